// xchg.cuh -- device side of the NVLink peer exchange (device_graph.h XDev): flags and the small push kernels that sit
// between separately launched kernels (linearize path).  The persistent PCG kernel uses the same primitives inline.
//
// Memory model: a producer writes its data into the consumer's region with plain stores, then makes them visible with
// __threadfence_system() and publishes a sequence number with st.release.sys; the consumer polls its OWN memory with
// ld.acquire.sys and only then reads the data.  Sums over ranks always run in rank order: deterministic.
//
// Included inside the anonymous namespace of kernels.cu.

__device__ __forceinline__ double* x_ptr(const Dev& d, uint32_t r, size_t off) { return reinterpret_cast<double*>(d.xc.peer[r] + off); }
__device__ __forceinline__ unsigned long long* x_flag(const Dev& d, uint32_t r, uint32_t point, uint32_t src) {
    return reinterpret_cast<unsigned long long*>(d.xc.peer[r] + d.xc.o_flags) + (size_t)point * kXMaxWorld + src;
}
// Called by a whole block: thread r < world publishes `v` at `point` to peer r (one remote store each, in parallel --
// a single thread looping over seven peers cost ~10 us per exchange point at 8 GPUs).  The data of the step must have
// been written, fenced at system scope and ordered before this call (grid barrier or stream order).
__device__ __forceinline__ void x_signal_block(const Dev& d, uint32_t point, unsigned long long v) {
    const uint32_t r = threadIdx.x;
    if (r < d.xc.world && r != d.xc.rank) {
        __threadfence_system();
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(x_flag(d, r, point, d.xc.rank)), "l"(v) : "memory");
    }
}
// Called by a whole block: thread r < world waits until peer r has published >= v at `point`; ends with a block barrier
__device__ __forceinline__ void x_wait_block(const Dev& d, uint32_t point, unsigned long long v) {
    const uint32_t r = threadIdx.x;
    if (r < d.xc.world && r != d.xc.rank) {
        const unsigned long long* f = x_flag(d, d.xc.rank, point, r);
        unsigned long long got;
        do { asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(got) : "l"(f) : "memory"); } while (got < v);
    }
    __syncthreads();
}

__global__ void k_x_signal(const Dev d, uint32_t point, unsigned long long v) { x_signal_block(d, point, v); }
__global__ void k_x_wait(const Dev d, uint32_t point, unsigned long long v) { x_wait_block(d, point, v); }

// boundary poses: entries of `vec` (3 per pose) that are ghosts on other ranks -> their ghost buffers
__global__ void k_x_push_ghost(const Dev d, const double* __restrict__ vec) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= d.xc.nExp) return;
    const uint32_t lp = d.xc.exp[3 * k], r = d.xc.exp[3 * k + 1], slot = d.xc.exp[3 * k + 2];
    double* o = x_ptr(d, r, d.xc.o_ghost) + 3ull * slot;
    o[0] = vec[3ull * lp]; o[1] = vec[3ull * lp + 1]; o[2] = vec[3ull * lp + 2];
}

// true in the last block of the grid to get here (all threads of that block); resets the counter for the next use.
// Every block fences its (remote) writes at system scope before it takes its ticket.
__device__ __forceinline__ bool x_last_block(uint32_t* counter) {
    __shared__ uint32_t s_last;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) { const uint32_t t = atomicAdd(counter, 1u); s_last = (t == gridDim.x - 1) ? 1u : 0u; if (s_last) *counter = 0; }
    __syncthreads();
    return s_last != 0;
}

// linearize, step 1 (after the slot partials exist): this rank's sum of every landmark it touches -> the owner's
// receive buffer rx[src = rank][l - first landmark of the owner] (5 doubles: Hll 00 01 11, b_l 0 1).  The last block to
// finish also sends this rank's chi2 to every rank and publishes the step (flag `v` at XP_LM5): no separate launches.
__global__ void __launch_bounds__(kThreads) k_x_lm5_push(const Dev d, unsigned long long v) {
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l < d.L) {
        uint32_t q = d.lm_part_ptr[l]; const uint32_t qe = d.lm_part_ptr[l + 1];
        if (q < qe) {                                               // touched here (lm_mask says so to the owner)
            double a0 = 0, a1 = 0, a2 = 0, a3 = 0, a4 = 0;
            for (; q < qe; ++q) { const double* u = d.lm_part + 5ull * d.lm_part_idx[q]; a0 += u[0]; a1 += u[1]; a2 += u[2]; a3 += u[3]; a4 += u[4]; }
            const uint32_t owner = l / d.xc.slice;
            double* o = x_ptr(d, owner, d.xc.o_lm5) + 5ull * ((size_t)d.xc.rank * d.xc.slice + (l - owner * d.xc.slice));
            o[0] = a0; o[1] = a1; o[2] = a2; o[3] = a3; o[4] = a4;
        }
    }
    if (x_last_block(d.xc.done)) {
        if (threadIdx.x < d.xc.world) x_ptr(d, threadIdx.x, d.xc.o_scal)[d.xc.rank] = d.scal[S_CHI2];      // slot 0: chi2 of this rank's edges
        x_signal_block(d, XP_LM5, v);
    }
}
// linearize, step 2: waits for every rank's step 1; the owner adds the contributions in rank order, finalises (prior,
// Hll, b_l, Hll^-1) and writes the result to every rank; chi2 = the ranks' parts in rank order; the last block
// publishes the step (flag `v` + 1 at XP_HINV).
__global__ void __launch_bounds__(kThreads) k_x_lm5_own(const Dev d, unsigned long long v) {
    x_wait_block(d, XP_LM5, v);
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t l = d.xc.rank * d.xc.slice + i;
    if (i < d.xc.slice && l < d.L) {
        const uint32_t mask = d.xc.lm_mask[l];
        double a0 = 0, a1 = 0, a2 = 0, a3 = 0, a4 = 0;
        for (uint32_t r = 0; r < d.xc.world; ++r) {
            if (!(mask >> r & 1u)) continue;
            const double* u = x_ptr(d, d.xc.rank, d.xc.o_lm5) + 5ull * ((size_t)r * d.xc.slice + i);
            a0 += u[0]; a1 += u[1]; a2 += u[2]; a3 += u[3]; a4 += u[4];
        }
        lm_finalize_one(d, l, a0, a1, a2, a3, a4);                  // writes this rank's hll / hinv / bl (they live in the region)
        const double h0 = d.hll[3ull * l], h1 = d.hll[3ull * l + 1], h2 = d.hll[3ull * l + 2];
        const double i0 = d.hinv[3ull * l], i1 = d.hinv[3ull * l + 1], i2 = d.hinv[3ull * l + 2];
        const double b0 = d.bl[2ull * l], b1 = d.bl[2ull * l + 1];
        for (uint32_t r = 0; r < d.xc.world; ++r) {
            if (r == d.xc.rank) continue;                            // every rank keeps the full arrays (read_system, back-substitution)
            double* hl = x_ptr(d, r, d.xc.o_hll) + 3ull * l; double* hi = x_ptr(d, r, d.xc.o_hinv) + 3ull * l; double* bl = x_ptr(d, r, d.xc.o_bl) + 2ull * l;
            hl[0] = h0; hl[1] = h1; hl[2] = h2; hi[0] = i0; hi[1] = i1; hi[2] = i2; bl[0] = b0; bl[1] = b1;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        double t = 0;
        for (uint32_t r = 0; r < d.xc.world; ++r) t += x_ptr(d, d.xc.rank, d.xc.o_scal)[r];
        d.scal[S_CHI2] = t;
    }
    if (x_last_block(d.xc.done + 1)) x_signal_block(d, XP_HINV, v + 1);
}
// scalar sum over ranks in rank order: every rank writes its value into slot `slot` of every region, then (after the
// flag) everybody adds the world values.  k_x_scal_push: one thread.
__global__ void k_x_scal_push(const Dev d, uint32_t slot, const double* __restrict__ src) {
    const double v = *src;
    for (uint32_t r = 0; r < d.xc.world; ++r) x_ptr(d, r, d.xc.o_scal)[(size_t)slot * kXMaxWorld + d.xc.rank] = v;
}
__global__ void k_x_scal_sum(const Dev d, uint32_t slot, double* __restrict__ dst) {
    double t = 0;
    for (uint32_t r = 0; r < d.xc.world; ++r) t += x_ptr(d, d.xc.rank, d.xc.o_scal)[(size_t)slot * kXMaxWorld + r];
    *dst = t;
}

// back-substitution across ranks: w = Hpl^T delta_p of this rank per touched landmark -> the owner; the owner forms
// delta_l = Hll^-1 (b_l - sum_r w_r) and pushes it to every rank (all ranks keep the full landmark state)
__global__ void __launch_bounds__(kThreads) k_x_w_push(const Dev d) {
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= d.L || d.lm_part_ptr[l] == d.lm_part_ptr[l + 1]) return;
    const double2 w = lm_w_sum(d, l);
    const uint32_t owner = l / d.xc.slice;
    reinterpret_cast<double2*>(x_ptr(d, owner, d.xc.o_bw))[(size_t)d.xc.rank * d.xc.slice + (l - owner * d.xc.slice)] = w;
}
__global__ void __launch_bounds__(kThreads) k_x_backsub_own(const Dev d) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t l = d.xc.rank * d.xc.slice + i;
    if (i >= d.xc.slice || l >= d.L) return;
    const uint32_t mask = d.xc.lm_mask[l];
    double w0 = 0, w1 = 0;
    for (uint32_t r = 0; r < d.xc.world; ++r) {
        if (!(mask >> r & 1u)) continue;
        const double2 w = reinterpret_cast<const double2*>(x_ptr(d, d.xc.rank, d.xc.o_bw))[(size_t)r * d.xc.slice + i];
        w0 += w.x; w1 += w.y;
    }
    const double t0 = d.bl[2ull * l] - w0, t1 = d.bl[2ull * l + 1] - w1;
    const double* hi = d.hinv + 3ull * l;
    const double2 dl = make_double2(hi[0] * t0 + hi[1] * t1, hi[1] * t0 + hi[2] * t1);
    for (uint32_t r = 0; r < d.xc.world; ++r) reinterpret_cast<double2*>(x_ptr(d, r, d.xc.o_dl))[l] = dl;
}

// ---- inside the persistent PCG kernel ------------------------------------------------------------------------------------
// grid barrier, then all-to-all over the peers: block 0 publishes this step, every block waits for the peers' flags.
// Blocks that wrote into peer regions fence at system scope before they arrive.
__device__ __forceinline__ void x_gsync(cg::grid_group& grid, const Dev& d, uint32_t& gtarget, uint32_t point, unsigned long long v) {
    __threadfence_system();
    res_gsync(grid, d.gbar, gtarget);
    if (blockIdx.x == 0) x_signal_block(d, point, v);
    x_wait_block(d, point, v);
}
// sum of one scalar per rank in rank order (every block computes the same `local`; block 0 publishes it): no grid barrier
// The exchange steps the persistent PCG kernel runs once per iteration are NOT inlined: inlined, their peer pointers and layout
// words were hoisted out of the iteration loop and stayed alive across the edge passes, and ptxas then put a dozen spill reloads
// into pass A's chunk loop (64-register cap) -- pass A took as long for half the chunks at 2 GPUs as for all of them at 1
// (profiles/r2c_pcg_phase_spread_n2.txt).  `d` is the kernel's __grid_constant__ parameter: the reference costs no copy.
__device__ __noinline__ double x_scalar_sum(const Dev& d, uint32_t slot, uint32_t point, unsigned long long v, double local) {
    if (blockIdx.x == 0) {
        const uint32_t r = threadIdx.x;
        if (r < d.xc.world && r != d.xc.rank) x_ptr(d, r, d.xc.o_scal)[(size_t)slot * kXMaxWorld + d.xc.rank] = local;
        x_signal_block(d, point, v);          // the same thread wrote the value: its fence orders it before the flag
    }
    x_wait_block(d, point, v);
    double t = 0;
    for (uint32_t r = 0; r < d.xc.world; ++r) t += r == d.xc.rank ? local : x_ptr(d, d.xc.rank, d.xc.o_scal)[(size_t)slot * kXMaxWorld + r];
    return t;
}
// Landmark phase inside the PCG, ONE exchange step (all-gather among the ranks that touch a landmark): every rank
// writes its partial w = Hpl^T p of a touched landmark into rx[src = me][l] of every OTHER rank that touches it (lm_mask)
// and keeps its own in the local region; after the flags every toucher adds the partials in rank order itself and forms
// u = Hll^-1 w (it holds Hll^-1 of every landmark it touches).  Same bits on every rank.  (Owner-computes -- partials to
// the owner, u back to the touchers -- needed two exchange steps per PCG iteration.)  Block 0 also sends the boundary
// poses' p to the ranks that hold them as ghosts.
__device__ __noinline__ void x_w_push(const Dev& d) {
    double2* mine = reinterpret_cast<double2*>(x_ptr(d, d.xc.rank, d.xc.o_w)) + (size_t)d.xc.rank * d.L;
    for (uint32_t l = 32 * (blockIdx.x + gridDim.x * (threadIdx.x >> 5)) + (threadIdx.x & 31); l < d.L; l += 32 * gridDim.x * (blockDim.x >> 5)) {      // warp-sized groups round-robin over blocks: see k_pcg_stream
        if (d.lm_part_ptr[l] == d.lm_part_ptr[l + 1]) continue;
        const double2 w = lm_w_sum(d, l);
        mine[l] = w;
        uint32_t mask = d.xc.lm_mask[l] & ~(1u << d.xc.rank);
        while (mask) {
            const uint32_t r = __ffs(mask) - 1; mask &= mask - 1;
            (reinterpret_cast<double2*>(x_ptr(d, r, d.xc.o_w)) + (size_t)d.xc.rank * d.L)[l] = w;
        }
    }
    if (blockIdx.x == 0)
        for (uint32_t k = threadIdx.x; k < d.xc.nExp; k += blockDim.x) {
            const uint32_t lp = d.xc.exp[3 * k], r = d.xc.exp[3 * k + 1], slot = d.xc.exp[3 * k + 2];
            double* o = x_ptr(d, r, d.xc.o_ghost) + 3ull * slot;
            o[0] = d.p[3ull * lp]; o[1] = d.p[3ull * lp + 1]; o[2] = d.p[3ull * lp + 2];
        }
}
// after the exchange: u of every touched landmark into d.w_sum (pass B gathers it), ghost p into the tail of d.p
__device__ __noinline__ void x_u_local(const Dev& d) {
    for (uint32_t i = threadIdx.x; i < 3 * d.nG; i += blockDim.x) d.p[3ull * d.nP + i] = x_ptr(d, d.xc.rank, d.xc.o_ghost)[i];   // same values from every block
    const double2* rx = reinterpret_cast<const double2*>(x_ptr(d, d.xc.rank, d.xc.o_w));
    for (uint32_t l = 32 * (blockIdx.x + gridDim.x * (threadIdx.x >> 5)) + (threadIdx.x & 31); l < d.L; l += 32 * gridDim.x * (blockDim.x >> 5)) {
        if (d.lm_part_ptr[l] == d.lm_part_ptr[l + 1]) continue;
        uint32_t mask = d.xc.lm_mask[l];
        double w0 = 0, w1 = 0;
        while (mask) {                                           // ascending rank
            const uint32_t r = __ffs(mask) - 1; mask &= mask - 1;
            const double2 w = rx[(size_t)r * d.L + l];
            w0 += w.x; w1 += w.y;
        }
        const double* hi = d.hinv + 3ull * l;
        reinterpret_cast<double2*>(d.w_sum)[l] = make_double2(hi[0] * w0 + hi[1] * w1, hi[1] * w0 + hi[2] * w1);
    }
    __syncthreads();
}
