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
// one thread: publish `v` at `point` to every peer (after all data writes of the step are done grid-wide)
__device__ __forceinline__ void x_signal_all(const Dev& d, uint32_t point, unsigned long long v) {
    __threadfence_system();
    for (uint32_t r = 0; r < d.xc.world; ++r) {
        if (r == d.xc.rank) continue;
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(x_flag(d, r, point, d.xc.rank)), "l"(v) : "memory");
    }
}
// one thread: wait until every peer has published >= v at `point`
__device__ __forceinline__ void x_wait_all(const Dev& d, uint32_t point, unsigned long long v) {
    for (uint32_t r = 0; r < d.xc.world; ++r) {
        if (r == d.xc.rank) continue;
        const unsigned long long* f = x_flag(d, d.xc.rank, point, r);
        unsigned long long got;
        do { asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(got) : "l"(f) : "memory"); } while (got < v);
    }
}

__global__ void k_x_signal(const Dev d, uint32_t point, unsigned long long v) { x_signal_all(d, point, v); }
__global__ void k_x_wait(const Dev d, uint32_t point, unsigned long long v) { x_wait_all(d, point, v); }

// boundary poses: entries of `vec` (3 per pose) that are ghosts on other ranks -> their ghost buffers
__global__ void k_x_push_ghost(const Dev d, const double* __restrict__ vec) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= d.xc.nExp) return;
    const uint32_t lp = d.xc.exp[3 * k], r = d.xc.exp[3 * k + 1], slot = d.xc.exp[3 * k + 2];
    double* o = x_ptr(d, r, d.xc.o_ghost) + 3ull * slot;
    o[0] = vec[3ull * lp]; o[1] = vec[3ull * lp + 1]; o[2] = vec[3ull * lp + 2];
}

// linearize, step 1 (after the slot partials exist): this rank's sum of every landmark it touches -> the owner's
// receive buffer rx[src = rank][l - first landmark of the owner] (5 doubles: Hll 00 01 11, b_l 0 1)
__global__ void __launch_bounds__(kThreads) k_x_lm5_push(const Dev d) {
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= d.L) return;
    uint32_t q = d.lm_part_ptr[l]; const uint32_t qe = d.lm_part_ptr[l + 1];
    if (q == qe) return;                                        // not touched here (lm_mask says so to the owner)
    double a0 = 0, a1 = 0, a2 = 0, a3 = 0, a4 = 0;
    for (; q < qe; ++q) { const double* u = d.lm_part + 5ull * d.lm_part_idx[q]; a0 += u[0]; a1 += u[1]; a2 += u[2]; a3 += u[3]; a4 += u[4]; }
    const uint32_t owner = l / d.xc.slice;
    double* o = x_ptr(d, owner, d.xc.o_lm5) + 5ull * ((size_t)d.xc.rank * d.xc.slice + (l - owner * d.xc.slice));
    o[0] = a0; o[1] = a1; o[2] = a2; o[3] = a3; o[4] = a4;
}
// linearize, step 2: the owner adds the contributions in rank order, finalises (prior, Hll, b_l, Hll^-1) and writes
// the result to every rank that touches the landmark (and to itself).  chi2 travels as a scalar the same way.
__global__ void __launch_bounds__(kThreads) k_x_lm5_own(const Dev d) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t l = d.xc.rank * d.xc.slice + i;
    if (i >= d.xc.slice || l >= d.L) return;
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
    reinterpret_cast<double2*>(x_ptr(d, owner, d.xc.o_w))[(size_t)d.xc.rank * d.xc.slice + (l - owner * d.xc.slice)] = w;
}
__global__ void __launch_bounds__(kThreads) k_x_backsub_own(const Dev d) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t l = d.xc.rank * d.xc.slice + i;
    if (i >= d.xc.slice || l >= d.L) return;
    const uint32_t mask = d.xc.lm_mask[l];
    double w0 = 0, w1 = 0;
    for (uint32_t r = 0; r < d.xc.world; ++r) {
        if (!(mask >> r & 1u)) continue;
        const double2 w = reinterpret_cast<const double2*>(x_ptr(d, d.xc.rank, d.xc.o_w))[(size_t)r * d.xc.slice + i];
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
    if (blockIdx.x == 0 && threadIdx.x == 0) x_signal_all(d, point, v);
    if (threadIdx.x == 0) x_wait_all(d, point, v);
    __syncthreads();
}
// sum of one scalar per rank in rank order (every block computes the same `local`; block 0 publishes it): no grid barrier
__device__ __forceinline__ double x_scalar_sum(const Dev& d, uint32_t slot, uint32_t point, unsigned long long v, double local) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        for (uint32_t r = 0; r < d.xc.world; ++r) if (r != d.xc.rank) x_ptr(d, r, d.xc.o_scal)[(size_t)slot * kXMaxWorld + d.xc.rank] = local;
        x_signal_all(d, point, v);
    }
    if (threadIdx.x == 0) x_wait_all(d, point, v);
    __syncthreads();
    double t = 0;
    for (uint32_t r = 0; r < d.xc.world; ++r) t += r == d.xc.rank ? local : x_ptr(d, d.xc.rank, d.xc.o_scal)[(size_t)slot * kXMaxWorld + r];
    return t;
}
// landmark phase, step 1: this rank's w = Hpl^T p per touched landmark -> the owner's receive buffer; block 0 also sends
// the boundary poses' p to the ranks that hold them as ghosts
__device__ __forceinline__ void x_w_push(const Dev& d, uint32_t gtid, uint32_t gsz) {
    for (uint32_t l = blockIdx.x + gridDim.x * threadIdx.x; l < d.L; l += gridDim.x * blockDim.x) {      // round-robin over blocks: see k_pcg_stream
        if (d.lm_part_ptr[l] == d.lm_part_ptr[l + 1]) continue;
        const double2 w = lm_w_sum(d, l);
        const uint32_t owner = l / d.xc.slice;
        reinterpret_cast<double2*>(x_ptr(d, owner, d.xc.o_w))[(size_t)d.xc.rank * d.xc.slice + (l - owner * d.xc.slice)] = w;
    }
    if (blockIdx.x == 0)
        for (uint32_t k = threadIdx.x; k < d.xc.nExp; k += blockDim.x) {
            const uint32_t lp = d.xc.exp[3 * k], r = d.xc.exp[3 * k + 1], slot = d.xc.exp[3 * k + 2];
            double* o = x_ptr(d, r, d.xc.o_ghost) + 3ull * slot;
            o[0] = d.p[3ull * lp]; o[1] = d.p[3ull * lp + 1]; o[2] = d.p[3ull * lp + 2];
        }
}
// landmark phase, step 2: owners add the contributions in rank order, u = Hll^-1 w, pushed to every rank that touches
// the landmark (pass B gathers it from d.w_sum == the own region's u array); ghost p into the tail of d.p
__device__ __forceinline__ void x_u_own(const Dev& d, uint32_t gtid, uint32_t gsz) {
    for (uint32_t i = threadIdx.x; i < 3 * d.nG; i += blockDim.x) d.p[3ull * d.nP + i] = x_ptr(d, d.xc.rank, d.xc.o_ghost)[i];   // same values from every block
    for (uint32_t i = blockIdx.x + gridDim.x * threadIdx.x; i < d.xc.slice; i += gridDim.x * blockDim.x) {
        const uint32_t l = d.xc.rank * d.xc.slice + i;
        if (l >= d.L) break;
        const uint32_t mask = d.xc.lm_mask[l];
        double w0 = 0, w1 = 0;
        for (uint32_t r = 0; r < d.xc.world; ++r) {
            if (!(mask >> r & 1u)) continue;
            const double2 w = reinterpret_cast<const double2*>(x_ptr(d, d.xc.rank, d.xc.o_w))[(size_t)r * d.xc.slice + i];
            w0 += w.x; w1 += w.y;
        }
        const double* hi = d.hinv + 3ull * l;
        const double2 u = make_double2(hi[0] * w0 + hi[1] * w1, hi[1] * w0 + hi[2] * w1);
        for (uint32_t r = 0; r < d.xc.world; ++r) if (mask >> r & 1u) reinterpret_cast<double2*>(x_ptr(d, r, d.xc.o_u))[l] = u;
    }
    __syncthreads();
}
