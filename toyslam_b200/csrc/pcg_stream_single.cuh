// pcg_stream_single.cuh -- the single-GPU instance of the streamed PCG kernel, kept as its own source.
//
// This is the streamed kernel as it stood before the sharded path's register work (plain ChunkDesc carriers, descriptor loads
// selected against a dummy, phase timers of block 0 only).  At the 64-register cap ptxas' allocation of this kernel is fragile:
// the variants that fix the spills of k_pcg_stream<true> (pcg_stream.cuh: packed descriptors, clamped loads, per-block timers,
// __grid_constant__) all compile the single-GPU instance to code that runs 2-5 % SLOWER on one B200 -- measured A/B on the same
// box, same inputs: C3 1 556 -> 1 605 / 1 626 / 1 633 ms per 50 GN iterations, C4 1 578 -> 1 613 ... 1 671 ms per GN iteration --
// while k_pcg_stream<true> needs them (2 GPUs, C4: 1.34 -> 1.03 s).  So each instance keeps the source that compiles best for it;
// the arithmetic and every summation order are the same (the 2-rank-vs-1-rank and streamed-vs-persistent tests compare them).
// Shared with pcg_stream.cuh: the stage / smem layout constants, ml_* and coarse_* phases, the grid barrier.
//
// Included inside the anonymous namespace of kernels.cu after pcg_stream.cuh.

// Producer (one thread): arm the stage's barrier and start the bulk copies of chunk `ch` for pass A (PASS 0) or B
template <int PASS>
__device__ __forceinline__ void ps1_issue(const Dev& d, const ChunkDesc& ch, unsigned char* stage, uint32_t bar) {
    const uint32_t n = ch.e1 - ch.e0, ns = ch.s1 - ch.s0;
    const uint32_t e4 = ch.e0 & ~3u, e8 = ch.e0 & ~7u, s4 = ch.s0 & ~3u, p4 = ch.p0 & ~3u;
    const uint32_t bPk = n ? (((ch.e1 - e4) + 3u) & ~3u) * 4u : 0u;
    const uint32_t bSe = (PASS == 0 && n) ? (((ch.e1 - e8) + 7u) & ~7u) * 2u : 0u;
    const uint32_t bPp = PASS == 1 ? (((ch.p1 + 1 - p4) + 3u) & ~3u) * 4u : 0u;
    const uint32_t bSp = PASS == 0 ? (((ch.s1 + 1 - s4) + 3u) & ~3u) * 4u : 0u;
    const uint32_t bSl = (PASS == 1 && ns) ? (((ch.s1 - s4) + 3u) & ~3u) * 4u : 0u;
    const uint32_t bH = n * 16u;
    const uint32_t s = smem_u32(stage);
    const double2* hp = hpl_row(d, ch.e0); const size_t hs = hpl_stride(d);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    mbar_expect_tx(bar, bPk + bSe + bPp + bSp + bSl + 3 * bH);
    if (bPk) bulk_g2s(s + kPsPk, d.lm_packed + e4, bPk, bar);
    if (bH) { bulk_g2s(s + kPsH, hp, bH, bar); bulk_g2s(s + kPsH + 16 * kChunkEdges, hp + hs, bH, bar); bulk_g2s(s + kPsH + 32 * kChunkEdges, hp + 2 * hs, bH, bar); }
    if (bSe) bulk_g2s(s + kPsIx1, d.slot_edge + e8, bSe, bar);
    if (bPp) bulk_g2s(s + kPsIx1, d.pose_ptr + p4, bPp, bar);
    if (bSp) bulk_g2s(s + kPsIx2, d.slot_ptr + s4, bSp, bar);
    if (bSl) bulk_g2s(s + kPsIx2, d.slot_lm + s4, bSl, bar);
}

// p_i = z_i (+ coarse correction) + beta p_i for component t of the poses starting at p0
__device__ __forceinline__ double ps1_pnew(const Dev& d, uint32_t p0, uint32_t t, double beta, bool coarse) {
    const size_t i = 3ull * p0 + t;
    double zz = d.z[i];
    if (coarse) zz += prolong(d, p0 + t / 3, t % 3);
    return zz + beta * d.p[i];
}

// Pass A over the block's chunks [cb, ce): p = z + beta p (fused), slot partials of Hpl^T p -> w_part
__device__ __forceinline__ void ps1_pass_a(const Dev& d, unsigned char* smem, uint32_t cb, uint32_t ce, uint32_t& kbase, double beta, bool coarse) {
    if (cb >= ce) return;
    double* rows = reinterpret_cast<double*>(smem + kPsRows);
    double* pvec = reinterpret_cast<double*>(smem + kPsAux);
    unsigned char* ring = smem + kPsRing;
    const uint32_t bar0 = smem_u32(smem + kPsBar), tid = threadIdx.x;
    const ChunkDesc zero = {0, 0, 0, 0, 0, 0, 0xFFFFFFFFu, 0};
    ChunkDesc cur = d.chunks[cb];
    ChunkDesc nx1 = cb + 1 < ce ? d.chunks[cb + 1] : zero;
    ChunkDesc nx2 = cb + 2 < ce ? d.chunks[cb + 2] : zero;
    if (tid == 0) {
        ps1_issue<0>(d, cur, ring + (kbase % kLsStages) * kPsStageBytes, bar0 + 8 * (kbase % kLsStages));
        if (cb + 1 < ce) ps1_issue<0>(d, nx1, ring + ((kbase + 1) % kLsStages) * kPsStageBytes, bar0 + 8 * ((kbase + 1) % kLsStages));
        if (cb + 2 < ce) ps1_issue<0>(d, nx2, ring + ((kbase + 2) % kLsStages) * kPsStageBytes, bar0 + 8 * ((kbase + 2) % kLsStages));
    }
    // p = z (+ coarse correction) + beta p for ALL poses of this block's chunk range, up front: the correction costs two
    // dependent global loads (aggregate -> yc), which inside the chunk loop stalled the block once per chunk
    {
        const uint32_t p_lo = cur.p0, p_hi = d.chunks[ce - 1].p1;
        if (coarse && d.mlLv) {         // multilevel: one thread per pose gathers the hat-weighted node corrections of every level
            for (uint32_t i = p_lo + tid; i < p_hi; i += kPsThreads) {
                const double4 ps = d.pose[i];
                double c0, c1, c2; ml_prolong(d, i, ps.x, ps.y, c0, c1, c2);
                double* pp = d.p + 3ull * i; const double* zz = d.z + 3ull * i;
                pp[0] = (zz[0] + c0) + beta * pp[0]; pp[1] = (zz[1] + c1) + beta * pp[1]; pp[2] = (zz[2] + c2) + beta * pp[2];
            }
        } else
        for (uint32_t i = tid; i < 3 * (p_hi - p_lo); i += kPsThreads) d.p[3ull * p_lo + i] = ps1_pnew(d, p_lo, i, beta, coarse);
    }
    __syncthreads();
    uint32_t vb = 0;
    if (tid < 3 * (cur.p1 - cur.p0)) pvec[tid] = d.p[3ull * cur.p0 + tid];
    __syncthreads();
    for (uint32_t c = cb; c < ce; ++c) {
        const uint32_t k = kbase + (c - cb), sidx = k % kLsStages;
        unsigned char* st = ring + sidx * kPsStageBytes;
        const uint32_t n = cur.e1 - cur.e0, ns = cur.s1 - cur.s0;
        const bool newunit = c + 1 < ce && nx1.block != cur.block;
        const ChunkDesc nx3 = c + 3 < ce ? d.chunks[c + 3] : zero;
        // p of the next chunk's poses when it starts a new unit: a raw load, in flight during this chunk
        const uint32_t tn = newunit ? 3 * (nx1.p1 - nx1.p0) : 0;
        double pn = 0;
        if (tid < tn) pn = d.p[3ull * nx1.p0 + tid];
        mbar_wait(bar0 + 8 * sidx, (k / kLsStages) & 1u);
        const uint32_t* pk = reinterpret_cast<const uint32_t*>(st + kPsPk) + (cur.e0 & 3u);
        const uint16_t* sedge = reinterpret_cast<const uint16_t*>(st + kPsIx1) + (cur.e0 & 7u);
        const uint32_t* sptr = reinterpret_cast<const uint32_t*>(st + kPsIx2) + (cur.s0 & 3u);
        const double2* hq = reinterpret_cast<const double2*>(st + kPsH);
        if (tid < n) {
            const double2 h0 = hq[tid], h1 = hq[kChunkEdges + tid], h2 = hq[2 * kChunkEdges + tid];
            const double* v = pvec + vb * kPsVecLen + 3 * (pk[tid] >> 16);
            const double v0 = v[0], v1 = v[1], v2 = v[2];
            rows[tid] = h0.x * v0 + h1.x * v1 + h2.x * v2; rows[kStageStride + tid] = h0.y * v0 + h1.y * v1 + h2.y * v2;
        }
        __syncthreads();
#pragma unroll 1
        for (uint32_t item = tid; item < 2 * ns; item += kPsThreads) {
            const uint32_t sl = item >> 1, comp = item & 1;
            d.w_part[2ull * (cur.s0 + sl) + comp] = list_sum_lean(rows + comp * kStageStride, sedge, sptr[sl] - cur.e0, sptr[sl + 1] - cur.e0);
        }
        if (tid < tn) pvec[(vb ^ 1u) * kPsVecLen + tid] = pn;
        if (newunit) vb ^= 1u;
        __syncthreads();
        if (tid == 0 && c + 3 < ce) ps1_issue<0>(d, nx3, st, bar0 + 8 * sidx);
        cur = nx1; nx1 = nx2; nx2 = nx3;
    }
    kbase += ce - cb;
}

// Pass B over the block's chunks: sp_i -= sum_e Hpl_e u_l(e), u = Hll^-1 w (in d.w_sum); returns this thread's part of p'Sp
__device__ __forceinline__ double ps1_pass_b(const Dev& d, unsigned char* smem, uint32_t cb, uint32_t ce, uint32_t& kbase) {
    if (cb >= ce) return 0.0;
    double* rows = reinterpret_cast<double*>(smem + kPsRows);
    double2* slot2 = reinterpret_cast<double2*>(smem + kPsAux);
    unsigned char* ring = smem + kPsRing;
    const uint32_t bar0 = smem_u32(smem + kPsBar), tid = threadIdx.x;
    const double2* usum = reinterpret_cast<const double2*>(d.w_sum);        // u_l = Hll^-1 w_l, written by the middle phase
    const ChunkDesc zero = {0, 0, 0, 0, 0, 0, 0xFFFFFFFFu, 0};
    ChunkDesc cur = d.chunks[cb];
    ChunkDesc nx1 = cb + 1 < ce ? d.chunks[cb + 1] : zero;
    ChunkDesc nx2 = cb + 2 < ce ? d.chunks[cb + 2] : zero;
    if (tid == 0) {
        ps1_issue<1>(d, cur, ring + (kbase % kLsStages) * kPsStageBytes, bar0 + 8 * (kbase % kLsStages));
        if (cb + 1 < ce) ps1_issue<1>(d, nx1, ring + ((kbase + 1) % kLsStages) * kPsStageBytes, bar0 + 8 * ((kbase + 1) % kLsStages));
        if (cb + 2 < ce) ps1_issue<1>(d, nx2, ring + ((kbase + 2) % kLsStages) * kPsStageBytes, bar0 + 8 * ((kbase + 2) % kLsStages));
    }
    {   // u of the first chunk's slots
        mbar_wait(bar0 + 8 * (kbase % kLsStages), (kbase / kLsStages) & 1u);
        const uint32_t* slm = reinterpret_cast<const uint32_t*>(ring + (kbase % kLsStages) * kPsStageBytes + kPsIx2) + (cur.s0 & 3u);
        if (tid < cur.s1 - cur.s0) slot2[tid] = usum[slm[tid]];
        __syncthreads();
    }
    double acc = 0, dot = 0;
    for (uint32_t c = cb; c < ce; ++c) {
        const uint32_t k = kbase + (c - cb), sidx = k % kLsStages;
        unsigned char* st = ring + sidx * kPsStageBytes;
        const uint32_t n = cur.e1 - cur.e0, np = cur.p1 - cur.p0;
        const bool last = nx1.block != cur.block;                    // `zero` ends the block's final unit
        const ChunkDesc nx3 = c + 3 < ce ? d.chunks[c + 3] : zero;
        // next chunk: u of its slots; this unit's sp and p if it ends here (global loads in flight during this chunk)
        double2 un = {};
        const uint32_t ns_next = nx1.s1 - nx1.s0;
        if (c + 1 < ce) {
            const uint32_t s2 = (k + 1) % kLsStages;
            mbar_wait(bar0 + 8 * s2, ((k + 1) / kLsStages) & 1u);
            const uint32_t* slm2 = reinterpret_cast<const uint32_t*>(ring + s2 * kPsStageBytes + kPsIx2) + (nx1.s0 & 3u);
            if (tid < ns_next) un = usum[slm2[tid]];            // raw load: nothing waits for it before the end of this chunk
        }
        const uint32_t nitems = 3 * np;
        double spv = 0, pv = 0;
        if (last && tid < nitems) { spv = d.sp[3ull * cur.p0 + tid]; pv = d.p[3ull * cur.p0 + tid]; }
        mbar_wait(bar0 + 8 * sidx, (k / kLsStages) & 1u);
        const uint32_t* pk = reinterpret_cast<const uint32_t*>(st + kPsPk) + (cur.e0 & 3u);
        const uint32_t* pptr = reinterpret_cast<const uint32_t*>(st + kPsIx1) + (cur.p0 & 3u);
        const double2* hq = reinterpret_cast<const double2*>(st + kPsH);
        if (tid < n) {
            const double2 h0 = hq[tid], h1 = hq[kChunkEdges + tid], h2 = hq[2 * kChunkEdges + tid];
            const double2 u = slot2[pk[tid] & 0xffffu];
            rows[tid] = h0.x * u.x + h0.y * u.y; rows[kStageStride + tid] = h1.x * u.x + h1.y * u.y;
            rows[2 * kStageStride + tid] = h2.x * u.x + h2.y * u.y;
        }
        __syncthreads();
        if (tid < nitems) {
            const uint32_t pl = tid / 3, row = tid - 3 * pl;
            const uint32_t b = max(pptr[pl], cur.e0) - cur.e0, e = min(pptr[pl + 1], cur.e1) - cur.e0;
            acc += seg_sum_lean(rows + row * kStageStride, b, e);
            if (last) { const double y = spv - acc; d.sp[3ull * cur.p0 + tid] = y; dot += pv * y; acc = 0; }
        }
        if (tid < ns_next) slot2[tid] = un;
        __syncthreads();
        if (tid == 0 && c + 3 < ce) ps1_issue<1>(d, nx3, st, bar0 + 8 * sidx);
        cur = nx1; nx1 = nx2; nx2 = nx3;
    }
    kbase += ce - cb;
    return dot;
}

// Pose-vector step of k_pcg_stream1: pose_step over this block's 512-pose windows.  With a large coarse space
// (nc > 3 kCoarseMaxAgg) whose aggregates nest in a window (aggK / 32 divides 16: C3 K = 64, C4 K = 512) the block also
// sums the segment partials it has just written into Z^T r of its aggregates (d.rcs[0 .. nc), same segment order as
// coarse_restrict), so that coarse_phase need not restrict 3 nSeg values once per block (C4: 296 x 750 KB of L2 reads
// per PCG iteration).  Returns whether d.rcs holds Z^T r.
template <bool INIT>
__device__ __forceinline__ bool ps1_pose_phase(const Dev& d, double alpha, bool coarse, double& dot) {
    const uint32_t nP32 = (d.nP + 31u) & ~31u, gsz = gridDim.x * kPsThreads, spa = d.aggK >> 5;
    const bool agg = coarse && d.nc > 3 * kCoarseMaxAgg && spa >= 1 && spa <= 16 && 16 % spa == 0;
    dot = 0;
    for (uint32_t base = blockIdx.x * kPsThreads; base < nP32; base += gsz) {
        const uint32_t i = base + threadIdx.x;
        if (i < nP32) dot += pose_step<INIT>(d, i, i < d.nP, alpha, coarse);
        if (agg) {
            __syncthreads();                                  // this block's rcp entries of the window are written
            const uint32_t seg0 = base >> 5, k = threadIdx.x;  // k = (aggregate in window, component)
            if (k < 3 * (16 / spa)) {
                const uint32_t a = seg0 / spa + k / 3, q = k % 3;
                if (a < d.nA) {
                    const uint32_t s0 = a * spa, s1 = min(d.nSeg, s0 + spa);
                    double t = 0;
                    for (uint32_t sg = s0; sg < s1; ++sg) t += d.rcp[3ull * sg + q];
                    d.rcs[3ull * a + q] = t;
                }
            }
        }
    }
    return agg;
}

// XP: compiled with the NVLink peer exchange (world > 1); the single-GPU instance carries none of that code
template <bool XP>
__global__ void __launch_bounds__(kPsThreads, 2) k_pcg_stream1(const Dev d) {
    extern __shared__ __align__(32) unsigned char smem_raw[];
    double* red = reinterpret_cast<double*>(smem_raw + kPsRed);
    double* rc_s = reinterpret_cast<double*>(smem_raw + kPsRows);
    cg::grid_group grid = cg::this_grid();
    uint32_t gtarget = 0, kbase = 0;
    const uint32_t tid = threadIdx.x, nb = gridDim.x, gtid = blockIdx.x * kPsThreads + tid, gsz = nb * kPsThreads;
    const uint32_t u_lo = (uint32_t)((uint64_t)d.nBlocks * blockIdx.x / nb), u_hi = (uint32_t)((uint64_t)d.nBlocks * (blockIdx.x + 1) / nb);
    const uint32_t cb = d.blk_chunk_ptr[u_lo], ce = d.blk_chunk_ptr[u_hi];
    if (tid == 0) {
        for (int s = 0; s < kLsStages; ++s) mbar_init(smem_u32(smem_raw + kPsBar) + 8 * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const bool coarse = d.flags[F_COARSE] != 0;
    const bool ml = coarse && d.mlLv > 0;                            // multilevel preconditioner (mlevel.cuh)
    const bool crow = coarse && coarse_row_owner(d);
    const uint32_t ncdot = ml ? d.mlCdot : d.nc;
    double* ml_sm = reinterpret_cast<double*>(smem_raw + kPsRing);   // the ring is idle during the pose phase
    double* rbuf[2] = {d.r, ml ? d.r2 : d.r};
    double dot = 0;
    bool rc_shared = false;
    if (ml) dot = ml_pose_phase<true>(d, ml_sm, 0.0, rbuf[0], rbuf[0]);
    else rc_shared = ps1_pose_phase<true>(d, 0.0, coarse, dot);
    double t = block_sum(dot, red);
    if (tid == 0) d.part[blockIdx.x] = t;
    // world > 1 (peer exchange, xchg.cuh): the ranks' parts of every scalar are added in rank order -- every rank sees
    // the same bits and takes the same branches.  With the hierarchy the top level is GLOBAL: every rank pushes the
    // residual of its top nodes to all ranks, the dense mat-vec and its part of r'z are replicated.
    constexpr bool xp = XP;
    unsigned long long xv = d.xc.seq;
    res_gsync(grid, d.gbar, gtarget);
    double vloc = 0;
    auto coarse_step = [&]() -> double {
        if (coarse) {
            if (ml && xp) {
                // one exchange step carries the residual of this rank's top nodes AND its part of r'z (fine level + its own
                // lower levels), which every block can form right after the grid barrier; block 0 publishes it with the flag
                ml_coarse_push(d, gtid, gsz);
                __threadfence_system();
                res_gsync(grid, d.gbar, gtarget);
                vloc = reduce_all(d.part, nb, red) + reduce_all(d.cdot + d.nc, ncdot - d.nc, red);
                ++xv;
                if (blockIdx.x == 0) {
                    if (tid < d.xc.world && tid != d.xc.rank) x_ptr(d, tid, d.xc.o_scal)[3 * kXMaxWorld + d.xc.rank] = vloc;
                    x_signal_block(d, XP_TOP, xv);
                }
                x_wait_block(d, XP_TOP, xv);
                if (crow) ml_coarse_top(d, gtid, gsz, rc_s);
            }
            else if (ml) ml_coarse_phase(d, gtid, gsz, rc_s, crow);
            else if (crow) coarse_phase(d, gtid >> 5, gsz >> 5, rc_s, rc_shared);
            res_gsync(grid, d.gbar, gtarget);
        }
        if (ml && xp) {          // ranks' parts in rank order, then the replicated top-level part
            double v = 0;
            for (uint32_t r = 0; r < d.xc.world; ++r) v += r == d.xc.rank ? vloc : x_ptr(d, d.xc.rank, d.xc.o_scal)[3 * kXMaxWorld + r];
            return v + reduce_all(d.cdot, d.nc, red);
        }
        double v = reduce_all(d.part, nb, red);
        if (coarse) v += reduce_all(d.cdot, ncdot, red);
        if (xp) v = x_scalar_sum(d, 3, XP_TOP, ++xv, v);
        return v;
    };
    double rz0 = coarse_step();
    double rz = rz0, beta = 0;
    const double tol2 = d.scal[S_TOL2]; const int maxit = d.flags[F_MAXIT];
    int it = 0;
    // optional phase timing (TSB_PCG_TIMING=1): SM cycles of block 0 per phase, barrier waits separate
    const bool timing = d.dbg != nullptr && blockIdx.x == 0 && tid == 0;
    long long tk = 0; long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#define TICK(slot) do { if (timing) { const long long now_ = clock64(); tacc[slot] += now_ - tk; tk = now_; } } while (0)
    if (rz0 > 0) {
        for (;;) {
            if (timing) tk = clock64();
            ps1_pass_a(d, smem_raw, cb, ce, kbase, beta, coarse);
            TICK(0);
            res_gsync(grid, d.gbar, gtarget);
            TICK(1);
            if (xp) {             // landmark sums across ranks: one all-gather step among the touchers (xchg.cuh)
                x_w_push(d);
                x_gsync(grid, d, gtarget, XP_W, ++xv);
                x_u_local(d);
                for (uint32_t p = gtid; p < d.nP; p += gsz) hpp_apply(d, p, d.sp + 3ull * p);
                TICK(2);
                res_gsync(grid, d.gbar, gtarget);
            } else {
            // landmarks are dealt round-robin over the BLOCKS (warp-sized groups, group g -> block g mod nb): with l = gtid the 20 k landmarks of
            // C3 all sat on the first 39 blocks while the other 257 waited at the barrier (35 k + 23 k cycles per iteration)
            // (groups of 32 consecutive landmarks, one per warp, so that the per-landmark arrays are still read coalesced)
            for (uint32_t l = 32 * (blockIdx.x + nb * (tid >> 5)) + (tid & 31); l < d.L; l += 32 * nb * (kPsThreads / 32)) {          // u_l = Hll^-1 w_l straight away: pass B then gathers it raw
                const double2 w = lm_w_sum(d, l); const double* hi = d.hinv + 3ull * l;
                reinterpret_cast<double2*>(d.w_sum)[l] = make_double2(hi[0] * w.x + hi[1] * w.y, hi[1] * w.x + hi[2] * w.y);
            }
            for (uint32_t p = gtid; p < d.nP; p += gsz) hpp_apply(d, p, d.sp + 3ull * p);
            TICK(2);
            res_gsync(grid, d.gbar, gtarget);
            }
            TICK(3);
            t = block_sum(ps1_pass_b(d, smem_raw, cb, ce, kbase), red);
            if (tid == 0) d.part2[blockIdx.x] = t;
            TICK(4);
            res_gsync(grid, d.gbar, gtarget);
            TICK(5);
            double psp = reduce_all(d.part2, nb, red);
            if (xp) psp = x_scalar_sum(d, 2, XP_PSP, ++xv, psp);
            const double alpha = rz / psp;
            if (ml) dot = ml_pose_phase<false>(d, ml_sm, alpha, rbuf[it & 1], rbuf[(it + 1) & 1]);
            else ps1_pose_phase<false>(d, alpha, coarse, dot);
            t = block_sum(dot, red);
            if (tid == 0) d.part[blockIdx.x] = t;
            TICK(6);
            res_gsync(grid, d.gbar, gtarget);
            const double rz_new = coarse_step();
            TICK(7);
            beta = rz_new / rz; rz = rz_new; ++it;
            if (!(rz_new > tol2 * rz0) || it >= maxit) break;
        }
    }
    if (blockIdx.x == 0 && tid == 0) {
        d.scal[S_RZ0] = rz0; d.scal[S_RZ] = rz; d.scal[S_BETA] = beta; d.flags[F_ITERS] = it; d.flags[F_DONE] = 1;
        if (timing) for (int q = 0; q < 8; ++q) d.dbg[q] = tacc[q];
    }
#undef TICK
}

