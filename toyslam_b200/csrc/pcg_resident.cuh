// pcg_resident.cuh -- K4 for graphs whose Hpl fits the shared memory of the whole GPU (C2-class: <= ~1280 LM
// edges per SM).  One block per SM, 1024 threads, one cooperative launch per solve.
//
// Why: at this size the PCG is latency-bound, not bandwidth-bound (profiles/r1_pcg_phase_cycles_c2.txt: five
// phases of 3-5 us each, most of it dependent L2 round trips, plus five grid barriers per iteration).  Here
//   * both orders of Hpl a block needs (its poses' edges in assembly order; its landmarks' edges in landmark-
//     major order) are loaded into shared memory once per solve, with the block's pose vectors (x, r, p, S p,
//     z), diagonal blocks, ODOM neighbour lists and landmark inverses: an iteration touches global memory only
//     for what crosses blocks (p and z of other poses, u of other landmarks, 2 x nb partial dot products, nc
//     coarse residuals);
//   * the landmark sum is computed landmark-major (one warp per landmark gathers p = z + beta p_old of its
//     edges' poses), so Hpl^T p needs no slot partials and no separate summation phase;
//   * the coarse residual is carried by recurrence, rc -= alpha Z^T(S p), every block redundantly, so that the
//     coarse solve of a block's own aggregates, the prolongation and r'z need no barrier of their own.
// Three grid barriers per iteration instead of five.  Every sum runs in a fixed order (deterministic).
//
// Included inside the anonymous namespace of kernels.cu.

constexpr int kResThreads = 1024;
constexpr uint32_t kResEC = kResEdgeCap, kResPC = kResPoseCap, kResLC = kResLmCap, kResAC = kResAggCap, kResOC = kResOdomCap;

struct ResSm {
    double* hplP; double* hplL; double* rows; double* x; double* r; double* p; double* y; double* zf;
    double* minv; double* hpp; double2* zoff; double* onb_h; double* onb_p; double* hinvL; double* rc; double* yc; double* red;
    uint32_t* elm; uint32_t* epos; uint32_t* eptr; uint32_t* lptr; uint32_t* pagg; uint32_t* onb_ptr; uint32_t* onb_idx; uint32_t* aptr;
};
constexpr size_t kResSmemBytes =
    sizeof(double) * (6 * kResEC * 2 + 3 * kResEC + 5 * 3 * kResPC + 2 * 6 * kResPC + 2 * kResPC + 6 * kResOC + 3 * kResLC +
                      3 * kCoarseMaxAgg + 3 * kResAC + 4 + 32 + 32 * 4) +
    sizeof(uint32_t) * (2 * kResEC + (kResPC + 4) + (kResLC + 4) + kResPC + (kResPC + 4) + kResOC + kResAC + 4);

__device__ __forceinline__ ResSm res_carve(unsigned char* b) {
    ResSm s; double* f = reinterpret_cast<double*>(b);
    s.hplP = f; f += 6 * kResEC; s.hplL = f; f += 6 * kResEC; s.rows = f; f += 3 * kResEC;
    s.x = f; f += 3 * kResPC; s.r = f; f += 3 * kResPC; s.p = f; f += 3 * kResPC; s.y = f; f += 3 * kResPC; s.zf = f; f += 3 * kResPC;
    s.minv = f; f += 6 * kResPC; s.hpp = f; f += 6 * kResPC; s.zoff = reinterpret_cast<double2*>(f); f += 2 * kResPC;
    s.onb_h = f; f += 3 * kResOC; s.onb_p = f; f += 3 * kResOC; s.hinvL = f; f += 3 * kResLC; s.rc = f; f += 3 * kCoarseMaxAgg; s.yc = f; f += 3 * kResAC + 4;
    s.red = f; f += 32 + 32 * 4;      // + partial row sums of the coarse mat-vec (one per warp)
    uint32_t* u = reinterpret_cast<uint32_t*>(f);
    s.elm = u; u += kResEC; s.epos = u; u += kResEC; s.eptr = u; u += kResPC + 4; s.lptr = u; u += kResLC + 4;
    s.pagg = u; u += kResPC; s.onb_ptr = u; u += kResPC + 4; s.onb_idx = u; u += kResOC; s.aptr = u;
    return s;
}

// component q of Z_i^T v for pose-local vector v (3 doubles)
__device__ __forceinline__ double res_zt(const double2 zo, const double* v, uint32_t q) {
    return q == 0 ? v[0] : q == 1 ? v[1] : zo.x * v[0] + zo.y * v[1] + v[2];
}

// Z^T v summed over the poses of each of this block's aggregates -> out[3 (a0 + a) + q]: warp a, lanes stride
// the aggregate's poses, fixed xor tree (deterministic).  aptr = local pose offsets of the aggregates.
__device__ __forceinline__ void res_restrict(const ResSm& s, const double* v, const uint32_t* aptr, uint32_t na, double* out) {
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp >= na) return;
    double t0 = 0, t1 = 0, t2 = 0;
    for (uint32_t i = aptr[warp] + lane; i < aptr[warp + 1]; i += 32) {
        const double2 zo = s.zoff[i];
        const double v0 = v[3 * i], v1 = v[3 * i + 1], v2 = v[3 * i + 2];
        t0 += v0; t1 += v1; t2 += zo.x * v0 + zo.y * v1 + v2;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        t0 += __shfl_xor_sync(0xffffffffu, t0, o); t1 += __shfl_xor_sync(0xffffffffu, t1, o); t2 += __shfl_xor_sync(0xffffffffu, t2, o);
    }
    if (lane == 0) { out[3 * warp] = t0; out[3 * warp + 1] = t1; out[3 * warp + 2] = t2; }
}

// total of the nb (<= 1024) per-block partial sums, every thread gets it: warp 0 loads them with all its loads in
// flight at once (lane-strided), adds in a fixed order and publishes the total through shared memory
__device__ __forceinline__ double res_total(const double* part, uint32_t nb, double* red) {
    if (threadIdx.x < 32) {
        double t = 0;
#pragma unroll 8
        for (uint32_t i = threadIdx.x; i < nb; i += 32) t += part[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (threadIdx.x == 0) red[0] = t;
    }
    __syncthreads();
    const double r = red[0];
    __syncthreads();
    return r;
}

// Coarse rows of this block's aggregates, prolongation, z = zj + Z yc (kept in zf, published to d.z) and the
// block's part of r'z -> d.part2[block].  zf holds zj = blockdiag(S)^-1 r on entry.
__device__ __forceinline__ void res_finish(const Dev& d, const ResSm& s, bool coarse, uint32_t p0, uint32_t np, uint32_t a0, uint32_t na) {
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nc = d.nc;
    if (coarse && na) {
        // each of the 3 na rows of Ai is split over `wpr` warps so that every lane has all its loads in flight at once
        const uint32_t rows = 3 * na, wpr = (kResThreads / 32) / rows;            // >= 2 (rows <= 12)
        double* part = s.red + 32;
        if (warp < rows * wpr) {
            const uint32_t rw = warp / wpr, sub = warp - rw * wpr;
            const double* row = d.Ai + (size_t)(3 * a0 + rw) * nc;
            double t = 0;
#pragma unroll 8
            for (uint32_t k = sub * 32 + lane; k < nc; k += 32 * wpr) t += row[k] * s.rc[k];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
            if (lane == 0) part[warp] = t;
        }
        __syncthreads();
    }
    double dot = 0;
    if (tid < 3 * np) {
        const uint32_t pl = tid / 3, q = tid - 3 * pl;
        double z = s.zf[tid];
        if (coarse) {
            // yc of the pose's aggregate: the partial row sums are added in warp order by every thread that needs them
            const uint32_t a = s.pagg[pl], wpr = (kResThreads / 32) / (3 * na);
            const double* part = s.red + 32 + 3 * a * wpr;
            double y0 = 0, y1 = 0, y2 = 0;
            for (uint32_t j = 0; j < wpr; ++j) { y0 += part[j]; y1 += part[wpr + j]; y2 += part[2 * wpr + j]; }
            z += q == 2 ? y2 : q == 0 ? y0 + s.zoff[pl].x * y2 : y1 + s.zoff[pl].y * y2;
        }
        s.zf[tid] = z; d.z[3ull * p0 + tid] = z;
        dot = s.r[tid] * z;
    }
    const double t = block_sum(dot, s.red);
    if (tid == 0) d.part2[blockIdx.x] = t;
}

// landmark-major pass over this block's landmarks: w_l = sum_e Hpl_e^T v_pose(e), v = va + beta vb gathered from
// global memory; MODE 0 stores u_l = Hll^-1 w_l to d.w_sum, MODE 1 stores delta_l = Hll^-1 (b_l - w_l) to d.dl
template <int MODE>
__device__ __forceinline__ void res_landmarks(const Dev& d, const ResSm& s, const double* va, const double* vb, double beta, uint32_t l0, uint32_t nl) {
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (uint32_t l = warp; l < nl; l += kResThreads / 32) {
        double w0 = 0, w1 = 0;
        for (uint32_t k = s.lptr[l] + lane; k < s.lptr[l + 1]; k += 32) {
            const uint32_t i = s.epos[k];
            double v0 = va[3ull * i], v1 = va[3ull * i + 1], v2 = va[3ull * i + 2];
            if (MODE == 0) { v0 += beta * vb[3ull * i]; v1 += beta * vb[3ull * i + 1]; v2 += beta * vb[3ull * i + 2]; }
            const double* h = s.hplL + 6 * k;
            w0 += h[0] * v0 + h[2] * v1 + h[4] * v2; w1 += h[1] * v0 + h[3] * v1 + h[5] * v2;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { w0 += __shfl_xor_sync(0xffffffffu, w0, o); w1 += __shfl_xor_sync(0xffffffffu, w1, o); }
        if (lane == 0) {
            const double* hi = s.hinvL + 3 * l;
            if (MODE == 0) { d.w_sum[2ull * (l0 + l)] = hi[0] * w0 + hi[1] * w1; d.w_sum[2ull * (l0 + l) + 1] = hi[1] * w0 + hi[2] * w1; }
            else { const double t0 = d.bl[2ull * (l0 + l)] - w0, t1 = d.bl[2ull * (l0 + l) + 1] - w1;
                   d.dl[2ull * (l0 + l)] = hi[0] * t0 + hi[1] * t1; d.dl[2ull * (l0 + l) + 1] = hi[1] * t0 + hi[2] * t1; }
        }
    }
}

// z_j = blockdiag_sb(S)^-1 r for the block's poses (sblock_smoother.cuh): dense symmetric mat-vec per sblock, the
// matrix streamed from L2 (column-major == row-major, so the threads of a warp -- consecutive rows -- read
// consecutive words).  Every row is split over G thread groups so that all 1024 threads have loads in flight; the
// partial sums are added in group order (deterministic).  `part` = G x 3 np doubles of scratch (s.rows).
__device__ __forceinline__ void res_smooth(const Dev& d, const ResSm& s, uint32_t np) {
    const uint32_t tid = threadIdx.x, nrows = 3 * np;
    if (nrows == 0) return;
    const uint32_t G = min(8u, (uint32_t)kResThreads / nrows);
    const uint32_t g = tid / nrows, row = tid - g * nrows;
    double* part = s.rows;
    if (g < G) {
        const uint32_t pl = row / 3;
        uint32_t ls, n; sb_range(np, 0, ls, n);
        uint32_t q = 0;
        if (pl >= n) { q = 1; sb_range(np, 1, ls, n); }
        const uint32_t dim = 3 * n, lr = row - 3 * ls;
        const uint32_t c0 = dim * g / G, c1 = dim * (g + 1) / G;
        const double* mat = d.sbinv + ((size_t)blockIdx.x * kSbPerBlock + q) * kSbLD * kSbLD + lr;
        const double* rv = s.r + 3 * ls;
        double t0 = 0, t1 = 0, t2 = 0, t3 = 0;
        uint32_t c = c0;
        for (; c + 4 <= c1; c += 4) {
            const double m0 = mat[(size_t)c * kSbLD], m1 = mat[(size_t)(c + 1) * kSbLD], m2 = mat[(size_t)(c + 2) * kSbLD], m3 = mat[(size_t)(c + 3) * kSbLD];
            t0 += m0 * rv[c]; t1 += m1 * rv[c + 1]; t2 += m2 * rv[c + 2]; t3 += m3 * rv[c + 3];
        }
        for (; c < c1; ++c) t0 += mat[(size_t)c * kSbLD] * rv[c];
        part[g * nrows + row] = (t0 + t1) + (t2 + t3);
    }
    __syncthreads();
    if (tid < nrows) {
        double z = 0;
        for (uint32_t k = 0; k < G; ++k) z += part[k * nrows + tid];
        s.zf[tid] = z;
    }
    __syncthreads();
}

// Grid barrier of the resident kernel (all blocks are co-resident: cooperative launch).  One counter that only
// grows: block b's thread 0 adds 1 with release semantics and polls with acquire loads until the counter reaches
// (barriers so far) x (blocks); bar.sync on both sides extends the ordering to the whole block.  About half the
// latency of cg::grid_group::sync() here (no second fence, no flip-bit arithmetic); d.gbar == nullptr falls back to it.
__device__ __forceinline__ void res_gsync(cg::grid_group& grid, uint32_t* ctr, uint32_t& target) {
    if (!ctr) { grid.sync(); return; }
    target += gridDim.x;
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(ctr), "r"(1u) : "memory");
        uint32_t v;
        do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory"); } while (v < target);
    }
    __syncthreads();
}

__global__ void __launch_bounds__(kResThreads, 1) k_pcg_resident(const Dev d) {
    extern __shared__ __align__(32) unsigned char smem_raw[];
    const ResSm s = res_carve(smem_raw);
    cg::grid_group grid = cg::this_grid();
    uint32_t gtarget = 0;
    const uint32_t tid = threadIdx.x, b = blockIdx.x, nb = gridDim.x;
    const uint32_t p0 = d.res_pose_cut[b], np = d.res_pose_cut[b + 1] - p0;
    const uint32_t l0 = d.res_lm_cut[b], nl = d.res_lm_cut[b + 1] - l0;
    const uint32_t a0 = d.res_agg_cut[b], na = d.res_agg_cut[b + 1] - a0;
    const uint32_t e0 = d.pose_ptr[p0], ne = d.pose_ptr[p0 + np] - e0;
    const uint32_t k0 = d.lmo_ptr[l0], nk = d.lmo_ptr[l0 + nl] - k0;
    const bool coarse = d.flags[F_COARSE] != 0;
    const uint32_t nc = d.nc;
    const double tol2 = d.scal[S_TOL2]; const int maxit = d.flags[F_MAXIT];
    double* pp[2] = {d.p, d.sp};                              // p of all poses, double-buffered across iterations

    // ---- once per solve: everything this block owns goes to shared memory ------------------------------------
    for (uint32_t i = tid; i < 6 * ne; i += kResThreads) { const uint32_t k = i / 6, q = i - 6 * k; s.hplP[i] = reinterpret_cast<const double*>(hpl_row(d, e0 + k) + (q >> 1) * hpl_stride(d))[q & 1u]; }
    for (uint32_t i = tid; i < ne; i += kResThreads) s.elm[i] = d.edge_lm[e0 + i];
    for (uint32_t i = tid; i < 6 * nk; i += kResThreads) { const uint32_t k = i / 6, q = i - 6 * k; s.hplL[i] = reinterpret_cast<const double*>(hpl_row(d, d.lmo_edge[k0 + k]) + (q >> 1) * hpl_stride(d))[q & 1u]; }
    for (uint32_t k = tid; k < nk; k += kResThreads) s.epos[k] = d.edge_pose[d.lmo_edge[k0 + k]];
    for (uint32_t i = tid; i <= np; i += kResThreads) s.eptr[i] = d.pose_ptr[p0 + i] - e0;
    for (uint32_t i = tid; i <= nl; i += kResThreads) s.lptr[i] = d.lmo_ptr[l0 + i] - k0;
    for (uint32_t i = tid; i <= na; i += kResThreads) s.aptr[i] = d.agg_ptr[a0 + i] - p0;
    for (uint32_t i = tid; i < 3 * nl; i += kResThreads) s.hinvL[i] = d.hinv[3ull * l0 + i];
    for (uint32_t i = tid; i < 6 * np; i += kResThreads) { s.minv[i] = d.minv[6ull * p0 + i]; s.hpp[i] = d.hpp[6ull * p0 + i]; }
    for (uint32_t i = tid; i < np; i += kResThreads) { s.zoff[i] = d.zoff[p0 + i]; s.pagg[i] = d.pose_agg[p0 + i] - a0; }
    for (uint32_t i = tid; i < 3 * np; i += kResThreads) {
        s.r[i] = d.rhs[3ull * p0 + i]; s.x[i] = 0; s.p[i] = 0; pp[0][3ull * p0 + i] = 0;
    }
    if (tid == 0) {                                           // ODOM neighbour lists of the block's poses (CSR)
        uint32_t n = 0;
        for (uint32_t i = 0; i < np; ++i) {
            s.onb_ptr[i] = n;
            n += (d.od_out_ptr[p0 + i + 1] - d.od_out_ptr[p0 + i]) + (d.od_in_ptr[p0 + i + 1] - d.od_in_ptr[p0 + i]);
        }
        s.onb_ptr[np] = n;
    }
    __syncthreads();
    if (tid < np) {
        uint32_t n = s.onb_ptr[tid];
        for (uint32_t k = d.od_out_ptr[p0 + tid]; k < d.od_out_ptr[p0 + tid + 1]; ++k, ++n) {
            s.onb_idx[n] = d.od_b[k]; s.onb_h[3 * n] = d.hoo[3ull * k]; s.onb_h[3 * n + 1] = d.hoo[3ull * k + 1]; s.onb_h[3 * n + 2] = d.hoo[3ull * k + 2];
        }
        for (uint32_t q = d.od_in_ptr[p0 + tid]; q < d.od_in_ptr[p0 + tid + 1]; ++q, ++n) {
            const uint32_t k = d.od_in_edge[q];
            s.onb_idx[n] = d.od_a[k]; s.onb_h[3 * n] = d.hoo[3ull * k]; s.onb_h[3 * n + 1] = d.hoo[3ull * k + 1]; s.onb_h[3 * n + 2] = d.hoo[3ull * k + 2];
        }
        if (!d.sbinv) sym3_mul(s.minv + 6 * tid, s.r[3 * tid], s.r[3 * tid + 1], s.r[3 * tid + 2], s.zf[3 * tid], s.zf[3 * tid + 1], s.zf[3 * tid + 2]);
    }
    __syncthreads();
    if (d.sbinv) res_smooth(d, s, np);
    // coarse residual Z^T r: this block's aggregates -> global, then every block keeps the whole vector
    if (coarse) res_restrict(s, s.r, s.aptr, na, d.yc + 3 * a0);
    res_gsync(grid, d.gbar, gtarget);
    if (coarse) { for (uint32_t k = tid; k < nc; k += kResThreads) s.rc[k] = d.yc[k]; __syncthreads(); }
    res_finish(d, s, coarse, p0, np, a0, na);
    res_gsync(grid, d.gbar, gtarget);

    double rz = 0, rz0 = 0, rz_last = 0;
    int it = 0, cur = 0;
    const bool timing = d.dbg != nullptr && b == 0 && tid == 0;      // TSB_PCG_TIMING=1: SM cycles of block 0 per phase
    long long tk = 0, tacc[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
#define RTICK(slot) do { if (timing) { const long long now_ = clock64(); tacc[slot] += now_ - tk; tk = now_; } } while (0)
    for (;;) {
        if (timing) tk = clock64();
        // ---- phase 1: beta, p = z + beta p (own poses), u_l = Hll^-1 Hpl^T p (own landmarks) -------------------
        const double rz_new = res_total(d.part2, nb, s.red);
        rz_last = rz_new;
        double beta = 0;
        if (it == 0) { rz0 = rz_new; if (!(rz0 > 0)) break; }
        else { beta = rz_new / rz; if (!(rz_new > tol2 * rz0) || it >= maxit) break; }
        rz = rz_new;
        const double* pold = pp[cur]; double* pnew = pp[cur ^ 1];
        if (tid < 3 * np) { const double v = fma(beta, s.p[tid], s.zf[tid]); s.p[tid] = v; pnew[3ull * p0 + tid] = v; }
        {   // same arithmetic as the owner's line above, on gathered z and p_old
            const uint32_t warp = tid >> 5, lane = tid & 31;
            for (uint32_t l = warp; l < nl; l += kResThreads / 32) {
                double w0 = 0, w1 = 0;
                for (uint32_t k = s.lptr[l] + lane; k < s.lptr[l + 1]; k += 32) {
                    const uint32_t i = s.epos[k];
                    const double v0 = fma(beta, pold[3ull * i], d.z[3ull * i]), v1 = fma(beta, pold[3ull * i + 1], d.z[3ull * i + 1]);
                    const double v2 = fma(beta, pold[3ull * i + 2], d.z[3ull * i + 2]);
                    const double* h = s.hplL + 6 * k;
                    w0 += h[0] * v0 + h[2] * v1 + h[4] * v2; w1 += h[1] * v0 + h[3] * v1 + h[5] * v2;
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) { w0 += __shfl_xor_sync(0xffffffffu, w0, o); w1 += __shfl_xor_sync(0xffffffffu, w1, o); }
                if (lane == 0) {
                    const double* hi = s.hinvL + 3 * l;
                    d.w_sum[2ull * (l0 + l)] = hi[0] * w0 + hi[1] * w1; d.w_sum[2ull * (l0 + l) + 1] = hi[1] * w0 + hi[2] * w1;
                }
            }
        }
        cur ^= 1;
        RTICK(0);
        res_gsync(grid, d.gbar, gtarget);
        RTICK(1);
        // ---- phase 2: y = S p for the own poses, p'y and Z^T y ---------------------------------------------------
        const uint32_t pl = tid / 3, q = tid - 3 * pl;
        {   // ODOM neighbours' p: one thread per (entry, component), all loads in flight at once
            const double* pg = pp[cur];
            const uint32_t nent = 3 * s.onb_ptr[np];
            for (uint32_t i = tid; i < nent; i += kResThreads) { const uint32_t n = i / 3; s.onb_p[i] = pg[3ull * s.onb_idx[n] + (i - 3 * n)]; }
        }
        for (uint32_t e = tid; e < ne; e += kResThreads) {
            const uint32_t l = s.elm[e];
            const double u0 = d.w_sum[2ull * l], u1 = d.w_sum[2ull * l + 1];
            const double* h = s.hplP + 6 * e;
            s.rows[e] = h[0] * u0 + h[1] * u1; s.rows[kResEC + e] = h[2] * u0 + h[3] * u1; s.rows[2 * kResEC + e] = h[4] * u0 + h[5] * u1;
        }
        RTICK(8);
        __syncthreads();
        RTICK(9);
        double dot = 0;
        if (tid < 3 * np) {
            const double acc = seg_sum_lean(s.rows + q * kResEC, s.eptr[pl], s.eptr[pl + 1]);
            const double* H = s.hpp + 6 * pl; const double* v = s.p + 3 * pl;
            const double hp = q == 0 ? H[0] * v[0] + H[1] * v[1] + H[2] * v[2] : q == 1 ? H[1] * v[0] + H[3] * v[1] + H[4] * v[2]
                                                                                        : H[2] * v[0] + H[4] * v[1] + H[5] * v[2];
            double yo = 0;                                         // ODOM blocks are diagonal: component-wise
            for (uint32_t n = s.onb_ptr[pl]; n < s.onb_ptr[pl + 1]; ++n) yo += s.onb_h[3 * n + q] * s.onb_p[3 * n + q];
            const double y = hp + yo - acc;
            s.y[tid] = y; dot = s.p[tid] * y;
        }
        RTICK(10);
        const double t = block_sum(dot, s.red);                    // (its barriers also publish s.y to the block)
        RTICK(11);
        if (tid == 0) d.part[b] = t;
        if (coarse) res_restrict(s, s.y, s.aptr, na, d.cdot + 3 * a0);
        RTICK(2);
        res_gsync(grid, d.gbar, gtarget);
        RTICK(3);
        // ---- phase 3: alpha, x, r, coarse residual by recurrence, z and r'z ----------------------------------------
        const double alpha = rz / res_total(d.part, nb, s.red);
        RTICK(12);
        if (coarse) for (uint32_t k = tid; k < nc; k += kResThreads) s.rc[k] -= alpha * d.cdot[k];
        if (tid < 3 * np) { s.x[tid] += alpha * s.p[tid]; s.r[tid] -= alpha * s.y[tid]; }
        __syncthreads();
        if (d.sbinv) res_smooth(d, s, np);
        else {
            if (tid < np) sym3_mul(s.minv + 6 * tid, s.r[3 * tid], s.r[3 * tid + 1], s.r[3 * tid + 2], s.zf[3 * tid], s.zf[3 * tid + 1], s.zf[3 * tid + 2]);
            __syncthreads();
        }
        RTICK(13);
        res_finish(d, s, coarse, p0, np, a0, na);
        ++it;
        RTICK(4);
        res_gsync(grid, d.gbar, gtarget);
        RTICK(5);
    }
    if (timing) for (int q = 0; q < 16; ++q) d.dbg[q] = tacc[q];
#undef RTICK
    // ---- results: delta_p, the PCG scalars, landmark back-substitution delta_l = Hll^-1 (b_l - Hpl^T delta_p) ----
    if (tid < 3 * np) d.x[3ull * p0 + tid] = s.x[tid];
    if (b == 0 && tid == 0) { d.scal[S_RZ0] = rz0; d.scal[S_RZ] = rz_last; d.flags[F_ITERS] = it; d.flags[F_DONE] = 1; }
    res_gsync(grid, d.gbar, gtarget);
    res_landmarks<1>(d, s, d.x, nullptr, 0.0, l0, nl);
}
