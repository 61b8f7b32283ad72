// sblock_smoother.cuh -- fine level of the two-level preconditioner of the shared-memory-resident PCG kernel:
//     M^-1 = blockdiag_sb(S)^-1 + Z (Z^T S Z)^-1 Z^T
// with "smoother blocks" (sblocks) of <= kSbCap consecutive poses instead of the 3x3 pose blocks of block-Jacobi.
// An sblock is the EXACT diagonal block of the Schur complement S = Hpp - Hpl Hll^-1 Hpl^T over its poses (consecutive
// poses see the same landmarks, so it is dense), inverted explicitly; the PCG applies it as a dense mat-vec.
// C2, first solve: 247 PCG iterations with 3x3 blocks -> 101 with sblocks (tools/precond_proto.py --schwarz predicted 102),
// at ~1.27x the time per iteration (24 MB of inverses stream from L2 per iteration) plus this kernel (0.26 ms); once the
// Huber weights have settled the iteration counts meet (113 vs 100 averaged over a 50-iteration request), so
// tsb_optimize starts a request with sblocks and drops them when a probe solve with 3x3 blocks is cheaper
// (TSB_SMOOTHER=1 / 0: always / never).  C2, 50 GN iterations: 420 -> 441 GN it/s.
//
// Resident block b (poses [res_pose_cut[b], res_pose_cut[b+1])) is cut into msb = ceil(np / kSbCap) <= kSbPerBlock
// equal sblocks; sblock q of block b is stored at sbinv + (b kSbPerBlock + q) kSbLD^2, leading dimension kSbLD,
// symmetric (row- and column-major are the same).
//
// k_sb_build: one thread block per sblock.
//   1. assembly in shared memory: thread per pose pair (i <= j): Hpp diagonal / ODOM blocks, minus the landmark
//      terms found by merging the two poses' edge lists (sorted by landmark).  One thread per entry block, fixed
//      order: deterministic, no atomics.
//   2. Gauss-Jordan without pivoting (the block is SPD) with the matrix in REGISTERS: thread (ty, tx) of 16 x 32 owns
//      rows ty + 16 a, columns tx + 32 b (9 x 5 tile); per pivot the owners publish row k (scaled) and column k through
//      double-buffered shared memory, one block barrier per pivot.  (The first version kept the matrix in shared
//      memory: 75 % of its 474 us were short-scoreboard stalls of the n^3 update loop, profiles/r1_sb_build_c2_source.txt.)
//      A non-positive pivot falls back to the 3x3 pose blocks of d.minv for this sblock.
//   3. symmetrised write to global memory.
// Included inside the anonymous namespace of kernels.cu.

constexpr uint32_t kSbLD = 3 * kSbCap;
constexpr int kSbThreads = 512;          // 16 x 32 threads, each owns a 9 x 5 register tile of the (padded) block during the inversion
constexpr uint32_t kSbColPad = 160;      // columns covered by 32 lanes x 5
constexpr size_t kSbSmemBytes = sizeof(double) * ((size_t)kSbLD * (kSbLD + 1) + 2 * kSbLD + 2 * kSbColPad + 8);
static_assert(kSbLD <= 16 * 9 && kSbLD <= kSbColPad, "register tile of the Gauss-Jordan step");
static_assert(kResPoseCap <= kSbPerBlock * kSbCap, "a resident block's poses must fit its sblocks");

// pose range [ls, ls + n) (block-local) of sblock q of a resident block with np poses
__device__ __forceinline__ void sb_range(uint32_t np, uint32_t q, uint32_t& ls, uint32_t& n) {
    const uint32_t msb = (np + kSbCap - 1) / kSbCap;
    if (q >= msb) { ls = np; n = 0; return; }
    ls = (uint32_t)((uint64_t)np * q / msb); n = (uint32_t)((uint64_t)np * (q + 1) / msb) - ls;
}

__global__ void __launch_bounds__(kSbThreads, 1) k_sb_build(const Dev d) {
    extern __shared__ __align__(32) unsigned char smem_raw[];
    double* A = reinterpret_cast<double*>(smem_raw);           // n x n, leading dimension ld = dim + 1 (odd: rows in different banks)
    const uint32_t b = blockIdx.x / kSbPerBlock, q = blockIdx.x % kSbPerBlock, tid = threadIdx.x;
    const uint32_t p0 = d.res_pose_cut[b], np = d.res_pose_cut[b + 1] - p0;
    uint32_t ls, n; sb_range(np, q, ls, n);
    if (n == 0) return;
    const uint32_t ps = p0 + ls, dim = 3 * n, ld = dim + 1;
    double* colk = A + (size_t)kSbLD * (kSbLD + 1); double* rowk = colk + 2 * kSbLD;      // two buffers each
    double* sh_ip = rowk + 2 * kSbColPad; int* bad = reinterpret_cast<int*>(sh_ip + 2);
    if (tid == 0) *bad = 0;
    // ---- 1. assembly ------------------------------------------------------------------------------------------
    // pose pairs i <= j: rows r and n - 1 - r of the upper triangle together hold n + 1 pairs
    for (uint32_t t = tid; t < ((n + 1) / 2) * (n + 1); t += kSbThreads) {
        const uint32_t r = t / (n + 1), c = t - r * (n + 1);
        uint32_t i, j;
        if (c < n - r) { i = r; j = r + c; }
        else { i = n - 1 - r; j = i + (c - (n - r)); if (i == r) continue; }
        const uint32_t pi = ps + i, pj = ps + j;
        double m[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        if (i == j) { const double* H = d.hpp + 6ull * pi; m[0] = H[0]; m[1] = m[3] = H[1]; m[2] = m[6] = H[2]; m[4] = H[3]; m[5] = m[7] = H[4]; m[8] = H[5]; }
        for (uint32_t k = d.od_out_ptr[pi]; k < d.od_out_ptr[pi + 1]; ++k)
            if (d.od_b[k] == pj) { m[0] += d.hoo[3ull * k]; m[4] += d.hoo[3ull * k + 1]; m[8] += d.hoo[3ull * k + 2]; }
        for (uint32_t k2 = d.od_in_ptr[pi]; k2 < d.od_in_ptr[pi + 1]; ++k2) {
            const uint32_t k = d.od_in_edge[k2];
            if (d.od_a[k] == pj) { m[0] += d.hoo[3ull * k]; m[4] += d.hoo[3ull * k + 1]; m[8] += d.hoo[3ull * k + 2]; }
        }
        uint32_t ei = d.pose_ptr[pi], ej = d.pose_ptr[pj];
        const uint32_t ei1 = d.pose_ptr[pi + 1], ej1 = d.pose_ptr[pj + 1];
        while (ei < ei1 && ej < ej1) {
            const uint32_t li = d.edge_lm[ei], lj = d.edge_lm[ej];
            if (li < lj) { ++ei; continue; }
            if (lj < li) { ++ej; continue; }
            // landmark li seen by both poses: sum each pose's blocks for it (a pose may see a landmark more than once)
            double a[6] = {0, 0, 0, 0, 0, 0}, c[6] = {0, 0, 0, 0, 0, 0};
            for (; ei < ei1 && d.edge_lm[ei] == li; ++ei) {
                const double2* h = hpl_row(d, ei); const double2 h0 = h[0], h1 = h[hpl_stride(d)], h2 = h[2 * hpl_stride(d)];
                a[0] += h0.x; a[1] += h0.y; a[2] += h1.x; a[3] += h1.y; a[4] += h2.x; a[5] += h2.y;
            }
            for (; ej < ej1 && d.edge_lm[ej] == li; ++ej) {
                const double2* h = hpl_row(d, ej); const double2 h0 = h[0], h1 = h[hpl_stride(d)], h2 = h[2 * hpl_stride(d)];
                c[0] += h0.x; c[1] += h0.y; c[2] += h1.x; c[3] += h1.y; c[4] += h2.x; c[5] += h2.y;
            }
            const double* hi = d.hinv + 3ull * li;
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const double t0 = a[2 * r] * hi[0] + a[2 * r + 1] * hi[1], t1 = a[2 * r] * hi[1] + a[2 * r + 1] * hi[2];
#pragma unroll
                for (int cc = 0; cc < 3; ++cc) m[3 * r + cc] -= t0 * c[2 * cc] + t1 * c[2 * cc + 1];
            }
        }
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int cc = 0; cc < 3; ++cc) {
                A[(size_t)(3 * i + r) * ld + 3 * j + cc] = m[3 * r + cc];
                if (i != j) A[(size_t)(3 * j + cc) * ld + 3 * i + r] = m[3 * r + cc];
            }
    }
    __syncthreads();
    // ---- 2. Gauss-Jordan, no pivoting, matrix in registers ------------------------------------------------------
    {
        const uint32_t ty = tid >> 5, tx = tid & 31;
        double R[9][5];
#pragma unroll
        for (int a = 0; a < 9; ++a)
#pragma unroll
            for (int b = 0; b < 5; ++b) {
                const uint32_t i = ty + 16 * a, j = tx + 32 * b;
                R[a][b] = (i < dim && j < dim) ? A[(size_t)i * ld + j] : (i == j ? 1.0 : 0.0);     // identity padding stays inert
            }
#pragma unroll
        for (int a0 = 0; a0 < 9; ++a0) {
            if (16u * a0 >= dim) break;
            const int b0 = a0 >> 1;                                  // k / 32 for k = 16 a0 + kk, kk < 16
#pragma unroll 1
            for (uint32_t kk = 0; kk < 16; ++kk) {
                const uint32_t k = 16u * a0 + kk;
                if (k >= dim) break;
                double* ck = colk + (k & 1u) * kSbLD; double* rk = rowk + (k & 1u) * kSbColPad;
                if (ty == kk) {                                      // this warp owns row k; lane k % 32 holds the pivot
                    const double piv = __shfl_sync(0xffffffffu, R[a0][b0], k & 31u);
                    const double ipv = 1.0 / piv;
                    if (tx == 0) { sh_ip[k & 1u] = ipv; if (!(piv > 0) || !(piv < 1e300)) *bad = 1; }
#pragma unroll
                    for (int b = 0; b < 5; ++b) rk[tx + 32 * b] = R[a0][b] * ipv;
                }
                if (tx == (k & 31u)) {
#pragma unroll
                    for (int a = 0; a < 9; ++a) ck[ty + 16 * a] = R[a][b0];
                }
                __syncthreads();
                const double ipv = sh_ip[k & 1u];
                double rj[5];
#pragma unroll
                for (int b = 0; b < 5; ++b) rj[b] = rk[tx + 32 * b];
#pragma unroll
                for (int a = 0; a < 9; ++a) {
                    if (16u * a >= dim) break;                       // tiles beyond the block are identity padding (block-uniform)
                    const uint32_t i = ty + 16 * a;
                    const double ci = ck[i];
#pragma unroll
                    for (int b = 0; b < 5; ++b) {
                        if (32u * b >= dim) break;
                        // row k lives in tile row a0 of warp kk, column k in tile column b0 of lane k % 32: a, a0, b, b0 are
                        // compile-time in the unrolled loops, so only 1/9 and 1/5 of the tile entries carry the selects
                        double v = R[a][b] - ci * rj[b];
                        if (b == b0 && tx == (k & 31u)) v = -ci * ipv;
                        if (a == a0 && ty == kk) v = (b == b0 && tx == (k & 31u)) ? ipv : rj[b];
                        R[a][b] = v;
                    }
                }
            }
        }
        __syncthreads();
#pragma unroll
        for (int a = 0; a < 9; ++a)
#pragma unroll
            for (int b = 0; b < 5; ++b) {
                const uint32_t i = ty + 16 * a, j = tx + 32 * b;
                if (i < dim && j < dim) A[(size_t)i * ld + j] = R[a][b];
            }
    }
    __syncthreads();
    // ---- 3. publish ------------------------------------------------------------------------------------------------
    double* out = d.sbinv + (size_t)blockIdx.x * kSbLD * kSbLD;
    if (*bad) {
        for (uint32_t t = tid; t < dim * dim; t += kSbThreads) {
            const uint32_t i = t / dim, j = t - i * dim;
            double v = 0;
            if (i / 3 == j / 3) {
                const double* M = d.minv + 6ull * (ps + i / 3); const uint32_t r = i % 3, c = j % 3;
                const uint32_t lo = r < c ? r : c, hi2 = r < c ? c : r;
                v = M[lo == 0 ? hi2 : (lo == 1 ? 2 + hi2 : 5)];
            }
            out[(size_t)i * kSbLD + j] = v;
        }
        return;
    }
    for (uint32_t i = tid >> 5; i < dim; i += kSbThreads / 32)
        for (uint32_t j = tid & 31; j < dim; j += 32)
            out[(size_t)i * kSbLD + j] = 0.5 * (A[(size_t)i * ld + j] + A[(size_t)j * ld + i]);
}
