// coarse_invert.cuh -- (Z^T S Z)^-1 in ONE cooperative launch.
//
// The coarse operator of the two-level preconditioner is a dense SPD matrix of order nc <= 1536 that is re-inverted
// every few GN iterations.  cuSOLVER potrf + potrs against the identity takes ~1.0 ms at nc = 939: ~54 launches of
// panel / trsm / gemm kernels that are each far too small for the machine.  Here: blocked Gauss-Jordan without
// pivoting (the matrix is SPD: every pivot block is SPD), tile 32, the whole grid in one kernel, one grid barrier
// per block step.  Step k reads X and writes Y (ping-pong, so no tile is read and written in the same step):
//     D = X[k,k]^-1      (32x32, every block inverts it redundantly in shared memory: cheaper than a barrier)
//     Y[k,k] = D,  Y[k,j] = D X[k,j],  Y[i,k] = -X[i,k] D,  Y[i,j] = X[i,j] - X[i,k] D X[k,j]      (i, j != k)
// A block owns column strip j and a slice of the rows: it forms P = D X[k,j] once and reuses it for its tiles.
// Fixed summation order everywhere: deterministic.  A non-positive pivot (not SPD / NaN) clears F_COARSE and the
// solve falls back to block-Jacobi, as with a failed potrf.
//
// Included inside the anonymous namespace of kernels.cu (after pcg_resident.cuh: uses res_gsync).

constexpr int kCiThreads = 1024, kCiSlots = 8;       // 8 tiles at a time, 128 threads (4 x 2 outputs each) per tile
constexpr int kCiLs = 36, kCiPs = 34;                // row strides (doubles) of the transposed-L and P tiles: 16-B aligned rows
constexpr uint32_t kCiOffD = 0;                                          // double[2][32][33]
constexpr uint32_t kCiOffP = kCiOffD + 2 * 32 * 33 * 8;                  // double[32][34]   P = D X[k,j]
constexpr uint32_t kCiOffB = kCiOffP + 32 * kCiPs * 8;                   // double[32][34]   X[k,j]
constexpr uint32_t kCiOffL = (kCiOffB + 32 * kCiPs * 8 + 15) & ~15u;     // double[8][32][36] L tiles, transposed
constexpr uint32_t kCiSmemBytes = kCiOffL + kCiSlots * 32 * kCiLs * 8;

// acc[r][c] = sum_q Lt[q][4 ty4 + r] * B[q][2 tx2 + c]: a 4 x 2 register tile per thread, 3 shared-memory wavefronts
// per 8 FMAs of a warp
__device__ __forceinline__ void ci_tile_mm(const double* __restrict__ Lt, const double* __restrict__ B, uint32_t ty4, uint32_t tx2,
                                           double acc[4][2]) {
#pragma unroll
    for (int r = 0; r < 4; ++r) { acc[r][0] = 0; acc[r][1] = 0; }
#pragma unroll 8
    for (int q = 0; q < 32; ++q) {
        const double2 a01 = *reinterpret_cast<const double2*>(Lt + q * kCiLs + 4 * ty4);
        const double2 a23 = *reinterpret_cast<const double2*>(Lt + q * kCiLs + 4 * ty4 + 2);
        const double2 b = *reinterpret_cast<const double2*>(B + q * kCiPs + 2 * tx2);
        acc[0][0] += a01.x * b.x; acc[0][1] += a01.x * b.y; acc[1][0] += a01.y * b.x; acc[1][1] += a01.y * b.y;
        acc[2][0] += a23.x * b.x; acc[2][1] += a23.x * b.y; acc[3][0] += a23.y * b.x; acc[3][1] += a23.y * b.y;
    }
}

__global__ void __launch_bounds__(kCiThreads, 1) k_coarse_invert(const Dev d) {
    extern __shared__ __align__(16) unsigned char ci_smem[];
    double (*sD)[32][33] = reinterpret_cast<double (*)[32][33]>(ci_smem + kCiOffD);
    double* sP = reinterpret_cast<double*>(ci_smem + kCiOffP);
    double* sB = reinterpret_cast<double*>(ci_smem + kCiOffB);
    double* sLt = reinterpret_cast<double*>(ci_smem + kCiOffL);
    cg::grid_group grid = cg::this_grid();
    const uint32_t n = d.nc, nb = (n + 31) / 32, G = gridDim.x;
    const uint32_t tx = threadIdx.x & 31, ty = threadIdx.x >> 5;        // element of a 32 x 32 tile (inversion, staging)
    const uint32_t slot = threadIdx.x >> 7, t = threadIdx.x & 127, ty4 = t >> 4, tx2 = t & 15;   // tile slot and 4 x 2 micro-tile
    const double* X = d.Ac; double* Y = d.Ai;
    uint32_t gtarget = 0;
    bool ok = true;
    const uint32_t slices = G / nb ? G / nb : 1;
    const uint32_t j = blockIdx.x % nb, sl = blockIdx.x / nb;
    const bool active = sl < slices;
    const uint32_t i_lo = nb * sl / slices, i_hi = active ? nb * (sl + 1) / slices : i_lo;
    // element (r, c) of tile (bi, bj); outside the matrix the padding is the identity
    auto ld = [&](const double* M, uint32_t bi, uint32_t bj, uint32_t r, uint32_t c) -> double {
        const uint32_t gr = 32 * bi + r, gc = 32 * bj + c;
        return (gr < n && gc < n) ? M[(size_t)gr * n + gc] : (gr == gc ? 1.0 : 0.0);
    };
    double* myL = sLt + slot * 32 * kCiLs;
    for (uint32_t k = 0; k < nb; ++k) {
        // ---- D = X[k,k]^-1: scalar Gauss-Jordan, one thread per element, one barrier per pivot ----------------
        const double xkk = ld(X, k, k, ty, tx);
        const double xkj = (active && j != k) ? ld(X, k, j, ty, tx) : 0.0;      // in flight during the inversion
        sD[0][ty][tx] = xkk;
        sB[ty * kCiPs + tx] = xkj;
        __syncthreads();
        int cur = 0;
#pragma unroll 1
        for (int p = 0; p < 32; ++p) {
            const double pv = sD[cur][p][p];
            if (!(pv > 0.0)) ok = false;
            const double piv = 1.0 / pv;
            const double rp = sD[cur][ty][p], pc = sD[cur][p][tx];
            double v;
            if ((int)ty == p) v = (int)tx == p ? piv : pc * piv;
            else if ((int)tx == p) v = -rp * piv;
            else v = sD[cur][ty][tx] - rp * (pc * piv);
            sD[cur ^ 1][ty][tx] = v;
            __syncthreads();
            cur ^= 1;
        }
        if (active) {
            // ---- P = D X[k,j]  (P = D for the pivot column) --------------------------------------------------
            if (j == k) sP[ty * kCiPs + tx] = sD[cur][ty][tx];
            else {
                sLt[tx * kCiLs + ty] = sD[cur][ty][tx];                 // slot 0 holds D transposed
                __syncthreads();
                if (slot == 0) {
                    double acc[4][2]; ci_tile_mm(sLt, sB, ty4, tx2, acc);
#pragma unroll
                    for (int r = 0; r < 4; ++r) { sP[(4 * ty4 + r) * kCiPs + 2 * tx2] = acc[r][0]; sP[(4 * ty4 + r) * kCiPs + 2 * tx2 + 1] = acc[r][1]; }
                }
            }
            __syncthreads();
            // ---- this block's tiles of column strip j, eight at a time -------------------------------------
#pragma unroll 1
            for (uint32_t i0 = i_lo; i0 < i_hi; i0 += kCiSlots) {
                const uint32_t i = i0 + slot;
                const bool mine = i < i_hi;
                double xo[4][2];
                if (mine && i != k) {
                    // L = X[i,k] -> transposed into this slot (coalesced global reads along a row)
#pragma unroll
                    for (int e = 0; e < 8; ++e) { const uint32_t idx = t + 128 * e, r = idx >> 5, q = idx & 31; myL[q * kCiLs + r] = ld(X, i, k, r, q); }
                    if (j != k) {
#pragma unroll
                        for (int r = 0; r < 4; ++r) { xo[r][0] = ld(X, i, j, 4 * ty4 + r, 2 * tx2); xo[r][1] = ld(X, i, j, 4 * ty4 + r, 2 * tx2 + 1); }
                    }
                }
                __syncthreads();
                if (mine) {
                    double acc[4][2];
                    if (i == k) {
#pragma unroll
                        for (int r = 0; r < 4; ++r) { acc[r][0] = sP[(4 * ty4 + r) * kCiPs + 2 * tx2]; acc[r][1] = sP[(4 * ty4 + r) * kCiPs + 2 * tx2 + 1]; }
                    } else {
                        ci_tile_mm(myL, sP, ty4, tx2, acc);
#pragma unroll
                        for (int r = 0; r < 4; ++r) { acc[r][0] = j == k ? -acc[r][0] : xo[r][0] - acc[r][0]; acc[r][1] = j == k ? -acc[r][1] : xo[r][1] - acc[r][1]; }
                    }
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        const uint32_t gr = 32 * i + 4 * ty4 + r, gc = 32 * j + 2 * tx2;
                        if (gr < n && gc < n) Y[(size_t)gr * n + gc] = acc[r][0];
                        if (gr < n && gc + 1 < n) Y[(size_t)gr * n + gc + 1] = acc[r][1];
                    }
                }
                __syncthreads();
            }
        }
        res_gsync(grid, d.cbar, gtarget);
        const double* tmp = X; X = Y; Y = const_cast<double*>(tmp);
    }
    // ---- the inverse is in X: symmetrise into Ai (pairwise, valid in place), publish the status ------------------
    const size_t tot = (size_t)n * n;
    for (size_t idx = (size_t)blockIdx.x * kCiThreads + threadIdx.x; idx < tot; idx += (size_t)G * kCiThreads) {
        const size_t r = idx / n, c = idx - r * n;
        if (r > c) continue;
        const double v = 0.5 * (X[r * n + c] + X[c * n + r]);
        d.Ai[r * n + c] = v; d.Ai[c * n + r] = v;
    }
    const int all_ok = __syncthreads_and(ok ? 1 : 0);
    if (blockIdx.x == 0 && threadIdx.x == 0) d.flags[F_COARSE] = all_ok;
}
