// mlevel.cuh -- multilevel preconditioner of the streamed PCG (HBM-sized graphs; replaces SolverEigen.h:20 together
// with pcg_stream.cuh).
//
//   M^-1 = blockdiag(S)^-1 + sum_{l < top} Z_l blockdiag(S_l)^-1 Z_l^T + Z_top S_top^-1 Z_top^T         (additive, "BPX")
//
// Level l has one node per `spacing` consecutive poses.  Z_l interpolates a node vector (dx, dy, dtheta: a rigid-body
// motion about the node's pose) LINEARLY along the trajectory: pose i between nodes q and q+1 takes (1-t) of node q
// and t of node q+1, t = (i mod spacing) / spacing.  Landmarks stay on every level, so S_l = Z_l^T S Z_l is
// App_l - C_l Hll^-1 C_l^T with C_l = Z_l^T Hpl held as one 3x2 block per distinct (node, landmark) pair, summed
// hierarchically from the level below (k_ml_C).  Only the diagonal blocks of S_l are used below the top level; the top
// level (<= a few thousand unknowns) is assembled densely (k_ml_top) and inverted by k_coarse_invert.
//   Measured on the CPU prototype (tools/mg_proto.py, first GN iteration, 1e-12): C3 550 -> 174 PCG iterations against
// the two-level method with piecewise-constant aggregates and the same dense size; the dense inverse shrinks 4x.
//
// Rigid-body transfer without per-pair offsets: with m_i = (r_x, r_y, r_t - y_i r_x + x_i r_y) (force and moment about the
// ORIGIN) the restriction is a scalar hat-weighted sum of m, corrected by the node's position when it is consumed;
// likewise a node correction e is prolongated as g = (e_x + Y e_t, e_y - X e_t, e_t) with scalar weights and turned
// into the pose's motion by (g_x - y_i g_t, g_y + x_i g_t, g_t).
// Every sum runs in a fixed order: deterministic, no atomics.
//
// Included inside the anonymous namespace of kernels.cu (after coarse_invert.cuh, before pcg_stream.cuh).

// hat weight of the child at offset j = child - ratio * K for parent K of a level with n nodes (the last node keeps
// weight 1 to its right: constant extrapolation)
__device__ __forceinline__ double ml_weight(int j, uint32_t ratio, uint32_t K, uint32_t n) {
    if (j >= 0) return (K + 1 == n) ? 1.0 : 1.0 - (double)j / (double)ratio;
    return (double)((int)ratio + j) / (double)ratio;
}
// weight of node K (level: spacing, n nodes) at pose / position `pos`
__device__ __forceinline__ double ml_hat_at(uint32_t pos, uint32_t K, uint32_t spacing, uint32_t n) {
    const uint32_t q = pos / spacing, rem = pos - q * spacing;
    if (q + 1 >= n || rem == 0) return K == q ? 1.0 : 0.0;
    const double t = (double)rem / (double)spacing;
    return K == q ? 1.0 - t : (K == q + 1 ? t : 0.0);
}

// ---- setup: C_l ---------------------------------------------------------------------------------------------------
// warp per node; a lane per coarse edge sums its child blocks (fixed order: ascending child edge index)
template <bool FIRST>
__global__ void __launch_bounds__(256) k_ml_C(const Dev d, const int l) {
    const MlDev& lv = d.ml[l];
    const uint32_t K = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (K >= lv.n) return;
    const uint32_t cs = FIRST ? 1u : d.ml[FIRST ? 0 : l - 1].spacing;
    const double2* ch = FIRST ? reinterpret_cast<const double2*>(d.hpl) : reinterpret_cast<const double2*>(d.ml[FIRST ? 0 : l - 1].C);
    const size_t chs = FIRST ? (size_t)d.nE : (size_t)d.ml[FIRST ? 0 : l - 1].nnz;
    const double4 pc = d.pose[(size_t)K * lv.spacing];
    double2* out = reinterpret_cast<double2*>(lv.C);
    const size_t os = lv.nnz;
    for (uint32_t e = lv.node_ptr[K] + lane; e < lv.node_ptr[K + 1]; e += 32) {
        double c0 = 0, c1 = 0, c2 = 0, c3 = 0, c4 = 0, c5 = 0;
        for (uint32_t q = lv.ref_ptr[e]; q < lv.ref_ptr[e + 1]; ++q) {
            const uint32_t rf = lv.ref[q], ce = rf >> 4;
            const int j = (int)(rf & 15u) - 7;
            const double w = ml_weight(j, lv.ratio, K, lv.n);
            const double4 pk = d.pose[(size_t)((int)(lv.ratio * K) + j) * cs];
            const double zx = -(pk.y - pc.y), zy = pk.x - pc.x;
            const double2 h0 = ch[ce], h1 = ch[chs + ce], h2 = ch[2 * chs + ce];
            c0 += w * h0.x; c1 += w * h0.y; c2 += w * h1.x; c3 += w * h1.y;
            c4 += w * (zx * h0.x + zy * h1.x + h2.x); c5 += w * (zx * h0.y + zy * h1.y + h2.y);
        }
        out[e] = make_double2(c0, c1); out[os + e] = make_double2(c2, c3); out[2 * os + e] = make_double2(c4, c5);
    }
}

// hat weight of local pose i for node K; support is ((K-1) s, (K+1) s), the last node extends to the end
__device__ __forceinline__ double ml_pose_weight(uint32_t i, uint32_t K, uint32_t s, uint32_t n) {
    const uint32_t c = K * s;
    if (i >= c) return (K + 1 == n) ? 1.0 : 1.0 - (double)(i - c) / (double)s;
    return (double)(i + s - c) / (double)s;
}
__device__ __forceinline__ void ml_support(const Dev& d, uint32_t K, uint32_t s, uint32_t n, uint32_t& lo, uint32_t& hi) {
    lo = K == 0 ? 0u : (K - 1) * s + 1;
    hi = (K + 1 == n) ? d.nP : min(d.nP, (K + 1) * s);          // exclusive
}

// ---- setup: blockdiag(S_l)^-1, warp per node ------------------------------------------------------------------------
//   D_K = sum_i w_i^2 R_i^T Hpp_ii R_i + sum_{ODOM (i,j)} w_i w_j (R_i^T Hoo R_j + transpose) - sum_e C_e Hll^-1 C_e^T
__global__ void __launch_bounds__(256) k_ml_D(const Dev d, const int l) {
    const MlDev& lv = d.ml[l];
    const uint32_t K = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (K >= lv.n) return;
    const uint32_t s = lv.spacing;
    const double4 pc = d.pose[(size_t)K * s];
    uint32_t lo, hi; ml_support(d, K, s, lv.n, lo, hi);
    double a00 = 0, a01 = 0, a02 = 0, a11 = 0, a12 = 0, a22 = 0;
    for (uint32_t i = lo + lane; i < hi; i += 32) {
        const double wi = ml_pose_weight(i, K, s, lv.n);
        const double4 pi = d.pose[i];
        const double zx = -(pi.y - pc.y), zy = pi.x - pc.x;
        const double* H = d.hpp + 6ull * i;
        const double v0 = H[0] * zx + H[1] * zy + H[2], v1 = H[1] * zx + H[3] * zy + H[4], v2 = H[2] * zx + H[4] * zy + H[5];
        const double w2 = wi * wi;
        a00 += w2 * H[0]; a01 += w2 * H[1]; a02 += w2 * v0; a11 += w2 * H[3]; a12 += w2 * v1; a22 += w2 * (zx * v0 + zy * v1 + v2);
        for (uint32_t k = d.od_out_ptr[i]; k < d.od_out_ptr[i + 1]; ++k) {
            const uint32_t j = d.od_b[k];
            if (j >= d.nP || j < lo || j >= hi) continue;
            const double ww = wi * ml_pose_weight(j, K, s, lv.n);
            const double4 pj = d.pose[j];
            const double yx = -(pj.y - pc.y), yy = pj.x - pc.x;
            const double* h = d.hoo + 3ull * k;
            a00 += ww * 2.0 * h[0]; a02 += ww * h[0] * (yx + zx); a11 += ww * 2.0 * h[1]; a12 += ww * h[1] * (yy + zy);
            a22 += ww * 2.0 * (zx * h[0] * yx + zy * h[1] * yy + h[2]);
        }
    }
    const double2* C = reinterpret_cast<const double2*>(lv.C);
    const size_t cs = lv.nnz;
    for (uint32_t e = lv.node_ptr[K] + lane; e < lv.node_ptr[K + 1]; e += 32) {
        const double* hi_ = d.hinv + 3ull * lv.edge_lm[e];
        const double i00 = hi_[0], i01 = hi_[1], i11 = hi_[2];
        const double2 r0 = C[e], r1 = C[cs + e], r2 = C[2 * cs + e];
        const double t00 = r0.x * i00 + r0.y * i01, t01 = r0.x * i01 + r0.y * i11;
        const double t10 = r1.x * i00 + r1.y * i01, t11 = r1.x * i01 + r1.y * i11;
        const double t20 = r2.x * i00 + r2.y * i01, t21 = r2.x * i01 + r2.y * i11;
        a00 -= t00 * r0.x + t01 * r0.y; a01 -= t00 * r1.x + t01 * r1.y; a02 -= t00 * r2.x + t01 * r2.y;
        a11 -= t10 * r1.x + t11 * r1.y; a12 -= t10 * r2.x + t11 * r2.y; a22 -= t20 * r2.x + t21 * r2.y;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a00 += __shfl_xor_sync(0xffffffffu, a00, o); a01 += __shfl_xor_sync(0xffffffffu, a01, o); a02 += __shfl_xor_sync(0xffffffffu, a02, o);
        a11 += __shfl_xor_sync(0xffffffffu, a11, o); a12 += __shfl_xor_sync(0xffffffffu, a12, o); a22 += __shfl_xor_sync(0xffffffffu, a22, o);
    }
    if (lane == 0) sym3_inverse(a00, a01, a02, a11, a12, a22, lv.dinv + 6ull * K);
}

// ---- setup: dense top operator, one block per top node K writes block row K of Ac ----------------------------------
// Row block (3 x nc) in shared memory.  Pose part: lane (r, c) of warp 0 walks the poses of the node's support (Hpp_ii
// couples the node with its chain neighbours through the overlapping hats; ODOM edges couple it with the nodes of the
// other end) -- one lane per entry, sequential: deterministic.  Landmark part: the node's coarse edges in list order;
// for each, the top-level edges of the same landmark (landmark-major CSR tl_ptr / tl_edge / tl_node); thread t owns the
// columns of the nodes K' = t (mod blockDim), so no two threads ever touch the same entry and every entry sees its
// landmarks in ascending order.
constexpr int kMlTopThreads = 256;
__global__ void __launch_bounds__(kMlTopThreads) k_ml_top(const Dev d, const uint32_t* __restrict__ tl_ptr, const uint32_t* __restrict__ tl_edge,
                                                          const uint32_t* __restrict__ tl_node) {
    extern __shared__ __align__(16) double ml_row[];         // [3][nc]
    const MlDev& lv = d.ml[d.mlLv - 1];
    const uint32_t K = blockIdx.x, nc = d.nc, tid = threadIdx.x, s = lv.spacing, n = lv.n, off = d.xc.top_off, Kg = off + K;
    const bool timing = d.dbg != nullptr && K == n / 2 && tid == 0;      // TSB_PCG_TIMING=1: cycles of one middle block per part
    long long tk0 = timing ? clock64() : 0;
    for (uint32_t k = tid; k < 3 * nc; k += kMlTopThreads) ml_row[k] = 0;
    __syncthreads();
    const double4 pc = d.pose[(size_t)K * s];
    // ---- pose part.  Hpp_ii couples node K with the chain neighbours its hat overlaps; an ODOM edge couples it with the
    // nodes of the other end.  For a chain (and any ODOM edge whose other end lies in K-1 .. K+1) a thread per support
    // pose adds into 3 x 9 register accumulators [Kp - K + 1][r][c]; warp xor tree, then the warps in order: deterministic.
    // ODOM edges that reach farther (loop-closure odometry) are walked sequentially by one lane per entry afterwards.
    __shared__ double ml_red[(kMlTopThreads / 32) * 27];
    uint32_t lo, hi; ml_support(d, K, s, n, lo, hi);
    double acc[3][9];
#pragma unroll
    for (int q = 0; q < 3; ++q)
#pragma unroll
        for (int e = 0; e < 9; ++e) acc[q][e] = 0;
    int far = 0;
    // acc[dk] += w * R(zx, zy)^T B R(yx, yy) with R(a, b) = [[1, 0, a], [0, 1, b], [0, 0, 1]], for the target node K + dk - 1
    auto add = [&](int dk, double w, double zx, double zy, const double* B, double yx, double yy) {
        double X[9];
#pragma unroll
        for (int r = 0; r < 3; ++r) { X[3 * r] = B[3 * r]; X[3 * r + 1] = B[3 * r + 1]; X[3 * r + 2] = yx * B[3 * r] + yy * B[3 * r + 1] + B[3 * r + 2]; }
#pragma unroll
        for (int q = 0; q < 3; ++q)
            if (dk == q) {
#pragma unroll
                for (int c = 0; c < 3; ++c) { acc[q][c] += w * X[c]; acc[q][3 + c] += w * X[3 + c]; acc[q][6 + c] += w * (zx * X[c] + zy * X[3 + c] + X[6 + c]); }
            }
    };
    for (uint32_t i = lo + tid; i < hi; i += kMlTopThreads) {
        const double wi = ml_pose_weight(i, K, s, n);
        const double4 pi = d.pose[i];
        const double zx = -(pi.y - pc.y), zy = pi.x - pc.x;
        {   // Hpp_ii against the (one or two) nodes pose i interpolates from
            const double* H = d.hpp + 6ull * i;
            const double B[9] = {H[0], H[1], H[2], H[1], H[3], H[4], H[2], H[4], H[5]};
            const uint32_t q = i / s, rem = i - q * s;
            const bool two = !(q + 1 >= n || rem == 0);
            for (int side = 0; side < (two ? 2 : 1); ++side) {
                const uint32_t Kp = q + side;
                const double wp = two ? (side == 0 ? 1.0 - (double)rem / (double)s : (double)rem / (double)s) : 1.0;
                const double4 pp = d.pose[(size_t)Kp * s];
                add((int)Kp - (int)K + 1, wi * wp, zx, zy, B, -(pi.y - pp.y), pi.x - pp.x);
            }
        }
        for (uint32_t pass = 0; pass < 2; ++pass) {        // ODOM edges of pose i, both directions: H_ij = diag(hoo)
            const uint32_t kb = pass == 0 ? d.od_out_ptr[i] : d.od_in_ptr[i], ke = pass == 0 ? d.od_out_ptr[i + 1] : d.od_in_ptr[i + 1];
            for (uint32_t kk = kb; kk < ke; ++kk) {
                const uint32_t k = pass == 0 ? kk : d.od_in_edge[kk];
                const uint32_t j = pass == 0 ? d.od_b[k] : d.od_a[k];
                if (j >= d.nP) continue;                              // pose of another rank: the cut edge stays out of the top operator
                const double* h = d.hoo + 3ull * k;
                const double B[9] = {h[0], 0, 0, 0, h[1], 0, 0, 0, h[2]};
                const double4 pj = d.pose[j];
                const uint32_t qj = j / s, remj = j - qj * s;
                const bool twoj = !(qj + 1 >= n || remj == 0);
                for (int side = 0; side < (twoj ? 2 : 1); ++side) {
                    const uint32_t Kp = qj + side;
                    if (Kp + 1 < K || Kp > K + 1) { far = 1; continue; }
                    const double wp = twoj ? (side == 0 ? 1.0 - (double)remj / (double)s : (double)remj / (double)s) : 1.0;
                    const double4 pp = d.pose[(size_t)Kp * s];
                    add((int)Kp - (int)K + 1, wi * wp, zx, zy, B, -(pj.y - pp.y), pj.x - pp.x);
                }
            }
        }
    }
#pragma unroll
    for (int q = 0; q < 3; ++q)
#pragma unroll
        for (int e = 0; e < 9; ++e) {
            double v = acc[q][e];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if ((tid & 31) == 0) ml_red[(tid >> 5) * 27 + q * 9 + e] = v;
        }
    far = __syncthreads_or(far);
    if (tid < 27) {
        const int dk = tid / 9, r = (tid % 9) / 3, c = tid % 3;
        const long long Kp = (long long)K + dk - 1;
        if (Kp >= 0 && Kp < (long long)n) {
            double t = 0;
            for (int wv = 0; wv < kMlTopThreads / 32; ++wv) t += ml_red[wv * 27 + tid];
            ml_row[(size_t)r * nc + 3 * (off + (uint32_t)Kp) + c] += t;
        }
    }
    __syncthreads();
    if (far && tid < 9) {          // far ODOM targets only: one lane per entry, sequential over the support (fixed order)
        const int r = tid / 3, c = tid % 3;
        double* rowp = ml_row + (size_t)r * nc + 3 * off;
        for (uint32_t i = lo; i < hi; ++i) {
            const double wi = ml_pose_weight(i, K, s, n);
            const double4 pi = d.pose[i];
            double a0, a1, a2; zcol(make_double2(-(pi.y - pc.y), pi.x - pc.x), r, a0, a1, a2);
            for (uint32_t pass = 0; pass < 2; ++pass) {
                const uint32_t kb = pass == 0 ? d.od_out_ptr[i] : d.od_in_ptr[i], ke = pass == 0 ? d.od_out_ptr[i + 1] : d.od_in_ptr[i + 1];
                for (uint32_t kk = kb; kk < ke; ++kk) {
                    const uint32_t k = pass == 0 ? kk : d.od_in_edge[kk];
                    const uint32_t j = pass == 0 ? d.od_b[k] : d.od_a[k];
                    if (j >= d.nP) continue;
                    const double* h = d.hoo + 3ull * k;
                    const double4 pj = d.pose[j];
                    const uint32_t qj = j / s, remj = j - qj * s;
                    const bool twoj = !(qj + 1 >= n || remj == 0);
                    for (int side = 0; side < (twoj ? 2 : 1); ++side) {
                        const uint32_t Kp = qj + side;
                        if (!(Kp + 1 < K || Kp > K + 1)) continue;
                        const double wp = twoj ? (side == 0 ? 1.0 - (double)remj / (double)s : (double)remj / (double)s) : 1.0;
                        const double4 pp = d.pose[(size_t)Kp * s];
                        double b0, b1, b2; zcol(make_double2(-(pj.y - pp.y), pj.x - pp.x), c, b0, b1, b2);
                        rowp[3 * Kp + c] += wi * wp * (a0 * h[0] * b0 + a1 * h[1] * b1 + a2 * h[2] * b2);
                    }
                }
            }
        }
    }
    __syncthreads();
    if (timing) { const long long t = clock64(); d.dbg[8] = t - tk0; tk0 = t; }
    // ---- landmark part (global lists: the partner node of a landmark may live on another rank).
    //   Ac[K, Kp] -= sum over landmarks l seen by both of C_{K,l} Hll^-1 C_{Kp,l}^T
    // Walking the node's coarse edges with four dependent global loads each (landmark -> list -> partner edge -> its C)
    // by every thread cost 3.4 M cycles per block at C3 (4.5 ms per launch).  Now the edges are staged 256 at a time by a
    // thread per edge (T = C Hll^-1, list offsets), the (edge, partner) pairs of the batch by a thread per pair, and only
    // then does the owner of column block Kp (thread Kp mod 256) walk the staged pairs -- reading shared memory, and global
    // memory only for its own pairs.  Every (row, column) entry still receives its landmarks in ascending order.
    constexpr uint32_t kPairCap = 3072;
    __shared__ double sT[kMlTopThreads][6];
    __shared__ uint32_t s_off[kMlTopThreads + 1], s_t0[kMlTopThreads], sp_e2[kPairCap];
    __shared__ uint16_t sp_node[kPairCap], sp_el[kPairCap];
    const double2* C = reinterpret_cast<const double2*>(d.gtC);
    const size_t cs = d.gt_nnz;
    const uint32_t e_lo = d.gt_node_ptr[Kg], e_hi = d.gt_node_ptr[Kg + 1];
    long long tq = 0, tacc[4] = {0, 0, 0, 0};
#define MLTICK(slot) do { if (timing) { const long long now_ = clock64(); tacc[slot] += now_ - tq; tq = now_; } } while (0)
    for (uint32_t e_base = e_lo; e_base < e_hi; e_base += kMlTopThreads) {
        const uint32_t ne = min((uint32_t)kMlTopThreads, e_hi - e_base);
        uint32_t len = 0;
        if (timing) tq = clock64();
        if (tid < ne) {
            const uint32_t e = e_base + tid, lm = d.gt_edge_lm[e];
            const uint32_t t0 = tl_ptr[lm]; len = tl_ptr[lm + 1] - t0; s_t0[tid] = t0;
            const double* hi_ = d.hinv + 3ull * lm;
            const double i00 = hi_[0], i01 = hi_[1], i11 = hi_[2];
            const double2 r0 = C[e], r1 = C[cs + e], r2 = C[2 * cs + e];
            sT[tid][0] = r0.x * i00 + r0.y * i01; sT[tid][1] = r0.x * i01 + r0.y * i11;
            sT[tid][2] = r1.x * i00 + r1.y * i01; sT[tid][3] = r1.x * i01 + r1.y * i11;
            sT[tid][4] = r2.x * i00 + r2.y * i01; sT[tid][5] = r2.x * i01 + r2.y * i11;
        }
        // exclusive prefix sum of the list lengths: s_off[t] = pairs before edge t of the batch, s_off[ne] = total
        s_off[tid + 1] = len;
        if (tid == 0) s_off[0] = 0;
        __syncthreads();
        if (tid == 0) for (uint32_t t = 0; t < ne; ++t) s_off[t + 1] += s_off[t];
        __syncthreads();
        const uint32_t total = s_off[ne];
        MLTICK(0);
        for (uint32_t pb = 0; pb < total; pb += kPairCap) {
            const uint32_t np = min(kPairCap, total - pb);
            for (uint32_t pr = tid; pr < np; pr += kMlTopThreads) {
                const uint32_t gp = pb + pr;
                uint32_t lo_ = 0, hi_ = ne;                  // last edge el with s_off[el] <= gp
                while (hi_ - lo_ > 1) { const uint32_t mid = (lo_ + hi_) >> 1; if (s_off[mid] <= gp) lo_ = mid; else hi_ = mid; }
                const uint32_t q = s_t0[lo_] + (gp - s_off[lo_]);
                sp_node[pr] = (uint16_t)tl_node[q]; sp_e2[pr] = tl_edge[q]; sp_el[pr] = (uint16_t)lo_;
            }
            __syncthreads();
            MLTICK(1);
            for (uint32_t pr = 0; pr < np; ++pr) {
                const uint32_t Kp = sp_node[pr];
                if ((Kp % kMlTopThreads) != tid) continue;
                const double* T = sT[sp_el[pr]];
                const uint32_t e2 = sp_e2[pr];
                const double2 s0 = C[e2], s1 = C[cs + e2], s2 = C[2 * cs + e2];
                double* o = ml_row + 3 * Kp;
                o[0] -= T[0] * s0.x + T[1] * s0.y; o[1] -= T[0] * s1.x + T[1] * s1.y; o[2] -= T[0] * s2.x + T[1] * s2.y;
                o[nc] -= T[2] * s0.x + T[3] * s0.y; o[nc + 1] -= T[2] * s1.x + T[3] * s1.y; o[nc + 2] -= T[2] * s2.x + T[3] * s2.y;
                o[2 * nc] -= T[4] * s0.x + T[5] * s0.y; o[2 * nc + 1] -= T[4] * s1.x + T[5] * s1.y; o[2 * nc + 2] -= T[4] * s2.x + T[5] * s2.y;
            }
            MLTICK(2);
            __syncthreads();
            MLTICK(3);
        }
    }
    if (timing) { d.dbg[11] = tacc[0]; d.dbg[12] = tacc[1]; d.dbg[13] = tacc[2]; d.dbg[14] = tacc[3]; d.dbg[15] = (long long)(e_hi - e_lo); }
#undef MLTICK
    __syncthreads();
    if (timing) { const long long t = clock64(); d.dbg[9] = t - tk0; tk0 = t; }
    double* out = d.Ac + (size_t)(3 * Kg) * nc;
    for (uint32_t k = tid; k < 3 * nc; k += kMlTopThreads) out[k] = ml_row[k];
    if (timing) d.dbg[10] = clock64() - tk0;
}

// ---- setup, sharded: ODOM edges across a rank cut in the replicated top operator -------------------------------------
// An ODOM edge (i on this rank, j = first pose of a higher rank) couples this rank's top nodes A that interpolate pose i with
// node B = node 0 of that rank (it sits AT pose j, weight 1, zero lever arm): (A, B) += w_A(i) R(z_iA)^T diag(hoo) and its
// transpose.  Without it the top operator is block-diagonal over the ranks on the odometry side and the PCG of a chain needs
// 1.3x (2 ranks) to 1.8x (4 ranks) the iterations.  The owner of the edge writes one record per edge; the records of all ranks
// are broadcast and every rank adds them to its copy of the operator in the same order (bit-identical copies).
__global__ void k_ml_cut_record(const Dev d) {
    const uint32_t q = threadIdx.x;
    if (q >= kXMaxCut) return;
    double* rec = d.xc.cutrec + (size_t)(d.xc.rank * kXMaxCut + q) * kXCutRec;
    if (q >= d.xc.nCut) { rec[0] = 0; return; }
    const MlDev& lv = d.ml[d.mlLv - 1];
    const uint32_t k = d.xc.cut[3 * q], i = d.xc.cut[3 * q + 1], B = d.xc.cut[3 * q + 2], s = lv.spacing, n = lv.n;
    const double* h = d.hoo + 3ull * k;
    const double4 pi = d.pose[i];
    const uint32_t qi = i / s, rem = i - qi * s;
    const bool two = !(qi + 1 >= n || rem == 0);
    rec[0] = 1; rec[1] = (double)B; rec[2] = two ? 2 : 1; rec[3] = (double)(d.xc.top_off + qi); rec[4] = (double)(d.xc.top_off + qi + 1); rec[5] = 0;
    for (int side = 0; side < (two ? 2 : 1); ++side) {
        const double w = two ? (side == 0 ? 1.0 - (double)rem / (double)s : (double)rem / (double)s) : 1.0;
        const double4 pc = d.pose[(size_t)(qi + side) * s];
        const double zx = -(pi.y - pc.y), zy = pi.x - pc.x;
        double* X = rec + 6 + 9 * side;
        X[0] = w * h[0]; X[1] = 0; X[2] = 0; X[3] = 0; X[4] = w * h[1]; X[5] = 0; X[6] = w * zx * h[0]; X[7] = w * zy * h[1]; X[8] = w * h[2];
    }
}
__global__ void k_ml_cut_apply(const Dev d) {
    if (threadIdx.x || blockIdx.x) return;
    const uint32_t nc = d.nc;
    for (uint32_t r = 0; r < d.xc.world; ++r)
        for (uint32_t q = 0; q < kXMaxCut; ++q) {
            const double* rec = d.xc.cutrec + (size_t)(r * kXMaxCut + q) * kXCutRec;
            if (rec[0] == 0) continue;
            const uint32_t B = (uint32_t)rec[1], nA = (uint32_t)rec[2];
            for (uint32_t a = 0; a < nA; ++a) {
                const uint32_t A = (uint32_t)rec[3 + a];
                const double* X = rec + 6 + 9 * a;
                for (uint32_t rr = 0; rr < 3; ++rr)
                    for (uint32_t cc = 0; cc < 3; ++cc) {
                        d.Ac[(size_t)(3 * A + rr) * nc + 3 * B + cc] += X[3 * rr + cc];
                        d.Ac[(size_t)(3 * B + cc) * nc + 3 * A + rr] += X[3 * rr + cc];
                    }
            }
        }
}

// ---- PCG phases -------------------------------------------------------------------------------------------------------
// correction of pose i (position x, y) from all levels: sum over levels of the hat-weighted node corrections g
__device__ __forceinline__ void ml_prolong(const Dev& d, uint32_t i, double px, double py, double& o0, double& o1, double& o2) {
    double gx = 0, gy = 0, gt = 0;
    const uint32_t top = d.mlLv - 1;
    for (uint32_t l = 0; l < top; ++l) {
        const MlDev& lv = d.ml[l];
        const uint32_t q = i / lv.spacing, rem = i - q * lv.spacing;
        const double* g = lv.g + 3ull * q;
        if (q + 1 >= lv.n || rem == 0) { gx += g[0]; gy += g[1]; gt += g[2]; }
        else { const double t = (double)rem / (double)lv.spacing, u = 1.0 - t;
            gx += u * g[0] + t * g[3]; gy += u * g[1] + t * g[4]; gt += u * g[2] + t * g[5]; }
    }
    {   // top level: e = yc, node position read from the pose array
        const MlDev& lv = d.ml[top];
        const uint32_t q = i / lv.spacing, rem = i - q * lv.spacing;
        const double* e = d.yc + 3ull * (d.xc.top_off + q);
        const double4 c0 = d.pose[(size_t)q * lv.spacing];
        const double g0x = e[0] + c0.y * e[2], g0y = e[1] - c0.x * e[2];
        if (q + 1 >= lv.n || rem == 0) { gx += g0x; gy += g0y; gt += e[2]; }
        else { const double t = (double)rem / (double)lv.spacing, u = 1.0 - t;
            const double4 c1 = d.pose[(size_t)(q + 1) * lv.spacing];
            gx += u * g0x + t * (e[3] + c1.y * e[5]); gy += u * g0y + t * (e[4] - c1.x * e[5]); gt += u * e[2] + t * e[5]; }
    }
    o0 = gx - py * gt; o1 = gy + px * gt; o2 = gt;
}

// node K of level l (not the top) has its restricted m: r = (m_x, m_y, m_t + Y m_x - X m_y), e = blockdiag(S_l)^-1 r,
// store g; returns r'e
__device__ __forceinline__ double ml_finalize(const Dev& d, const MlDev& lv, uint32_t K, double mx, double my, double mt) {
    const double4 pc = d.pose[(size_t)K * lv.spacing];
    const double r2 = mt + pc.y * mx - pc.x * my;
    double e0, e1, e2; sym3_mul(lv.dinv + 6ull * K, mx, my, r2, e0, e1, e2);
    double* g = lv.g + 3ull * K;
    g[0] = e0 + pc.y * e2; g[1] = e1 - pc.x * e2; g[2] = e2;
    return mx * e0 + my * e1 + r2 * e2;
}

// seven-tap restriction of one level inside a window: parent v (window-local) from the children 4v-3 .. 4v+3
__device__ __forceinline__ void ml_tap7(const double* __restrict__ child, uint32_t nchild, double* __restrict__ out, uint32_t v,
                                        long long K, uint32_t n) {
    double a0 = 0, a1 = 0, a2 = 0;
    if (K >= 0 && K < (long long)n) {
#pragma unroll
        for (int j = -3; j <= 3; ++j) {
            const int cidx = 4 * (int)v + j;
            if (cidx < 0 || cidx >= (int)nchild) continue;
            const double w = ml_weight(j, 4, (uint32_t)K, n);
            a0 += w * child[3 * cidx]; a1 += w * child[3 * cidx + 1]; a2 += w * child[3 * cidx + 2];
        }
    }
    out[3 * v] = a0; out[3 * v + 1] = a1; out[3 * v + 2] = a2;
}

// Pose-vector step of the streamed PCG with the multilevel preconditioner.  A block takes windows of kMlWindow poses
// (thread per pose, plus kMlHalo poses to the left whose new residual is recomputed -- r ping-pongs between two buffers
// so that a neighbour's in-place update cannot be seen):
//   x += alpha p, r_new = r - alpha Sp, z = blockdiag(S)^-1 r_new, r'z;
//   m of the window-local levels (ratio 4, spacing <= 64) by seven-tap sums in shared memory, their corrections g;
//   for the levels above: the window's partial sums per node the window touches (summed over windows in ml_upper).
// `sm` needs (512 + 128 + 32 + 8) * 3 doubles.  Returns this thread's part of r'z (fine level and local levels).
template <bool INIT>
__device__ __forceinline__ double ml_pose_phase(const Dev& d, double* __restrict__ sm, const double alpha,
                                                const double* __restrict__ rd, double* __restrict__ wr) {
    double* m0 = sm; double* m1 = m0 + 3 * 512; double* m2 = m1 + 3 * 128; double* m3 = m2 + 3 * 32;
    const uint32_t tid = threadIdx.x, nloc = d.mlLocal, Lv = d.mlLv, top = Lv - 1;
    double dot = 0;
    for (uint32_t w = blockIdx.x; w < d.mlWin; w += gridDim.x) {
        const long long il = (long long)w * kMlWindow + (long long)tid - (long long)kMlHalo;
        const bool own = tid >= kMlHalo;
        double mx = 0, my = 0, mt = 0;
        if (il >= 0 && il < (long long)d.nP && (own || nloc > 0)) {
            const size_t i = (size_t)il;
            double r0, r1, r2;
            if (INIT) { r0 = d.rhs[3 * i]; r1 = d.rhs[3 * i + 1]; r2 = d.rhs[3 * i + 2]; }
            else { r0 = rd[3 * i] - alpha * d.sp[3 * i]; r1 = rd[3 * i + 1] - alpha * d.sp[3 * i + 1]; r2 = rd[3 * i + 2] - alpha * d.sp[3 * i + 2]; }
            if (own) {
                if (INIT) { d.x[3 * i] = 0; d.x[3 * i + 1] = 0; d.x[3 * i + 2] = 0; d.p[3 * i] = 0; d.p[3 * i + 1] = 0; d.p[3 * i + 2] = 0; }
                else { d.x[3 * i] += alpha * d.p[3 * i]; d.x[3 * i + 1] += alpha * d.p[3 * i + 1]; d.x[3 * i + 2] += alpha * d.p[3 * i + 2]; }
                wr[3 * i] = r0; wr[3 * i + 1] = r1; wr[3 * i + 2] = r2;
                double z0, z1, z2; sym3_mul(d.minv + 6ull * i, r0, r1, r2, z0, z1, z2);
                d.z[3 * i] = z0; d.z[3 * i + 1] = z1; d.z[3 * i + 2] = z2;
                dot += r0 * z0 + r1 * z1 + r2 * z2;
            }
            const double4 ps = d.pose[i];
            mx = r0; my = r1; mt = r2 - ps.y * r0 + ps.x * r1;
        }
        m0[3 * tid] = mx; m0[3 * tid + 1] = my; m0[3 * tid + 2] = mt;
        __syncthreads();
        // window-local levels: level 1 nodes at window-local poses 4v (128), level 2 at 16v (32), level 3 at 64v (8)
        const long long kb = (long long)w * kMlWindow - (long long)kMlHalo;       // pose of window-local index 0 (multiple of 64)
        if (nloc >= 1) { if (tid < 128) ml_tap7(m0, 512, m1, tid, kb / 4 + (long long)tid, d.ml[0].n); __syncthreads(); }
        if (nloc >= 2) { if (tid < 32) ml_tap7(m1, 128, m2, tid, kb / 16 + (long long)tid, d.ml[1].n); __syncthreads(); }
        if (nloc >= 3) { if (tid < 8) ml_tap7(m2, 32, m3, tid, kb / 64 + (long long)tid, d.ml[2].n); __syncthreads(); }
        // own nodes of the local levels: store m (top) or finalize (g, r'e).  Threads [0,128) level 1, [128,160) level 2, [160,168) level 3
        {
            int lvl = -1; uint32_t v = 0; const double* src = nullptr; uint32_t first_own = 0;
            if (tid < 128) { lvl = 0; v = tid; src = m1; first_own = 16; }
            else if (tid < 160) { lvl = 1; v = tid - 128; src = m2; first_own = 4; }
            else if (tid < 168) { lvl = 2; v = tid - 160; src = m3; first_own = 1; }
            if (lvl >= 0 && (uint32_t)lvl < nloc && v >= first_own) {
                const MlDev& lv = d.ml[lvl];
                const long long K = kb / (long long)lv.spacing + (long long)v;
                if (K >= 0 && K < (long long)lv.n) {
                    if ((uint32_t)lvl == top) { double* o = lv.m + 3ull * K; o[0] = src[3 * v]; o[1] = src[3 * v + 1]; o[2] = src[3 * v + 2]; }
                    else dot += ml_finalize(d, lv, (uint32_t)K, src[3 * v], src[3 * v + 1], src[3 * v + 2]);
                }
            }
        }
        // levels above the local ones: partial sums of this window, from the own nodes of the last local level
        if (nloc < Lv) {
            const uint32_t sB = nloc == 0 ? 1u : d.ml[nloc - 1].spacing, nb = kMlWindow / sB;
            const double* base = nloc == 0 ? m0 + 3 * kMlHalo : nloc == 1 ? m1 + 3 * 16 : nloc == 2 ? m2 + 3 * 4 : m3 + 3;
            const uint32_t pos0 = w * kMlWindow;
            uint32_t item = tid;
            for (uint32_t l = nloc; l < Lv; ++l) {
                const MlDev& lv = d.ml[l];
                if (item < lv.slots) {
                    const uint32_t K = pos0 / lv.spacing + item;
                    double a0 = 0, a1 = 0, a2 = 0;
                    if (K < lv.n) {
                        for (uint32_t kk = 0; kk < nb; ++kk) {
                            const uint32_t pos = pos0 + kk * sB;
                            if (pos >= d.nP) break;
                            const double wt = ml_hat_at(pos, K, lv.spacing, lv.n);
                            a0 += wt * base[3 * kk]; a1 += wt * base[3 * kk + 1]; a2 += wt * base[3 * kk + 2];
                        }
                    }
                    double* o = lv.wpart + 3ull * ((size_t)w * lv.slots + item);
                    o[0] = a0; o[1] = a1; o[2] = a2;
                    item = 0xFFFFFFFFu;
                } else item -= lv.slots;
            }
        }
        __syncthreads();
    }
    return dot;
}

// m of node K of an upper level: its windows' partial sums in window order
__device__ __forceinline__ void ml_upper_sum(const Dev& d, const MlDev& lv, uint32_t K, double& mx, double& my, double& mt) {
    uint32_t lo, hi; ml_support(d, K, lv.spacing, lv.n, lo, hi);
    const uint32_t w_lo = lo / kMlWindow, w_hi = (hi - 1) / kMlWindow;
    double a0 = 0, a1 = 0, a2 = 0;
    for (uint32_t w = w_lo; w <= w_hi; ++w) {
        const uint32_t slot = K - (w * kMlWindow) / lv.spacing;
        if (slot >= lv.slots) continue;
        const double* p = lv.wpart + 3ull * ((size_t)w * lv.slots + slot);
        a0 += p[0]; a1 += p[1]; a2 += p[2];
    }
    mx = a0; my = a1; mt = a2;
}

// Coarse phase after the grid barrier that follows ml_pose_phase:
//   upper levels below the top: m from the window partials, g, r'e into cdot (thread per node over the whole grid);
//   top level: every block that owns rows of the dense inverse gathers the corrected top residual into rc_s, then
//   yc = Ai rc and cdot (coarse_matvec).
__device__ __forceinline__ void ml_coarse_phase(const Dev& d, uint32_t gtid, uint32_t gsz, double* rc_smem, bool crow) {
    const uint32_t Lv = d.mlLv, top = Lv - 1, nloc = d.mlLocal;
    for (uint32_t l = nloc; l < top; ++l) {
        const MlDev& lv = d.ml[l];
        for (uint32_t K = gtid; K < lv.n; K += gsz) {
            double mx, my, mt; ml_upper_sum(d, lv, K, mx, my, mt);
            d.cdot[lv.cdot_off + K] = ml_finalize(d, lv, K, mx, my, mt);
        }
    }
    if (!crow) return;
    const MlDev& lv = d.ml[top];
    const uint32_t nc = d.nc;
    double* rc_s = nc <= 3 * kCoarseMaxAgg ? rc_smem : d.rcs + (size_t)blockIdx.x * nc;
    for (uint32_t K = threadIdx.x; K < lv.n; K += blockDim.x) {
        double mx, my, mt;
        if (top < nloc) { const double* m = lv.m + 3ull * K; mx = m[0]; my = m[1]; mt = m[2]; }
        else ml_upper_sum(d, lv, K, mx, my, mt);
        const double4 pc = d.pose[(size_t)K * lv.spacing];
        rc_s[3 * K] = mx; rc_s[3 * K + 1] = my; rc_s[3 * K + 2] = mt + pc.y * mx - pc.x * my;
    }
    __syncthreads();
    coarse_matvec(d, gtid >> 5, gsz >> 5, rc_s, rc_smem);
}

// Sharded (peer exchange): the same in two steps with the exchange between them.  Step 1: the upper levels of this rank
// and the corrected top-level residual of this rank's top nodes, pushed into every rank's copy of the global top residual.
__device__ __noinline__ void ml_coarse_push(const Dev& d, uint32_t gtid, uint32_t gsz) {
    const uint32_t Lv = d.mlLv, top = Lv - 1, nloc = d.mlLocal;
    for (uint32_t l = nloc; l < top; ++l) {
        const MlDev& lv = d.ml[l];
        for (uint32_t K = gtid; K < lv.n; K += gsz) {
            double mx, my, mt; ml_upper_sum(d, lv, K, mx, my, mt);
            d.cdot[lv.cdot_off + K] = ml_finalize(d, lv, K, mx, my, mt);
        }
    }
    const MlDev& lv = d.ml[top];
    for (uint32_t K = gtid; K < lv.n; K += gsz) {
        double mx, my, mt;
        if (top < nloc) { const double* m = lv.m + 3ull * K; mx = m[0]; my = m[1]; mt = m[2]; }
        else ml_upper_sum(d, lv, K, mx, my, mt);
        const double4 pc = d.pose[(size_t)K * lv.spacing];
        const double r2 = mt + pc.y * mx - pc.x * my;
        for (uint32_t r = 0; r < d.xc.world; ++r) {
            double* o = x_ptr(d, r, d.xc.o_top) + 3ull * (d.xc.top_off + K);
            o[0] = mx; o[1] = my; o[2] = r2;
        }
    }
}
// Step 2 (after the exchange): yc = Ai r_top with the global top residual read from the own region; the same on every rank
__device__ __forceinline__ void ml_coarse_top(const Dev& d, uint32_t gtid, uint32_t gsz, double* rc_smem) {
    coarse_matvec(d, gtid >> 5, gsz >> 5, x_ptr(d, d.xc.rank, d.xc.o_top), rc_smem);
}
