// kernels.cu -- hand-written sm_100a kernels of the Gauss-Newton inner loop.
//
// Math restated from the reference's CPU pipeline (the parity target), not from remote/cuda:
//   K1  linearize+assemble  EdgeSe2.h:24-38, EdgeSe2Point2d.h:28-70, OptimizerCpu.h:36-46,88-138
//   K2  landmark blocks     OptimizerCpu.h:109-118 (the H_jj / b_j part for Point2 vertices)
//   K3/K4 Schur + PCG       replaces SolverEigen.h:20 (dense pivoted QR of a dim x dim H)
//   K5  update              VertexSe2.h:16-27, Vertex2d.h:16-19, OptimizerCpu.h:159-165,173
// All reductions run in a fixed order (segmented sums over a precomputed pattern): no floating
// point atomics anywhere, so results are bit-reproducible run to run.
//
// Memory-bound design (HBM roofline): every per-edge array is read or written once per pass with
// unit-stride accesses in assembly order; vertex blocks (poses of the unit, landmarks of the
// chunk's slots) are staged in shared memory; per-edge products are staged in shared memory and
// reduced along precomputed row segments (pose side) and slot lists (landmark side).
//
// A "unit" is one entry of blk_chunk_ptr: a run of whole poses with <= kChunkEdges LM edges (or one
// heavy pose split into several chunks).  Every chunked kernel is written as a device function over
// one unit so that the same code serves the one-launch-per-phase kernels and the persistent
// cooperative PCG kernel.
#include "device_graph.h"
#include "../../include/toyslam_b200.h"

#include <cooperative_groups.h>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <mutex>
namespace cg = cooperative_groups;

namespace tsb {
namespace {

constexpr size_t align32(size_t x) { return (x + 31) & ~size_t(31); }

// Shared-memory layout of a chunked kernel, sized per kernel (occupancy is smem-limited).
template <int ROWS, bool SLOT2, bool SLOT3, bool POSE, bool VEC, bool SEDGE>
struct Layout {   // every layout also holds the unit's row pointers and the chunk's slot pointers
    static constexpr size_t oStage = 0;
    static constexpr size_t oSlot2 = oStage + align32(sizeof(double) * ROWS * kStageStride);
    static constexpr size_t oSlot3 = oSlot2 + (SLOT2 ? sizeof(double2) * kChunkEdges : 0);
    static constexpr size_t oPose = oSlot3 + (SLOT3 ? sizeof(double) * 3 * kChunkEdges : 0);
    static constexpr size_t oVec = oPose + (POSE ? sizeof(double4) * kBlockPoses : 0);
    static constexpr size_t oSedge = oVec + (VEC ? sizeof(double) * 3 * kBlockPoses : 0);
    static constexpr size_t oPptr = oSedge + (SEDGE ? align32(sizeof(uint16_t) * kChunkEdges) : 0);
    static constexpr size_t oSptr = oPptr + align32(sizeof(uint32_t) * (kBlockPoses + 1));
    static constexpr size_t oRed = oSptr + align32(sizeof(uint32_t) * (kChunkEdges + 1));
    static constexpr size_t bytes = oRed + sizeof(double) * 32;
};
using LayLin = Layout<9, true, false, true, false, true>;
using LaySchur = Layout<9, true, true, false, false, false>;
using LayA = Layout<2, false, false, false, true, true>;
using LayB = Layout<3, true, false, false, false, false>;
using LayPcg = Layout<3, true, false, false, true, true>;     // persistent kernel: pass A and pass B

struct Sm {
    double* stage; double* acc; double2* slot2; double* slot3; double4* pose; double* vec; uint16_t* sedge; double* red;
    uint32_t* pptr; uint32_t* sptr;    // pose_ptr[p0..p0+np] and slot_ptr[s0..s1] of the current unit / chunk
    template <class L> __device__ static Sm make(unsigned char* b) {
        Sm s;
        s.stage = reinterpret_cast<double*>(b + L::oStage); s.acc = s.stage;      // acc aliases stage (used after the last chunk)
        s.slot2 = reinterpret_cast<double2*>(b + L::oSlot2); s.slot3 = reinterpret_cast<double*>(b + L::oSlot3);
        s.pose = reinterpret_cast<double4*>(b + L::oPose); s.vec = reinterpret_cast<double*>(b + L::oVec);
        s.sedge = reinterpret_cast<uint16_t*>(b + L::oSedge); s.red = reinterpret_cast<double*>(b + L::oRed);
        s.pptr = reinterpret_cast<uint32_t*>(b + L::oPptr); s.sptr = reinterpret_cast<uint32_t*>(b + L::oSptr);
        return s;
    }
};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

// deterministic block sum; result valid in thread 0 (all threads must call)
__device__ __forceinline__ double block_sum(double v, double* red) {
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0;
    if (threadIdx.x == 0) { for (uint32_t i = 0; i < blockDim.x / 32; ++i) t += red[i]; }
    __syncthreads();
    return t;
}

// Hpl of LM edge e (3x2, A^T W B): row r is the double2 at hpl_row(d, e)[r * d.nE] -- three planes, so that the
// threads of a warp, which hold consecutive edges, read and write consecutive 16-byte words
__device__ __forceinline__ double2* hpl_row(const Dev& d, uint64_t e) { return reinterpret_cast<double2*>(d.hpl) + (TSB_HPL_PLANES ? e : 3 * e); }
__device__ __forceinline__ size_t hpl_stride(const Dev& d) { return TSB_HPL_PLANES ? (size_t)d.nE : 1; }

// sums `n` values in a fixed order with one block; every thread gets the total
__device__ __forceinline__ double reduce_all(const double* part, uint32_t n, double* red) {
    double t = 0;
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) t += part[i];
    t = warp_sum(t);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = t;
    __syncthreads();
    double r = 0;
    for (uint32_t i = 0; i < blockDim.x / 32; ++i) r += red[i];
    __syncthreads();
    return r;
}

// guarded inverse of a symmetric 3x3 block (00 01 02 11 12 22); a singular / non-finite block gives the zero block:
// that vertex then takes no correction from this level (the dense QR of the reference leaves such unknowns at zero)
__device__ __forceinline__ bool sym3_inverse(const double s00, const double s01, const double s02, const double s11, const double s12,
                                             const double s22, double* M) {
    const double c00 = s11 * s22 - s12 * s12, c01 = s02 * s12 - s01 * s22, c02 = s01 * s12 - s02 * s11;
    const double det = s00 * c00 + s01 * c01 + s02 * c02;
    const double scale = fabs(s00 * s11 * s22);
    if (!(det > 1e-13 * scale) || !isfinite(det)) { M[0] = M[1] = M[2] = M[3] = M[4] = M[5] = 0; return false; }
    const double id = 1.0 / det;
    M[0] = c00 * id; M[1] = c01 * id; M[2] = c02 * id;
    M[3] = (s00 * s22 - s02 * s02) * id; M[4] = (s01 * s02 - s00 * s12) * id; M[5] = (s00 * s11 - s01 * s01) * id;
    return true;
}

// Huber on chi2 with delta = 1.5 (OptimizerCpu.h:36-46)
__device__ __forceinline__ void robustify(double chi2, double& rho, double& w) {
    if (chi2 <= 2.25) { rho = chi2; w = 1.0; }
    else { const double sq = sqrt(chi2); rho = 3.0 * sq - 2.25; w = 1.5 / sq; }
}

// ODOM edge (EdgeSe2.h:24-38): e = (D02, D12, atan2(D10, D00)), D = Zinv * (Xa^-1 * Xb); A = -I, B = +I.
// Returns the weighted information W = huber_w * diag(w) and W*e.
__device__ __forceinline__ void odom_edge(const double4 a, const double4 b, const double* __restrict__ zi,
                                          const double* __restrict__ w, double W[3], double We[3], double& rho) {
    // Xa^-1 = [[c, s, -(c x + s y)], [-s, c, s x - c y]]
    const double i02 = -(a.z * a.x + a.w * a.y), i12 = a.w * a.x - a.z * a.y;
    const double p00 = a.z * b.z + a.w * b.w, p10 = -a.w * b.z + a.z * b.w;       // (Xa^-1 Xb) first column
    const double p02 = a.z * b.x + a.w * b.y + i02, p12 = -a.w * b.x + a.z * b.y + i12;
    const double d00 = zi[0] * p00 + zi[1] * p10, d10 = zi[3] * p00 + zi[4] * p10;
    const double e0 = zi[0] * p02 + zi[1] * p12 + zi[2];
    const double e1 = zi[3] * p02 + zi[4] * p12 + zi[5];
    const double e2 = atan2(d10, d00);
    const double chi2 = e0 * (w[0] * e0) + e1 * (w[1] * e1) + e2 * (w[2] * e2);
    double hw; robustify(chi2, rho, hw);
    W[0] = w[0] * hw; W[1] = w[1] * hw; W[2] = w[2] * hw;
    We[0] = W[0] * e0; We[1] = W[1] * e1; We[2] = W[2] * e2;
}

// fixed-order sum of s[b..e) with four independent partial sums (shortens the dependent chain)
__device__ __forceinline__ double seg_sum(const double* s, uint32_t b, uint32_t e) {
    double t0 = 0, t1 = 0, t2 = 0, t3 = 0;
    uint32_t k = b;
    for (; k + 3 < e; k += 4) { t0 += s[k]; t1 += s[k + 1]; t2 += s[k + 2]; t3 += s[k + 3]; }
    for (; k < e; ++k) t0 += s[k];
    return (t0 + t1) + (t2 + t3);
}
__device__ __forceinline__ double list_sum(const double* s, const uint16_t* idx, uint32_t b, uint32_t e) {
    double t0 = 0, t1 = 0, t2 = 0, t3 = 0;
    uint32_t k = b;
    for (; k + 3 < e; k += 4) { t0 += s[idx[k]]; t1 += s[idx[k + 1]]; t2 += s[idx[k + 2]]; t3 += s[idx[k + 3]]; }
    for (; k < e; ++k) t0 += s[idx[k]];
    return (t0 + t1) + (t2 + t3);
}

// pose side: items (pose, row); sums stage[row][.] over the pose's edges that fall in the chunk
template <int ROWS>
__device__ __forceinline__ void pose_reduce(const uint32_t* pptr, const ChunkDesc& ch, const double* stage, uint32_t np,
                                            double (&acc)[ROWS]) {
#pragma unroll
    for (int it = 0; it < ROWS; ++it) {
        const uint32_t item = threadIdx.x + it * kThreads;
        if (item < np * ROWS) {
            const uint32_t pl = item / ROWS, row = item % ROWS;
            const uint32_t b = max(pptr[pl], ch.e0) - ch.e0;
            const uint32_t e = min(pptr[pl + 1], ch.e1) - ch.e0;
            acc[it] += seg_sum(stage + row * kStageStride, b, e);
        }
    }
}

template <int ROWS>
__device__ __forceinline__ void acc_to_smem(const double (&acc)[ROWS], double* sacc, uint32_t np) {
#pragma unroll
    for (int it = 0; it < ROWS; ++it) {
        const uint32_t item = threadIdx.x + it * kThreads;
        if (item < np * ROWS) sacc[item] = acc[it];
    }
}

// one LM edge: residual, Jacobian products, Hpl block, staged per-edge products (EdgeSe2Point2d.h:28-70)
__device__ __forceinline__ double lm_edge(const double4 ps, const double2 l, const double2 z, const double2 w,
                                          double2* __restrict__ h, size_t hs, double* __restrict__ st) {
    const double c_ = ps.z, s_ = ps.w;
    const double e0 = c_ * l.x + s_ * l.y - (c_ * ps.x + s_ * ps.y) - z.x;       // e = X^-1 l - z_loc (:34-38)
    const double e1 = -s_ * l.x + c_ * l.y + (s_ * ps.x - c_ * ps.y) - z.y;
    const double q0 = c_ * l.y - s_ * l.x - c_ * ps.y + s_ * ps.x;              // A = [[-c,-s,q0],[s,-c,q1]] (:55-61)
    const double q1 = -s_ * l.y - c_ * l.x + s_ * ps.y + c_ * ps.x;             // B = [[c,s],[-s,c]]        (:63-67)
    double rho, hw; robustify(e0 * (w.x * e0) + e1 * (w.y * e1), rho, hw);
    const double i0 = w.x * hw, i1 = w.y * hw, f0 = i0 * e0, f1 = i1 * e1;
    const double m00 = c_ * i0 * c_ + s_ * i1 * s_;          // B^T W B
    const double m01 = c_ * i0 * s_ - s_ * i1 * c_;
    const double m11 = s_ * i0 * s_ + c_ * i1 * c_;
    const double t0 = c_ * i0 * q0 - s_ * i1 * q1;           // B^T W q  ( = -(A^T W A)[0:2,2] )
    const double t1 = s_ * i0 * q0 + c_ * i1 * q1;
    // Hpl = A^T W B = [[-m00,-m01],[-m01,-m11],[t0,t1]]
    h[0] = make_double2(-m00, -m01); h[hs] = make_double2(-m01, -m11); h[2 * hs] = make_double2(t0, t1);
    st[0 * kStageStride] = m00; st[1 * kStageStride] = m01; st[2 * kStageStride] = m11;
    st[3 * kStageStride] = -t0; st[4 * kStageStride] = -t1; st[5 * kStageStride] = q0 * i0 * q0 + q1 * i1 * q1;
    st[6 * kStageStride] = c_ * f0 - s_ * f1; st[7 * kStageStride] = s_ * f0 + c_ * f1;   // B^T W e
    st[8 * kStageStride] = q0 * f0 + q1 * f1;
    return rho;
}

// ------------------------------------------------------------------------------------------------
// K1a: ODOM edges, pose-parallel.  Each pose sums its out-edges (id1 == pose) and in-edges
// (id2 == pose); an edge is therefore evaluated twice, which costs less than a per-edge record
// round trip through HBM, and the dependent index loads are hidden by P-way parallelism instead
// of stalling a unit's 256 threads.  Writes pose_od (H diag 3, b 3), hoo and chi2 partials.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) k_odom(const Dev d) {
    __shared__ double red[32];
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    double rho_sum = 0;
    if (p < d.nP) {
        const double4 me = d.pose[p];
        double h0 = 0, h1 = 0, h2 = 0, b0 = 0, b1 = 0, b2 = 0;
        for (uint32_t k = d.od_out_ptr[p]; k < d.od_out_ptr[p + 1]; ++k) {
            double W[3], We[3], rho;
            odom_edge(me, d.pose[d.od_b[k]], d.od_zinv + 6ull * k, d.od_w + 3ull * k, W, We, rho);
            rho_sum += rho;
            h0 += W[0]; h1 += W[1]; h2 += W[2];               // A^T W A, A = -I
            b0 += We[0]; b1 += We[1]; b2 += We[2];            // b -= A^T W e
            d.hoo[3ull * k] = -W[0]; d.hoo[3ull * k + 1] = -W[1]; d.hoo[3ull * k + 2] = -W[2];
        }
        for (uint32_t q = d.od_in_ptr[p]; q < d.od_in_ptr[p + 1]; ++q) {
            const uint32_t k = d.od_in_edge[q];
            double W[3], We[3], rho;
            odom_edge(d.pose[d.od_a[k]], me, d.od_zinv + 6ull * k, d.od_w + 3ull * k, W, We, rho);
            h0 += W[0]; h1 += W[1]; h2 += W[2];               // B^T W B, B = +I
            b0 -= We[0]; b1 -= We[1]; b2 -= We[2];
            if (k >= d.nOwned) { d.hoo[3ull * k] = -W[0]; d.hoo[3ull * k + 1] = -W[1]; d.hoo[3ull * k + 2] = -W[2]; }
        }
        const double prior = 1e6 * (double)d.fixed_pose_count[p];     // OptimizerCpu.h:132-138
        double* o = d.pose_od + 6ull * p;
        o[0] = h0 + prior; o[1] = h1 + prior; o[2] = h2 + prior; o[3] = b0; o[4] = b1; o[5] = b2;
    }
    const double t = block_sum(rho_sum, red);
    if (threadIdx.x == 0) d.part2[blockIdx.x] = t;
}

// stages the unit's row pointers (pose_ptr[p0 .. p0+np]) in shared memory
__device__ __forceinline__ void stage_pptr(const Dev& d, const Sm& sm, uint32_t p0, uint32_t np) {
    for (uint32_t i = threadIdx.x; i <= np; i += kThreads) sm.pptr[i] = d.pose_ptr[p0 + i];
}
// stages the chunk's slot pointers, rebased to the chunk
__device__ __forceinline__ void stage_sptr(const Dev& d, const Sm& sm, const ChunkDesc& ch) {
    for (uint32_t i = threadIdx.x; i <= ch.s1 - ch.s0; i += kThreads) sm.sptr[i] = d.slot_ptr[ch.s0 + i] - ch.e0;
}

// ------------------------------------------------------------------------------------------------
// K1b unit: LM edges -- linearize + assemble.  All global loads of a chunk are issued up front
// (edge data, row/slot pointers, slot landmarks); the reductions then run out of shared memory.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void unit_linearize(const Dev& d, const uint32_t unit, const Sm& sm) {
    const uint32_t c0 = d.blk_chunk_ptr[unit], c1 = d.blk_chunk_ptr[unit + 1];
    const uint32_t p0 = d.chunks[c0].p0, np = d.chunks[c0].p1 - p0;
    for (uint32_t i = threadIdx.x; i < np; i += kThreads) sm.pose[i] = d.pose[p0 + i];
    stage_pptr(d, sm, p0, np);
    double acc[kStageRows];
#pragma unroll
    for (int i = 0; i < kStageRows; ++i) acc[i] = 0;
    double rho_sum = 0;
    for (uint32_t c = c0; c < c1; ++c) {
        const ChunkDesc ch = d.chunks[c];
        const uint32_t n = ch.e1 - ch.e0, ns = ch.s1 - ch.s0;
        const uint32_t k0 = threadIdx.x, k1 = threadIdx.x + kThreads;
        uint32_t pk0 = 0, pk1 = 0; double2 z0 = {}, w0 = {}, z1 = {}, w1 = {};
        if (k0 < n) { pk0 = d.lm_packed[ch.e0 + k0]; z0 = d.le_z[ch.e0 + k0]; w0 = d.le_w[ch.e0 + k0]; }
        if (k1 < n) { pk1 = d.lm_packed[ch.e0 + k1]; z1 = d.le_z[ch.e0 + k1]; w1 = d.le_w[ch.e0 + k1]; }
        for (uint32_t s = threadIdx.x; s < ns; s += kThreads) sm.slot2[s] = d.lm[d.slot_lm[ch.s0 + s]];
        for (uint32_t k = threadIdx.x; k < n; k += kThreads) sm.sedge[k] = d.slot_edge[ch.e0 + k];
        stage_sptr(d, sm, ch);
        __syncthreads();
        if (k0 < n) rho_sum += lm_edge(sm.pose[pk0 >> 16], sm.slot2[pk0 & 0xffffu], z0, w0,
                                       hpl_row(d, ch.e0 + k0), hpl_stride(d), sm.stage + k0);
        if (k1 < n) rho_sum += lm_edge(sm.pose[pk1 >> 16], sm.slot2[pk1 & 0xffffu], z1, w1,
                                       hpl_row(d, ch.e0 + k1), hpl_stride(d), sm.stage + k1);
        __syncthreads();
        pose_reduce<kStageRows>(sm.pptr, ch, sm.stage, np, acc);
        // landmark side: one partial (M00 M01 M11 g0 g1) per slot, edges of the slot in list order
        for (uint32_t item = threadIdx.x; item < ns * 5; item += kThreads) {
            const uint32_t sl = item / 5, comp = item % 5;
            const double* s = sm.stage + (comp < 3 ? comp : comp + 3) * kStageStride;
            d.lm_part[5ull * (ch.s0 + sl) + comp] = list_sum(s, sm.sedge, sm.sptr[sl], sm.sptr[sl + 1]);
        }
        __syncthreads();
    }
    acc_to_smem<kStageRows>(acc, sm.acc, np);
    __syncthreads();
    // pose finalize: LM sums + the ODOM / prior part prepared by k_odom
    for (uint32_t i = threadIdx.x; i < np; i += kThreads) {
        const double* a = sm.acc + i * kStageRows;
        const uint32_t p = p0 + i;
        const double* o = d.pose_od + 6ull * p;
        double* H = d.hpp + 6ull * p;
        H[0] = a[0] + o[0]; H[1] = a[1]; H[2] = a[3]; H[3] = a[2] + o[1]; H[4] = a[4]; H[5] = a[5] + o[2];
        double* B = d.bp + 3ull * p;
        B[0] = a[6] + o[3]; B[1] = a[7] + o[4]; B[2] = o[5] - a[8];       // b -= A^T W e
    }
    const double t = block_sum(rho_sum, sm.red);
    if (threadIdx.x == 0) d.part[unit] = t;
}

__global__ void __launch_bounds__(kThreads, 4) k_linearize(const Dev d) {
    extern __shared__ __align__(32) unsigned char smem_raw[];
    const Sm sm = Sm::make<LayLin>(smem_raw);
    unit_linearize(d, blockIdx.x, sm);
}

#include "linearize_stream.cuh"

// ------------------------------------------------------------------------------------------------
// K2a: landmark-major sum of the slot partials (5 values per landmark), fixed order.
// K2b: prior, Hll, b_l, Hll^-1 (closed-form 2x2).
// ------------------------------------------------------------------------------------------------
// prior, Hll, b_l and the closed-form inverse of the 2x2 landmark block from the five landmark sums
__device__ __forceinline__ void lm_finalize_one(const Dev& d, uint32_t l, double a0, double a1, double a2, double a3, double a4) {
    const double prior = 1e6 * (double)d.fixed_lm_count[l];
    const double h00 = a0 + prior, h01 = a1, h11 = a2 + prior;
    d.hll[3ull * l] = h00; d.hll[3ull * l + 1] = h01; d.hll[3ull * l + 2] = h11;
    d.bl[2ull * l] = -a3; d.bl[2ull * l + 1] = -a4;               // b_l -= B^T W e
    // a landmark without (effective) observations has a singular block: it keeps delta = 0, as the rank-revealing QR of
    // the reference leaves an unconstrained unknown (SolverEigen.h:20); counted in tsb_report.singular_blocks
    const double det = h00 * h11 - h01 * h01;
    if (!(det > 1e-13 * fabs(h00 * h11)) || !isfinite(det)) {
        d.hinv[3ull * l] = 0; d.hinv[3ull * l + 1] = 0; d.hinv[3ull * l + 2] = 0; d.flags[F_SING] = 1; return; }
    const double id = 1.0 / det;
    d.hinv[3ull * l] = h11 * id; d.hinv[3ull * l + 1] = -h01 * id; d.hinv[3ull * l + 2] = h00 * id;
}

// Thread per landmark: the slot partials of the landmark (five doubles each, one contiguous 40-byte read) are added
// in list order, two partials in flight per step.  FIN (world == 1): finalize in place.  (A warp per landmark -- six
// slots x five components across the lanes -- was latency-bound: ~10 partials per landmark, three dependent loads.)
template <bool FIN>
__global__ void __launch_bounds__(kThreads) k_lm_sum(const Dev d) {
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= d.L) return;
    uint32_t q = d.lm_part_ptr[l]; const uint32_t qe = d.lm_part_ptr[l + 1];
    double a0 = 0, a1 = 0, a2 = 0, a3 = 0, a4 = 0;
    for (; q + 2 <= qe; q += 2) {
        const double* u = d.lm_part + 5ull * d.lm_part_idx[q]; const double* v = d.lm_part + 5ull * d.lm_part_idx[q + 1];
        const double u0 = u[0], u1 = u[1], u2 = u[2], u3 = u[3], u4 = u[4], v0 = v[0], v1 = v[1], v2 = v[2], v3 = v[3], v4 = v[4];
        a0 += u0; a1 += u1; a2 += u2; a3 += u3; a4 += u4; a0 += v0; a1 += v1; a2 += v2; a3 += v3; a4 += v4;
    }
    if (q < qe) { const double* u = d.lm_part + 5ull * d.lm_part_idx[q]; a0 += u[0]; a1 += u[1]; a2 += u[2]; a3 += u[3]; a4 += u[4]; }
    if (FIN) lm_finalize_one(d, l, a0, a1, a2, a3, a4);
    else { double* o = d.lm_sum + 5ull * l; o[0] = a0; o[1] = a1; o[2] = a2; o[3] = a3; o[4] = a4; }
}

__global__ void k_lm_finalize(const Dev d) {
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l == 0 && d.lm_sum_chi2) d.scal[S_CHI2] = d.lm_sum[5ull * d.L];
    if (l >= d.L) return;
    const double* a = d.lm_sum + 5ull * l;
    lm_finalize_one(d, l, a[0], a[1], a[2], a[3], a[4]);
}

__global__ void k_reduce2_to(const double* a, uint32_t na, const double* b, uint32_t nb, double* out) {
    __shared__ double red[32];
    const double r = reduce_all(a, na, red) + reduce_all(b, nb, red);
    if (threadIdx.x == 0) *out = r;
}
__global__ void k_reduce_to(const double* part, uint32_t n, double* out) {
    __shared__ double red[32];
    const double r = reduce_all(part, n, red);
    if (threadIdx.x == 0) *out = r;
}

// ------------------------------------------------------------------------------------------------
// K3: Schur preparation.  diag(S)_i = Hpp_ii - sum_e Hpl_e Hll^-1 Hpl_e^T,  rhs_i = b_i - sum_e Hpl_e Hll^-1 b_l
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void schur_edge(const double2* __restrict__ hp, size_t hs, const double* __restrict__ hi, const double2 bl,
                                           double* __restrict__ st) {
    const double2 r0 = hp[0], r1 = hp[hs], r2 = hp[2 * hs];
    const double i00 = hi[0], i01 = hi[1], i11 = hi[2];
    const double t00 = r0.x * i00 + r0.y * i01, t01 = r0.x * i01 + r0.y * i11;     // T = Hpl * Hll^-1 (3x2)
    const double t10 = r1.x * i00 + r1.y * i01, t11 = r1.x * i01 + r1.y * i11;
    const double t20 = r2.x * i00 + r2.y * i01, t21 = r2.x * i01 + r2.y * i11;
    st[0 * kStageStride] = t00 * r0.x + t01 * r0.y;      // (T Hpl^T) 00 01 02 11 12 22
    st[1 * kStageStride] = t00 * r1.x + t01 * r1.y;
    st[2 * kStageStride] = t00 * r2.x + t01 * r2.y;
    st[3 * kStageStride] = t10 * r1.x + t11 * r1.y;
    st[4 * kStageStride] = t10 * r2.x + t11 * r2.y;
    st[5 * kStageStride] = t20 * r2.x + t21 * r2.y;
    st[6 * kStageStride] = t00 * bl.x + t01 * bl.y;      // T b_l
    st[7 * kStageStride] = t10 * bl.x + t11 * bl.y;
    st[8 * kStageStride] = t20 * bl.x + t21 * bl.y;
}

__global__ void __launch_bounds__(kThreads, 3) k_schur_prep(const Dev d) {
    extern __shared__ __align__(32) unsigned char smem_raw[];
    const Sm sm = Sm::make<LaySchur>(smem_raw);
    const uint32_t unit = blockIdx.x;
    const uint32_t c0 = d.blk_chunk_ptr[unit], c1 = d.blk_chunk_ptr[unit + 1];
    const uint32_t p0 = d.chunks[c0].p0, np = d.chunks[c0].p1 - p0;
    stage_pptr(d, sm, p0, np);
    double acc[kStageRows];
#pragma unroll
    for (int i = 0; i < kStageRows; ++i) acc[i] = 0;
    for (uint32_t c = c0; c < c1; ++c) {
        const ChunkDesc ch = d.chunks[c];
        const uint32_t n = ch.e1 - ch.e0, ns = ch.s1 - ch.s0;
        for (uint32_t s = threadIdx.x; s < ns; s += kThreads) {
            const uint32_t l = d.slot_lm[ch.s0 + s];
            sm.slot2[s] = make_double2(d.bl[2ull * l], d.bl[2ull * l + 1]);
            sm.slot3[3 * s] = d.hinv[3ull * l]; sm.slot3[3 * s + 1] = d.hinv[3ull * l + 1]; sm.slot3[3 * s + 2] = d.hinv[3ull * l + 2];
        }
        __syncthreads();
        for (uint32_t k = threadIdx.x; k < n; k += kThreads) {
            const uint32_t sl = d.lm_packed[ch.e0 + k] & 0xffffu;
            schur_edge(hpl_row(d, ch.e0 + k), hpl_stride(d), sm.slot3 + 3 * sl, sm.slot2[sl], sm.stage + k);
        }
        __syncthreads();
        pose_reduce<kStageRows>(sm.pptr, ch, sm.stage, np, acc);
        __syncthreads();
    }
    acc_to_smem<kStageRows>(acc, sm.acc, np);
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < np; i += kThreads) {
        const double* a = sm.acc + i * kStageRows;
        const uint32_t p = p0 + i;
        const double* H = d.hpp + 6ull * p;
        const double s00 = H[0] - a[0], s01 = H[1] - a[1], s02 = H[2] - a[2], s11 = H[3] - a[3], s12 = H[4] - a[4], s22 = H[5] - a[5];
        // symmetric 3x3 inverse by cofactors; an unconstrained pose (no edge, not fixed) gets the zero block: delta = 0
        if (!sym3_inverse(s00, s01, s02, s11, s12, s22, d.minv + 6ull * p)) d.flags[F_SING] = 1;
        d.rhs[3ull * p] = d.bp[3ull * p] - a[6]; d.rhs[3ull * p + 1] = d.bp[3ull * p + 1] - a[7]; d.rhs[3ull * p + 2] = d.bp[3ull * p + 2] - a[8];
    }
}

// ------------------------------------------------------------------------------------------------
// K4: block-Jacobi PCG on the reduced pose system, matrix-free:
//     S v = Hpp v - Hpl (Hll^-1 (Hpl^T v))
//   pass A : t_e = Hpl_e^T v_pose(e)  -> per-slot partial sums            (one stream over Hpl)
//   w_sum  : landmark-major sum of the slot partials
//   pass B : u_l = Hll^-1 w_l;  (S v)_i = Hpp_ii v_i + sum_odom Hoo v_j - sum_e Hpl_e u_l(e); v'Sv partials
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void sym3_mul(const double* M, double a, double b, double c, double& o0, double& o1, double& o2) {
    o0 = M[0] * a + M[1] * b + M[2] * c; o1 = M[1] * a + M[3] * b + M[4] * c; o2 = M[2] * a + M[4] * b + M[5] * c;
}

// MODE 0: v = p after the fused update p = z + beta p;  MODE 1: v = x
// z_i + Z_i yc[agg(i)] : the prolongated coarse correction added to the block-Jacobi part (component q of pose p)
__device__ __forceinline__ double prolong(const Dev& d, uint32_t p, uint32_t q) {
    const uint32_t a = d.pose_agg[p];
    const double y2 = d.yc[3 * a + 2];
    if (q == 2) return y2;
    const double2 zo = d.zoff[p];
    return q == 0 ? d.yc[3 * a] + zo.x * y2 : d.yc[3 * a + 1] + zo.y * y2;
}

template <int MODE>
__device__ __forceinline__ void unit_pass_a(const Dev& d, const uint32_t unit, const Sm& sm, const double beta, const bool coarse) {
    double* v = sm.vec;
    const uint32_t c0 = d.blk_chunk_ptr[unit], c1 = d.blk_chunk_ptr[unit + 1];
    const uint32_t p0 = d.chunks[c0].p0, np = d.chunks[c0].p1 - p0;
    if (MODE == 0) {
        for (uint32_t i = threadIdx.x; i < 3 * np; i += kThreads) {
            double zz = d.z[3ull * p0 + i];
            if (coarse) zz += prolong(d, p0 + i / 3, i % 3);
            const double t = zz + beta * d.p[3ull * p0 + i];
            d.p[3ull * p0 + i] = t; v[i] = t;
        }
    } else {
        for (uint32_t i = threadIdx.x; i < 3 * np; i += kThreads) v[i] = d.x[3ull * p0 + i];
    }
    for (uint32_t c = c0; c < c1; ++c) {
        const ChunkDesc ch = d.chunks[c];
        const uint32_t n = ch.e1 - ch.e0, ns = ch.s1 - ch.s0;
        const uint32_t k0 = threadIdx.x, k1 = threadIdx.x + kThreads;
        uint32_t pl0 = 0, pl1 = 0; double2 a0 = {}, a1 = {}, a2 = {}, b0 = {}, b1 = {}, b2 = {};
        if (k0 < n) { const double2* hp = hpl_row(d, ch.e0 + k0);
            pl0 = d.lm_packed[ch.e0 + k0] >> 16; a0 = hp[0]; a1 = hp[hpl_stride(d)]; a2 = hp[2 * hpl_stride(d)]; }
        if (k1 < n) { const double2* hp = hpl_row(d, ch.e0 + k1);
            pl1 = d.lm_packed[ch.e0 + k1] >> 16; b0 = hp[0]; b1 = hp[hpl_stride(d)]; b2 = hp[2 * hpl_stride(d)]; }
        for (uint32_t k = threadIdx.x; k < n; k += kThreads) sm.sedge[k] = d.slot_edge[ch.e0 + k];
        stage_sptr(d, sm, ch);
        __syncthreads();
        if (k0 < n) { const double v0 = v[3 * pl0], v1 = v[3 * pl0 + 1], v2 = v[3 * pl0 + 2];
            sm.stage[k0] = a0.x * v0 + a1.x * v1 + a2.x * v2; sm.stage[kStageStride + k0] = a0.y * v0 + a1.y * v1 + a2.y * v2; }
        if (k1 < n) { const double v0 = v[3 * pl1], v1 = v[3 * pl1 + 1], v2 = v[3 * pl1 + 2];
            sm.stage[k1] = b0.x * v0 + b1.x * v1 + b2.x * v2; sm.stage[kStageStride + k1] = b0.y * v0 + b1.y * v1 + b2.y * v2; }
        __syncthreads();
        for (uint32_t item = threadIdx.x; item < ns * 2; item += kThreads) {
            const uint32_t sl = item >> 1, comp = item & 1;
            d.w_part[2ull * (ch.s0 + sl) + comp] = list_sum(sm.stage + comp * kStageStride, sm.sedge, sm.sptr[sl], sm.sptr[sl + 1]);
        }
        __syncthreads();
    }
}

// sp_i = Hpp_ii p_i + sum over ODOM edges of Hoo p_j   (the part of S p that needs no landmark data)
__device__ __forceinline__ void hpp_apply(const Dev& d, uint32_t p, double* out) {
    const double v0 = d.p[3ull * p], v1 = d.p[3ull * p + 1], v2 = d.p[3ull * p + 2];
    double y0, y1, y2; sym3_mul(d.hpp + 6ull * p, v0, v1, v2, y0, y1, y2);
    const uint4 nb = d.od_nb[p];
    if (nb.x != 0xFFFFFFFEu) {                 // chain pose: both neighbours known after ONE load (same order of additions)
        if (nb.x != 0xFFFFFFFFu) { const double* h = d.hoo + 3ull * nb.y; const double* q = d.p + 3ull * nb.x;
            y0 += h[0] * q[0]; y1 += h[1] * q[1]; y2 += h[2] * q[2]; }
        if (nb.z != 0xFFFFFFFFu) { const double* h = d.hoo + 3ull * nb.w; const double* q = d.p + 3ull * nb.z;
            y0 += h[0] * q[0]; y1 += h[1] * q[1]; y2 += h[2] * q[2]; }
        out[0] = y0; out[1] = y1; out[2] = y2;
        return;
    }
    for (uint32_t k = d.od_out_ptr[p]; k < d.od_out_ptr[p + 1]; ++k) {
        const double* h = d.hoo + 3ull * k; const double* q = d.p + 3ull * d.od_b[k];
        y0 += h[0] * q[0]; y1 += h[1] * q[1]; y2 += h[2] * q[2];
    }
    for (uint32_t j = d.od_in_ptr[p]; j < d.od_in_ptr[p + 1]; ++j) {
        const uint32_t k = d.od_in_edge[j];
        const double* h = d.hoo + 3ull * k; const double* q = d.p + 3ull * d.od_a[k];
        y0 += h[0] * q[0]; y1 += h[1] * q[1]; y2 += h[2] * q[2];
    }
    out[0] = y0; out[1] = y1; out[2] = y2;
}

// Pass B over one unit; returns the unit's contribution to p'Sp in thread 0.  w_sum holds Hpl^T p reduced
// over slots (and over ranks when world > 1); sp holds the Hpp part of S p from the middle phase.
// (Summing the slot partials inside this pass instead -- one grid barrier less -- was measured slower:
// landmarks shared by many units make the redundant gathers the critical path.)
__device__ __forceinline__ double unit_pass_b(const Dev& d, const uint32_t unit, const Sm& sm) {
    const uint32_t c0 = d.blk_chunk_ptr[unit], c1 = d.blk_chunk_ptr[unit + 1];
    const uint32_t p0 = d.chunks[c0].p0, np = d.chunks[c0].p1 - p0;
    stage_pptr(d, sm, p0, np);
    double acc[3] = {0, 0, 0};
    for (uint32_t c = c0; c < c1; ++c) {
        const ChunkDesc ch = d.chunks[c];
        const uint32_t n = ch.e1 - ch.e0, ns = ch.s1 - ch.s0;
        const uint32_t k0 = threadIdx.x, k1 = threadIdx.x + kThreads;
        uint32_t sl0 = 0, sl1 = 0; double2 a0 = {}, a1 = {}, a2 = {}, b0 = {}, b1 = {}, b2 = {};
        if (k0 < n) { const double2* hp = hpl_row(d, ch.e0 + k0);
            sl0 = d.lm_packed[ch.e0 + k0] & 0xffffu; a0 = hp[0]; a1 = hp[hpl_stride(d)]; a2 = hp[2 * hpl_stride(d)]; }
        if (k1 < n) { const double2* hp = hpl_row(d, ch.e0 + k1);
            sl1 = d.lm_packed[ch.e0 + k1] & 0xffffu; b0 = hp[0]; b1 = hp[hpl_stride(d)]; b2 = hp[2 * hpl_stride(d)]; }
        for (uint32_t s = threadIdx.x; s < ns; s += kThreads) {
            const uint32_t l = d.slot_lm[ch.s0 + s];
            const double2 w = reinterpret_cast<const double2*>(d.w_sum)[l];
            const double* hi = d.hinv + 3ull * l;
            sm.slot2[s] = make_double2(hi[0] * w.x + hi[1] * w.y, hi[1] * w.x + hi[2] * w.y);
        }
        __syncthreads();
        if (k0 < n) { const double2 u = sm.slot2[sl0];
            sm.stage[k0] = a0.x * u.x + a0.y * u.y; sm.stage[kStageStride + k0] = a1.x * u.x + a1.y * u.y;
            sm.stage[2 * kStageStride + k0] = a2.x * u.x + a2.y * u.y; }
        if (k1 < n) { const double2 u = sm.slot2[sl1];
            sm.stage[k1] = b0.x * u.x + b0.y * u.y; sm.stage[kStageStride + k1] = b1.x * u.x + b1.y * u.y;
            sm.stage[2 * kStageStride + k1] = b2.x * u.x + b2.y * u.y; }
        __syncthreads();
        pose_reduce<3>(sm.pptr, ch, sm.stage, np, acc);
        __syncthreads();
    }
    acc_to_smem<3>(acc, sm.acc, np);
    __syncthreads();
    double dot = 0;
    for (uint32_t i = threadIdx.x; i < 3 * np; i += kThreads) {
        const double y = d.sp[3ull * p0 + i] - sm.acc[i];
        d.sp[3ull * p0 + i] = y;
        dot += d.p[3ull * p0 + i] * y;
    }
    return block_sum(dot, sm.red);
}

// landmark-major sum of the slot partials of w = Hpl^T v, one THREAD per landmark, partials added in list order
// (deterministic).  A landmark has about ten partials at C3/C4: a warp per landmark left two thirds of its lanes
// idle and put three dependent global loads (pointer, index, partial) in series per landmark and warp -- 260 us of
// the 1.26 ms C4 iteration; with a thread per landmark all of them are in flight at once.
__device__ __forceinline__ double2 lm_w_sum(const Dev& d, uint32_t l) {
    uint32_t q = d.lm_part_ptr[l]; const uint32_t qe = d.lm_part_ptr[l + 1];
    const double2* wp = reinterpret_cast<const double2*>(d.w_part);
    double a0 = 0, a1 = 0;
    for (; q + 4 <= qe; q += 4) {
        const uint32_t i0 = d.lm_part_idx[q], i1 = d.lm_part_idx[q + 1], i2 = d.lm_part_idx[q + 2], i3 = d.lm_part_idx[q + 3];
        const double2 r0 = wp[i0], r1 = wp[i1], r2 = wp[i2], r3 = wp[i3];
        a0 += r0.x; a1 += r0.y; a0 += r1.x; a1 += r1.y; a0 += r2.x; a1 += r2.y; a0 += r3.x; a1 += r3.y;
    }
    for (; q < qe; ++q) { const double2 r = wp[d.lm_part_idx[q]]; a0 += r.x; a1 += r.y; }
    return make_double2(a0, a1);
}

// One PCG vector step for pose i (all 32 lanes of a warp call it; `valid` masks the tail):
//   init : x = 0, r = rhs            else : x += alpha p, r -= alpha Sp
//   z = blockdiag(S)^-1 r (the block-Jacobi part of the preconditioner); returns r'z of the pose.
// With the coarse level on, the warp also restricts: rcp[segment] = sum over its 32 poses of Z_i^T r_i
// (fixed xor tree -> deterministic).
template <bool INIT>
__device__ __forceinline__ double pose_step(const Dev& d, uint32_t i, bool valid, double alpha, bool coarse) {
    double r[3] = {0, 0, 0}, dot = 0;
    if (valid) {
        double z[3];
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            if (INIT) { r[q] = d.rhs[3ull * i + q]; d.x[3ull * i + q] = 0; d.p[3ull * i + q] = 0; }
            else { d.x[3ull * i + q] += alpha * d.p[3ull * i + q]; r[q] = d.r[3ull * i + q] - alpha * d.sp[3ull * i + q]; }
            d.r[3ull * i + q] = r[q];
        }
        sym3_mul(d.minv + 6ull * i, r[0], r[1], r[2], z[0], z[1], z[2]);
        d.z[3ull * i] = z[0]; d.z[3ull * i + 1] = z[1]; d.z[3ull * i + 2] = z[2];
        dot = r[0] * z[0] + r[1] * z[1] + r[2] * z[2];
    }
    if (coarse) {
        double2 zo = make_double2(0, 0);
        if (valid) zo = d.zoff[i];
        double c0 = r[0], c1 = r[1], c2 = zo.x * r[0] + zo.y * r[1] + r[2];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            c0 += __shfl_xor_sync(0xffffffffu, c0, o); c1 += __shfl_xor_sync(0xffffffffu, c1, o); c2 += __shfl_xor_sync(0xffffffffu, c2, o);
        }
        if ((threadIdx.x & 31) == 0 && (i >> 5) < d.nSeg) { double* o = d.rcp + 3ull * (i >> 5); o[0] = c0; o[1] = c1; o[2] = c2; }
    }
    return dot;
}

// Coarse level: yc = Ac^-1 (Z^T r) with Z^T r summed from the segment partials, and cdot_j = (Z^T r)_j yc_j.
// Rows are dealt to warps (`gwarp` of `nwarps`); rc_smem is block shared memory for 3 kCoarseMaxAgg doubles; larger
// coarse spaces (HBM-sized graphs: up to kCoarseMaxAggLarge aggregates) keep the block's copy of Z^T r in its own
// slice of d.rcs (global, L1/L2-resident: written and read by this block only).
// (Z^T r)_k for coarse unknown k: the segment partials of its aggregate added in segment order
__device__ __forceinline__ double coarse_restrict(const Dev& d, uint32_t k) {
    const uint32_t spa = d.aggK >> 5, a = k / 3, q = k % 3, s0 = a * spa, s1 = min(d.nSeg, s0 + spa);
    double t = 0;
#pragma unroll 8
    for (uint32_t sg = s0; sg < s1; ++sg) t += d.rcp[3ull * sg + q];
    return t;
}
// `shared_rc`: Z^T r was left in d.rcs[0 .. nc) by a preceding launch (k_pcg_restrict, phased path, large coarse spaces)
// does this block take part in the dense coarse mat-vec?  Large coarse spaces: every block that gets at least one group
// of four rows; small ones: blocks whose first warp has a row
__device__ __forceinline__ bool coarse_row_owner(const Dev& d) {
    return d.nc > 3 * kCoarseMaxAgg ? 4 * blockIdx.x < d.nc : blockIdx.x * (blockDim.x >> 5) < d.nc;
}
// yc = Ai rc_s and cdot_j = yc_j rc_s_j; rc_s holds the coarse residual (shared memory for nc <= 3 kCoarseMaxAgg, else a
// global scratch row); rc_smem: the block's shared scratch (reduction space in the large mode)
__device__ __forceinline__ void coarse_matvec(const Dev& d, uint32_t gwarp, uint32_t nwarps, const double* rc_s, double* rc_smem) {
    const uint32_t nc = d.nc;
    const uint32_t lane = threadIdx.x & 31;
    if (nc > 3 * kCoarseMaxAgg) {
        // Large coarse space (Ai = 100-300 MB, streamed from HBM every iteration): a warp per row leaves one 47 KB row
        // behind a handful of loads in flight (C4: 190 us for 275 MB).  Here a whole block takes four rows at a time,
        // every thread with 4 x nc / blockDim independent loads; warp xor tree, then the warps' partials in warp order.
        const uint32_t wpb = blockDim.x >> 5, warp = threadIdx.x >> 5;
        const uint32_t nblk = min(nwarps / wpb, (nc + 3) / 4), bidx = gwarp / wpb;      // blocks that own coarse rows: coarse_row_owner()
        double* red = rc_smem;                                   // shared memory is free in this mode (rc_s is global)
        for (uint32_t j0 = 4 * bidx; j0 < nc; j0 += 4 * nblk) {
            const uint32_t nr = min(4u, nc - j0);
            const double* r0 = d.Ai + (size_t)j0 * nc;
            const double* r1 = r0 + (nr > 1 ? nc : 0); const double* r2 = r0 + (nr > 2 ? 2 * (size_t)nc : 0); const double* r3 = r0 + (nr > 3 ? 3 * (size_t)nc : 0);
            double t0 = 0, t1 = 0, t2 = 0, t3 = 0;
#pragma unroll 2
            for (uint32_t k = threadIdx.x; k < nc; k += blockDim.x) {
                const double c = rc_s[k];
                t0 += r0[k] * c; t1 += r1[k] * c; t2 += r2[k] * c; t3 += r3[k] * c;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                t0 += __shfl_xor_sync(0xffffffffu, t0, o); t1 += __shfl_xor_sync(0xffffffffu, t1, o);
                t2 += __shfl_xor_sync(0xffffffffu, t2, o); t3 += __shfl_xor_sync(0xffffffffu, t3, o);
            }
            if (lane == 0) { red[4 * warp] = t0; red[4 * warp + 1] = t1; red[4 * warp + 2] = t2; red[4 * warp + 3] = t3; }
            __syncthreads();
            if (threadIdx.x < nr) {
                double t = 0;
                for (uint32_t w = 0; w < wpb; ++w) t += red[4 * w + threadIdx.x];
                d.yc[j0 + threadIdx.x] = t; d.cdot[j0 + threadIdx.x] = t * rc_s[j0 + threadIdx.x];
            }
            __syncthreads();
        }
        return;
    }
    for (uint32_t j = gwarp; j < nc; j += nwarps) {
        const double* row = d.Ai + (size_t)j * nc;
        double t0 = 0, t1 = 0;
        uint32_t k = lane;
#pragma unroll 4
        for (; k + 32 < nc; k += 64) { t0 += row[k] * rc_s[k]; t1 += row[k + 32] * rc_s[k + 32]; }
        if (k < nc) t0 += row[k] * rc_s[k];
        double t = t0 + t1;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (lane == 0) { d.yc[j] = t; d.cdot[j] = t * rc_s[j]; }
    }
    __syncthreads();
}
__device__ __forceinline__ void coarse_phase(const Dev& d, uint32_t gwarp, uint32_t nwarps, double* rc_smem, bool shared_rc = false) {
    const uint32_t nc = d.nc;
    double* rc_s = nc <= 3 * kCoarseMaxAgg ? rc_smem : shared_rc ? d.rcs : d.rcs + (size_t)blockIdx.x * nc;
    if (!(shared_rc && nc > 3 * kCoarseMaxAgg)) {
        for (uint32_t k = threadIdx.x; k < nc; k += blockDim.x) rc_s[k] = coarse_restrict(d, k);
        __syncthreads();
    }
    coarse_matvec(d, gwarp, nwarps, rc_s, rc_smem);
}

// ------------------------------------------------------------------------------------------------
// Coarse operator of the two-level preconditioner: Ac = Z^T S Z (dense nc x nc, nc = 3 aggregates),
//   Z_i = [[1,0,zo.x],[0,1,zo.y],[0,0,1]]: translation x, translation y and rigid rotation of aggregate
//   agg(i) about its centroid -- the near-null vectors of S (a rigid motion of an aggregate together
//   with its landmarks leaves every LM residual unchanged).  Every (a,b) block is accumulated by one
//   thread in a fixed order: deterministic, no atomics.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) k_coarse_cen(const Dev d) {      // block per aggregate
    __shared__ double red[32];
    __shared__ double cxy[2];
    const uint32_t a = blockIdx.x, p0 = d.agg_ptr[a], p1 = d.agg_ptr[a + 1];
    double sx = 0, sy = 0;
    for (uint32_t i = p0 + threadIdx.x; i < p1; i += kThreads) { const double4 ps = d.pose[i]; sx += ps.x; sy += ps.y; }
    const double tx = block_sum(sx, red), ty = block_sum(sy, red);
    if (threadIdx.x == 0) { cxy[0] = tx / (double)(p1 - p0); cxy[1] = ty / (double)(p1 - p0); }
    __syncthreads();
    for (uint32_t i = p0 + threadIdx.x; i < p1; i += kThreads) { const double4 ps = d.pose[i]; d.zoff[i] = make_double2(-(ps.y - cxy[1]), ps.x - cxy[0]); }
}

__global__ void k_coarse_C(const Dev d) {                                  // thread per aslot: C = sum Hpl_e^T Z_i
    const uint32_t sidx = blockIdx.x * blockDim.x + threadIdx.x;
    if (sidx >= d.nAS) return;
    double c[6] = {0, 0, 0, 0, 0, 0};
    for (uint32_t q = d.aslot_ptr[sidx]; q < d.aslot_ptr[sidx + 1]; ++q) {
        const uint32_t e = d.aslot_edge[q];
        const double2 zo = d.zoff[d.edge_pose[e]];
        const double2* h = hpl_row(d, e);
        const double2 h0 = h[0], h1 = h[hpl_stride(d)], h2 = h[2 * hpl_stride(d)];
        c[0] += h0.x; c[1] += h1.x; c[2] += zo.x * h0.x + zo.y * h1.x + h2.x;
        c[3] += h0.y; c[4] += h1.y; c[5] += zo.x * h0.y + zo.y * h1.y + h2.y;
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) d.cC[6ull * sidx + k] = c[k];
}

// column c of Z_i as a 3-vector
__device__ __forceinline__ void zcol(const double2 zo, int c, double& v0, double& v1, double& v2) {
    v0 = c == 0 ? 1.0 : (c == 2 ? zo.x : 0.0); v1 = c == 1 ? 1.0 : (c == 2 ? zo.y : 0.0); v2 = c == 2 ? 1.0 : 0.0;
}

// pose-diagonal and ODOM part: warp per aggregate, lane (r,c) of 9 owns entry [r][c] of every block of block-row a
__global__ void k_coarse_pose(const Dev d) {
    const uint32_t a = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (a >= d.nA || lane >= 9) return;
    const int r = lane / 3, c = lane % 3;
    const uint32_t p0 = d.agg_ptr[a], p1 = d.agg_ptr[a + 1], nc = d.nc;
    double* rowp = d.Ac + (size_t)(3 * a + r) * nc;
    double diag = 0;
    for (uint32_t i = p0; i < p1; ++i) {
        const double2 zi = d.zoff[i];
        double a0, a1, a2, b0, b1, b2; zcol(zi, r, a0, a1, a2); zcol(zi, c, b0, b1, b2);
        double y0, y1, y2; sym3_mul(d.hpp + 6ull * i, b0, b1, b2, y0, y1, y2);
        diag += a0 * y0 + a1 * y1 + a2 * y2;
        for (uint32_t k = d.od_out_ptr[i]; k < d.od_out_ptr[i + 1]; ++k) {
            const uint32_t j = d.od_b[k];
            if (j >= d.nP) continue;                                   // pose of another rank: not in this rank's coarse space
            const double* h = d.hoo + 3ull * k; zcol(d.zoff[j], c, b0, b1, b2);
            const double v = a0 * h[0] * b0 + a1 * h[1] * b1 + a2 * h[2] * b2;
            const uint32_t b = d.pose_agg[j];
            if (b == a) diag += v; else rowp[3 * b + c] += v;
        }
        for (uint32_t q = d.od_in_ptr[i]; q < d.od_in_ptr[i + 1]; ++q) {
            const uint32_t k = d.od_in_edge[q], j = d.od_a[k];
            if (j >= d.nP) continue;
            const double* h = d.hoo + 3ull * k; zcol(d.zoff[j], c, b0, b1, b2);
            const double v = a0 * h[0] * b0 + a1 * h[1] * b1 + a2 * h[2] * b2;
            const uint32_t b = d.pose_agg[j];
            if (b == a) diag += v; else rowp[3 * b + c] += v;
        }
    }
    rowp[3 * a + c] += diag;
}

// landmark (Schur) part: thread per block pair (a,b): Ac[a,b] -= sum_l C_{l,a}^T Hll^-1 C_{l,b}
__global__ void k_coarse_lm(const Dev d) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= d.nPairs) return;
    double acc[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (uint32_t q = d.cpair_ptr[t]; q < d.cpair_ptr[t + 1]; ++q) {
        const uint32_t sa = d.ctrip[2ull * q], sb = d.ctrip[2ull * q + 1];
        const double* hi = d.hinv + 3ull * d.aslot_lm[sa];
        const double* ca = d.cC + 6ull * sa; const double* cb = d.cC + 6ull * sb;
        const double i00 = hi[0], i01 = hi[1], i11 = hi[2];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const double t0 = i00 * ca[r] + i01 * ca[3 + r], t1 = i01 * ca[r] + i11 * ca[3 + r];
#pragma unroll
            for (int c = 0; c < 3; ++c) acc[3 * r + c] += t0 * cb[c] + t1 * cb[3 + c];
        }
    }
    double* blk = d.Ac + (size_t)(3 * d.cpair_a[t]) * d.nc + 3 * d.cpair_b[t];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) blk[(size_t)r * d.nc + c] -= acc[3 * r + c];
}

// ---- one launch per phase (used when collectives sit between the phases, world > 1) -------------------
__global__ void k_pcg_init(const Dev d) {
    __shared__ double red[32];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;        // grid covers nP rounded up to whole warps
    const double t = block_sum(pose_step<true>(d, i, i < d.nP, 0.0, d.flags[F_COARSE] != 0), red);
    if (threadIdx.x == 0) d.part[blockIdx.x] = t;
}
// yc = Ac^-1 Z^T r and the coarse part of r'z (phased path; the persistent kernel calls coarse_phase itself)
__global__ void __launch_bounds__(kThreads) k_pcg_coarse(const Dev d, int init) {
    if (!d.flags[F_COARSE] || (!init && d.flags[F_DONE])) return;
    __shared__ double rc_s[3 * kCoarseMaxAgg];
    coarse_phase(d, (blockIdx.x * blockDim.x + threadIdx.x) >> 5, (gridDim.x * blockDim.x) >> 5, rc_s, true);
}
// large coarse spaces, phased path: Z^T r once into d.rcs instead of once per block of k_pcg_coarse
__global__ void k_pcg_restrict(const Dev d, int init) {
    if (!d.flags[F_COARSE] || (!init && d.flags[F_DONE])) return;
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < d.nc) d.rcs[k] = coarse_restrict(d, k);
}
__global__ void k_pcg_init_scal(const Dev d, uint32_t nparts) {
    __shared__ double red[32];
    double rz = reduce_all(d.part, nparts, red);
    if (d.flags[F_COARSE]) rz += reduce_all(d.cdot, d.nc, red);
    if (threadIdx.x == 0) {
        d.scal[S_RZ0] = rz; d.scal[S_RZ] = rz; d.scal[S_BETA] = 0; d.scal[S_ALPHA] = 0;
        d.flags[F_ITERS] = 0; d.flags[F_DONE] = (rz > 0) ? 0 : 1;
    }
}
template <int MODE>
__global__ void __launch_bounds__(kThreads) k_pcg_pass_a(const Dev d) {
    if (MODE == 0 && d.flags[F_DONE]) return;
    extern __shared__ __align__(32) unsigned char smem_raw[];
    const Sm sm = Sm::make<LayA>(smem_raw);
    unit_pass_a<MODE>(d, blockIdx.x, sm, MODE == 0 ? d.scal[S_BETA] : 0.0, MODE == 0 && d.flags[F_COARSE] != 0);
}
// landmark-major sum of the slot partials (thread per landmark); used where w must exist as an array:
// before the all-reduce over ranks (world > 1) and before the back-substitution
__global__ void k_pcg_mid(const Dev d, int mode) {
    if (mode == 0 && d.flags[F_DONE]) return;
    const uint32_t gtid = blockIdx.x * blockDim.x + threadIdx.x, gsz = gridDim.x * blockDim.x;
    for (uint32_t l = gtid; l < d.L; l += gsz) reinterpret_cast<double2*>(d.w_sum)[l] = lm_w_sum(d, l);
    if (mode == 0) for (uint32_t p = gtid; p < d.nP; p += gsz) hpp_apply(d, p, d.sp + 3ull * p);
}
__global__ void __launch_bounds__(kThreads) k_pcg_pass_b(const Dev d) {
    if (d.flags[F_DONE]) return;
    extern __shared__ __align__(32) unsigned char smem_raw[];
    const Sm sm = Sm::make<LayB>(smem_raw);
    const double t = unit_pass_b(d, blockIdx.x, sm);
    if (threadIdx.x == 0) d.part[blockIdx.x] = t;
}
__global__ void k_pcg_psp(const Dev d, uint32_t nparts) {   // local part of p'Sp (all-reduced over ranks before k_pcg_alpha)
    if (d.flags[F_DONE]) return;
    __shared__ double red[32];
    const double psp = reduce_all(d.part, nparts, red);
    if (threadIdx.x == 0) d.scal[S_PSP] = psp;
}
__global__ void k_pcg_alpha(const Dev d) {
    if (d.flags[F_DONE]) return;
    d.scal[S_ALPHA] = d.scal[S_RZ] / d.scal[S_PSP];
}
__global__ void k_pcg_hpp(const Dev d) {
    if (d.flags[F_DONE]) return;
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < d.nP) hpp_apply(d, p, d.sp + 3ull * p);
}
__global__ void k_pack_boundary(const Dev d, const double* vec) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= d.nBnd) return;
    const uint32_t lp = d.bnd_local[k];
#pragma unroll
    for (int q = 0; q < 3; ++q) d.bnd[3ull * k + q] = lp != 0xFFFFFFFFu ? vec[3ull * lp + q] : 0.0;
}
__global__ void k_unpack_ghost(const Dev d, double* vec) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= d.nG) return;
    const uint32_t k = d.ghost_bnd_slot[g];
#pragma unroll
    for (int q = 0; q < 3; ++q) vec[3ull * (d.nP + g) + q] = d.bnd[3ull * k + q];
}
__global__ void k_pcg_update(const Dev d) {
    if (d.flags[F_DONE]) return;
    __shared__ double red[32];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const double t = block_sum(pose_step<false>(d, i, i < d.nP, d.scal[S_ALPHA], d.flags[F_COARSE] != 0), red);
    if (threadIdx.x == 0) d.part[blockIdx.x] = t;
}
__global__ void k_pcg_rznew(const Dev d, uint32_t nparts) {   // local part of r'z (all-reduced over ranks before k_pcg_beta)
    if (d.flags[F_DONE]) return;
    __shared__ double red[32];
    double rz_new = reduce_all(d.part, nparts, red);
    if (d.flags[F_COARSE]) rz_new += reduce_all(d.cdot, d.nc, red);
    if (threadIdx.x == 0) d.scal[S_RZNEW] = rz_new;
}
__global__ void k_pcg_beta(const Dev d) {
    if (d.flags[F_DONE]) return;
    const double rz_new = d.scal[S_RZNEW];
    d.scal[S_BETA] = rz_new / d.scal[S_RZ]; d.scal[S_RZ] = rz_new;
    const int it = d.flags[F_ITERS] + 1; d.flags[F_ITERS] = it;
    if (!(rz_new > d.scal[S_TOL2] * d.scal[S_RZ0]) || it >= d.flags[F_MAXIT]) d.flags[F_DONE] = 1;
}
// after the all-reduce of scal[S_RZ] at the start of a phased solve
__global__ void k_pcg_init_fix(const Dev d) {
    const double rz = d.scal[S_RZ];
    d.scal[S_RZ0] = rz; d.flags[F_DONE] = (rz > 0) ? 0 : 1;
}

// ---- persistent cooperative PCG: the whole solve in one launch (world == 1) ------------------------------
// Four grid-wide barriers per iteration; alpha, beta and the stop test are computed redundantly by every
// block from the same partial sums in the same order, so all blocks take the same branch.
__global__ void __launch_bounds__(kThreads, 3) k_pcg_persistent(const Dev d) {
    extern __shared__ __align__(32) unsigned char smem_raw[];
    const Sm sm = Sm::make<LayPcg>(smem_raw);
    cg::grid_group grid = cg::this_grid();
    const uint32_t nb = gridDim.x, gtid = blockIdx.x * kThreads + threadIdx.x, gsz = nb * kThreads;
    const bool coarse = d.flags[F_COARSE] != 0;
    const uint32_t nP32 = (d.nP + 31u) & ~31u;                      // whole warps take part in the restriction
    const bool crow = coarse && coarse_row_owner(d);   // this block owns coarse rows
    double dot = 0;
    for (uint32_t i = gtid; i < nP32; i += gsz) dot += pose_step<true>(d, i, i < d.nP, 0.0, coarse);
    double t = block_sum(dot, sm.red);
    if (threadIdx.x == 0) d.part[blockIdx.x] = t;
    grid.sync();
    if (coarse) {
        if (crow) coarse_phase(d, gtid >> 5, gsz >> 5, sm.stage);
        grid.sync();
    }
    double rz0 = reduce_all(d.part, nb, sm.red);
    if (coarse) rz0 += reduce_all(d.cdot, d.nc, sm.red);
    double rz = rz0, beta = 0;
    const double tol2 = d.scal[S_TOL2]; const int maxit = d.flags[F_MAXIT];
    int it = 0;
    // optional phase timing (TSB_PCG_TIMING=1): SM cycles of block 0 per phase, barrier waits separate
    const bool timing = d.dbg != nullptr && blockIdx.x == 0 && threadIdx.x == 0;
    long long tk = 0; long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#define TICK(slot) do { if (timing) { const long long now_ = clock64(); tacc[slot] += now_ - tk; tk = now_; } } while (0)
    if (rz0 > 0) {
        for (;;) {
            if (timing) tk = clock64();
            for (uint32_t u = blockIdx.x; u < d.nBlocks; u += nb) unit_pass_a<0>(d, u, sm, beta, coarse);
            TICK(0);
            grid.sync();
            TICK(1);
            for (uint32_t l = gtid; l < d.L; l += gsz) reinterpret_cast<double2*>(d.w_sum)[l] = lm_w_sum(d, l);
            for (uint32_t p = gtid; p < d.nP; p += gsz) hpp_apply(d, p, d.sp + 3ull * p);
            TICK(2);
            grid.sync();
            TICK(3);
            double bsum = 0;
            for (uint32_t u = blockIdx.x; u < d.nBlocks; u += nb) bsum += unit_pass_b(d, u, sm);
            if (threadIdx.x == 0) d.part2[blockIdx.x] = bsum;
            TICK(4);
            grid.sync();
            TICK(5);
            const double alpha = rz / reduce_all(d.part2, nb, sm.red);
            dot = 0;
            for (uint32_t i = gtid; i < nP32; i += gsz) dot += pose_step<false>(d, i, i < d.nP, alpha, coarse);
            t = block_sum(dot, sm.red);
            if (threadIdx.x == 0) d.part[blockIdx.x] = t;
            TICK(6);
            grid.sync();
            if (coarse) {
                if (crow) coarse_phase(d, gtid >> 5, gsz >> 5, sm.stage);
                grid.sync();
            }
            TICK(7);
            double rz_new = reduce_all(d.part, nb, sm.red);
            if (coarse) rz_new += reduce_all(d.cdot, d.nc, sm.red);
            beta = rz_new / rz; rz = rz_new; ++it;
            if (!(rz_new > tol2 * rz0) || it >= maxit) break;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        d.scal[S_RZ0] = rz0; d.scal[S_RZ] = rz; d.scal[S_BETA] = beta; d.flags[F_ITERS] = it; d.flags[F_DONE] = 1;
        if (timing) for (int q = 0; q < 8; ++q) d.dbg[q] = tacc[q];
    }
#undef TICK
    // landmark back-substitution: delta_l = Hll^-1 (b_l - Hpl^T delta_p)
    grid.sync();
    for (uint32_t u = blockIdx.x; u < d.nBlocks; u += nb) unit_pass_a<1>(d, u, sm, 0.0, false);
    grid.sync();
    for (uint32_t l = gtid; l < d.L; l += gsz) {
        const double2 w = lm_w_sum(d, l);
        const double t0 = d.bl[2ull * l] - w.x, t1 = d.bl[2ull * l + 1] - w.y;
        const double* hi = d.hinv + 3ull * l;
        d.dl[2ull * l] = hi[0] * t0 + hi[1] * t1; d.dl[2ull * l + 1] = hi[1] * t0 + hi[2] * t1;
    }
}

#include "sblock_smoother.cuh"
#include "pcg_resident.cuh"
#include "coarse_invert.cuh"
#include "xchg.cuh"
#include "mlevel.cuh"
#include "pcg_stream.cuh"
#include "pcg_stream_single.cuh"

// delta_l = Hll^-1 (b_l - Hpl^T delta_p); w_sum holds Hpl^T delta_p
__global__ void k_backsub(const Dev d) {
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= d.L) return;
    const double t0 = d.bl[2ull * l] - d.w_sum[2ull * l], t1 = d.bl[2ull * l + 1] - d.w_sum[2ull * l + 1];
    const double* hi = d.hinv + 3ull * l;
    d.dl[2ull * l] = hi[0] * t0 + hi[1] * t1; d.dl[2ull * l + 1] = hi[1] * t0 + hi[2] * t1;
}

// ------------------------------------------------------------------------------------------------
// K5: x <- x [+] step * delta  (VertexSe2.h:16-27, Vertex2d.h:16-19) and ||delta||^2 partials
// ------------------------------------------------------------------------------------------------
__global__ void k_update_state(const Dev d, double step) {
    __shared__ double red[32];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    double np2 = 0, nl2 = 0;
    if (i < d.nP) {
        const double d0 = d.x[3ull * i], d1 = d.x[3ull * i + 1], d2 = d.x[3ull * i + 2];
        double4 ps = d.pose[i];
        const double th = atan2(ps.w, ps.z) + d2 * step;
        double s, c; sincos(th, &s, &c);
        ps.x += d0 * step; ps.y += d1 * step; ps.z = c; ps.w = s;
        d.pose[i] = ps;
        np2 = d0 * d0 + d1 * d1 + d2 * d2;
    }
    if (i < d.L) {
        const double d0 = d.dl[2ull * i], d1 = d.dl[2ull * i + 1];
        double2 l = d.lm[i]; l.x += d0 * step; l.y += d1 * step; d.lm[i] = l;
        nl2 = d0 * d0 + d1 * d1;
    }
    if (i < d.nG) {        // ghost copy of a pose owned elsewhere: same delta, same arithmetic, same bits as its owner
        const double* dd = d.bnd + 3ull * d.ghost_bnd_slot[i];
        double4 ps = d.pose[d.nP + i];
        const double th = atan2(ps.w, ps.z) + dd[2] * step;
        double s, c; sincos(th, &s, &c);
        ps.x += dd[0] * step; ps.y += dd[1] * step; ps.z = c; ps.w = s;
        d.pose[d.nP + i] = ps;
    }
    const double t = block_sum(np2, red);
    const double u = block_sum(nl2, red);
    if (threadIdx.x == 0) { d.part[blockIdx.x] = t; d.part2[blockIdx.x] = u; }
}

__global__ void k_init_pose(const double* xyt, uint32_t n, double4* out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // CreateTransform2d (Helper.h:6-18): rotation through the unit quaternion (cos t/2, 0, 0, sin t/2)
    const double th = xyt[3ull * i + 2];
    double sh, ch; sincos(0.5 * th, &sh, &ch);
    const double tz = 2.0 * sh;
    out[i] = make_double4(xyt[3ull * i], xyt[3ull * i + 1], 1.0 - tz * sh, tz * ch);
}

__global__ void k_read_pose(const double4* pose, uint32_t n, double* xyt) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double4 p = pose[i];
    xyt[3ull * i] = p.x; xyt[3ull * i + 1] = p.y; xyt[3ull * i + 2] = atan2(p.w, p.z);   // SerializeGraphFuncCpu.h:28
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
std::atomic<uint64_t> g_launches{0};
static inline uint32_t cdiv(uint32_t a, uint32_t b) { return (a + b - 1) / b; }
size_t linearize_smem_bytes() { return LayLin::bytes; }

// Per-DEVICE launch parameters.  cudaFuncSetAttribute (the > 48 KB shared-memory opt-in) and the occupancy-derived grid
// sizes belong to a device, not to the process: one server process may drive several GPUs (one context per GPU, one
// host thread each), and a context on device 1 must not find the kernels configured for device 0 only.
struct DevInfo {
    bool init = false;
    int coop_blocks = 0;       // co-resident blocks of k_pcg_persistent
    int stream_blocks = 0;     // grid of k_linearize_stream: resident blocks per SM x SMs
    int pcg_stream_blocks = 0; // grid of k_pcg_stream (0: no cooperative launch)
    int ci_blocks = 0;         // grid of k_coarse_invert (SM count, 0: unavailable)
    int res_blocks = 0;        // SM count if k_pcg_resident can be co-scheduled with one block per SM
    int sb_ok = 0;             // k_sb_build configured
    int ps_b_blocks = 0;       // grid of the stand-alone streamed pass B
    size_t top_smem = 0;       // dynamic shared memory k_ml_top may use
};
constexpr int kMaxDevices = 64;
constexpr size_t kMlTopSmemMax = 150 * 1024;      // + ~41 KB static < 227 KB: nc up to 6 400
static DevInfo g_dev[kMaxDevices];
static std::mutex g_dev_mu;
static DevInfo& set_attrs() {
    int dev = 0;
    cudaGetDevice(&dev);
    DevInfo& di = g_dev[dev < 0 || dev >= kMaxDevices ? 0 : dev];
    if (di.init) return di;
    std::lock_guard<std::mutex> lock(g_dev_mu);
    if (di.init) return di;
    int sms = 0, per_sm = 0, coop = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    auto opt_in = [](const void* fn, size_t bytes) { if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes) != cudaSuccess) { cudaGetLastError(); return false; } return true; };
    opt_in((const void*)k_linearize, LayLin::bytes); opt_in((const void*)k_schur_prep, LaySchur::bytes);
    opt_in((const void*)k_pcg_pass_a<0>, LayA::bytes); opt_in((const void*)k_pcg_pass_a<1>, LayA::bytes);
    opt_in((const void*)k_pcg_pass_b, LayB::bytes); opt_in((const void*)k_pcg_persistent, LayPcg::bytes);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_pcg_persistent, kThreads, LayPcg::bytes);
    di.coop_blocks = coop ? sms * per_sm : 0;
    if (opt_in((const void*)k_linearize_stream, kLsSmemBytes)) { per_sm = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_linearize_stream, kLsThreads, kLsSmemBytes); di.stream_blocks = sms * per_sm; }
    if (opt_in((const void*)k_pcg_stream1<false>, kPsSmemBytes) && opt_in((const void*)k_pcg_stream<true>, kPsSmemBytes)) {
        int a = 0, b = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&a, k_pcg_stream1<false>, kPsThreads, kPsSmemBytes);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, k_pcg_stream<true>, kPsThreads, kPsSmemBytes);
        di.pcg_stream_blocks = coop ? sms * (a < b ? a : b) : 0;
    }
    if (opt_in((const void*)k_coarse_invert, kCiSmemBytes)) { per_sm = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_coarse_invert, kCiThreads, kCiSmemBytes); di.ci_blocks = (coop && per_sm >= 1) ? sms : 0; }
    if (opt_in((const void*)k_pcg_resident, kResSmemBytes)) { per_sm = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_pcg_resident, kResThreads, kResSmemBytes); di.res_blocks = (coop && per_sm >= 1) ? sms : 0; }
    di.sb_ok = opt_in((const void*)k_sb_build, kSbSmemBytes) ? 1 : 0;
    if (opt_in((const void*)k_ps_pass_b, kPsSmemBytes)) { per_sm = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_ps_pass_b, kPsThreads, kPsSmemBytes); di.ps_b_blocks = sms * per_sm; }
    di.top_smem = opt_in((const void*)k_ml_top, kMlTopSmemMax) ? kMlTopSmemMax : 48 * 1024;
    di.init = true;
    return di;
}

void launch_linearize(const Dev& d, cudaStream_t s) {
    const int g_stream_blocks = set_attrs().stream_blocks;
    const uint32_t nob = cdiv(d.nP, kThreads);
    if (nob) { k_odom<<<nob, kThreads, 0, s>>>(d); ++g_launches; }
    // HBM-sized graphs stream their chunks through persistent blocks (2 per SM); small ones take one unit per block
    uint32_t nparts = d.nBlocks;
    uint32_t min_units = 4u * (uint32_t)g_stream_blocks;
    bool no_stream = false;
    { const char* e = getenv("TSB_STREAM_MIN_UNITS"); if (e) min_units = (uint32_t)atoi(e);      // tests force either path
      e = getenv("TSB_NO_STREAM"); no_stream = e && e[0] == '1'; }
    if (g_stream_blocks > 0 && !no_stream && d.nBlocks >= min_units && d.nBlocks) {
        k_linearize_stream<<<g_stream_blocks, kLsThreads, kLsSmemBytes, s>>>(d); nparts = (uint32_t)g_stream_blocks;
        ++g_launches;
    } else if (d.nBlocks) { k_linearize<<<d.nBlocks, kThreads, LayLin::bytes, s>>>(d); ++g_launches; }
    // sharded: this rank's chi2 rides at the end of lm_sum through the landmark all-reduce
    k_reduce2_to<<<1, 1024, 0, s>>>(d.part2, nob, d.part, nparts, d.lm_sum_chi2 ? d.lm_sum + 5ull * d.L : d.scal + S_CHI2); ++g_launches;
}
void launch_lm_sum(const Dev& d, bool finalize, cudaStream_t s) {
    if (!d.L) return;
    if (finalize) k_lm_sum<true><<<cdiv(d.L, kThreads), kThreads, 0, s>>>(d);
    else k_lm_sum<false><<<cdiv(d.L, kThreads), kThreads, 0, s>>>(d);
    ++g_launches;
}
void launch_lm_finalize(const Dev& d, cudaStream_t s) {
    if (d.L || d.lm_sum_chi2) { k_lm_finalize<<<d.L ? cdiv(d.L, kThreads) : 1u, kThreads, 0, s>>>(d); ++g_launches; }
}
void launch_schur_prep(const Dev& d, cudaStream_t s) {
    set_attrs();
    if (d.nBlocks) { k_schur_prep<<<d.nBlocks, kThreads, LaySchur::bytes, s>>>(d); ++g_launches; }
}
// whole PCG + back-substitution in one cooperative launch; returns false if the device cannot co-schedule it
bool launch_pcg_persistent(const Dev& d, cudaStream_t s) {
    const int g_coop_blocks = set_attrs().coop_blocks;
    if (g_coop_blocks <= 0) return false;
    uint32_t want = d.nBlocks > cdiv(d.nP, kThreads) ? d.nBlocks : cdiv(d.nP, kThreads);
    if (cdiv(d.L, kThreads) > want) want = cdiv(d.L, kThreads);
    uint32_t grid = want < (uint32_t)g_coop_blocks ? want : (uint32_t)g_coop_blocks;
    if (grid == 0) grid = 1;
    Dev dd = d; void* args[] = {&dd};
    const cudaError_t e = cudaLaunchCooperativeKernel((void*)k_pcg_persistent, dim3(grid), dim3(kThreads), args, LayPcg::bytes, s);
    if (e != cudaSuccess) { cudaGetLastError(); return false; }
    ++g_launches;
    return true;
}
// HBM-sized graphs on one GPU: the streamed variant of the persistent kernel (>= 4 units per resident block);
// leaves the landmark back-substitution to the caller (launch_pcg_pass_a<1>, launch_pcg_mid, launch_backsub)
bool launch_pcg_stream(const Dev& d, cudaStream_t s) {
    const int blocks = set_attrs().pcg_stream_blocks;
    uint32_t min_units = 4u * (uint32_t)blocks;
    { const char* e = getenv("TSB_STREAM_MIN_UNITS"); if (e) min_units = (uint32_t)atoi(e);
      e = getenv("TSB_NO_STREAM"); if (e && e[0] == '1' && !d.mlLv && d.xc.world <= 1) return false; }
    if (blocks <= 0 || !d.gbar || !d.nBlocks || (d.nBlocks < min_units && !d.mlLv && d.xc.world <= 1)) return false;     // the multilevel hierarchy and the peer exchange always stream
    Dev dd = d; void* args[] = {&dd};
    cudaMemsetAsync(dd.gbar, 0, sizeof(uint32_t), s);
    const uint32_t grid = d.nBlocks < (uint32_t)blocks ? d.nBlocks : (uint32_t)blocks;
    // one source per instance: see pcg_stream_single.cuh
    const void* fn = d.xc.world > 1 ? (const void*)k_pcg_stream<true> : (const void*)k_pcg_stream1<false>;
    if (cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kPsThreads), args, kPsSmemBytes, s) != cudaSuccess) { cudaGetLastError(); return false; }
    ++g_launches;
    return true;
}
void launch_coarse_assemble(const Dev& d, cudaStream_t s) {
    if (!d.nA) return;
    cudaMemsetAsync(d.Ac, 0, sizeof(double) * (size_t)d.nc * d.nc, s);
    k_coarse_cen<<<d.nA, kThreads, 0, s>>>(d); ++g_launches;
    if (d.nAS) { k_coarse_C<<<cdiv(d.nAS, kThreads), kThreads, 0, s>>>(d); ++g_launches; }
    k_coarse_pose<<<cdiv(d.nA * 32, kThreads), kThreads, 0, s>>>(d); ++g_launches;
    if (d.nPairs) { k_coarse_lm<<<cdiv(d.nPairs, 128), 128, 0, s>>>(d); ++g_launches; }
}
// multilevel setup (mlevel.cuh)
void launch_ml_levels(const Dev& d, cudaStream_t s) {
    for (uint32_t l = 0; l < d.mlLv; ++l) {
        const uint32_t nb = cdiv(d.ml[l].n * 32, 256);
        if (l == 0) k_ml_C<true><<<nb, 256, 0, s>>>(d, 0); else k_ml_C<false><<<nb, 256, 0, s>>>(d, (int)l);
        ++g_launches;
        if (l + 1 < d.mlLv) { k_ml_D<<<nb, 256, 0, s>>>(d, (int)l); ++g_launches; }
    }
}
void launch_ml_cut_record(const Dev& d, cudaStream_t s) { k_ml_cut_record<<<1, 32, 0, s>>>(d); ++g_launches; }
void launch_ml_cut_apply(const Dev& d, cudaStream_t s) { k_ml_cut_apply<<<1, 32, 0, s>>>(d); ++g_launches; }
void launch_ml_top(const Dev& d, cudaStream_t s) {
    if (!d.mlLv) return;
    const size_t smem = sizeof(double) * 3 * d.nc;
    if (smem > set_attrs().top_smem) return;              // cannot happen for nc <= 32 x 148 (k_coarse_invert's own limit)
    k_ml_top<<<d.ml[d.mlLv - 1].n, kMlTopThreads, smem, s>>>(d, d.tl_ptr, d.tl_edge, d.tl_node); ++g_launches;
}
// Ai = Ac^-1 and F_COARSE in one cooperative launch; false: not launchable here
bool launch_coarse_invert(const Dev& d, cudaStream_t s) {
    const int blocks = set_attrs().ci_blocks;
    const uint32_t nb = (d.nc + 31) / 32;
    if (!blocks || !d.cbar || !nb || nb > (uint32_t)blocks) return false;
    Dev dd = d; void* args[] = {&dd};
    cudaMemsetAsync(dd.cbar, 0, sizeof(uint32_t), s);
    if (cudaLaunchCooperativeKernel((void*)k_coarse_invert, dim3(blocks), dim3(kCiThreads), args, kCiSmemBytes, s) != cudaSuccess) { cudaGetLastError(); return false; }
    ++g_launches;
    return true;
}
void launch_pcg_coarse(const Dev& d, int init, cudaStream_t s) {
    if (!d.nc) return;
    if (d.nc > 3 * kCoarseMaxAgg) {          // block-per-4-rows mat-vec: one block per SM-slot is enough, the rows stream from HBM
        k_pcg_restrict<<<cdiv(d.nc, kThreads), kThreads, 0, s>>>(d, init); ++g_launches;
        k_pcg_coarse<<<min(cdiv(d.nc, 4u), 148u * 4u), kThreads, 0, s>>>(d, init); ++g_launches;   // grid <= groups of four rows: see nblk in coarse_matvec
        return;
    }
    k_pcg_coarse<<<cdiv(d.nc, kThreads / 32), kThreads, 0, s>>>(d, init); ++g_launches;
}
int resident_blocks_available() { return set_attrs().res_blocks; }
bool launch_sb_build(const Dev& d, cudaStream_t s) {
    if (!set_attrs().sb_ok) return false;
    if (!d.res_nb || !d.sbinv) return false;
    k_sb_build<<<d.res_nb * kSbPerBlock, kSbThreads, kSbSmemBytes, s>>>(d); ++g_launches;
    return true;
}
bool launch_pcg_resident(const Dev& d, cudaStream_t s) {
    if (!d.res_nb || (int)d.res_nb != resident_blocks_available()) return false;
    Dev dd = d; void* args[] = {&dd};
    if (dd.gbar) cudaMemsetAsync(dd.gbar, 0, sizeof(uint32_t), s);
    const cudaError_t e = cudaLaunchCooperativeKernel((void*)k_pcg_resident, dim3(d.res_nb), dim3(kResThreads), args, kResSmemBytes, s);
    if (e != cudaSuccess) { cudaGetLastError(); return false; }
    ++g_launches;
    return true;
}
void launch_pcg_init(const Dev& d, cudaStream_t s) {
    const uint32_t nb = cdiv(d.nP, kThreads);
    if (nb) { k_pcg_init<<<nb, kThreads, 0, s>>>(d); ++g_launches; }
    launch_pcg_coarse(d, 1, s);
    k_pcg_init_scal<<<1, 1024, 0, s>>>(d, nb); ++g_launches;
}
// grid of the stand-alone streamed pass B (2 blocks per SM) when this graph is large enough for it, else 0
static uint32_t ps_phase_grid(const Dev& d) {
    const int blocks = set_attrs().ps_b_blocks;
    uint32_t min_units = 4u * (uint32_t)(blocks > 0 ? blocks : 0);
    { const char* e = getenv("TSB_STREAM_MIN_UNITS"); if (e) min_units = (uint32_t)atoi(e);
      e = getenv("TSB_NO_STREAM"); if (e && e[0] == '1') return 0; }
    if (blocks <= 0 || !d.nBlocks || d.nBlocks < min_units) return 0;
    return d.nBlocks < (uint32_t)blocks ? d.nBlocks : (uint32_t)blocks;
}
void launch_pcg_pass_a(const Dev& d, int mode, cudaStream_t s) {
    set_attrs();
    if (!d.nBlocks) return;
    if (mode == 0) k_pcg_pass_a<0><<<d.nBlocks, kThreads, LayA::bytes, s>>>(d);
    else k_pcg_pass_a<1><<<d.nBlocks, kThreads, LayA::bytes, s>>>(d);
    ++g_launches;
}
void launch_pcg_mid(const Dev& d, int mode, cudaStream_t s) {
    uint32_t nb = cdiv(d.L, kThreads);
    if (mode == 0 && cdiv(d.nP, kThreads) > nb) nb = cdiv(d.nP, kThreads);
    if (nb > 148 * 8) nb = 148 * 8;
    if (nb) { k_pcg_mid<<<nb, kThreads, 0, s>>>(d, mode); ++g_launches; }
}
// returns the number of p'Sp partials it leaves in d.part (for launch_pcg_psp)
uint32_t launch_pcg_pass_b(const Dev& d, cudaStream_t s) {
    set_attrs();
    if (!d.nBlocks) return 0;
    if (const uint32_t g = ps_phase_grid(d)) {
        if (d.L) { k_ps_w_to_u<<<cdiv(d.L, kThreads), kThreads, 0, s>>>(d); ++g_launches; }
        k_ps_pass_b<<<g, kPsThreads, kPsSmemBytes, s>>>(d); ++g_launches;
        return g;
    }
    k_pcg_pass_b<<<d.nBlocks, kThreads, LayB::bytes, s>>>(d); ++g_launches;
    return d.nBlocks;
}
void launch_pcg_hpp(const Dev& d, cudaStream_t s) {
    if (d.nP) { k_pcg_hpp<<<cdiv(d.nP, kThreads), kThreads, 0, s>>>(d); ++g_launches; }
}
void launch_pack_boundary(const Dev& d, const double* vec, cudaStream_t s) {
    if (d.nBnd) { k_pack_boundary<<<cdiv(d.nBnd, kThreads), kThreads, 0, s>>>(d, vec); ++g_launches; }
}
void launch_unpack_ghost(const Dev& d, double* vec, cudaStream_t s) {
    if (d.nG) { k_unpack_ghost<<<cdiv(d.nG, kThreads), kThreads, 0, s>>>(d, vec); ++g_launches; }
}
void launch_pcg_psp(const Dev& d, uint32_t nparts, cudaStream_t s) { k_pcg_psp<<<1, 1024, 0, s>>>(d, nparts); ++g_launches; }
void launch_pcg_alpha(const Dev& d, cudaStream_t s) { k_pcg_alpha<<<1, 1, 0, s>>>(d); ++g_launches; }
void launch_pcg_update(const Dev& d, cudaStream_t s) {
    const uint32_t nb = cdiv(d.nP, kThreads);
    if (nb) { k_pcg_update<<<nb, kThreads, 0, s>>>(d); ++g_launches; }
}
void launch_pcg_rznew(const Dev& d, cudaStream_t s) { launch_pcg_coarse(d, 0, s); k_pcg_rznew<<<1, 1024, 0, s>>>(d, cdiv(d.nP, kThreads)); ++g_launches; }
void launch_pcg_beta(const Dev& d, cudaStream_t s) { k_pcg_beta<<<1, 1, 0, s>>>(d); ++g_launches; }
void launch_pcg_init_fix(const Dev& d, cudaStream_t s) { k_pcg_init_fix<<<1, 1, 0, s>>>(d); ++g_launches; }
void launch_backsub(const Dev& d, cudaStream_t s) {
    if (d.L) { k_backsub<<<cdiv(d.L, kThreads), kThreads, 0, s>>>(d); ++g_launches; }
}
void launch_update_state(const Dev& d, double step, cudaStream_t s) {
    uint32_t nmax = d.nP > d.L ? d.nP : d.L; if (d.nG > nmax) nmax = d.nG;
    const uint32_t nb = cdiv(nmax, kThreads);
    if (!nb) return;
    k_update_state<<<nb, kThreads, 0, s>>>(d, step); ++g_launches;
    k_reduce_to<<<1, 1024, 0, s>>>(d.part, nb, d.scal + S_DELTA2); ++g_launches;
    k_reduce_to<<<1, 1024, 0, s>>>(d.part2, nb, d.scal + S_DELTA2L); ++g_launches;
}
void launch_x_signal_wait(const Dev& d, uint32_t point, uint64_t value, cudaStream_t s) {
    k_x_signal<<<1, 32, 0, s>>>(d, point, (unsigned long long)value); ++g_launches;
    k_x_wait<<<1, 32, 0, s>>>(d, point, (unsigned long long)value); ++g_launches;
}
// fused landmark exchange of the linearisation: push (+ chi2, flag v at XP_LM5), own (waits v, flag v + 1 at XP_HINV), wait v + 1
void launch_x_lm5_exchange(const Dev& d, uint64_t v, cudaStream_t s) {
    k_x_lm5_push<<<cdiv(d.L ? d.L : 1u, kThreads), kThreads, 0, s>>>(d, (unsigned long long)v); ++g_launches;
    k_x_lm5_own<<<cdiv(d.xc.slice, kThreads), kThreads, 0, s>>>(d, (unsigned long long)v); ++g_launches;
    k_x_wait<<<1, 32, 0, s>>>(d, XP_HINV, (unsigned long long)(v + 1)); ++g_launches;
}
void launch_x_push_ghost(const Dev& d, const double* vec, cudaStream_t s) {
    if (d.xc.nExp) { k_x_push_ghost<<<cdiv(d.xc.nExp, kThreads), kThreads, 0, s>>>(d, vec); ++g_launches; }
}
void launch_x_w_push(const Dev& d, cudaStream_t s) { if (d.L) { k_x_w_push<<<cdiv(d.L, kThreads), kThreads, 0, s>>>(d); ++g_launches; } }
void launch_x_backsub_own(const Dev& d, cudaStream_t s) { k_x_backsub_own<<<cdiv(d.xc.slice, kThreads), kThreads, 0, s>>>(d); ++g_launches; }
void launch_x_scal_push(const Dev& d, uint32_t slot, const double* src, cudaStream_t s) { k_x_scal_push<<<1, 1, 0, s>>>(d, slot, src); ++g_launches; }
void launch_x_scal_sum(const Dev& d, uint32_t slot, double* dst, cudaStream_t s) { k_x_scal_sum<<<1, 1, 0, s>>>(d, slot, dst); ++g_launches; }
void launch_init_pose(const double* xyt, uint32_t n, double4* out, cudaStream_t s) {
    if (n) { k_init_pose<<<cdiv(n, kThreads), kThreads, 0, s>>>(xyt, n, out); ++g_launches; }
}
void launch_read_pose(const double4* pose, uint32_t n, double* xyt, cudaStream_t s) {
    if (n) { k_read_pose<<<cdiv(n, kThreads), kThreads, 0, s>>>(pose, n, xyt); ++g_launches; }
}

}  // namespace tsb

// ------------------------------------------------------------------------------------------------
// tsb_odom_linearize_noalloc (include/toyslam_b200.h): the ODOM edge kernel for callers that must not allocate.
// No counterpart in the reference.
//
// k_odom leaves one chi2 partial per thread block in context scratch (Dev::part2, from the context's arena) and a
// second launch adds them in a fixed order.  This variant keeps that scratch in a file-scope device array of
// kNaMaxBlocks doubles (32 KB, held by every GPU that loads the library until the process exits), so the call needs no
// context and performs no cudaMalloc / cudaFree: the caller passes device pointers for every input and output.
// Edge-parallel (one thread per edge, same odom_edge() arithmetic as k_odom); chi2 = block partials added in block order.
//
// The array is shared by every call in the process and keeps the previous call's contents.  A launch writes partials
// [0, blocks) before the reduction reads exactly those, so nothing has to start at zero and stale entries beyond
// `blocks` are never read.  Calls that overlap on the GPU (different streams, or several host threads) RACE on it:
// serialise them.
// ------------------------------------------------------------------------------------------------
namespace tsb {
constexpr uint32_t kNaMaxBlocks = 4096;                                   // x kThreads edges = TSB_NOALLOC_MAX_ODOM_EDGES
constexpr int kNaMaxDevices = 64;
}
static __device__ __align__(256) double g_na_scratch[tsb::kNaMaxBlocks];
static_assert(tsb::kNaMaxBlocks * tsb::kThreads == TSB_NOALLOC_MAX_ODOM_EDGES, "header constant");

namespace tsb {
namespace {

__global__ void __launch_bounds__(kThreads) k_odom_noalloc(uint32_t nP, uint32_t nE, const double* __restrict__ xyt,
                                                           const uint32_t* __restrict__ ij, const double* __restrict__ zinv,
                                                           const double* __restrict__ w, double* __restrict__ hoff,
                                                           double* __restrict__ we, double* __restrict__ part) {
    __shared__ double red[32];
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    double rho = 0;
    if (k < nE) {
        const uint32_t i = ij[2ull * k], j = ij[2ull * k + 1];
        double W[3] = {0, 0, 0}, We[3] = {0, 0, 0};
        if (i < nP && j < nP) {                                           // an edge with an id out of range contributes nothing
            double4 a, b;
            {   // CreateTransform2d (Helper.h:6-18), as k_init_pose
                double sh, ch; sincos(0.5 * xyt[3ull * i + 2], &sh, &ch); const double tz = 2.0 * sh;
                a = make_double4(xyt[3ull * i], xyt[3ull * i + 1], 1.0 - tz * sh, tz * ch);
                sincos(0.5 * xyt[3ull * j + 2], &sh, &ch); const double ty = 2.0 * sh;
                b = make_double4(xyt[3ull * j], xyt[3ull * j + 1], 1.0 - ty * sh, ty * ch);
            }
            odom_edge(a, b, zinv + 6ull * k, w + 3ull * k, W, We, rho);
        }
        hoff[3ull * k] = -W[0]; hoff[3ull * k + 1] = -W[1]; hoff[3ull * k + 2] = -W[2];     // H_ij = A^T W B, A = -I, B = +I
        we[3ull * k] = We[0]; we[3ull * k + 1] = We[1]; we[3ull * k + 2] = We[2];
    }
    const double t = block_sum(rho, red);
    if (threadIdx.x == 0) part[blockIdx.x] = t;
}
__global__ void k_odom_noalloc_sum(const double* __restrict__ part, uint32_t n, double* __restrict__ chi2) {
    __shared__ double red[32];
    const double r = reduce_all(part, n, red);
    if (threadIdx.x == 0) *chi2 = r;
}

}  // namespace
}  // namespace tsb

extern "C" int tsb_odom_linearize_noalloc(int32_t device, void* stream, uint32_t n_pose, uint32_t n_edge,
                                          const double* pose_xytheta, const uint32_t* edge_ij, const double* edge_zinv,
                                          const double* edge_w, double* hpp_off, double* w_e, double* chi2,
                                          char* err, uint32_t err_cap) {
    using namespace tsb;
    static double* scratch_of[kNaMaxDevices] = {};       // device address of g_na_scratch, per device ordinal
    auto fail = [&](int code, const char* msg, const char* detail) {
        if (err && err_cap) std::snprintf(err, err_cap, "tsb_odom_linearize_noalloc: %s%s%s", msg, detail ? ": " : "", detail ? detail : "");
        return code;
    };
    if (err && err_cap) err[0] = 0;
    if (n_edge > TSB_NOALLOC_MAX_ODOM_EDGES) {
        char d[96]; std::snprintf(d, sizeof d, "%u edges, the static scratch holds partials for at most %u", n_edge, (unsigned)TSB_NOALLOC_MAX_ODOM_EDGES);
        return fail(1 /* TSB_EINVAL */, "too many ODOM edges", d);
    }
    if (!chi2 || (n_edge && (!pose_xytheta || !edge_ij || !edge_zinv || !edge_w || !hpp_off || !w_e))) return fail(1, "null pointer", nullptr);
    if (device < 0 || device >= kNaMaxDevices) return fail(1, "bad device ordinal", nullptr);
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail(3 /* TSB_ECUDA */, "cudaSetDevice", cudaGetErrorString(e));
    if (!scratch_of[device]) {
        // g_na_scratch names a host-side stand-in here: the device address comes from the runtime, once per device
        double* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_na_scratch);
        if (e != cudaSuccess || !p) return fail(3, "cudaGetSymbolAddress", cudaGetErrorString(e));
        scratch_of[device] = p;
    }
    double* part = scratch_of[device];
    cudaStream_t s = (cudaStream_t)stream;
    const uint32_t blocks = (n_edge + kThreads - 1) / kThreads;
    if (blocks) {
        k_odom_noalloc<<<blocks, kThreads, 0, s>>>(n_pose, n_edge, pose_xytheta, edge_ij, edge_zinv, edge_w, hpp_off, w_e, part); ++g_launches;
    }
    k_odom_noalloc_sum<<<1, 1024, 0, s>>>(part, blocks, chi2); ++g_launches;
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(3, "kernel launch", cudaGetErrorString(e));
    return 0;
}
