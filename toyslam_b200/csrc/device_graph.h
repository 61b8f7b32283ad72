// device_graph.h -- device-resident graph, assembly structure and linear system (fp64 SoA in HBM).
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstdint>
#include "host_graph.h"

namespace tsb {

constexpr int kThreads = 256;          // threads per block of the chunked kernels
#ifndef TSB_HPL_PLANES
#define TSB_HPL_PLANES 1              // Hpl as three planes of nE double2 (1) or 3x2 per edge (0)
#endif
constexpr uint32_t kCoarseScratchBlocks = 1024;   // upper bound on the blocks of a persistent PCG kernel (each owns a scratch row of nc doubles in Dev::rcs)
constexpr int kStageRows = 9;          // values staged per LM edge
constexpr int kStageStride = kChunkEdges + 1;   // odd stride in doubles: rows land in different banks

// scalars kept on the device (index into Dev::scal)
enum Scal { S_CHI2 = 0, S_RZ0, S_RZ, S_PSP, S_ALPHA, S_BETA, S_DELTA2, S_DELTA2L, S_TOL2, S_RZNEW, S_COUNT };
// integer flags (index into Dev::flags)
enum Flag { F_DONE = 0, F_ITERS, F_MAXIT, F_COARSE, F_SING, F_COUNT };   // F_COARSE: 1 = coarse inverse valid for this solve; F_SING: a singular vertex block was zeroed

// Peer exchange over NVLink (world > 1, one process per GPU).  Every rank owns an "exchange region" (one cudaMalloc,
// opened on the peers through CUDA IPC) with the SAME layout on all ranks; a rank WRITES into its peers' regions and only
// ever READS its own (push model: every wait polls local memory).  Flags are 64-bit sequence numbers, one row per
// exchange point and source rank; the data of a step is written first, fenced at system scope, then the flag.
constexpr uint32_t kXMaxWorld = 8, kXPoints = 16, kXMaxGhost = 16;
enum XPoint { XP_LM5 = 0, XP_HINV, XP_GHOST, XP_W, XP_U, XP_PSP, XP_TOP, XP_MISC, XP_BW, XP_BU, XP_COUNT };   // XP_W .. XP_TOP: inside the PCG kernel
struct XDev {
    uint32_t world, rank, slice, nExp;     // slice: landmarks owned per rank (owner(l) = l / slice); nExp: export entries
    uint64_t seq;                          // flag value base of this launch (the host advances it identically on all ranks)
    char* peer[kXMaxWorld];                // region of every rank as mapped here (peer[rank] = own region)
    size_t o_flags, o_scal, o_ghost, o_lm5, o_hll, o_hinv, o_bl, o_w, o_bw, o_u, o_dl, o_top;   // byte offsets inside a region
    uint32_t* done;                        // two "blocks finished" counters of the fused push / own kernels (local, zero between uses)
    const uint32_t* exp;                   // 3 per export entry: local pose, destination rank, ghost slot there
    const uint8_t* lm_mask;                // per landmark: bit r set = rank r has LM edges to it
    uint32_t top_off, top_total;           // this rank's first node / all nodes of the global top level
    // ODOM edges that cross a rank cut and end on the FIRST pose of a higher rank (node 0 of that rank's top level, weight 1):
    // their off-diagonal blocks join the replicated top operator (k_ml_cut_record / k_ml_cut_apply).  3 words per edge:
    // local ODOM edge, local pose of the near end, global top node of the far end; records of kXCutRec doubles per edge.
    uint32_t nCut; const uint32_t* cut; double* cutrec;
};

// one level of the multilevel hierarchy (host_graph.h MlLevel) on the device
constexpr uint32_t kXMaxCut = 8;        // cut edges per rank that enter the top operator
constexpr uint32_t kXCutRec = 24;       // doubles per record: valid, B, nA, A0, A1, pad, X0[9], X1[9]

struct MlDev {
    uint32_t n, ratio, spacing, nnz;
    uint32_t slots, cdot_off;                 // upper levels: partial sums per window; first entry of the level in Dev::cdot
    const uint32_t* node_ptr; const uint32_t* edge_lm; const uint32_t* ref_ptr; const uint32_t* ref;
    double* C;                                // C_l = Z^T Hpl: three planes of nnz double2, like Dev::hpl
    double* dinv;                             // 6 per node: inverse of the diagonal block of S_l (not on the top level)
    double* m;                                // 3 per node: restricted residual as (force x, force y, moment about the origin)
    double* g;                                // 3 per node: correction (e_x + Y e_t, e_y - X e_t, e_t), prolongated with scalar hat weights
    double* wpart;                            // upper levels: nWin x slots x 3 partial sums of m
};

struct Dev {
    uint32_t nP, nG, L, nE, nO, nOwned, nChunks, nBlocks, nSlots;
    // state
    double4* pose;              // nP + nG : x, y, cos(theta), sin(theta)
    double2* lm;                // L
    // LM edges, assembly order
    const uint32_t* lm_packed;  // (pose - chunk.p0) << 16 | slot - chunk.s0
    const double2* le_z;        // landmark in the sensor frame (r cos b, r sin b)
    const double2* le_w;        // information diagonal
    uint32_t* gbar;             // grid-barrier counter of the resident PCG kernel (zeroed before each launch); null: cg grid.sync
    double* hpl;                // A^T W B (3x2) per LM edge, as three planes of nE double2: row r of edge e at 2*(r*nE+e)
    const uint32_t* pose_ptr;   // nP + 1
    const ChunkDesc* chunks;
    const uint32_t* blk_chunk_ptr;
    const uint32_t* slot_lm;
    const uint32_t* slot_ptr;
    const uint16_t* slot_edge;
    const uint32_t* lm_part_ptr;
    const uint32_t* lm_part_idx;
    // ODOM edges (owned first, then ghost)
    const uint32_t* od_a; const uint32_t* od_b;
    const double* od_zinv;      // 6 per edge
    const double* od_w;         // 3 per edge
    const uint32_t* od_out_ptr; const uint32_t* od_in_ptr; const uint32_t* od_in_edge;
    const uint4* od_nb;         // per pose with <= 1 out- and <= 1 in-edge (a chain): (out pose, out edge, in pose, in edge), 0xFFFFFFFF = none;
                                // x == 0xFFFFFFFE: general pose, walk the CSR lists.  Saves two dependent loads per pose in S p
    double* hoo;                // 3 per edge: diagonal of the (i,j) block = -W
    const uint32_t* fixed_pose_count; const uint32_t* fixed_lm_count;
    // assembled system
    double* pose_od;            // 6 per pose: ODOM + prior part of the diagonal block (3) and of b (3), from k_odom
    double* hpp;                // 6 per pose (00 01 02 11 12 22), prior included
    double* bp;                 // 3 per pose
    double* lm_part;            // 5 per slot: partial (Hll 00 01 11, b_l 0 1)
    uint32_t lm_sum_chi2;       // world > 1: chi2 is written to lm_sum[5 L] (one all-reduce for both)
    double* lm_sum;             // 5 per landmark: sum over this rank's slots (all-reduced when world > 1)
    double* hll;                // 3 per landmark
    double* hinv;               // 3 per landmark: inverse of Hll (prior included)
    double* bl;                 // 2 per landmark
    // Schur / PCG
    double* minv;               // 6 per pose: inverse of the diagonal block of S
    double* rhs;                // 3 per pose
    double* x; double* r; double* z; double* p; double* sp;   // 3 per pose (p: nP + nG)
    double* w_part;             // 2 per slot
    double* w_sum;              // 2 per landmark (+ tail used by the multi-GPU exchange)
    double* dl;                 // 2 per landmark: delta of the landmarks
    double* part;               // nBlocks partial sums (one reduction at a time)
    double* part2;              // second set (chi2 of ODOM etc.)
    double* scal;               // S_COUNT
    int* flags;                 // F_COUNT
    // world > 1: boundary poses exchanged through the tail of the w_sum all-reduce buffer
    uint32_t nBnd;
    const uint32_t* bnd_local;  // per boundary slot: local pose index, or 0xFFFFFFFF if owned elsewhere
    const uint32_t* ghost_bnd_slot;   // per ghost pose: boundary slot
    double* bnd;                // 3 per boundary slot; == w_sum + 2 L (one contiguous all-reduce)
    // two-level preconditioner: M^-1 = blockdiag(S)^-1 + Z (Z^T S Z)^-1 Z^T, Z = rigid-body modes of pose aggregates
    uint32_t aggK, nA, nc, nAS, nPairs, nSeg;     // nc = 3 nA; nSeg = ceil(nP / 32) restriction segments
    const uint32_t* edge_pose; const uint32_t* aslot_lm; const uint32_t* aslot_agg; const uint32_t* aslot_ptr;
    const uint32_t* aslot_edge; const uint32_t* cpair_a; const uint32_t* cpair_b; const uint32_t* cpair_ptr; const uint32_t* ctrip;
    const uint32_t* agg_ptr;    // nA + 1 pose offsets of the aggregates; pose_agg: aggregate of each pose
    const uint32_t* pose_agg;
    // resident plan (res_nb > 0): block b owns poses [res_pose_cut[b], +1), landmarks [res_lm_cut[b], +1), aggregates [res_agg_cut[b], +1)
    uint32_t res_nb;
    const uint32_t* res_pose_cut; const uint32_t* res_lm_cut; const uint32_t* res_agg_cut;
    const uint32_t* lmo_ptr; const uint32_t* lmo_edge; const uint32_t* edge_lm;
    double* sbinv;              // res_nb x kSbPerBlock dense symmetric inverses (kSbLD^2 each) of the smoother blocks; null: 3x3 pose blocks (minv)
    double2* zoff;              // per pose: (-(y - cy), x - cx) about its aggregate's centroid -> third column of Z_i
    double* cC;                 // 6 per aslot: sum of Hpl_e^T Z_i (2x3 row-major)
    double* Ac;                 // nc x nc: Z^T S Z, overwritten by its Cholesky factor
    double* Ai;                 // nc x nc: (Z^T S Z)^-1, applied as a dense mat-vec by the PCG
    double* rcp;                // 3 per segment: Z^T r summed over the segment's 32 poses
    double* rcs;                // nc > 3 kCoarseMaxAgg only: kCoarseScratchBlocks x nc, block b's copy of Z^T r in coarse_phase
    double* yc;                 // nc: coarse correction Ac^-1 (Z^T r)
    double* cdot;               // nc: (Z^T r)_j * yc_j  (the coarse part of r'z)
    uint32_t* cbar;             // grid-barrier counter of k_coarse_invert
    int* cinfo;                 // potrf / potrs status words
    long long* dbg;             // optional: 8 phase-cycle counters of the persistent PCG kernel (TSB_PCG_TIMING=1)
    // multilevel preconditioner of the streamed PCG (mlLv > 0): M^-1 = blockdiag(S)^-1 + sum_l Z_l blockdiag(S_l)^-1 Z_l^T
    //   + Z_top S_top^-1 Z_top^T; the top level uses Ac / Ai / yc / cdot above with nc = 3 * ml[mlLv - 1].n
    uint32_t mlLv, mlLocal, mlWin, mlCdot;   // levels, window-local levels, restriction windows, entries of cdot
    const uint32_t* tl_ptr; const uint32_t* tl_edge; const uint32_t* tl_node;   // landmark-major CSR over the top level's coarse edges
    // the top level as k_ml_top sees it: world == 1 the last level's own arrays; sharded: the top levels of all ranks
    // concatenated (node index = xc.top_off + local node), C gathered from the peers before every assembly
    const uint32_t* gt_node_ptr; const uint32_t* gt_edge_lm; double* gtC; uint32_t gt_nnz;
    XDev xc;                    // peer exchange (xc.world <= 1: none)
    double* r2;                 // second residual buffer (the restriction recomputes halo poses: r ping-pongs between r and r2)
    MlDev ml[kMlMaxLevels];
};

// kernel launchers (kernels.cu).  All take the launching stream.
void launch_linearize(const Dev& d, cudaStream_t s);            // K1: residuals, Jacobians, Hpl, pose-side H/b, slot partials
void launch_lm_sum(const Dev& d, bool finalize, cudaStream_t s);   // K2a: landmark-major sum of slot partials (+ K2b when finalize)
void launch_lm_finalize(const Dev& d, cudaStream_t s);          // K2b: Hll, b_l, prior, Hll^-1; chi2
void launch_schur_prep(const Dev& d, cudaStream_t s);           // K3: diag(S)^-1 and reduced rhs
void launch_coarse_assemble(const Dev& d, cudaStream_t s);      // zoff, C, Ac = Z^T S Z (dense, deterministic)
void launch_ml_levels(const Dev& d, cudaStream_t s);            // multilevel: C_l and blockdiag(S_l)^-1 level by level
void launch_ml_top(const Dev& d, cudaStream_t s);               // this rank's block rows of the dense top operator into Ac
void launch_ml_cut_record(const Dev& d, cudaStream_t s);        // sharded: records of this rank's ODOM edges across a cut (xc.cutrec)
void launch_ml_cut_apply(const Dev& d, cudaStream_t s);         // sharded: all ranks' records added to the replicated top operator
bool launch_pcg_stream(const Dev& d, cudaStream_t s);           // streamed persistent PCG (no back-substitution)
bool launch_coarse_invert(const Dev& d, cudaStream_t s);        // Ai = Ac^-1 + F_COARSE, one cooperative launch (blocked Gauss-Jordan)
void launch_pcg_coarse(const Dev& d, int init, cudaStream_t s);           // yc = Ac^-1 Z^T r, cdot (phased path)
bool launch_pcg_persistent(const Dev& d, cudaStream_t s);     // whole PCG + back-substitution, one cooperative launch
bool launch_sb_build(const Dev& d, cudaStream_t s);           // smoother-block inverses of the resident kernel (needs minv, hinv: after launch_schur_prep)
bool launch_pcg_resident(const Dev& d, cudaStream_t s);       // same, Hpl resident in shared memory (needs d.res_nb > 0)
int resident_blocks_available();                              // SM count if the resident kernel can be co-scheduled, else 0
void launch_pcg_init(const Dev& d, cudaStream_t s);
void launch_pcg_pass_a(const Dev& d, int mode, cudaStream_t s); // mode 0: v = p (with p = z + beta p fused), 1: v = x
void launch_pcg_mid(const Dev& d, int mode, cudaStream_t s);   // w = sum of slot partials; mode 0 also sp = Hpp p + ODOM part
uint32_t launch_pcg_pass_b(const Dev& d, cudaStream_t s);      // returns the number of p'Sp partials left in d.part
void launch_pcg_hpp(const Dev& d, cudaStream_t s);              // sp = Hpp p + ODOM part (pose-parallel)
void launch_pack_boundary(const Dev& d, const double* vec, cudaStream_t s);   // own boundary entries of vec -> d.bnd, zeros elsewhere
void launch_unpack_ghost(const Dev& d, double* vec, cudaStream_t s);          // d.bnd -> ghost entries of vec
void launch_pcg_psp(const Dev& d, uint32_t nparts, cudaStream_t s);              // local p'Sp -> scal[S_PSP]
void launch_pcg_alpha(const Dev& d, cudaStream_t s);
void launch_pcg_update(const Dev& d, cudaStream_t s);
void launch_pcg_rznew(const Dev& d, cudaStream_t s);            // local r'z -> scal[S_RZNEW]
void launch_pcg_beta(const Dev& d, cudaStream_t s);
void launch_pcg_init_fix(const Dev& d, cudaStream_t s);          // rz0 / done flag from the all-reduced scal[S_RZ]
void launch_backsub(const Dev& d, cudaStream_t s);              // delta_l = Hll^-1 (b_l - Hpl^T delta_p)
void launch_update_state(const Dev& d, double step, cudaStream_t s);   // K5 + ||delta||^2 (ghost poses take delta from d.bnd)
void launch_init_pose(const double* xyt, uint32_t n, double4* out, cudaStream_t s);
void launch_read_pose(const double4* pose, uint32_t n, double* xyt, cudaStream_t s);
// peer exchange between separately launched kernels (xchg.cuh)
void launch_x_signal_wait(const Dev& d, uint32_t point, uint64_t value, cudaStream_t s);   // publish `value`, then wait for every peer's
void launch_x_lm5_exchange(const Dev& d, uint64_t v, cudaStream_t s);   // landmark sums -> owners (flag v), owners finalize and push Hll / Hll^-1 / b_l to all (flag v + 1), wait
void launch_x_push_ghost(const Dev& d, const double* vec, cudaStream_t s);   // boundary-pose entries of vec -> ghost buffers of the peers
void launch_x_w_push(const Dev& d, cudaStream_t s);            // back-substitution: this rank's Hpl^T delta_p per touched landmark -> owners
void launch_x_backsub_own(const Dev& d, cudaStream_t s);       // owners: delta_l = Hll^-1 (b_l - sum over ranks), pushed to every rank
void launch_x_scal_push(const Dev& d, uint32_t slot, const double* src, cudaStream_t s);
void launch_x_scal_sum(const Dev& d, uint32_t slot, double* dst, cudaStream_t s);
size_t linearize_smem_bytes();
extern std::atomic<uint64_t> g_launches;   // kernels launched by this library since load (all contexts, all threads)

}  // namespace tsb
