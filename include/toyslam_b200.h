/* toyslam_b200.h -- C ABI of the B200-native (sm_100a) ToySlam graph-optimisation path.
 *
 * Drop-in boundary for the reference's PIPELINE=gpu branch.  The reference has no FFI of its
 * own (its CUDA classes are C++ templates selected in remote/app/GraphManager.h:73-122); the
 * entry points below are what `GraphGpu<T>` / `OptimizerGpu<T>` shims bind instead:
 *
 *   tsb_create / tsb_destroy   <- per-request CreateSolver/CreateOptimizer/CreateGraph
 *                                 (remote/app/ConnectionHandler.h:18-20); here one context
 *                                 persists across requests
 *   tsb_load_wire              <- DeserializeGraph<T,GraphGpu,..>  (remote/serialization/DeserializeGraph.h:9-174)
 *                                 + GraphGpu::ToDevice             (remote/cuda/graph/GraphGpu.h:80-187)
 *   tsb_load_soa               <- the per-object AddVertex/AddEdge/FixVertex path (GraphGpu.h:44-58)
 *   tsb_optimize               <- IOptimizer<T>::Optimize(IGraph*)  (remote/optimizer/IOptimizer.h:14-21),
 *                                 semantics of OptimizerCpu<T>::Optimize (remote/optimizer/OptimizerCpu.h:25-183)
 *   tsb_linearize              <- one pass of the edge loop + fixed prior (OptimizerCpu.h:82-138)
 *   tsb_solve                  <- ISolver<T>::Solve (remote/solver/ISolver.h:10, SolverEigen.h:11-21)
 *   tsb_read_state             <- GraphGpu::ToHost (GraphGpu.h:190-224), fp64, before the wire cast
 *   tsb_reply_size / tsb_write_reply <- SerializeGraph<T,GraphGpu,..> (remote/serialization/SerializeGraph.h:8-72,
 *                                 SerializeGraphFuncCpu.h:10-65); payload only, the caller adds the
 *                                 4-byte size prefix (ConnectionHandlerGraph.h:16-29)
 *
 * Conventions: every function returns 0 on success or a TSB_E* code (message: tsb_last_error);
 * nothing throws, exits or aborts (reference: remote/cuda/Helper.h:5-16 exits the process).
 * The caller owns every host buffer it passes; the context owns all device memory.  One context,
 * one host thread.  All floating point in the context is fp64; the wire stays fp32 (main.cpp:40).
 * There is no CPU fallback: without a CUDA device tsb_create fails with TSB_ECUDA.
 */
#ifndef TOYSLAM_B200_H
#define TOYSLAM_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define TSB_OK 0
#define TSB_EINVAL 1   /* bad argument / call order */
#define TSB_EPARSE 2   /* malformed or unsupported wire payload */
#define TSB_ECUDA 3    /* CUDA runtime failure (or no device) */
#define TSB_ENCCL 4    /* NCCL failure */
#define TSB_ENOMEM 5

typedef struct tsb_ctx tsb_ctx;

typedef struct tsb_config {
    int32_t device;            /* CUDA device ordinal of this rank                                   */
    int32_t rank, world;       /* one process per GPU; world == 1: no communicator                    */
    const void* nccl_id;       /* world > 1: the 128-byte ncclUniqueId shared by all ranks            */
    double pcg_tol;            /* stop when r'M^-1r <= tol^2 * (r'M^-1r)_0;  <= 0 selects 1e-12       */
    int32_t pcg_max_iters;     /* <= 0 selects 4000                                                   */
    int32_t enable_exits;      /* 1: the three early exits of OptimizerCpu.h:140-179; 0: run all iterations */
    int32_t verbose;           /* 1: print the reference's stdout lines ("Summary() error = ...")     */
} tsb_config;

/* Graph as structure-of-arrays (fp32 values exactly as they would travel on the wire). */
typedef struct tsb_graph_soa {
    uint32_t n_pose, n_lm, n_odom, n_lmedge, n_fixed;
    const uint32_t* pose_id;  const float* pose_xyt;   /* n_pose x (x, y, theta)                      */
    const uint32_t* lm_id;    const float* lm_xy;      /* n_lm x (x, y)                               */
    const uint32_t* odom_ij;  const float* odom_Z;  const float* odom_w;  /* ids (2), 3x3 row-major (9), info diag (3) */
    const uint32_t* lmedge_ij; const float* lmedge_meas; const float* lmedge_w; /* ids (2), (range, bearing), info diag (2) */
    const uint32_t* fixed_id;
} tsb_graph_soa;

typedef struct tsb_counts {
    uint32_t n_pose, n_lm, n_odom, n_lmedge, n_fixed, dim;   /* global graph                          */
    uint32_t pose_begin, pose_end;                            /* canonical pose range owned by this rank */
    uint32_t n_lmedge_local, n_odom_local;                    /* edges linearised by this rank          */
    uint32_t n_chunks, n_blocks, n_slots;                     /* assembly structure of this rank        */
    uint32_t pcg_path;                                        /* 0 persistent kernel, 1 shared-memory-resident kernel, 2 one launch per phase (+ NCCL), 3 streamed persistent kernel */
    uint32_t replicated;                                      /* 1: world > 1 but the graph is solved whole on every rank (it fits one GPU's resident kernel) */
    uint32_t n_aggregates;                                    /* coarse space of the two-level preconditioner (this rank) */
    uint32_t n_ml_levels, n_ml_top;                           /* multilevel hierarchy of the streamed PCG: levels, nodes of the dense top level */
    uint32_t structure_reuses;                                /* loads on this context that found the previous request's topology (same integer part of the
                                                                 payload) and kept its assembly structure and device index arrays; only values were refreshed */
} tsb_counts;

typedef struct tsb_iter_stats {
    double chi2;               /* robustified error sum `err` (OptimizerCpu.h:117)                    */
    double delta_norm;         /* ||delta||_2 of the undamped step (OptimizerCpu.h:173)               */
    int32_t pcg_iters;
    double pcg_rel_residual;   /* sqrt(r'M^-1r / (r'M^-1r)_0) reached                                  */
    float ms_linearize;        /* device time of linearise+assemble (K1+K2), CUDA events              */
    float ms_solve;            /* device time of Schur prep + PCG + back-substitution                 */
} tsb_iter_stats;

typedef struct tsb_report {
    int32_t iterations_run;    /* iterations entered (== chi2 values written)                         */
    int32_t exit_code;         /* 0 all iterations, 1 "Error is getting worse", 2 "Plateau", 3 "CONVERGED" */
    double final_chi2;
    int64_t pcg_iters_total;
    int64_t kernel_launches;   /* kernels this library launched during the call                       */
    float ms_total;            /* device time of the whole call (CUDA events)                         */
    float ms_linearize_total, ms_solve_total;
    float ms_pcg_total;        /* part of ms_solve_total spent in the PCG itself (the persistent kernel when world == 1) */
    int32_t coarse_refreshes;  /* how many solves rebuilt the coarse operator of the two-level preconditioner */
    int32_t pcg_not_converged; /* solves that hit pcg_max_iters or ended with a non-finite residual (the step is then inexact) */
    int32_t coarse_failures;   /* solves whose dense coarse operator could not be inverted (they ran with block-Jacobi only) */
    int32_t singular_blocks;   /* GN iterations in which an unconstrained vertex (singular 2x2 / 3x3 block) was held at delta = 0 */
} tsb_report;

/* world > 1: rank 0 calls this once and ships the 128 bytes to every rank (e.g. over torch.distributed or MPI) */
int tsb_nccl_unique_id(void* out128);
int tsb_create(const tsb_config* cfg, tsb_ctx** out);
void tsb_destroy(tsb_ctx* ctx);
const char* tsb_last_error(const tsb_ctx* ctx);     /* ctx may be NULL: error of the last failed tsb_create */

int tsb_load_wire(tsb_ctx* ctx, const void* payload, uint32_t nbytes);
int tsb_load_soa(tsb_ctx* ctx, const tsb_graph_soa* g);
int tsb_get_counts(const tsb_ctx* ctx, tsb_counts* out);
/* host-only planning (no GPU): rank `rank` of `world`'s share of a payload; structure_hash covers the
 * rank-independent parts (pose cuts, boundary-pose list) and must be equal on every rank */
int tsb_host_plan(const void* payload, uint32_t nbytes, int32_t rank, int32_t world, tsb_counts* out,
                  uint32_t* n_ghost, uint32_t* n_boundary, uint64_t* structure_hash);
/* host-only: rank `rank`'s integer structure (same arrays as tsb_read_structure; sizes from tsb_host_plan) -- lets the
 * "per-rank index map == the 1-rank map restricted to the rank's pose range" test run without a GPU
 * (replaces the per-rank part of GraphGpu::ToDevice, remote/cuda/graph/GraphGpu.h:80-187) */
int tsb_host_structure(const void* payload, uint32_t nbytes, int32_t rank, int32_t world, uint32_t* lm_perm,
                       uint32_t* pose_ptr, uint32_t* chunk_desc, uint32_t* slot_lm, uint32_t* lm_packed);
int tsb_reset_state(tsb_ctx* ctx);                  /* state <- the values loaded (device-to-device)  */
int tsb_set_state(tsb_ctx* ctx, const double* pose_xytheta, const double* lm_xy);

int tsb_linearize(tsb_ctx* ctx, tsb_iter_stats* stats);
int tsb_solve(tsb_ctx* ctx, tsb_iter_stats* stats); /* needs a preceding tsb_linearize; leaves delta on the device */
int tsb_optimize(tsb_ctx* ctx, uint32_t iterations, tsb_report* report,
                 tsb_iter_stats* per_iter /* NULL or `iterations` entries */);

/* canonical order: poses by ascending id, then landmarks by ascending id (all ranks return the full state) */
int tsb_read_ids(const tsb_ctx* ctx, uint32_t* pose_id, uint32_t* lm_id);
int tsb_read_state(tsb_ctx* ctx, double* pose_xytheta, double* lm_xy);
int tsb_read_delta(tsb_ctx* ctx, double* delta_pose /* 3 per local pose */, double* delta_lm /* 2 per landmark */);

/* The assembled system of the last tsb_linearize, block by block (this rank's part):
 *   hpp_diag   6 per local pose  (00 01 02 11 12 22), prior included
 *   b_pose     3 per local pose
 *   hpl        6 per local LM edge, 3x2 row-major, in *request order* of the local LM edges
 *   hpp_off    3 per local ODOM edge: diagonal of the (i,j) block (A=-I, B=+I, diagonal information)
 *   hll        3 per landmark (00 01 11), b_lm 2 per landmark (reduced over ranks)
 * Any pointer may be NULL. */
int tsb_read_system(tsb_ctx* ctx, double* hpp_diag, double* b_pose, double* hpl, double* hpp_off,
                    double* hll, double* b_lm);

/* Integer structure ("edge -> block index map"), this rank's part.  Any pointer may be NULL.
 *   lm_perm[k]     request-order index (among local LM edges) of the k-th edge in assembly order
 *   lm_packed[k]   (pose - chunk_pose0) << 16 | slot-in-chunk   for the k-th edge in assembly order
 *   chunk_desc     8 x u32 per chunk: edge0, edge1, pose0, pose1, slot0, slot1, block, reserved
 *   slot_lm[s]     landmark index of partial-sum slot s
 *   pose_ptr       n_local_pose + 1 offsets into assembly order
 *   lm_part_ptr / lm_part_idx   landmark-major CSR over slots */
int tsb_read_structure(tsb_ctx* ctx, uint32_t* lm_perm, uint32_t* lm_packed, uint32_t* chunk_desc,
                       uint32_t* slot_lm, uint32_t* pose_ptr, uint32_t* lm_part_ptr, uint32_t* lm_part_idx);

/* host<->device bytes copied since the last tsb_load_* (for the end-to-end bench accounting) */
int tsb_get_transfer_bytes(const tsb_ctx* ctx, uint64_t* h2d, uint64_t* d2h);

int tsb_reply_size(tsb_ctx* ctx, uint32_t* nbytes);
int tsb_write_reply(tsb_ctx* ctx, void* out, uint32_t capacity);

/* ODOM edges without any allocation.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Evaluates every ODOM edge at the given state with the arithmetic of the pose-parallel ODOM kernel behind
 * tsb_linearize (EdgeSe2.h:24-38 residual, Huber weight of OptimizerCpu.h:36-46):
 *   hpp_off[3k..]  diagonal of the (i,j) Hessian block of edge k, = what tsb_read_system returns as hpp_off
 *   w_e[3k..]      W e of edge k (b_i += W e, b_j -= W e)
 *   chi2[0]        sum of the robustified edge errors (== tsb_iter_stats.chi2 of an ODOM-only graph)
 * Every pointer is a DEVICE pointer owned by the caller (poses 3 per pose as x, y, theta; edge_ij 2 per edge, indices
 * into the pose array; edge_zinv 6 per edge, the top two rows of inverse(Z); edge_w 3 per edge, information diagonal).
 * `stream` is a cudaStream_t (NULL: the default stream); the call enqueues two kernels and returns without
 * synchronising.  No context is needed and nothing is allocated or freed: the per-block chi2 partials live in a static
 * device array of TSB_NOALLOC_MAX_ODOM_EDGES / 256 doubles (32 KB per GPU that loads the library, until exit).  More
 * edges are rejected with TSB_EINVAL.  The array is shared by all calls in the process and keeps the last call's
 * contents (every launch overwrites what it reads): calls that overlap on the GPU race on it -- serialise them.
 * `err` (may be NULL) receives a message of at most err_cap bytes on failure. */
#define TSB_NOALLOC_MAX_ODOM_EDGES 1048576
int tsb_odom_linearize_noalloc(int32_t device, void* stream, uint32_t n_pose, uint32_t n_edge,
                               const double* pose_xytheta, const uint32_t* edge_ij, const double* edge_zinv,
                               const double* edge_w, double* hpp_off, double* w_e, double* chi2,
                               char* err, uint32_t err_cap);

#ifdef __cplusplus
}
#endif
#endif
