// api.cu -- C ABI (include/toyslam_b200.h): context, upload, Gauss-Newton driver.
// The driver restates the control flow of OptimizerCpu<T>::Optimize (remote/optimizer/OptimizerCpu.h:70-182):
// linearize -> "getting worse" test -> solve -> damped update (0.2) -> plateau / convergence tests.
#include "../../include/toyslam_b200.h"
#include "device_graph.h"

#include <nccl.h>

#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <type_traits>
#include <vector>

using namespace tsb;

static std::string g_create_error;

struct tsb_ctx {
    tsb_config cfg{};
    std::string err;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // [4],[5]: around the PCG proper
    HostGraph g;
    Structure st;
    int world = 1, rank = 0;                              // effective for the loaded graph: (1, 0) when replicated
    ResidentPlan rp;                                      // rp.nb > 0: this graph runs the shared-memory-resident PCG kernel
    Dev d{};
    // device memory of the loaded graph: bump-allocated from a few large blocks that outlive the graph, so a server
    // that optimises one request after another stops calling cudaMalloc / cudaFree after the first request
    std::vector<std::pair<char*, size_t>> arena;          // (base, capacity)
    size_t arena_blk = 0, arena_off = 0, arena_used = 0;
    double4* pose0 = nullptr; double2* lm0 = nullptr;     // state as loaded (tsb_reset_state)
    double* tmp_xyt = nullptr;                            // 3 * (nP + nG) scratch
    double* h_scal = nullptr; int* h_flags = nullptr;     // pinned
    bool loaded = false, linearized = false, solved = false;
    uint32_t part_cap = 0;
    bool pcg_streamed = false;                            // the last solve ran k_pcg_stream
    bool force_phased = false;                            // TSB_PCG_PHASED=1: one launch per PCG phase (the multi-GPU path)
    uint64_t h2d = 0, d2h = 0;                            // bytes copied since the last tsb_load_*
    ncclComm_t comm = nullptr;                            // world > 1: one communicator, collectives on `stream`
    // peer exchange over NVLink (device_graph.h XDev): this rank's region, the peers' regions opened through CUDA IPC
    char* xreg = nullptr; size_t xbytes = 0; char* xpeer[kXMaxWorld] = {}; bool xpeer_ipc[kXMaxWorld] = {}; uint64_t xseq = 0, xlaunch = 0;
    std::vector<uint32_t> gt_ecnt, gt_eoff;               // coarse edges of every rank's top level (global top lists)
    // structure reuse: topology of the graph whose structure / index arrays are on the device
    uint64_t topo_hash = 0; uint32_t topo_counts[6] = {0, 0, 0, 0, 0, 0}; bool topo_valid = false, reuse_ints = false;
    pvec<uint8_t> soa_wire;                                   // tsb_load_soa: the request bytes laid out from the SoA arrays (kept across requests)
    int structure_reuses = 0, st_world = 1, st_rank = 0;      // st_*: effective world / rank the kept structure was built for
    bool use_p2p = true;                                  // TSB_P2P=0: NCCL collectives between phase launches (round-1 path)
    // coarse levels of the preconditioner (two-level aggregates for the resident kernel, multilevel hierarchy for the
    // streamed one): the dense top operator is inverted by k_coarse_invert
    bool use_coarse = true;                               // TSB_COARSE=0: block-Jacobi only
    uint32_t ml_top = kMlTopDefault;                      // TSB_ML_TOP: nodes of the hierarchy's top level; TSB_ML=0: two-level aggregates everywhere
    // the coarse inverse is reused across GN iterations while it keeps working (any SPD M is a valid
    // preconditioner; a stale one only costs iterations): refreshed on the first solve of a request, when the
    // PCG needs > 1.25x the iterations it took right after the last refresh, or after `coarse_max_age` solves
    int sb_mode = 0;                                      // smoother blocks: 0 never, 1 always, 2 adaptive (default)
    double* sbinv_buf = nullptr;                          // Dev::sbinv is this buffer or null, solve by solve
    int coarse_age = -1, coarse_ref_iters = 0, coarse_max_age = 16, coarse_refreshes = 0, coarse_hold = 0, coarse_backoff = 1;
};

namespace {

#define CU(call)                                                                                   \
    do { cudaError_t e_ = (call); if (e_ != cudaSuccess) {                                         \
        c->err = std::string(#call) + ": " + cudaGetErrorString(e_); return TSB_ECUDA; } } while (0)

void release_arena(tsb_ctx* c) {
    for (auto& b : c->arena) cudaFree(b.first);
    c->arena.clear(); c->arena_blk = c->arena_off = 0;
}
void free_graph(tsb_ctx* c) {
    // a graph that needed several blocks: give them back so that the next one gets a single block of the full size
    if (c->arena.size() > 1) { cudaStreamSynchronize(c->stream); release_arena(c); }
    c->arena_blk = c->arena_off = 0;
    c->d = Dev{}; c->loaded = c->linearized = c->solved = false;
    c->pose0 = nullptr; c->lm0 = nullptr; c->tmp_xyt = nullptr; c->coarse_age = -1; c->sbinv_buf = nullptr;
}

template <typename T> int dalloc(tsb_ctx* c, T*& out, size_t n) {
    // 64 B of slack: the streaming kernels copy from addresses aligned down / sizes rounded up to 16 B
    const size_t bytes = (std::max<size_t>(n, 1) * sizeof(T) + 64 + 255) & ~size_t(255);
    while (c->arena_blk < c->arena.size() && c->arena_off + bytes > c->arena[c->arena_blk].second) { ++c->arena_blk; c->arena_off = 0; }
    if (c->arena_blk == c->arena.size()) {
        // first block: what the previous graph used (plus headroom); later blocks double the total
        size_t total = 0; for (auto& b : c->arena) total += b.second;
        const size_t cap = std::max(bytes, std::max<size_t>(c->arena.empty() ? c->arena_used + c->arena_used / 8 : total, size_t(16) << 20));
        void* p = nullptr;
        if (cudaMalloc(&p, cap) != cudaSuccess) {
            cudaGetLastError(); c->err = "cudaMalloc failed (" + std::to_string(cap) + " bytes)"; return TSB_ENOMEM; }
        c->arena.emplace_back((char*)p, cap); c->arena_off = 0;
    }
    out = (T*)(c->arena[c->arena_blk].first + c->arena_off); c->arena_off += bytes;
    { size_t used = c->arena_off; for (size_t i = 0; i < c->arena_blk; ++i) used += c->arena[i].second; c->arena_used = std::max(c->arena_used, used); }
    return TSB_OK;
}
template <typename T, typename U, typename A> int upload(tsb_ctx* c, const U*& out, const std::vector<T, A>& v) {
    static_assert(sizeof(T) == sizeof(U) || sizeof(U) % sizeof(T) == 0, "layout");
    T* p = nullptr;
    if (int rc = dalloc(c, p, v.size())) return rc;
    // same topology as the previous request: the bump allocator hands out the same addresses in the same order and the
    // integer (structure) arrays are still there -- only floating-point values travel again
    if (c->reuse_ints && std::is_integral<T>::value) { out = reinterpret_cast<const U*>(p); return TSB_OK; }
    if (!v.empty()) CU(cudaMemcpyAsync(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, c->stream));
    c->h2d += v.size() * sizeof(T);
    out = reinterpret_cast<const U*>(p); return TSB_OK;
}
#define TRY(x) do { if (int rc_ = (x)) return rc_; } while (0)

int x_ensure(tsb_ctx* c, size_t bytes);
int x_gatherv_u32(tsb_ctx* c, const std::vector<uint32_t>& mine, std::vector<std::vector<uint32_t>>& all);

int upload_graph(tsb_ctx* c) {
    const HostGraph& g = c->g; const Structure& s = c->st; Dev& d = c->d;
    d.nP = s.nP; d.nG = s.nG; d.L = g.L; d.nE = (uint32_t)s.lm_perm.size(); d.nO = (uint32_t)s.od_local.size();
    d.nOwned = s.n_od_owned; d.nChunks = (uint32_t)s.chunks.size(); d.nBlocks = (uint32_t)s.blk_chunk_ptr.size() - 1;
    d.nSlots = (uint32_t)s.slot_lm.size();
    // peer exchange (sharded, one process per GPU): region layout -- identical on every rank -- and the peers' mappings
    d.xc = XDev{};
    if (c->world > 1 && c->use_p2p) {
        if (s.nG > kXMaxGhost) { c->err = "peer exchange: too many ghost poses (general ODOM graphs across a cut need TSB_P2P=0)"; return TSB_EINVAL; }
        XDev& x = d.xc;
        x.world = (uint32_t)c->world; x.rank = (uint32_t)c->rank; x.slice = (g.L + x.world - 1) / x.world; if (!x.slice) x.slice = 1;
        auto al = [](size_t v) { return (v + 255) & ~size_t(255); };
        size_t o = 0;
        x.o_flags = o; o = al(o + 8ull * kXPoints * kXMaxWorld);
        x.o_scal = o; o = al(o + 8ull * 16 * kXMaxWorld);
        x.o_ghost = o; o = al(o + 8ull * 3 * kXMaxGhost);
        x.o_lm5 = o; o = al(o + 8ull * 5 * x.world * x.slice);
        x.o_hll = o; o = al(o + 8ull * 3 * g.L + 64);
        x.o_hinv = o; o = al(o + 8ull * 3 * g.L + 64);
        x.o_bl = o; o = al(o + 8ull * 2 * g.L + 64);
        x.o_w = o; o = al(o + 8ull * 2 * x.world * g.L + 64);          // rx[src][landmark] of the PCG's all-gather
        x.o_bw = o; o = al(o + 8ull * 2 * x.world * x.slice);
        x.o_u = o; o = al(o + 8ull * 2 * g.L + 64);
        x.o_dl = o; o = al(o + 8ull * 2 * g.L + 64);
        x.o_top = o; o = al(o + 8ull * (3 * 4096 + 4 * kXMaxWorld));
        TRY(x_ensure(c, o));
        for (int r = 0; r < c->world; ++r) x.peer[r] = c->xpeer[r];
        x.nExp = (uint32_t)(s.x_exp.size() / 3);
        TRY(upload(c, x.exp, s.x_exp)); TRY(upload(c, x.lm_mask, s.lm_mask));
        TRY(dalloc(c, x.done, (size_t)2)); CU(cudaMemsetAsync(x.done, 0, 8, c->stream));
    }
    const bool xp = d.xc.world > 1;
    // per-edge data gathered into assembly order
    pvec<double> z, w; z.resize(2ull * d.nE); w.resize(2ull * d.nE);
#pragma omp parallel for schedule(static)
    for (long long kk = 0; kk < (long long)d.nE; ++kk) {
        const uint32_t k = (uint32_t)kk, e = s.le_local[s.lm_perm[k]];
        z[2ull * k] = g.le_z[2ull * e]; z[2ull * k + 1] = g.le_z[2ull * e + 1];
        w[2ull * k] = g.le_w[2ull * e]; w[2ull * k + 1] = g.le_w[2ull * e + 1];
    }
    std::vector<double> oz(6ull * d.nO), ow(3ull * d.nO);
    for (uint32_t k = 0; k < d.nO; ++k) {
        const uint32_t e = s.od_local[k];
        std::memcpy(&oz[6ull * k], &g.od_zinv[6ull * e], 48); std::memcpy(&ow[3ull * k], &g.od_w[3ull * e], 24);
    }
    TRY(upload(c, d.lm_packed, s.lm_packed)); TRY(upload(c, d.le_z, z)); TRY(upload(c, d.le_w, w));
    TRY(upload(c, d.pose_ptr, s.pose_ptr)); TRY(upload(c, d.blk_chunk_ptr, s.blk_chunk_ptr));
    { std::vector<uint32_t> raw(8ull * s.chunks.size());
      if (!s.chunks.empty()) std::memcpy(raw.data(), s.chunks.data(), raw.size() * 4);
      TRY(upload(c, d.chunks, raw)); }
    TRY(upload(c, d.slot_lm, s.slot_lm)); TRY(upload(c, d.slot_ptr, s.slot_ptr)); TRY(upload(c, d.slot_edge, s.slot_edge));
    TRY(upload(c, d.lm_part_ptr, s.lm_part_ptr)); TRY(upload(c, d.lm_part_idx, s.lm_part_idx));
    TRY(upload(c, d.od_a, s.od_a)); TRY(upload(c, d.od_b, s.od_b)); TRY(upload(c, d.od_zinv, oz)); TRY(upload(c, d.od_w, ow));
    TRY(upload(c, d.od_out_ptr, s.od_out_ptr)); TRY(upload(c, d.od_in_ptr, s.od_in_ptr)); TRY(upload(c, d.od_in_edge, s.od_in_edge));
    { std::vector<uint32_t> nb(4ull * d.nP, 0xFFFFFFFFu);
      for (uint32_t p = 0; p < d.nP; ++p) {
          const uint32_t no = s.od_out_ptr[p + 1] - s.od_out_ptr[p], ni = s.od_in_ptr[p + 1] - s.od_in_ptr[p];
          if (no > 1 || ni > 1) { nb[4ull * p] = 0xFFFFFFFEu; continue; }
          if (no) { const uint32_t k = s.od_out_ptr[p]; nb[4ull * p] = s.od_b[k]; nb[4ull * p + 1] = k; }
          if (ni) { const uint32_t k = s.od_in_edge[s.od_in_ptr[p]]; nb[4ull * p + 2] = s.od_a[k]; nb[4ull * p + 3] = k; }
      }
      TRY(upload(c, d.od_nb, nb)); }
    TRY(upload(c, d.fixed_pose_count, s.fixed_pose_count)); TRY(upload(c, d.fixed_lm_count, s.fixed_lm_count));
    // state: local poses then ghosts
    const uint32_t nPG = d.nP + d.nG;
    std::vector<double> xyt(3ull * nPG);
    if (d.nP) std::memcpy(xyt.data(), &g.pose_xyt[3ull * s.pose_begin], 24ull * d.nP);
    for (uint32_t q = 0; q < d.nG; ++q) std::memcpy(&xyt[3ull * (d.nP + q)], &g.pose_xyt[3ull * s.ghost_pose[q]], 24);
    TRY(dalloc(c, c->tmp_xyt, 3ull * std::max(nPG, g.P)));
    TRY(dalloc(c, d.pose, nPG)); TRY(dalloc(c, c->pose0, nPG)); TRY(dalloc(c, d.lm, g.L)); TRY(dalloc(c, c->lm0, g.L));
    if (nPG) CU(cudaMemcpyAsync(c->tmp_xyt, xyt.data(), 24ull * nPG, cudaMemcpyHostToDevice, c->stream));
    launch_init_pose(c->tmp_xyt, nPG, c->pose0, c->stream);
    if (g.L) CU(cudaMemcpyAsync(c->lm0, g.lm_xy.data(), 16ull * g.L, cudaMemcpyHostToDevice, c->stream));
    c->h2d += 24ull * nPG + 16ull * g.L;
    // system and solver vectors
    TRY(dalloc(c, d.hpl, 6ull * d.nE)); TRY(dalloc(c, d.hoo, 3ull * d.nO));
    TRY(dalloc(c, d.hpp, 6ull * d.nP)); TRY(dalloc(c, d.bp, 3ull * d.nP)); TRY(dalloc(c, d.pose_od, 6ull * d.nP));
    TRY(dalloc(c, d.lm_part, 5ull * d.nSlots)); TRY(dalloc(c, d.lm_sum, 5ull * g.L + 1));     // + chi2, so that one all-reduce carries both when world > 1
    if (xp) {     // written by the landmark's owner rank: they live in the exchange region
        d.hll = reinterpret_cast<double*>(c->xreg + d.xc.o_hll); d.hinv = reinterpret_cast<double*>(c->xreg + d.xc.o_hinv); d.bl = reinterpret_cast<double*>(c->xreg + d.xc.o_bl);
    } else { TRY(dalloc(c, d.hll, 3ull * g.L)); TRY(dalloc(c, d.hinv, 3ull * g.L)); TRY(dalloc(c, d.bl, 2ull * g.L)); }
    TRY(dalloc(c, d.minv, 6ull * d.nP)); TRY(dalloc(c, d.rhs, 3ull * d.nP));
    TRY(dalloc(c, d.x, 3ull * d.nP)); TRY(dalloc(c, d.r, 3ull * d.nP)); TRY(dalloc(c, d.z, 3ull * d.nP));
    TRY(dalloc(c, d.p, 3ull * nPG)); TRY(dalloc(c, d.sp, 3ull * d.nP));
    TRY(dalloc(c, d.w_part, 2ull * d.nSlots)); d.nBnd = (uint32_t)s.boundary.size();
    if (xp) {     // u = Hll^-1 w is pushed by the owners, ghost-pose values by their owners
        d.w_sum = reinterpret_cast<double*>(c->xreg + d.xc.o_u); d.bnd = reinterpret_cast<double*>(c->xreg + d.xc.o_ghost); d.nBnd = d.nG;
        std::vector<uint32_t> ident(d.nG); for (uint32_t q = 0; q < d.nG; ++q) ident[q] = q;
        TRY(upload(c, d.bnd_local, ident)); TRY(upload(c, d.ghost_bnd_slot, ident));
    } else {
    TRY(dalloc(c, d.w_sum, 2ull * g.L + 3ull * d.nBnd + 8)); d.bnd = d.w_sum + 2ull * g.L;
    { std::vector<uint32_t> bl(d.nBnd, 0xFFFFFFFFu);
      for (size_t k = 0; k < s.own_bnd_slot.size(); ++k) bl[s.own_bnd_slot[k]] = s.own_bnd_pose[k];
      TRY(upload(c, d.bnd_local, bl)); TRY(upload(c, d.ghost_bnd_slot, s.ghost_bnd_slot)); }
    }
    if (xp) d.dl = reinterpret_cast<double*>(c->xreg + d.xc.o_dl); else TRY(dalloc(c, d.dl, 2ull * g.L));
    // coarse space
    d.aggK = s.agg_K; d.nA = s.nA; d.nc = 3 * s.nA; d.nAS = (uint32_t)s.aslot_lm.size(); d.nPairs = (uint32_t)s.cpair_a.size();
    d.nSeg = (d.nP + 31) / 32;
    TRY(upload(c, d.agg_ptr, s.agg_ptr)); TRY(upload(c, d.pose_agg, s.pose_agg));
    d.res_nb = c->rp.nb;
    d.lm_sum_chi2 = (c->world > 1 && !xp) ? 1u : 0u;
    { const char* e = std::getenv("TSB_RES_CGSYNC");                  // 1: cooperative-groups grid.sync() in the resident kernel
      if (!(e && e[0] == '1')) TRY(dalloc(c, d.gbar, (size_t)1)); }
    if (d.res_nb) {
        TRY(upload(c, d.res_pose_cut, c->rp.pose_cut)); TRY(upload(c, d.res_lm_cut, c->rp.lm_cut)); TRY(upload(c, d.res_agg_cut, c->rp.agg_cut));
        TRY(upload(c, d.lmo_ptr, c->rp.lmo_ptr)); TRY(upload(c, d.lmo_edge, c->rp.lmo_edge)); TRY(upload(c, d.edge_lm, c->rp.edge_lm));
        // Dense smoother blocks (sblock_smoother.cuh) instead of the 3x3 pose blocks as the fine level of the resident
        // kernel's preconditioner: 2.4x fewer PCG iterations on the first solves of C2 but ~1.27x the time per iteration
        // (the inverses stream from L2), and little gain once the Huber weights have settled.  Default: adaptive per
        // request (tsb_optimize); TSB_SMOOTHER=1 always, TSB_SMOOTHER=0 never.
        { const char* e = std::getenv("TSB_SMOOTHER");
          c->sb_mode = e && e[0] == '0' ? 0 : e && e[0] == '1' ? 1 : 2;
          if (c->sb_mode) TRY(dalloc(c, c->sbinv_buf, (size_t)d.res_nb * kSbPerBlock * (3 * kSbCap) * (3 * kSbCap)));
          d.sbinv = c->sb_mode ? c->sbinv_buf : nullptr; }
    }
    TRY(upload(c, d.edge_pose, s.edge_pose)); TRY(upload(c, d.aslot_lm, s.aslot_lm)); TRY(upload(c, d.aslot_agg, s.aslot_agg));
    TRY(upload(c, d.aslot_ptr, s.aslot_ptr)); TRY(upload(c, d.aslot_edge, s.aslot_edge));
    TRY(upload(c, d.cpair_a, s.cpair_a)); TRY(upload(c, d.cpair_b, s.cpair_b)); TRY(upload(c, d.cpair_ptr, s.cpair_ptr));
    TRY(upload(c, d.ctrip, s.ctrip));
    // multilevel hierarchy (streamed PCG): structure arrays and per-level vectors
    d.mlLv = (uint32_t)s.ml.size(); d.mlLocal = s.ml_local; d.mlWin = (d.nP + kMlWindow - 1) / kMlWindow; d.mlCdot = d.nc;
    if (d.mlLv) {
        // the exactly-solved top level: this rank's last level, or (sharded) the last levels of all ranks concatenated
        const MlLevel& t = s.ml.back();
        uint32_t top_total = t.n; d.xc.top_off = 0;
        if (xp) { top_total = 0; for (int r = 0; r < c->world; ++r) { if (r == c->rank) d.xc.top_off = top_total; top_total += s.top_nodes[r]; } }
        d.xc.top_total = top_total;
        if (top_total > 4096) { c->err = "multilevel hierarchy: global top level too large"; return TSB_EINVAL; }
        d.nc = 3 * top_total; d.mlCdot = d.nc;
        for (uint32_t l = 0; l < d.mlLv; ++l) {
            const MlLevel& h = s.ml[l]; MlDev& m = d.ml[l];
            m.n = h.n; m.ratio = h.ratio; m.spacing = h.spacing; m.nnz = (uint32_t)h.edge_lm.size();
            m.slots = (kMlWindow + h.spacing - 1) / h.spacing + 2; m.cdot_off = 0;
            TRY(upload(c, m.node_ptr, h.node_ptr)); TRY(upload(c, m.edge_lm, h.edge_lm)); TRY(upload(c, m.ref_ptr, h.ref_ptr)); TRY(upload(c, m.ref, h.ref));
            TRY(dalloc(c, m.C, 6ull * m.nnz)); TRY(dalloc(c, m.dinv, 6ull * m.n)); TRY(dalloc(c, m.m, 3ull * m.n)); TRY(dalloc(c, m.g, 3ull * m.n + 3));
            if (l >= d.mlLocal) TRY(dalloc(c, m.wpart, 3ull * d.mlWin * m.slots));
            if (l >= d.mlLocal && l + 1 < d.mlLv) { m.cdot_off = d.mlCdot; d.mlCdot += m.n; }
        }
        // global top lists: per node the landmark-sorted coarse edges; sharded: every rank's (counts, landmarks) gathered
        std::vector<uint32_t> gptr, glm;
        c->gt_ecnt.assign(c->world, 0); c->gt_eoff.assign(c->world + 1, 0);
        if (!xp) { gptr = t.node_ptr; glm = t.edge_lm; c->gt_ecnt[0] = (uint32_t)glm.size(); c->gt_eoff[1] = c->gt_ecnt[0]; }
        else {
            std::vector<uint32_t> cnt(t.n);
            for (uint32_t K = 0; K < t.n; ++K) cnt[K] = t.node_ptr[K + 1] - t.node_ptr[K];
            std::vector<std::vector<uint32_t>> acnt, alm;
            TRY(x_gatherv_u32(c, cnt, acnt)); TRY(x_gatherv_u32(c, t.edge_lm, alm));
            gptr.push_back(0);
            for (int r = 0; r < c->world; ++r) {
                if (acnt[r].size() != s.top_nodes[r]) { c->err = "multilevel hierarchy: ranks disagree on the top level"; return TSB_EINVAL; }
                for (uint32_t v : acnt[r]) gptr.push_back(gptr.back() + v);
                glm.insert(glm.end(), alm[r].begin(), alm[r].end());
                c->gt_ecnt[r] = (uint32_t)alm[r].size(); c->gt_eoff[r + 1] = c->gt_eoff[r] + c->gt_ecnt[r];
            }
        }
        d.gt_nnz = (uint32_t)glm.size();
        // landmark-major CSR over the (global) top-level coarse edges (k_ml_top)
        std::vector<uint32_t> tp(g.L + 1, 0), te(glm.size()), tn(glm.size());
        for (uint32_t l : glm) tp[l + 1]++;
        for (uint32_t l = 0; l < g.L; ++l) tp[l + 1] += tp[l];
        { std::vector<uint32_t> c2(tp.begin(), tp.end() - 1);
          for (uint32_t K = 0; K < top_total; ++K) for (uint32_t e = gptr[K]; e < gptr[K + 1]; ++e) { const uint32_t at = c2[glm[e]]++; te[at] = e; tn[at] = K; } }
        TRY(upload(c, d.tl_ptr, tp)); TRY(upload(c, d.tl_edge, te)); TRY(upload(c, d.tl_node, tn));
        if (xp) { TRY(upload(c, d.gt_node_ptr, gptr)); TRY(upload(c, d.gt_edge_lm, glm)); TRY(dalloc(c, d.gtC, 6ull * d.gt_nnz)); }
        d.xc.nCut = 0; d.xc.cut = nullptr; d.xc.cutrec = nullptr;
        if (xp) {     // owned ODOM edges whose far end is the first pose of a higher rank: they join the top operator (mlevel.cuh)
            std::vector<uint32_t> cut;
            std::vector<uint32_t> toff(c->world + 1, 0);
            for (int r = 0; r < c->world; ++r) toff[r + 1] = toff[r] + s.top_nodes[r];
            for (uint32_t k = 0; k < s.n_od_owned && cut.size() < 3 * kXMaxCut; ++k) {
                if (s.od_a[k] >= d.nP || s.od_b[k] < d.nP) continue;
                const uint32_t gp = s.ghost_pose[s.od_b[k] - d.nP];
                const uint32_t ro = (uint32_t)(std::upper_bound(s.cuts.begin(), s.cuts.end(), gp) - s.cuts.begin()) - 1;
                if ((int)ro > c->rank && gp == s.cuts[ro]) { cut.push_back(k); cut.push_back(s.od_a[k]); cut.push_back(toff[ro]); }
            }
            d.xc.nCut = (uint32_t)(cut.size() / 3);
            cut.resize(3 * kXMaxCut, 0);
            TRY(upload(c, d.xc.cut, cut));
            TRY(dalloc(c, d.xc.cutrec, (size_t)c->world * kXMaxCut * kXCutRec));
        }
        else { d.gt_node_ptr = d.ml[d.mlLv - 1].node_ptr; d.gt_edge_lm = d.ml[d.mlLv - 1].edge_lm; d.gtC = d.ml[d.mlLv - 1].C; }
        TRY(dalloc(c, d.r2, 3ull * d.nP));
    }
    TRY(dalloc(c, d.zoff, d.nP)); TRY(dalloc(c, d.cC, 6ull * d.nAS)); TRY(dalloc(c, d.Ac, (size_t)d.nc * d.nc)); TRY(dalloc(c, d.Ai, (size_t)d.nc * d.nc));
    if (d.nc > 3 * kCoarseMaxAgg) TRY(dalloc(c, d.rcs, (size_t)kCoarseScratchBlocks * d.nc));
    TRY(dalloc(c, d.rcp, 3ull * d.nSeg)); TRY(dalloc(c, d.yc, d.nc + 3ull)); TRY(dalloc(c, d.cdot, (size_t)std::max(d.mlCdot, d.nc))); TRY(dalloc(c, d.cinfo, (size_t)2)); TRY(dalloc(c, d.cbar, (size_t)1));
    CU(cudaMemsetAsync(d.cinfo, 0, 8, c->stream));
    // partial-sum scratch: one entry per unit, per 256 vertices, or per block of a persistent kernel (<= 8 per SM)
    c->part_cap = std::max<uint32_t>(std::max<uint32_t>(d.nBlocks, (std::max(d.nP, g.L) + kThreads - 1) / kThreads) + 1, 4096);
    TRY(dalloc(c, d.part, c->part_cap)); TRY(dalloc(c, d.part2, c->part_cap));
    TRY(dalloc(c, d.scal, (size_t)S_COUNT)); TRY(dalloc(c, d.flags, (size_t)F_COUNT));
    { const char* e = std::getenv("TSB_PCG_TIMING"); if (e && e[0] == '1') { TRY(dalloc(c, d.dbg, (size_t)16 + 9 * 1024)); CU(cudaMemsetAsync(d.dbg, 0, 8 * (16 + 9 * 1024), c->stream)); } }
    CU(cudaMemsetAsync(d.scal, 0, sizeof(double) * S_COUNT, c->stream));
    CU(cudaMemsetAsync(d.flags, 0, sizeof(int) * F_COUNT, c->stream));
    CU(cudaMemsetAsync(d.p, 0, 24ull * nPG, c->stream));
    CU(cudaMemsetAsync(d.x, 0, 24ull * d.nP, c->stream));
    CU(cudaMemsetAsync(d.dl, 0, 16ull * g.L, c->stream));
    const double tol = c->cfg.pcg_tol > 0 ? c->cfg.pcg_tol : 1e-12;
    const double tol2 = tol * tol; const int maxit = c->cfg.pcg_max_iters > 0 ? c->cfg.pcg_max_iters : 4000;
    CU(cudaMemcpyAsync(d.scal + S_TOL2, &tol2, 8, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d.flags + F_MAXIT, &maxit, 4, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d.pose, c->pose0, sizeof(double4) * nPG, cudaMemcpyDeviceToDevice, c->stream));
    CU(cudaMemcpyAsync(d.lm, c->lm0, sizeof(double2) * g.L, cudaMemcpyDeviceToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->loaded = true; c->linearized = c->solved = false;
    return TSB_OK;
}

// in-place sum over ranks on the compute stream; no-op for world == 1
int allreduce(tsb_ctx* c, double* buf, size_t n) {
    if (c->world <= 1 || n == 0) return TSB_OK;
    const ncclResult_t r = ncclAllReduce(buf, buf, n, ncclDouble, ncclSum, c->comm, c->stream);
    if (r != ncclSuccess) { c->err = std::string("ncclAllReduce: ") + ncclGetErrorString(r); return TSB_ENCCL; }
    return TSB_OK;
}

// Exchange region of at least `bytes` on every rank (collective: all ranks pass the same size).  Allocated with
// cudaMalloc (CUDA IPC needs a plain allocation), zeroed, and opened on the peers: handles travel through NCCL.
int x_ensure(tsb_ctx* c, size_t bytes) {
    if (bytes <= c->xbytes) return TSB_OK;
    CU(cudaStreamSynchronize(c->stream));
    for (int r = 0; r < c->cfg.world; ++r) { if (r != c->cfg.rank && c->xpeer[r] && c->xpeer_ipc[r]) cudaIpcCloseMemHandle(c->xpeer[r]); c->xpeer[r] = nullptr; }
    if (c->xreg) {      // peers may still be reading the old region: everybody reaches this point before anybody frees
        int* dz = nullptr; CU(cudaMalloc((void**)&dz, 4)); CU(cudaMemsetAsync(dz, 0, 4, c->stream));
        ncclAllReduce(dz, dz, 1, ncclInt, ncclSum, c->comm, c->stream); CU(cudaStreamSynchronize(c->stream)); cudaFree(dz);
        cudaFree(c->xreg); c->xreg = nullptr; c->xbytes = 0;
    }
    bytes = std::max(bytes + bytes / 4, size_t(32) << 20);
    CU(cudaMalloc((void**)&c->xreg, bytes));
    CU(cudaMemsetAsync(c->xreg, 0, bytes, c->stream));
    // what every rank tells the others about its region: the CUDA-IPC handle (ranks in other processes open it), and
    // process id + device + raw address (ranks that are threads of THIS process -- graph_optimizer driving several GPUs --
    // enable peer access and use the address directly: cudaIpcOpenMemHandle refuses handles of its own process)
    struct XInfo { cudaIpcMemHandle_t h; long long pid; unsigned long long addr; int device; int pad; };
    static_assert(sizeof(cudaIpcMemHandle_t) == 64 && sizeof(XInfo) == 88, "exchange record");
    XInfo mine{}; mine.pid = (long long)getpid(); mine.addr = (unsigned long long)(uintptr_t)c->xreg; mine.device = c->cfg.device;
    CU(cudaIpcGetMemHandle(&mine.h, c->xreg));
    char* hb = nullptr;
    CU(cudaMalloc((void**)&hb, sizeof(XInfo) * (size_t)c->cfg.world));
    CU(cudaMemcpyAsync(hb + sizeof(XInfo) * (size_t)c->cfg.rank, &mine, sizeof(XInfo), cudaMemcpyHostToDevice, c->stream));
    const ncclResult_t r = ncclAllGather(hb + sizeof(XInfo) * (size_t)c->cfg.rank, hb, sizeof(XInfo), ncclChar, c->comm, c->stream);
    if (r != ncclSuccess) { cudaFree(hb); c->err = std::string("ncclAllGather: ") + ncclGetErrorString(r); return TSB_ENCCL; }
    std::vector<XInfo> all(c->cfg.world);
    CU(cudaMemcpyAsync(all.data(), hb, sizeof(XInfo) * (size_t)c->cfg.world, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    cudaFree(hb);
    for (int q = 0; q < c->cfg.world; ++q) {
        c->xpeer_ipc[q] = false;
        if (q == c->cfg.rank) { c->xpeer[q] = c->xreg; continue; }
        if (all[q].pid == mine.pid) {
            const cudaError_t e = cudaDeviceEnablePeerAccess(all[q].device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { c->err = std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e); return TSB_ECUDA; }
            cudaGetLastError();
            c->xpeer[q] = (char*)(uintptr_t)all[q].addr;
            continue;
        }
        const cudaError_t e = cudaIpcOpenMemHandle((void**)&c->xpeer[q], all[q].h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) { c->err = std::string("cudaIpcOpenMemHandle (the peer exchange needs GPUs with peer access): ") + cudaGetErrorString(e); return TSB_ECUDA; }
        c->xpeer_ipc[q] = true;
    }
    c->xbytes = bytes; c->xseq = 0; c->xlaunch = 0;           // flags are zero everywhere: every rank memset before it entered the all-gather
    return TSB_OK;
}
// load-time gather of one uint32 list per rank (collective; NCCL on the compute stream)
int x_gatherv_u32(tsb_ctx* c, const std::vector<uint32_t>& mine, std::vector<std::vector<uint32_t>>& all) {
    const int W = c->cfg.world, R = c->cfg.rank;
    unsigned long long* dcnt = nullptr;
    CU(cudaMalloc((void**)&dcnt, 8 * (size_t)W));
    const unsigned long long n = mine.size();
    CU(cudaMemcpyAsync(dcnt + R, &n, 8, cudaMemcpyHostToDevice, c->stream));
    if (ncclAllGather(dcnt + R, dcnt, 1, ncclUint64, c->comm, c->stream) != ncclSuccess) { cudaFree(dcnt); c->err = "ncclAllGather failed"; return TSB_ENCCL; }
    std::vector<unsigned long long> cnt(W);
    CU(cudaMemcpyAsync(cnt.data(), dcnt, 8 * (size_t)W, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    cudaFree(dcnt);
    std::vector<size_t> off(W + 1, 0);
    for (int r = 0; r < W; ++r) off[r + 1] = off[r] + cnt[r];
    uint32_t* buf = nullptr;
    CU(cudaMalloc((void**)&buf, 4 * std::max<size_t>(off[W], 1)));
    if (n) CU(cudaMemcpyAsync(buf + off[R], mine.data(), 4 * n, cudaMemcpyHostToDevice, c->stream));
    for (int r = 0; r < W; ++r)
        if (cnt[r] && ncclBroadcast(buf + off[r], buf + off[r], cnt[r], ncclUint32, r, c->comm, c->stream) != ncclSuccess) { cudaFree(buf); c->err = "ncclBroadcast failed"; return TSB_ENCCL; }
    std::vector<uint32_t> flat(off[W]);
    if (off[W]) CU(cudaMemcpyAsync(flat.data(), buf, 4 * off[W], cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    cudaFree(buf);
    all.assign(W, {});
    for (int r = 0; r < W; ++r) all[r].assign(flat.begin() + off[r], flat.begin() + off[r + 1]);
    return TSB_OK;
}
inline bool p2p(const tsb_ctx* c) { return c->world > 1 && c->d.xc.world > 1; }
// all-to-all "step done" over the peers' flag rows: publish after everything enqueued so far, wait for every peer
void x_sync(tsb_ctx* c, uint32_t point) {
    ++c->xseq;
    launch_x_signal_wait(c->d, point, c->xseq, c->stream);
}

int fetch_scalars(tsb_ctx* c) {
    CU(cudaMemcpyAsync(c->h_scal, c->d.scal, sizeof(double) * S_COUNT, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(c->h_flags, c->d.flags, sizeof(int) * F_COUNT, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->d2h += sizeof(double) * S_COUNT + sizeof(int) * F_COUNT;
    return TSB_OK;
}

// K1 + K2.  world > 1: every rank linearises the LM edges of its poses; the per-landmark partial
// (Hll, b_l) sums and chi2 are summed over ranks before the landmark blocks are inverted.
int enqueue_linearize(tsb_ctx* c) {
    static const bool timing = [] { const char* e = std::getenv("TSB_LIN_TIMING"); return e && e[0] == '1'; }();
    cudaEvent_t ev[5] = {};
    auto mark = [&](int i) { if (timing) { cudaEventCreate(&ev[i]); cudaEventRecord(ev[i], c->stream); } };
    mark(0);
    launch_linearize(c->d, c->stream);
    mark(1);
    if (c->world <= 1) launch_lm_sum(c->d, true, c->stream);          // sum + finalize in one kernel
    else if (p2p(c)) {
        // owner-computes over NVLink: every rank pushes the sums of the landmarks it touches to their owners (+ its chi2),
        // the owner adds them in rank order, inverts and pushes Hll / Hll^-1 / b_l back to every rank
        mark(2);
        launch_x_lm5_exchange(c->d, c->xseq + 1, c->stream); c->xseq += 2;
        mark(3);
    } else {
        launch_lm_sum(c->d, false, c->stream);
        mark(2);
        TRY(allreduce(c, c->d.lm_sum, 5ull * c->g.L + 1));     // landmark sums and chi2 (k_lm_finalize moves it to scal)
        mark(3);
        launch_lm_finalize(c->d, c->stream);
    }
    mark(4);
    if (timing) {
        cudaEventSynchronize(ev[4]);
        float a = 0, b = 0, x = 0, f = 0;
        cudaEventElapsedTime(&a, ev[0], ev[1]);
        if (c->world > 1) { cudaEventElapsedTime(&b, ev[1], ev[2]); cudaEventElapsedTime(&x, ev[2], ev[3]); cudaEventElapsedTime(&f, ev[3], ev[4]); }
        else cudaEventElapsedTime(&b, ev[1], ev[4]);
        std::fprintf(stderr, "[tsb] rank %d linearize: odom+edges+chi2 %.1f us, landmark sums %.1f us, all-reduce %.1f us, finalize %.1f us\n",
                     c->rank, 1e3 * a, 1e3 * b, 1e3 * x, 1e3 * f);
        for (auto& e : ev) if (e) cudaEventDestroy(e);
    }
    return TSB_OK;
}

// Schur preparation + PCG + landmark back-substitution; leaves delta in d.x / d.dl
int run_solve(tsb_ctx* c, bool refresh_coarse = true, bool smoother = true) {
    c->d.sbinv = (smoother && c->d.res_nb && c->sb_mode) ? c->sbinv_buf : nullptr;      // this solve's fine level
    const Dev& d = c->d;
    launch_schur_prep(d, c->stream);
    if (d.res_nb && d.sbinv && !launch_sb_build(d, c->stream)) { c->err = "smoother-block kernel could not be launched"; return TSB_ECUDA; }
    if (c->use_coarse && d.nc && (refresh_coarse || c->coarse_age < 0)) {
        c->coarse_age = 0; ++c->coarse_refreshes;
        // coarse operator (two-level: Ac = Z^T S Z of the aggregates; multilevel: C_l, blockdiag(S_l)^-1 and the dense top
        // operator), inverted explicitly so that the PCG applies it as a dense mat-vec; a failed inversion (non-positive
        // pivot) turns the coarse levels off for this solve on the device (F_COARSE = 0)
        cudaEvent_t te[3] = {nullptr, nullptr, nullptr};
        const bool timing = d.dbg != nullptr;
        if (timing) { for (auto& e : te) cudaEventCreate(&e); cudaEventRecord(te[0], c->stream); }
        if (d.mlLv) {
            launch_ml_levels(d, c->stream);
            if (p2p(c)) {      // every rank's top-level C into the global planes (stride gt_nnz), each rank's segment broadcast by its owner
                const MlDev& t = d.ml[d.mlLv - 1];
                for (int q = 0; q < 3; ++q)
                    if (t.nnz) CU(cudaMemcpyAsync(d.gtC + 2ull * ((size_t)q * d.gt_nnz + c->gt_eoff[c->rank]), t.C + 2ull * (size_t)q * t.nnz, 16ull * t.nnz, cudaMemcpyDeviceToDevice, c->stream));
                for (int r = 0; r < c->world; ++r) for (int q = 0; q < 3; ++q) {
                    double* seg = d.gtC + 2ull * ((size_t)q * d.gt_nnz + c->gt_eoff[r]);
                    if (c->gt_ecnt[r] && ncclBroadcast(seg, seg, 2ull * c->gt_ecnt[r], ncclDouble, r, c->comm, c->stream) != ncclSuccess) { c->err = "ncclBroadcast failed"; return TSB_ENCCL; }
                }
            }
            launch_ml_top(d, c->stream);
            if (p2p(c)) {      // block rows of the dense top operator: every rank assembled its own nodes' rows
                size_t row = 0;
                for (int r = 0; r < c->world; ++r) {
                    const size_t cnt = 3ull * c->st.top_nodes[r] * d.nc;
                    if (cnt && ncclBroadcast(d.Ac + row, d.Ac + row, cnt, ncclDouble, r, c->comm, c->stream) != ncclSuccess) { c->err = "ncclBroadcast failed"; return TSB_ENCCL; }
                    row += cnt;
                }
                // ODOM edges across the cuts: every rank's records to every rank, then added to the replicated operator
                launch_ml_cut_record(d, c->stream);
                for (int r = 0; r < c->world; ++r) {
                    double* seg = d.xc.cutrec + (size_t)r * kXMaxCut * kXCutRec;
                    if (ncclBroadcast(seg, seg, (size_t)kXMaxCut * kXCutRec, ncclDouble, r, c->comm, c->stream) != ncclSuccess) { c->err = "ncclBroadcast failed"; return TSB_ENCCL; }
                }
                launch_ml_cut_apply(d, c->stream);
            }
        } else launch_coarse_assemble(d, c->stream);
        if (timing) cudaEventRecord(te[1], c->stream);
        if (!launch_coarse_invert(d, c->stream)) { c->err = "coarse inverse kernel could not be launched (nc = " + std::to_string(d.nc) + ")"; return TSB_ECUDA; }
        if (timing) {
            cudaEventRecord(te[2], c->stream); cudaEventSynchronize(te[2]);
            float a = 0, b = 0; cudaEventElapsedTime(&a, te[0], te[1]); cudaEventElapsedTime(&b, te[1], te[2]);
            std::fprintf(stderr, "[tsb] coarse setup nc=%u (%u levels): assemble %.3f ms, k_coarse_invert %.3f ms\n", d.nc, d.mlLv, a, b);
            if (d.mlLv) { long long t[8] = {}; cudaMemcpy(t, d.dbg + 8, sizeof(t), cudaMemcpyDeviceToHost);
                std::fprintf(stderr, "[tsb] k_ml_top, cycles of a middle block (%lld coarse edges): pose part %lld, landmark part %lld (stage edges %lld, stage pairs %lld, "
                             "own walk of thread 0 %lld, wait for the other owners %lld), write-out %lld\n", t[7], t[0], t[1], t[3], t[4], t[5], t[6], t[2]); }
            for (auto& ev : te) cudaEventDestroy(ev);
        }
    }
    cudaEventRecord(c->ev[4], c->stream);
    if (d.res_nb) {            // small graph: Hpl stays in shared memory for the whole solve, three grid barriers per iteration
        if (!launch_pcg_resident(d, c->stream)) { c->err = "resident PCG kernel could not be launched"; return TSB_ECUDA; }
        if (d.dbg) {
            long long t[16]; int its = 0;
            CU(cudaMemcpyAsync(t, d.dbg, sizeof(t), cudaMemcpyDeviceToHost, c->stream));
            CU(cudaMemcpyAsync(&its, d.flags + F_ITERS, 4, cudaMemcpyDeviceToHost, c->stream));
            CU(cudaStreamSynchronize(c->stream));
            const char* nm[14] = {"phase1", "sync", "phase2(rest)", "sync", "phase3(finish)", "sync", "", "", "p2:gather+products", "p2:bar", "p2:segsum", "p2:blocksum", "p3:alpha", "p3:update"};
            std::fprintf(stderr, "[tsb] resident PCG, %d iterations, SM cycles per iteration (block 0):", its);
            for (int q = 0; q < 14; ++q) if (nm[q][0]) std::fprintf(stderr, " %s=%lld", nm[q], its ? t[q] / its : 0);
            std::fprintf(stderr, "\n");
        }
        cudaEventRecord(c->ev[5], c->stream);
        return TSB_OK;
    }
    // single GPU, HBM-sized graph: streamed persistent kernel, then the back-substitution by the phased kernels
    if (p2p(c)) c->d.xc.seq = (uint64_t)(++c->xlaunch) << 32;      // flag values of this launch: the same on every rank
    if ((c->world <= 1 || p2p(c)) && !c->force_phased && launch_pcg_stream(d, c->stream)) {
        c->pcg_streamed = true;
        if (d.dbg) {
            long long t[8]; int its = 0;
            CU(cudaMemcpyAsync(t, d.dbg, sizeof(t), cudaMemcpyDeviceToHost, c->stream));
            CU(cudaMemcpyAsync(&its, d.flags + F_ITERS, 4, cudaMemcpyDeviceToHost, c->stream));
            CU(cudaStreamSynchronize(c->stream));
            const char* nm[8] = {"pass_a", "sync", "mid", "sync", "pass_b", "sync", "update", "sync+coarse"};
            std::fprintf(stderr, "[tsb] streamed PCG, %d iterations, SM cycles per iteration (block 0):", its);
            for (int q = 0; q < 8; ++q) std::fprintf(stderr, " %s=%lld", nm[q], its ? t[q] / its : 0);
            std::fprintf(stderr, "\n");
            if (std::getenv("TSB_PCG_TIMING_BLOCKS") && its) {       // spread over the blocks: which phase waits for whom
                std::vector<long long> all(9 * 1024);
                CU(cudaMemcpy(all.data(), d.dbg + 16, 8 * all.size(), cudaMemcpyDeviceToHost));
                uint32_t nb = 0; while (nb < 1024 && all[9 * nb + 8]) ++nb;
                if (!nb) std::fprintf(stderr, "[tsb]   (per-block counters are kept by the sharded instance of the kernel only)\n");
                if (nb)
                for (int q : {0, 2, 4, 6}) {
                    std::vector<std::pair<long long, uint32_t>> v;
                    for (uint32_t b = 0; b < nb; ++b) v.emplace_back(all[9 * b + q] / its, b);
                    std::sort(v.begin(), v.end());
                    std::fprintf(stderr, "[tsb]   %-7s over %u blocks: min %lld  p10 %lld  median %lld  p90 %lld  max %lld | slowest:", nm[q], nb, v[0].first, v[nb / 10].first,
                                 v[nb / 2].first, v[nb * 9 / 10].first, v[nb - 1].first);
                    for (uint32_t k = 0; k < 4 && k < nb; ++k) { const uint32_t b = v[nb - 1 - k].second;
                        std::fprintf(stderr, " b%u/sm%lld/%lldch=%lld", b, all[9 * b + 8] & 0xffffffff, all[9 * b + 8] >> 32, v[nb - 1 - k].first); }
                    std::fprintf(stderr, "\n");
                }
            }
        }
        launch_pcg_pass_a(d, 1, c->stream);
        if (p2p(c)) {          // Hpl^T delta_p summed over ranks by the landmark's owner, delta_l pushed back to every rank
            launch_x_w_push(d, c->stream); x_sync(c, XP_BW);
            launch_x_backsub_own(d, c->stream); x_sync(c, XP_BU);
        } else {
        launch_pcg_mid(d, 1, c->stream);
        launch_backsub(d, c->stream);
        }
        cudaEventRecord(c->ev[5], c->stream);
        return TSB_OK;
    }
    c->pcg_streamed = false;
    if (p2p(c)) { c->err = "streamed PCG kernel could not be launched (peer exchange)"; return TSB_ECUDA; }
    if (d.mlLv) { c->err = "streamed PCG kernel could not be launched"; return TSB_ECUDA; }      // the hierarchy has no other consumer
    // single GPU: the whole PCG and the back-substitution run in one persistent cooperative kernel
    if (c->world <= 1 && !c->force_phased && launch_pcg_persistent(d, c->stream)) {
        if (d.dbg) {
            long long t[8]; int its = 0;
            CU(cudaMemcpyAsync(t, d.dbg, sizeof(t), cudaMemcpyDeviceToHost, c->stream));
            CU(cudaMemcpyAsync(&its, d.flags + F_ITERS, 4, cudaMemcpyDeviceToHost, c->stream));
            CU(cudaStreamSynchronize(c->stream));
            const char* nm[8] = {"pass_a", "sync", "mid", "sync", "pass_b", "sync", "update", "sync"};
            std::fprintf(stderr, "[tsb] persistent PCG, %d iterations, SM cycles per iteration (block 0):", its);
            for (int q = 0; q < 8; ++q) std::fprintf(stderr, " %s=%lld", nm[q], its ? t[q] / its : 0);
            std::fprintf(stderr, "\n");
        }
        cudaEventRecord(c->ev[5], c->stream);
        return TSB_OK;
    }
    // one launch per phase, collectives between the phases (the multi-GPU path; also TSB_PCG_PHASED=1)
    launch_pcg_init(d, c->stream);
    TRY(allreduce(c, d.scal + S_RZ, 1));
    launch_pcg_init_fix(d, c->stream);
    const int maxit = c->cfg.pcg_max_iters > 0 ? c->cfg.pcg_max_iters : 4000;
    const int batch = 16;
    const size_t nx = 2ull * c->g.L + 3ull * d.nBnd;          // [w | boundary p] travels as one all-reduce
    for (int it = 0; it < maxit + batch; it += batch) {
        for (int b = 0; b < batch; ++b) {
            launch_pcg_pass_a(d, 0, c->stream);                // p = z + beta p; slot partials of Hpl^T p
            launch_pcg_mid(d, 1, c->stream);                   // w = sum of this rank's slot partials
            launch_pack_boundary(d, d.p, c->stream);
            TRY(allreduce(c, d.w_sum, nx));
            launch_unpack_ghost(d, d.p, c->stream);
            launch_pcg_hpp(d, c->stream);                      // sp = Hpp p + ODOM part (needs the ghost p)
            const uint32_t nparts = launch_pcg_pass_b(d, c->stream);
            launch_pcg_psp(d, nparts, c->stream);
            TRY(allreduce(c, d.scal + S_PSP, 1));
            launch_pcg_alpha(d, c->stream);
            launch_pcg_update(d, c->stream);
            launch_pcg_rznew(d, c->stream);
            TRY(allreduce(c, d.scal + S_RZNEW, 1));
            launch_pcg_beta(d, c->stream);
        }
        CU(cudaMemcpyAsync(c->h_flags, d.flags, sizeof(int) * F_COUNT, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        c->d2h += sizeof(int) * F_COUNT;
        if (c->h_flags[F_DONE]) break;
    }
    launch_pcg_pass_a(d, 1, c->stream);
    launch_pcg_mid(d, 1, c->stream);
    TRY(allreduce(c, d.w_sum, 2ull * c->g.L));
    launch_backsub(d, c->stream);
    cudaEventRecord(c->ev[5], c->stream);
    return TSB_OK;
}

// K5: damped update of poses and landmarks; ghost poses follow their owners
int enqueue_update(tsb_ctx* c) {
    const Dev& d = c->d;
    if (p2p(c)) {
        launch_x_push_ghost(d, d.x, c->stream);                // delta of the boundary poses -> the ranks that hold them as ghosts
        x_sync(c, XP_GHOST);
        launch_update_state(d, 0.2, c->stream);
        launch_x_scal_push(d, 1, d.scal + S_DELTA2, c->stream);    // ||delta_pose||^2 summed over ranks in rank order
        x_sync(c, XP_MISC);
        launch_x_scal_sum(d, 1, d.scal + S_DELTA2, c->stream);
        return TSB_OK;
    }
    launch_pack_boundary(d, d.x, c->stream);
    TRY(allreduce(c, d.bnd, 3ull * d.nBnd));
    launch_update_state(d, 0.2, c->stream);                    // OptimizerCpu.h:159-165
    TRY(allreduce(c, d.scal + S_DELTA2, 1));
    return TSB_OK;
}

}  // namespace

extern "C" {

const char* tsb_last_error(const tsb_ctx* c) { return c ? c->err.c_str() : g_create_error.c_str(); }

int tsb_create(const tsb_config* cfg, tsb_ctx** out) {
    if (!cfg || !out) { g_create_error = "tsb_create: null argument"; return TSB_EINVAL; }
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        g_create_error = "no CUDA device: the PIPELINE=gpu path has no CPU fallback"; return TSB_ECUDA; }
    if (cfg->device < 0 || cfg->device >= ndev) { g_create_error = "tsb_create: bad device ordinal"; return TSB_EINVAL; }
    if (cfg->world > 1 && (!cfg->nccl_id || cfg->rank < 0 || cfg->rank >= cfg->world)) { g_create_error = "tsb_create: world > 1 needs nccl_id and 0 <= rank < world"; return TSB_EINVAL; }
    tsb_ctx* c = new tsb_ctx();
    c->cfg = *cfg; if (c->cfg.world < 1) { c->cfg.world = 1; c->cfg.rank = 0; }
    { const char* e = std::getenv("TSB_PCG_PHASED"); c->force_phased = e && e[0] == '1'; }
    { const char* e = std::getenv("TSB_COARSE"); c->use_coarse = !(e && e[0] == '0'); }
    { const char* e = std::getenv("TSB_COARSE_MAX_AGE"); if (e && std::atoi(e) > 0) c->coarse_max_age = std::atoi(e); }
    auto fail = [&](const char* what, cudaError_t e) { g_create_error = std::string(what) + ": " + cudaGetErrorString(e); delete c; return TSB_ECUDA; };
    cudaError_t e;
    if ((e = cudaSetDevice(cfg->device)) != cudaSuccess) return fail("cudaSetDevice", e);
    if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
    for (auto& ev : c->ev) if ((e = cudaEventCreate(&ev)) != cudaSuccess) return fail("cudaEventCreate", e);
    if ((e = cudaMallocHost((void**)&c->h_scal, sizeof(double) * S_COUNT)) != cudaSuccess) return fail("cudaMallocHost", e);
    if ((e = cudaMallocHost((void**)&c->h_flags, sizeof(int) * F_COUNT)) != cudaSuccess) return fail("cudaMallocHost", e);
    { const char* e = std::getenv("TSB_P2P"); c->use_p2p = !(e && e[0] == '0'); }
    if (cfg->world > (int)kXMaxWorld) { g_create_error = "tsb_create: world > 8 is not supported"; delete c; return TSB_EINVAL; }
    { const char* e = std::getenv("TSB_ML"); if (e && e[0] == '0') c->ml_top = 0; }
    { const char* e = std::getenv("TSB_ML_TOP"); if (e && std::atoi(e) >= 8) c->ml_top = std::min<uint32_t>((uint32_t)std::atoi(e), kMlTopMax); }
    if (c->cfg.world > 1) {
        ncclUniqueId id; std::memcpy(&id, cfg->nccl_id, sizeof(id));
        const ncclResult_t r = ncclCommInitRank(&c->comm, c->cfg.world, id, c->cfg.rank);
        if (r != ncclSuccess) { g_create_error = std::string("ncclCommInitRank: ") + ncclGetErrorString(r); delete c; return TSB_ENCCL; }
    }
    *out = c;
    return TSB_OK;
}

int tsb_nccl_unique_id(void* out128) {
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
    if (!out128) return TSB_EINVAL;
    ncclUniqueId id;
    if (ncclGetUniqueId(&id) != ncclSuccess) { g_create_error = "ncclGetUniqueId failed"; return TSB_ENCCL; }
    std::memcpy(out128, &id, sizeof(id));
    return TSB_OK;
}

void tsb_destroy(tsb_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->cfg.device);
    free_graph(c); release_arena(c);
    for (int r = 0; r < c->cfg.world; ++r) if (r != c->cfg.rank && c->xpeer[r] && c->xpeer_ipc[r]) cudaIpcCloseMemHandle(c->xpeer[r]);
    if (c->xreg) cudaFree(c->xreg);
    if (c->comm) ncclCommDestroy(c->comm);
    if (c->h_scal) cudaFreeHost(c->h_scal);
    if (c->h_flags) cudaFreeHost(c->h_flags);
    for (auto& ev : c->ev) if (ev) cudaEventDestroy(ev);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int tsb_load_wire(tsb_ctx* c, const void* payload, uint32_t nbytes) {
    if (!c || (!payload && nbytes)) return TSB_EINVAL;
    CU(cudaSetDevice(c->cfg.device));
    free_graph(c); c->h2d = c->d2h = 0;
    if (!parse_wire((const uint8_t*)payload, nbytes, c->g, c->err)) { c->topo_valid = false; return TSB_EPARSE; }
    if (c->g.reply_bytes > 0x7FFFFFFFu) { c->topo_valid = false; c->err = "reply would exceed the signed 32-bit frame size of the wire protocol"; return TSB_EPARSE; }
    // Structure reuse: a request whose INTEGER part (ids, types, edge lists, fixed ids) hashes like the previous one on
    // this context keeps the assembly structure and every index array already on the device; only values are refreshed.
    // (The reference rebuilds graph, index map and device buffers per request: ConnectionHandler.h:18-20, GraphGpu.h:80-187.)
    {
        // the switches that decide which structure is built are part of the key
        auto on = [](const char* n) { const char* e = std::getenv(n); return (e && e[0] == '1') ? 1u : 0u; };
        const uint32_t mode = on("TSB_FORCE_SHARD") | on("TSB_NO_RESIDENT") << 1 | on("TSB_NO_STREAM") << 2 | (c->force_phased ? 8u : 0u) | c->ml_top << 4;
        const uint32_t cnt[6] = {c->g.P, c->g.L, c->g.Eo, c->g.El, (uint32_t)c->g.fixed.size() ^ (mode << 8), nbytes};
        const char* nr = std::getenv("TSB_NO_REUSE");
        c->reuse_ints = c->topo_valid && c->topo_hash == c->g.topo_hash && std::memcmp(cnt, c->topo_counts, sizeof(cnt)) == 0 &&
                        c->arena.size() == 1 && !(nr && nr[0] == '1');
        c->topo_hash = c->g.topo_hash; std::memcpy(c->topo_counts, cnt, sizeof(cnt)); c->topo_valid = false;
        if (c->reuse_ints) {
            ++c->structure_reuses;
            c->world = c->st_world; c->rank = c->st_rank;
            const int rc = upload_graph(c);
            c->reuse_ints = false; c->topo_valid = rc == TSB_OK;
            return rc;
        }
    }
    // Sharding decision.  A graph that fits the shared-memory-resident kernel of ONE GPU is latency-bound, not
    // bandwidth-bound: splitting it over NVLink only adds collective latency to every PCG phase, so with world > 1
    // every rank solves it whole (replicated, no collective; results are bit-identical on all ranks).  Larger
    // graphs -- and TSB_FORCE_SHARD=1 -- are edge-sharded over the ranks.
    uint32_t res_nb = 0;
    { const char* e = std::getenv("TSB_NO_RESIDENT");
      if (c->use_coarse && !c->force_phased && !(e && e[0] == '1')) res_nb = (uint32_t)resident_blocks_available(); }
    const char* fs = std::getenv("TSB_FORCE_SHARD");
    c->world = c->cfg.world; c->rank = c->cfg.rank;
    if (c->cfg.world > 1 && res_nb > 0 && !(fs && fs[0] == '1')) {
        build_structure(c->g, 0, 1, c->st, res_nb, &c->rp);
        if (c->rp.nb > 0) { c->world = 1; c->rank = 0; c->st_world = 1; c->st_rank = 0; const int rc = upload_graph(c); c->topo_valid = rc == TSB_OK; return rc; }
    }
    // graphs beyond the resident kernel run the streamed PCG with the multilevel hierarchy (one GPU; the sharded path
    // and TSB_PCG_PHASED=1 / TSB_ML=0 keep the two-level aggregates)
    const char* ns = std::getenv("TSB_NO_STREAM");
    // top level: 1024 nodes, 1536 for the million-pose graphs (their PCG iteration moves 2.5 GB, the 170 MB dense inverse is
    // cheap beside it and every halving of the top spacing saves ~30 % of the iterations); TSB_ML_TOP overrides
    const uint32_t top_default = (c->ml_top == kMlTopDefault && c->g.P >= 500000u) ? kMlTopMax : c->ml_top;
    const uint32_t ml_top = ((c->world <= 1 || c->use_p2p) && c->use_coarse && !c->force_phased && !(ns && ns[0] == '1')) ? top_default : 0;
    build_structure(c->g, (uint32_t)c->rank, (uint32_t)c->world, c->st, c->world <= 1 ? res_nb : 0, &c->rp, ml_top);
    c->st_world = c->world; c->st_rank = c->rank;
    {   const bool tm = std::getenv("TSB_HOST_TIMING") != nullptr;        // parse_wire and build_structure print their own sections
        const auto t0 = std::chrono::steady_clock::now();
        const int rc = upload_graph(c); c->topo_valid = rc == TSB_OK;
        if (tm) { cudaStreamSynchronize(c->stream);
            std::fprintf(stderr, "[tsb] upload_graph (gathers + host-to-device copies, synchronised) %8.1f ms\n",
                         std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count()); }
        return rc; }
}

int tsb_load_soa(tsb_ctx* c, const tsb_graph_soa* s) {
    if (!c || !s) return TSB_EINVAL;
    // The request bytes are needed in any case (the reply echoes the edges from them, SerializeGraph.h:30-63), so the SoA
    // arrays are laid out as exactly what the reference client would send (graph_to_bytes.py:32-67) -- fixed-size records
    // at known offsets, written by a parallel loop into a buffer the context keeps -- and take the wire path, whose decode
    // is parallel as well
    const uint64_t oV = 4, oL = oV + 20ull * s->n_pose, oE = oL + 16ull * s->n_lm, oO = oE + 4, oM = oO + 76ull * s->n_odom,
                   oF = oM + 44ull * s->n_lmedge, total = oF + 4 + 4ull * s->n_fixed;
    if (total > 0xFFFFFFFFull) { c->err = "graph too large for the 32-bit wire size prefix"; return TSB_EINVAL; }
    if ((s->n_pose && (!s->pose_id || !s->pose_xyt)) || (s->n_lm && (!s->lm_id || !s->lm_xy)) || (s->n_odom && (!s->odom_ij || !s->odom_Z || !s->odom_w)) ||
        (s->n_lmedge && (!s->lmedge_ij || !s->lmedge_meas || !s->lmedge_w)) || (s->n_fixed && !s->fixed_id)) { c->err = "tsb_load_soa: null array"; return TSB_EINVAL; }
    pvec<uint8_t>& buf = c->soa_wire; buf.resize(total);
    uint8_t* o = buf.data();
    auto pu = [](uint8_t* p, uint32_t v) { std::memcpy(p, &v, 4); };
    pu(o, s->n_pose + s->n_lm); pu(o + oE, s->n_odom + s->n_lmedge); pu(o + oF, s->n_fixed);
#pragma omp parallel
    {
#pragma omp for schedule(static) nowait
        for (long long i = 0; i < (long long)s->n_pose; ++i) { uint8_t* p = o + oV + 20ull * i; pu(p, s->pose_id[i]); pu(p + 4, 0); std::memcpy(p + 8, s->pose_xyt + 3ull * i, 12); }
#pragma omp for schedule(static) nowait
        for (long long i = 0; i < (long long)s->n_lm; ++i) { uint8_t* p = o + oL + 16ull * i; pu(p, s->lm_id[i]); pu(p + 4, 1); std::memcpy(p + 8, s->lm_xy + 2ull * i, 8); }
#pragma omp for schedule(static) nowait
        for (long long i = 0; i < (long long)s->n_odom; ++i) {
            uint8_t* p = o + oO + 76ull * i;
            pu(p, 0); std::memcpy(p + 4, s->odom_ij + 2ull * i, 8); pu(p + 12, 3); pu(p + 16, 3); std::memcpy(p + 20, s->odom_Z + 9ull * i, 36);
            pu(p + 56, 0); pu(p + 60, 3); std::memcpy(p + 64, s->odom_w + 3ull * i, 12);
        }
#pragma omp for schedule(static) nowait
        for (long long i = 0; i < (long long)s->n_lmedge; ++i) {
            uint8_t* p = o + oM + 44ull * i;
            pu(p, 1); std::memcpy(p + 4, s->lmedge_ij + 2ull * i, 8); pu(p + 12, 0); pu(p + 16, 2); std::memcpy(p + 20, s->lmedge_meas + 2ull * i, 8);
            pu(p + 28, 0); pu(p + 32, 2); std::memcpy(p + 36, s->lmedge_w + 2ull * i, 8);
        }
#pragma omp for schedule(static)
        for (long long i = 0; i < (long long)s->n_fixed; ++i) pu(o + oF + 4 + 4ull * i, s->fixed_id[i]);
    }
    return tsb_load_wire(c, buf.data(), (uint32_t)buf.size());
}

// Host-only: parse a payload and build rank `rank`'s assembly structure without touching a GPU.
// Lets the partition logic (pose cuts, edge ownership, ghosts, boundary list) be tested on CPU.
int tsb_host_plan(const void* payload, uint32_t nbytes, int32_t rank, int32_t world, tsb_counts* o,
                  uint32_t* n_ghost, uint32_t* n_boundary, uint64_t* structure_hash) {
    if (!payload || !o || world < 1 || rank < 0 || rank >= world) return TSB_EINVAL;
    HostGraph g; Structure st; std::string err;
    if (!parse_wire((const uint8_t*)payload, nbytes, g, err)) { g_create_error = err; return TSB_EPARSE; }
    build_structure(g, (uint32_t)rank, (uint32_t)world, st, 0, nullptr, world <= 1 ? kMlTopDefault : 0);
    std::memset(o, 0, sizeof(*o));
    o->n_pose = g.P; o->n_lm = g.L; o->n_odom = g.Eo; o->n_lmedge = g.El; o->n_fixed = (uint32_t)g.fixed.size();
    o->dim = 3 * g.P + 2 * g.L; o->pose_begin = st.pose_begin; o->pose_end = st.pose_end;
    o->n_lmedge_local = (uint32_t)st.lm_perm.size(); o->n_odom_local = st.n_od_owned;
    o->n_chunks = (uint32_t)st.chunks.size(); o->n_blocks = (uint32_t)st.blk_chunk_ptr.size() - 1; o->n_slots = (uint32_t)st.slot_lm.size();
    o->n_aggregates = st.nA;
    o->n_ml_levels = (uint32_t)st.ml.size(); o->n_ml_top = st.ml.empty() ? 0 : st.ml.back().n;
    if (n_ghost) *n_ghost = st.nG;
    if (n_boundary) *n_boundary = (uint32_t)st.boundary.size();
    if (structure_hash) {            // FNV-1a over the rank-independent parts (cuts, boundary list): must agree on all ranks
        uint64_t h = 1469598103934665603ull;
        auto mix = [&](const std::vector<uint32_t>& v) { for (uint32_t x : v) { h ^= x; h *= 1099511628211ull; } };
        mix(st.cuts); mix(st.boundary);
        *structure_hash = h;
    }
    return TSB_OK;
}

// Host-only: the integer structure arrays of rank `rank` (sizes from tsb_host_plan: n_lmedge_local, pose range,
// n_chunks, n_slots).  lm_perm holds REQUEST-order LM edge indices in this rank's assembly order.
int tsb_host_structure(const void* payload, uint32_t nbytes, int32_t rank, int32_t world, uint32_t* lm_perm, uint32_t* pose_ptr,
                       uint32_t* chunk_desc, uint32_t* slot_lm, uint32_t* lm_packed) {
    if (!payload || world < 1 || rank < 0 || rank >= world) return TSB_EINVAL;
    HostGraph g; Structure st; std::string err;
    if (!parse_wire((const uint8_t*)payload, nbytes, g, err)) { g_create_error = err; return TSB_EPARSE; }
    build_structure(g, (uint32_t)rank, (uint32_t)world, st, 0, nullptr, 0);
    if (lm_perm) for (size_t k = 0; k < st.lm_perm.size(); ++k) lm_perm[k] = st.le_local[st.lm_perm[k]];
    if (pose_ptr) std::memcpy(pose_ptr, st.pose_ptr.data(), 4 * st.pose_ptr.size());
    if (chunk_desc && !st.chunks.empty()) std::memcpy(chunk_desc, st.chunks.data(), 32 * st.chunks.size());
    if (slot_lm && !st.slot_lm.empty()) std::memcpy(slot_lm, st.slot_lm.data(), 4 * st.slot_lm.size());
    if (lm_packed && !st.lm_packed.empty()) std::memcpy(lm_packed, st.lm_packed.data(), 4 * st.lm_packed.size());
    return TSB_OK;
}

int tsb_get_counts(const tsb_ctx* c, tsb_counts* o) {
    if (!c || !o || !c->loaded) return TSB_EINVAL;
    o->n_pose = c->g.P; o->n_lm = c->g.L; o->n_odom = c->g.Eo; o->n_lmedge = c->g.El; o->n_fixed = (uint32_t)c->g.fixed.size();
    o->dim = 3 * c->g.P + 2 * c->g.L; o->pose_begin = c->st.pose_begin; o->pose_end = c->st.pose_end;
    o->n_lmedge_local = c->d.nE; o->n_odom_local = c->d.nOwned;
    o->n_chunks = c->d.nChunks; o->n_blocks = c->d.nBlocks; o->n_slots = c->d.nSlots;
    o->pcg_path = c->d.res_nb ? 1u : (c->pcg_streamed || c->d.mlLv || c->d.xc.world > 1) ? 3u : (c->world <= 1 && !c->force_phased ? 0u : 2u);
    o->replicated = (c->cfg.world > 1 && c->world == 1) ? 1u : 0u; o->n_aggregates = c->d.nA;
    o->n_ml_levels = c->d.mlLv; o->n_ml_top = c->d.mlLv ? c->d.ml[c->d.mlLv - 1].n : 0;
    o->structure_reuses = (uint32_t)c->structure_reuses;
    return TSB_OK;
}

int tsb_reset_state(tsb_ctx* c) {
    if (!c || !c->loaded) return TSB_EINVAL;
    CU(cudaSetDevice(c->cfg.device));
    CU(cudaMemcpyAsync(c->d.pose, c->pose0, sizeof(double4) * (c->d.nP + c->d.nG), cudaMemcpyDeviceToDevice, c->stream));
    CU(cudaMemcpyAsync(c->d.lm, c->lm0, sizeof(double2) * c->g.L, cudaMemcpyDeviceToDevice, c->stream));
    c->linearized = c->solved = false;
    return TSB_OK;
}

int tsb_set_state(tsb_ctx* c, const double* pose_xyt, const double* lm_xy) {
    if (!c || !c->loaded || !pose_xyt || (!lm_xy && c->g.L)) return TSB_EINVAL;
    CU(cudaSetDevice(c->cfg.device));
    const Structure& s = c->st; const uint32_t nPG = c->d.nP + c->d.nG;
    std::vector<double> xyt(3ull * nPG);
    if (c->d.nP) std::memcpy(xyt.data(), pose_xyt + 3ull * s.pose_begin, 24ull * c->d.nP);
    for (uint32_t q = 0; q < c->d.nG; ++q) std::memcpy(&xyt[3ull * (c->d.nP + q)], pose_xyt + 3ull * s.ghost_pose[q], 24);
    if (nPG) CU(cudaMemcpyAsync(c->tmp_xyt, xyt.data(), 24ull * nPG, cudaMemcpyHostToDevice, c->stream));
    launch_init_pose(c->tmp_xyt, nPG, c->d.pose, c->stream);
    if (c->g.L) CU(cudaMemcpyAsync(c->d.lm, lm_xy, 16ull * c->g.L, cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->linearized = c->solved = false;
    return TSB_OK;
}

int tsb_linearize(tsb_ctx* c, tsb_iter_stats* st) {
    if (!c || !c->loaded) return TSB_EINVAL;
    CU(cudaSetDevice(c->cfg.device));
    CU(cudaEventRecord(c->ev[0], c->stream));
    TRY(enqueue_linearize(c));
    CU(cudaEventRecord(c->ev[1], c->stream));
    TRY(fetch_scalars(c));
    CU(cudaGetLastError());
    c->linearized = true; c->solved = false;
    if (st) { std::memset(st, 0, sizeof(*st)); st->chi2 = c->h_scal[S_CHI2]; CU(cudaEventElapsedTime(&st->ms_linearize, c->ev[0], c->ev[1])); }
    return TSB_OK;
}

int tsb_solve(tsb_ctx* c, tsb_iter_stats* st) {
    if (!c || !c->loaded || !c->linearized) { if (c) c->err = "tsb_solve needs a preceding tsb_linearize"; return TSB_EINVAL; }
    CU(cudaSetDevice(c->cfg.device));
    CU(cudaEventRecord(c->ev[1], c->stream));
    TRY(run_solve(c));
    CU(cudaEventRecord(c->ev[2], c->stream));
    TRY(fetch_scalars(c));
    CU(cudaGetLastError());
    c->solved = true;
    if (st) { st->pcg_iters = c->h_flags[F_ITERS];
        st->pcg_rel_residual = c->h_scal[S_RZ0] > 0 ? std::sqrt(c->h_scal[S_RZ] / c->h_scal[S_RZ0]) : 0;
        CU(cudaEventElapsedTime(&st->ms_solve, c->ev[1], c->ev[2])); }
    return TSB_OK;
}

int tsb_optimize(tsb_ctx* c, uint32_t iterations, tsb_report* rep, tsb_iter_stats* per_iter) {
    if (!c || !c->loaded) return TSB_EINVAL;
    CU(cudaSetDevice(c->cfg.device));
    const uint64_t launches0 = g_launches;
    const bool exits = c->cfg.enable_exits != 0;
    double prevErr = -1, err = 0; int penalty = 0, code = 0; uint32_t done = 0;
    // every request starts from a fresh coarse operator and refreshes it on its first solves: far from the optimum the
    // Huber weights move so much per step that a one-step-old operator costs ~1.7x the iterations (C2: 240 -> 415)
    c->coarse_age = -1; c->coarse_hold = 4; c->coarse_backoff = 1;
    int64_t pcg_total = 0; float ms_lin = 0, ms_sol = 0, ms_pcg = 0; const int refreshes0 = c->coarse_refreshes;
    bool sb_on = true; int sb_last_iters = 0;
    int not_conv = 0, coarse_fail = 0, singular = 0;
    const int maxit = c->cfg.pcg_max_iters > 0 ? c->cfg.pcg_max_iters : 4000;
    CU(cudaMemsetAsync(c->d.flags + F_SING, 0, sizeof(int), c->stream));
    CU(cudaEventRecord(c->ev[3], c->stream));
    for (uint32_t it = 0; it < iterations; ++it) {
        CU(cudaEventRecord(c->ev[0], c->stream));
        TRY(enqueue_linearize(c));
        CU(cudaEventRecord(c->ev[1], c->stream));
        if (exits) {                                     // OptimizerCpu.h:140-153 needs err before the solve
            TRY(fetch_scalars(c)); err = c->h_scal[S_CHI2];
            if (prevErr > 0 && err > prevErr) {
                if (++penalty > 2) {
                    code = 1;
                    if (per_iter) { std::memset(&per_iter[it], 0, sizeof(tsb_iter_stats)); per_iter[it].chi2 = err; }
                    ++done;
                    if (c->cfg.verbose) std::printf("Error is getting worse\n");
                    break;
                }
            } else penalty = 0;
        }
        // Refresh policy (iteration counts only: deterministic and identical on every rank).  A stale operator is
        // probed after each refresh; if the very first stale solve already costs > 15 % more iterations the system
        // is still moving and the next `backoff` solves refresh unconditionally (backoff doubles up to 16, halves
        // when a stale operator lasted); otherwise it is kept until it degrades or reaches coarse_max_age.
        bool refresh = c->coarse_age < 0 || c->coarse_age >= c->coarse_max_age;
        if (!refresh && c->coarse_hold > 0) { refresh = true; --c->coarse_hold; }
        if (!refresh && c->coarse_age >= 2 && 20 * c->h_flags[F_ITERS] > 23 * c->coarse_ref_iters + 100) {
            refresh = true;
            if (c->coarse_age == 2) { c->coarse_hold = c->coarse_backoff; c->coarse_backoff = std::min(2 * c->coarse_backoff, 16); }
            else c->coarse_backoff = std::max(1, c->coarse_backoff / 2);
        }
        // Fine level of the resident kernel, adaptive mode: a request starts with the dense smoother blocks; every sixth
        // solve is a probe with the 3x3 pose blocks (and a fresh coarse operator), and once the probe is cheaper -- an
        // iteration with smoother blocks costs ~1.27x one without and k_sb_build ~17.5 iterations (C2, measured) -- the
        // request stays with the 3x3 blocks.
        // Iteration counts only: deterministic.
        bool use_sb = c->sb_mode == 1 || (c->sb_mode == 2 && sb_on);
        const bool sb_probe = c->sb_mode == 2 && sb_on && c->d.res_nb && it > 0 && it % 6 == 0;
        if (sb_probe) { use_sb = false; refresh = true; }
        TRY(run_solve(c, refresh, use_sb));
        CU(cudaEventRecord(c->ev[2], c->stream));
        TRY(enqueue_update(c));
        TRY(fetch_scalars(c));
        CU(cudaGetLastError());
        err = c->h_scal[S_CHI2];
        // the solve is only as good as its PCG: count the ones that stopped at the cap or lost the residual
        const bool bad_pcg = c->h_flags[F_ITERS] >= maxit || !std::isfinite(c->h_scal[S_RZ]);
        if (bad_pcg) { ++not_conv; if (c->cfg.verbose) std::printf("warning: PCG stopped after %d iterations, residual %g\n", c->h_flags[F_ITERS], c->h_scal[S_RZ]); }
        if (c->use_coarse && c->d.nc && !c->h_flags[F_COARSE]) ++coarse_fail;
        if (c->h_flags[F_SING]) { ++singular; CU(cudaMemsetAsync(c->d.flags + F_SING, 0, sizeof(int), c->stream)); }
        const double dnorm = std::sqrt(c->h_scal[S_DELTA2] + c->h_scal[S_DELTA2L]);
        float a = 0, b = 0;
        CU(cudaEventElapsedTime(&a, c->ev[0], c->ev[1])); CU(cudaEventElapsedTime(&b, c->ev[1], c->ev[2]));
        { float p_ = 0; CU(cudaEventElapsedTime(&p_, c->ev[4], c->ev[5])); ms_pcg += p_; }
        ms_lin += a; ms_sol += b; pcg_total += c->h_flags[F_ITERS]; ++done;
        if (sb_probe && 100 * c->h_flags[F_ITERS] < 127 * sb_last_iters + 1750) sb_on = false;
        if (use_sb) sb_last_iters = c->h_flags[F_ITERS];
        if (c->coarse_age == 0) c->coarse_ref_iters = c->h_flags[F_ITERS];
        ++c->coarse_age;
        if (per_iter) { tsb_iter_stats& s = per_iter[it]; s.chi2 = err; s.delta_norm = dnorm; s.pcg_iters = c->h_flags[F_ITERS];
            s.pcg_rel_residual = c->h_scal[S_RZ0] > 0 ? std::sqrt(c->h_scal[S_RZ] / c->h_scal[S_RZ0]) : 0; s.ms_linearize = a; s.ms_solve = b; }
        if (exits && std::fabs(err - prevErr) < 0.001) { code = 2; if (c->cfg.verbose) std::printf("Plateau: NO MORE OPT\n"); break; }
        if (exits && dnorm < 0.001) { code = 3; if (c->cfg.verbose) std::printf("CONVERGED\n"); break; }
        prevErr = err;
    }
    CU(cudaEventRecord(c->ev[2], c->stream));
    CU(cudaEventSynchronize(c->ev[2]));
    if (c->cfg.verbose) std::printf("Summary() error = %g\n", err);
    c->linearized = c->solved = false;
    if (rep) { rep->iterations_run = (int32_t)done; rep->exit_code = code; rep->final_chi2 = err; rep->pcg_iters_total = pcg_total;
        rep->kernel_launches = (int64_t)(g_launches - launches0); rep->ms_linearize_total = ms_lin; rep->ms_solve_total = ms_sol;
        rep->ms_pcg_total = ms_pcg; rep->coarse_refreshes = c->coarse_refreshes - refreshes0;
        rep->pcg_not_converged = not_conv; rep->coarse_failures = coarse_fail; rep->singular_blocks = singular;
        CU(cudaEventElapsedTime(&rep->ms_total, c->ev[3], c->ev[2])); }
    return TSB_OK;
}

}  // extern "C"

namespace {
template <typename T> int download(tsb_ctx* c, std::vector<T>& out, const void* dev, size_t n) {
    out.resize(n);
    if (n) CU(cudaMemcpyAsync(out.data(), dev, n * sizeof(T), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return TSB_OK;
}
}  // namespace

extern "C" {

int tsb_read_ids(const tsb_ctx* c, uint32_t* pose_id, uint32_t* lm_id) {
    if (!c || !c->loaded) return TSB_EINVAL;
    if (pose_id) std::memcpy(pose_id, c->g.pose_id.data(), 4ull * c->g.P);
    if (lm_id) std::memcpy(lm_id, c->g.lm_id.data(), 4ull * c->g.L);
    return TSB_OK;
}

int tsb_read_state(tsb_ctx* c, double* pose_xyt, double* lm_xy) {
    if (!c || !c->loaded) return TSB_EINVAL;
    CU(cudaSetDevice(c->cfg.device));
    if (pose_xyt && c->world > 1) {
        // every rank returns the full state: local ranges summed into a zeroed 3P buffer (sum with zeros = gather)
        CU(cudaMemsetAsync(c->tmp_xyt, 0, 24ull * c->g.P, c->stream));
        launch_read_pose(c->d.pose, c->d.nP, c->tmp_xyt + 3ull * c->st.pose_begin, c->stream);
        TRY(allreduce(c, c->tmp_xyt, 3ull * c->g.P));
        if (c->g.P) CU(cudaMemcpyAsync(pose_xyt, c->tmp_xyt, 24ull * c->g.P, cudaMemcpyDeviceToHost, c->stream));
    } else if (pose_xyt) {
        launch_read_pose(c->d.pose, c->d.nP, c->tmp_xyt, c->stream);
        if (c->d.nP) CU(cudaMemcpyAsync(pose_xyt + 3ull * c->st.pose_begin, c->tmp_xyt, 24ull * c->d.nP, cudaMemcpyDeviceToHost, c->stream));
    }
    if (lm_xy && c->g.L) CU(cudaMemcpyAsync(lm_xy, c->d.lm, 16ull * c->g.L, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->d2h += (pose_xyt ? 24ull * c->d.nP : 0) + (lm_xy ? 16ull * c->g.L : 0);
    return TSB_OK;
}

int tsb_read_delta(tsb_ctx* c, double* delta_pose, double* delta_lm) {
    if (!c || !c->loaded || !c->solved) { if (c) c->err = "tsb_read_delta needs a preceding tsb_solve"; return TSB_EINVAL; }
    CU(cudaSetDevice(c->cfg.device));
    if (delta_pose && c->d.nP) CU(cudaMemcpyAsync(delta_pose, c->d.x, 24ull * c->d.nP, cudaMemcpyDeviceToHost, c->stream));
    if (delta_lm && c->g.L) CU(cudaMemcpyAsync(delta_lm, c->d.dl, 16ull * c->g.L, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return TSB_OK;
}

int tsb_read_system(tsb_ctx* c, double* hpp_diag, double* b_pose, double* hpl, double* hpp_off, double* hll, double* b_lm) {
    if (!c || !c->loaded || !c->linearized) { if (c) c->err = "tsb_read_system needs a preceding tsb_linearize"; return TSB_EINVAL; }
    CU(cudaSetDevice(c->cfg.device));
    const Dev& d = c->d; const Structure& s = c->st;
    if (hpp_diag && d.nP) CU(cudaMemcpyAsync(hpp_diag + 6ull * s.pose_begin, d.hpp, 48ull * d.nP, cudaMemcpyDeviceToHost, c->stream));
    if (b_pose && d.nP) CU(cudaMemcpyAsync(b_pose + 3ull * s.pose_begin, d.bp, 24ull * d.nP, cudaMemcpyDeviceToHost, c->stream));
    if (hll && d.L) CU(cudaMemcpyAsync(hll, d.hll, 24ull * d.L, cudaMemcpyDeviceToHost, c->stream));
    if (b_lm && d.L) CU(cudaMemcpyAsync(b_lm, d.bl, 16ull * d.L, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (hpl) { std::vector<double> t; TRY(download(c, t, d.hpl, 6ull * d.nE));
        for (uint32_t k = 0; k < d.nE; ++k) for (int r = 0; r < 3; ++r)      // device: three planes of double2; out: 3x2 per edge
            std::memcpy(hpl + 6ull * s.le_local[s.lm_perm[k]] + 2 * r, &t[TSB_HPL_PLANES ? 2ull * ((uint64_t)r * d.nE + k) : 6ull * k + 2 * r], 16); }
    if (hpp_off) { std::vector<double> t; TRY(download(c, t, d.hoo, 3ull * d.nOwned));
        for (uint32_t k = 0; k < d.nOwned; ++k) std::memcpy(hpp_off + 3ull * s.od_local[k], &t[3ull * k], 24); }
    return TSB_OK;
}

int tsb_read_structure(tsb_ctx* c, uint32_t* lm_perm, uint32_t* lm_packed, uint32_t* chunk_desc, uint32_t* slot_lm,
                       uint32_t* pose_ptr, uint32_t* lm_part_ptr, uint32_t* lm_part_idx) {
    if (!c || !c->loaded) return TSB_EINVAL;
    CU(cudaSetDevice(c->cfg.device));
    const Dev& d = c->d; const Structure& s = c->st;
    // lm_perm is host-side knowledge (request index of each assembly position); the rest comes back from the device
    if (lm_perm) for (uint32_t k = 0; k < d.nE; ++k) lm_perm[k] = s.le_local[s.lm_perm[k]];
    auto get = [&](uint32_t* out, const void* dev, size_t n) -> int {
        if (out && n) CU(cudaMemcpyAsync(out, dev, 4 * n, cudaMemcpyDeviceToHost, c->stream)); return TSB_OK; };
    TRY(get(lm_packed, d.lm_packed, d.nE)); TRY(get(chunk_desc, d.chunks, 8ull * d.nChunks)); TRY(get(slot_lm, d.slot_lm, d.nSlots));
    TRY(get(pose_ptr, d.pose_ptr, d.nP + 1ull)); TRY(get(lm_part_ptr, d.lm_part_ptr, d.L + 1ull)); TRY(get(lm_part_idx, d.lm_part_idx, d.nSlots));
    CU(cudaStreamSynchronize(c->stream));
    return TSB_OK;
}

int tsb_get_transfer_bytes(const tsb_ctx* c, uint64_t* h2d, uint64_t* d2h) {
    if (!c) return TSB_EINVAL;
    if (h2d) *h2d = c->h2d;
    if (d2h) *d2h = c->d2h;
    return TSB_OK;
}

int tsb_reply_size(tsb_ctx* c, uint32_t* nbytes) {
    if (!c || !c->loaded || !nbytes) return TSB_EINVAL;
    *nbytes = c->g.reply_bytes; return TSB_OK;
}

int tsb_write_reply(tsb_ctx* c, void* out, uint32_t capacity) {
    if (!c || !c->loaded || !out) return TSB_EINVAL;
    if (capacity < c->g.reply_bytes) { c->err = "tsb_write_reply: buffer too small"; return TSB_EINVAL; }
    std::vector<double> pose(3ull * c->g.P), lm(2ull * c->g.L);
    TRY(tsb_read_state(c, pose.data(), lm.data()));
    encode_reply(c->g, pose.data(), lm.data(), (uint8_t*)out);
    return TSB_OK;
}

}  // extern "C"
