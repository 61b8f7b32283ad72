// host_graph.h -- host-side graph in structure-of-arrays form + the assembly structure.
// Replaces the per-object containers of remote/graph/GraphCpu.h and remote/cuda/graph/GraphGpu.h
// (unique_ptr per vertex/edge) with flat arrays parsed straight from the wire payload.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include <memory>
#include <utility>

namespace tsb {

// vector whose resize() leaves new elements uninitialised: the big per-edge arrays are filled by a parallel loop, which
// then also takes the first-touch page faults in parallel (zero-filling ~2 GB from one thread cost more than the parse)
template <class T> struct NoInitAlloc : std::allocator<T> {
    template <class U> struct rebind { using other = NoInitAlloc<U>; };
    NoInitAlloc() = default;
    template <class U> NoInitAlloc(const NoInitAlloc<U>&) {}
    template <class U, class... A> void construct(U* p, A&&... a) {
        if constexpr (sizeof...(A) == 0) ::new ((void*)p) U; else ::new ((void*)p) U(std::forward<A>(a)...);
    }
};
template <class T> using pvec = std::vector<T, NoInitAlloc<T>>;

struct HostGraph {
    uint32_t P = 0, L = 0, Eo = 0, El = 0;
    // canonical vertex order: poses by ascending id, then landmarks by ascending id
    std::vector<uint32_t> pose_id, lm_id;
    std::vector<double> pose_xyt;     // 3 per pose (x, y, theta) widened from fp32
    std::vector<double> lm_xy;        // 2 per landmark
    // edges in request order, split by type
    pvec<uint32_t> od_i, od_j;        // canonical pose indices
    pvec<double> od_zinv;             // 6 per edge: top two rows of inverse(Z) (EdgeSe2.h:32), fp64
    pvec<double> od_w;                // 3 per edge: information diagonal
    pvec<uint32_t> le_i, le_l;        // pose index, landmark index
    pvec<double> le_z;                // 2 per edge: (r cos b, r sin b)  (EdgeSe2Point2d.h:34-35)
    pvec<double> le_w;                // 2 per edge
    std::vector<uint32_t> fixed;      // canonical vertex index (pose k -> k, landmark k -> P + k), repeats kept
    // reply echo
    pvec<uint8_t> request;            // copy of the payload
    std::vector<uint32_t> req_vertex; // per request vertex: canonical index | type << 31; 0xFFFFFFFF = dropped duplicate
    uint32_t edges_offset = 0;        // byte offset of the edge count in `request`
    uint32_t reply_bytes = 0;
    // per request edge: byte offset of its record and its index among the edges of its type (ODOM / LM): what lets the
    // value decode and the reply encode run in parallel over edges (records have no fixed stride in general)
    pvec<uint32_t> edge_off, edge_slot;
    uint32_t fixed_offset = 0;        // byte offset of the fixed-id count
    // 64-bit hash of the INTEGER part of the payload (ids, types, shapes, fixed ids -- everything but the float values):
    // two requests with the same hash and counts have the same assembly structure (tsb_load_wire reuses it)
    uint64_t topo_hash = 0;
    // empties the graph but keeps every buffer: the next request of a persistent context does not page-fault them in again
    void reset() {
        P = L = Eo = El = 0; pose_id.clear(); lm_id.clear(); pose_xyt.clear(); lm_xy.clear(); od_i.clear(); od_j.clear(); od_zinv.clear();
        od_w.clear(); le_i.clear(); le_l.clear(); le_z.clear(); le_w.clear(); fixed.clear(); request.clear(); req_vertex.clear();
        edges_offset = reply_bytes = fixed_offset = 0; edge_off.clear(); edge_slot.clear(); topo_hash = 0;
    }
};

// DeserializeGraph.h:9-174, bounds-checked.  Returns false and sets `err` on malformed input.
bool parse_wire(const uint8_t* buf, uint32_t nbytes, HostGraph& g, std::string& err);
// SerializeGraph.h:8-72 + SerializeGraphFuncCpu.h:10-65 (payload, no size prefix).
// pose_xyt / lm_xy: current fp64 state in canonical order.
void encode_reply(const HostGraph& g, const double* pose_xyt, const double* lm_xy, uint8_t* out);

struct ChunkDesc { uint32_t e0, e1, p0, p1, s0, s1, block, flags; };

// ---- multilevel coarse hierarchy of the streamed PCG (HBM-sized graphs) -----------------------------------------
// Level l >= 1 has one node per `spacing` consecutive local poses (node K sits at local pose K * spacing); a vector on
// level l is interpolated to the poses LINEARLY along the trajectory (hat functions) as rigid-body motions about the
// node's pose.  The landmarks are kept on every level, so S_l = App_l - C_l Hll^-1 C_l^T is the Galerkin operator
// Z^T S Z and C_l (one 3x2 block per distinct (node, landmark) pair: a "coarse edge") is a weighted sum of the blocks
// of the level below.  All integer structure; built once per request.
constexpr uint32_t kMlMaxLevels = 10;
constexpr uint32_t kMlWindow = 448;     // poses whose residual one block restricts at a time (7 x 64) + 64 halo poses = 512 threads
constexpr uint32_t kMlHalo = 64;
struct MlLevel {
    uint32_t n = 0, ratio = 4, spacing = 4;      // nodes, nodes of the level below per node interval, poses per node interval
    std::vector<uint32_t> node_ptr;              // n + 1 offsets into edge_lm: coarse edges of a node, landmark ascending
    std::vector<uint32_t> edge_lm;               // landmark of each coarse edge
    pvec<uint32_t> ref_ptr;                      // nnz + 1 offsets into ref
    pvec<uint32_t> ref;                          // child edge index (level below; level 1: LM edge in assembly order) << 4 | (child - ratio * K + 7)
};

// Assembly structure of one rank (all integer; compared bit-exactly in tests).
struct Structure {
    uint32_t pose_begin = 0, pose_end = 0;   // canonical pose range owned by this rank
    uint32_t nP = 0, nG = 0;                 // local poses, ghost poses (ODOM neighbours owned elsewhere)
    std::vector<uint32_t> ghost_pose;        // canonical index of each ghost pose
    // world > 1: poses that are a ghost on some rank ("boundary poses"), identical list on every rank.
    // Their vector entries travel in the tail of the per-iteration all-reduce (sum with zeros = gather).
    std::vector<uint32_t> cuts;              // world + 1 pose cut points
    std::vector<uint32_t> boundary;          // canonical pose index, ascending
    std::vector<uint32_t> own_bnd_slot, own_bnd_pose;   // slots of `boundary` owned here and their local pose index
    std::vector<uint32_t> ghost_bnd_slot;    // per ghost pose: its slot in `boundary`
    // peer exchange (world > 1): poses of this rank that are ghosts elsewhere, and which ranks touch each landmark
    std::vector<uint32_t> x_exp;             // 3 per entry: local pose, destination rank, ghost slot on that rank
    std::vector<uint8_t> lm_mask;            // L: bit r set = rank r owns a pose with an LM edge to the landmark
    std::vector<uint32_t> top_nodes;         // world > 1: nodes of every rank's top level (global top = concatenation)
    // LM edges of local poses, assembly order = sorted by (pose, landmark, request index)
    std::vector<uint32_t> le_local;          // request-order indices (into HostGraph::le_*) of the local LM edges
    std::vector<uint32_t> lm_perm;           // assembly k -> index into le_local
    std::vector<uint32_t> lm_packed;         // (pose - chunk.p0) << 16 | slot - chunk.s0
    std::vector<uint32_t> pose_ptr;          // nP + 1
    std::vector<ChunkDesc> chunks;
    std::vector<uint32_t> blk_chunk_ptr;     // n_blocks + 1
    std::vector<uint32_t> slot_lm;           // landmark index per slot (one partial sum per slot)
    std::vector<uint32_t> slot_ptr;          // n_slots + 1, offsets into slot_edge (== assembly positions)
    std::vector<uint16_t> slot_edge;         // per assembly position: edge index within its chunk, grouped by slot
    std::vector<uint32_t> lm_part_ptr, lm_part_idx;   // landmark-major CSR over slots
    // ODOM edges touching a local pose.  Owned (id1 local) edges first, sorted by id1; then ghost
    // edges (id2 local, id1 elsewhere).  Pose indices are local (0..nP) or ghost (nP..nP+nG).
    std::vector<uint32_t> od_local;          // request-order index (into HostGraph::od_*)
    uint32_t n_od_owned = 0;
    std::vector<uint32_t> od_a, od_b;        // local/ghost pose index of id1, id2
    std::vector<uint32_t> od_out_ptr;        // nP + 1 : owned edges by id1 (contiguous)
    std::vector<uint32_t> od_in_ptr, od_in_edge;   // nP + 1 : edges (owned or ghost) by id2
    std::vector<uint32_t> fixed_pose_count;  // nP : how many times the pose is listed as fixed
    std::vector<uint32_t> fixed_lm_count;    // L
    // ---- coarse space of the two-level PCG preconditioner (this rank's poses only) -------------------
    // Aggregate a = local poses [a*agg_K, (a+1)*agg_K); agg_K is a multiple of 32 (one warp-wide restriction
    // segment).  An "aslot" is a distinct (aggregate, landmark) pair; C_aslot = sum of Hpl_e^T Z_pose(e).
    uint32_t agg_K = 32, nA = 0;             // agg_K == 0: aggregates of varying size (resident plan), see agg_ptr
    std::vector<uint32_t> agg_ptr;           // nA + 1 pose offsets; pose_agg: nP, aggregate of each pose
    std::vector<uint32_t> pose_agg;
    std::vector<uint32_t> edge_pose;         // nE : local pose of each LM edge in assembly order
    std::vector<uint32_t> aslot_lm, aslot_agg;      // per aslot (sorted by aggregate, then landmark)
    std::vector<uint32_t> aslot_ptr, aslot_edge;    // nAS + 1 offsets into aslot_edge (assembly edge index, ascending)
    std::vector<uint32_t> cpair_a, cpair_b, cpair_ptr;   // coarse block pairs (a, b) sharing a landmark; offsets into ctrip
    std::vector<uint32_t> ctrip;             // 2 per triple: (aslot of a, aslot of b) with the same landmark, landmark ascending
    // ---- multilevel hierarchy (replaces the aggregates above for graphs that run the streamed PCG) -------------------
    std::vector<MlLevel> ml;                 // ml[0] = level 1; the last level is solved exactly (dense inverse)
    uint32_t ml_local = 0;                   // levels 1..ml_local are restricted inside a window (ratio 4, spacing <= 64)
};
// ---- resident plan (world == 1, graphs whose Hpl fits the shared memory of the whole GPU) -----------------
// One block per SM owns a contiguous pose range with all its LM edges (assembly order) and a contiguous landmark
// range with all its LM edges in landmark-major order; both copies of Hpl stay in shared memory for the whole PCG.
struct ResidentPlan {
    uint32_t nb = 0;                          // blocks; 0 = not resident
    std::vector<uint32_t> pose_cut, lm_cut;   // nb + 1 each
    std::vector<uint32_t> agg_cut;            // nb + 1: aggregates owned by each block (aggregates never straddle blocks)
    std::vector<uint32_t> lmo_ptr;            // L + 1: landmark-major CSR over LM edges
    std::vector<uint32_t> lmo_edge;           // nE: assembly index of each landmark-major position (ascending inside a landmark)
    std::vector<uint32_t> edge_lm;            // nE: landmark of each LM edge in assembly order
};
constexpr uint32_t kResEdgeCap = 1280;    // LM edges per block, either order
constexpr uint32_t kResPoseCap = 96;      // poses per block
constexpr uint32_t kResLmCap = 64;        // landmarks per block
constexpr uint32_t kResAggCap = 4;        // aggregates per block
constexpr uint32_t kResOdomCap = 384;     // ODOM neighbour entries per block
constexpr uint32_t kSbCap = 48;           // poses per smoother block of the resident kernel (dense 144 x 144 inverse)
constexpr uint32_t kSbPerBlock = 2;       // smoother blocks per resident block
constexpr uint32_t kCoarseMaxAgg = 512;   // aggregates whose coarse residual (3 doubles each) fits the shared memory the PCG kernels set aside for it
constexpr uint32_t kCoarseMaxAggLarge = 768;    // upper bound on aggregates per rank of the two-level method (dense coarse matrix, inverted by
                                                // k_coarse_invert: nc <= 4736); above kCoarseMaxAgg the coarse residual lives in per-block global
                                                // scratch (Dev::rcs).  HBM-sized graphs use the multilevel hierarchy (MlLevel) instead
constexpr uint32_t kMlTopDefault = 1024;        // nodes of the exactly-solved top level of the hierarchy (nc = 3072: 75 MB inverse)
constexpr uint32_t kMlTopMax = 1536;            // k_coarse_invert handles nc <= 32 * 148

constexpr uint32_t kChunkEdges = 512;   // LM edges staged per chunk
constexpr uint32_t kBlockPoses = 128;   // poses per thread block (upper bound)

// resident_blocks > 0 (world == 1): try to lay the graph out for the shared-memory-resident PCG kernel with that many
// blocks; `rp.nb` stays 0 when it does not fit and the uniform aggregates (agg_K) are used instead
// ml_top > 0: a graph that does not fit the resident plan gets the multilevel hierarchy (top level <= ml_top nodes)
// instead of the uniform aggregates
void build_structure(const HostGraph& g, uint32_t rank, uint32_t world, Structure& s, uint32_t resident_blocks = 0,
                     ResidentPlan* rp = nullptr, uint32_t ml_top = 0);

}  // namespace tsb
