// structure.cpp -- integer assembly structure ("edge -> block index map").
// Replaces the reference's per-edge hash lookups into a dense H (OptimizerCpu.h:48-68,96-99;
// GraphGpu.h:95-111 `indexesInH`) by a precomputed block pattern:
//   * LM edges sorted by (pose, landmark): pose-side sums are contiguous row segments;
//   * chunks of <= kChunkEdges edges covering whole poses; inside a chunk every distinct landmark
//     gets a slot, the chunk emits one partial (Hll, b_l) sum per slot, and a landmark-major CSR
//     over slots finishes the sum.  Fixed summation order everywhere -> deterministic, no atomics.
#include "host_graph.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <numeric>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace tsb {

namespace {

// TSB_HOST_TIMING=1: wall time of each section of build_structure on stderr
struct SectionTimer {
    bool on = std::getenv("TSB_HOST_TIMING") != nullptr;
    std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now();
    void lap(const char* what) {
        if (!on) return;
        const auto n = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[tsb] build_structure %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(n - t).count());
        t = n;
    }
};

// One level of the hierarchy from the level below (child nodes with their landmark-sorted edge lists).
// Parent K interpolates to the children ratio*K - (ratio-1) .. ratio*K + (ratio-1); its coarse edges are the distinct
// landmarks of those children, each with the list of child edges it sums (ascending child edge index).
void build_ml_level(const std::vector<uint32_t>& cptr, const std::vector<uint32_t>& clm, uint32_t ratio, uint32_t spacing, MlLevel& lv) {
    const uint32_t nc = (uint32_t)cptr.size() - 1, n = (nc + ratio - 1) / ratio;
    lv.n = n; lv.ratio = ratio; lv.spacing = spacing;
    lv.node_ptr.assign(n + 1, 0);
    int nt = 1;
#ifdef _OPENMP
    nt = omp_get_max_threads();
#endif
    nt = std::max(1, std::min<int>(nt, (int)(n / 64 + 1)));
    auto k_lo_of = [&](uint32_t K) { return ratio * K >= ratio - 1 ? ratio * K - (ratio - 1) : 0u; };
    auto k_hi_of = [&](uint32_t K) { return (uint32_t)std::min<uint64_t>(nc - 1, (uint64_t)ratio * K + ratio - 1); };
    // every child edge of a node's children is one entry of `ref`: the node's offset into it is known up front, so the
    // threads write their part of `ref` in place; the thread ranges are balanced by that count
    std::vector<uint64_t> roff(n + 1, 0);
    for (uint32_t K = 0; K < n; ++K) roff[K + 1] = roff[K] + (cptr[k_hi_of(K) + 1] - cptr[k_lo_of(K)]);
    lv.ref.resize(roff[n]);
    std::vector<std::vector<uint32_t>> t_lm(nt), t_cnt(nt);
    std::vector<uint32_t> t_lo(nt + 1);
    for (int t = 0; t <= nt; ++t) t_lo[t] = t == nt ? n : (uint32_t)(std::lower_bound(roff.begin(), roff.end(), roff[n] * (uint64_t)t / nt) - roff.begin());
#pragma omp parallel for num_threads(nt) schedule(static, 1)
    for (int t = 0; t < nt; ++t) {
        std::vector<uint64_t> tmp, tmp2;
        uint32_t run[2 * 8 + 2];
        auto& olm = t_lm[t]; auto& ocnt = t_cnt[t];
        const size_t guess = (size_t)((roff[t_lo[t + 1]] - roff[t_lo[t]]) / 2 + 64);
        olm.reserve(guess); ocnt.reserve(guess);
        for (uint32_t K = t_lo[t]; K < t_lo[t + 1]; ++K) {
            const uint32_t k_lo = k_lo_of(K), k_hi = k_hi_of(K);
            // keys (landmark, child edge, child offset); every child's list is landmark-sorted and the child edge indices
            // ascend from child to child, so the sorted order is a merge of <= 7 sorted runs (a third of std::sort's time)
            tmp.clear();
            uint32_t nrun = 0;
            for (uint32_t k = k_lo; k <= k_hi; ++k) {
                run[nrun++] = (uint32_t)tmp.size();
                for (uint32_t e = cptr[k]; e < cptr[k + 1]; ++e)
                    tmp.push_back(((uint64_t)clm[e] << 32) | ((uint64_t)e << 4) | (uint64_t)(k + 7 - ratio * K));
            }
            run[nrun] = (uint32_t)tmp.size();
            tmp2.resize(tmp.size());
            while (nrun > 1) {
                uint32_t nout = 0;
                for (uint32_t q = 0; q < nrun; q += 2) {
                    const uint32_t lo = run[q], mid = run[q + 1], hi = q + 2 <= nrun ? run[std::min(q + 2, nrun)] : mid;
                    const uint64_t *a = tmp.data() + lo, *ae = tmp.data() + mid, *b = ae, *be = tmp.data() + (q + 1 < nrun ? hi : mid);
                    uint64_t* o = tmp2.data() + lo;
                    while (a < ae && b < be) { const bool tb = *b < *a; *o++ = tb ? *b : *a; a += !tb; b += tb; }
                    while (a < ae) *o++ = *a++;
                    while (b < be) *o++ = *b++;
                    run[nout++] = lo;
                }
                run[nout] = (uint32_t)tmp.size();
                nrun = nout; tmp.swap(tmp2);
            }
            uint32_t ne = 0;
            uint32_t* oref = lv.ref.data() + roff[K];
            for (size_t q = 0; q < tmp.size(); ++q) {
                const uint32_t l = (uint32_t)(tmp[q] >> 32);
                if (q == 0 || l != (uint32_t)(tmp[q - 1] >> 32)) { olm.push_back(l); ocnt.push_back(0); ++ne; }
                ocnt.back()++; oref[q] = (uint32_t)tmp[q];
            }
            lv.node_ptr[K + 1] = ne;
        }
    }
    for (uint32_t K = 0; K < n; ++K) lv.node_ptr[K + 1] += lv.node_ptr[K];
    const uint32_t nnz = lv.node_ptr[n];
    lv.edge_lm.resize(nnz); lv.ref_ptr.resize((size_t)nnz + 1); lv.ref_ptr[0] = 0;
    std::vector<size_t> e_off(nt + 1, 0);
    for (int t = 0; t < nt; ++t) e_off[t + 1] = e_off[t] + t_lm[t].size();
#pragma omp parallel for num_threads(nt) schedule(static, 1)
    for (int t = 0; t < nt; ++t) {
        std::copy(t_lm[t].begin(), t_lm[t].end(), lv.edge_lm.begin() + e_off[t]);
        size_t r = roff[t_lo[t]];
        for (size_t q = 0; q < t_cnt[t].size(); ++q) { r += t_cnt[t][q]; lv.ref_ptr[e_off[t] + q + 1] = (uint32_t)r; }
    }
}

// Hierarchy over the nP local poses: ratio 4 per level (a final ratio of 2 where 4 would leave the exactly-solved top
// level with fewer than top_max / 2 nodes); no level at all for graphs with <= 64 poses (block-Jacobi is enough there).
// level plan of nP consecutive poses: (ratio, spacing) per level; returns the nodes of the top level (0: no hierarchy)
uint32_t ml_plan(uint32_t nP, uint32_t top_max, std::vector<std::pair<uint32_t, uint32_t>>* levels = nullptr) {
    if (levels) levels->clear();
    if (nP <= 64 || top_max < 8) return 0;
    uint32_t n = nP, spacing = 1, cnt = 0;
    bool first = true;
    while ((first || n > top_max) && cnt < kMlMaxLevels) {
        uint32_t ratio = 4;                       // the last step takes the smallest ratio that fits the top level: the finer it is, the fewer PCG iterations
        if (!first) { if ((n + 1) / 2 <= top_max) ratio = 2; else if ((n + 2) / 3 <= top_max) ratio = 3; }
        spacing *= ratio; n = (n + ratio - 1) / ratio; ++cnt; first = false;
        if (levels) levels->emplace_back(ratio, spacing);
    }
    return n;
}
void build_hierarchy(Structure& s, const std::vector<uint32_t>& elm, uint32_t top_max) {
    s.ml.clear(); s.ml_local = 0;
    SectionTimer tm;
    s.ml.reserve(kMlMaxLevels);                  // the loop keeps pointers into the previous level
    std::vector<std::pair<uint32_t, uint32_t>> plan;
    if (!ml_plan(s.nP, top_max, &plan)) return;
    const std::vector<uint32_t>* cptr = &s.pose_ptr; const std::vector<uint32_t>* clm = &elm;
    for (const auto& lv : plan) {
        s.ml.emplace_back();
        build_ml_level(*cptr, *clm, lv.first, lv.second, s.ml.back());
        if (std::getenv("TSB_HOST_TIMING")) std::fprintf(stderr, "[tsb]   level spacing %u: n %u nnz %zu refs %zu\n", lv.second, s.ml.back().n, s.ml.back().edge_lm.size(), s.ml.back().ref.size());
        if (lv.first == 4 && lv.second <= kMlHalo && s.ml_local + 1 == s.ml.size()) s.ml_local = (uint32_t)s.ml.size();
        cptr = &s.ml.back().node_ptr; clm = &s.ml.back().edge_lm;
        tm.lap("  hierarchy level");
    }
}

}  // namespace

void build_structure(const HostGraph& g, uint32_t rank, uint32_t world, Structure& s, uint32_t resident_blocks, ResidentPlan* rp,
                     uint32_t ml_top) {
    s = Structure();
    SectionTimer tm;
    const uint32_t P = g.P, L = g.L;
    // ---- pose partition: contiguous ranges balanced by edge work --------------------------------
    std::vector<uint32_t> deg(P + 1, 0);
    for (uint32_t e = 0; e < g.El; ++e) deg[g.le_i[e]]++;
    s.cuts.assign(world < 1 ? 2 : world + 1, 0);
    if (world <= 1) { s.pose_begin = 0; s.pose_end = P; s.cuts = {0, P}; }
    else {
        std::vector<uint64_t> wsum(P + 1, 0);
        std::vector<uint32_t> odeg(P, 0);
        for (uint32_t e = 0; e < g.Eo; ++e) odeg[g.od_i[e]]++;
        for (uint32_t p = 0; p < P; ++p) wsum[p + 1] = wsum[p] + deg[p] + 2ull * odeg[p] + 1;
        for (uint32_t r = 0; r <= world; ++r) {
            if (r == 0) s.cuts[r] = 0;
            else if (r == world) s.cuts[r] = P;
            else s.cuts[r] = (uint32_t)(std::lower_bound(wsum.begin(), wsum.end(), wsum[P] * r / world) - wsum.begin());
        }
        s.pose_begin = s.cuts[rank]; s.pose_end = s.cuts[rank + 1];
    }
    const uint32_t pb = s.pose_begin, pe = s.pose_end, nP = pe - pb;
    s.nP = nP;
    // ---- local LM edges in assembly order --------------------------------------------------------
    s.pose_ptr.assign(nP + 1, 0);
    for (uint32_t p = 0; p < nP; ++p) s.pose_ptr[p + 1] = s.pose_ptr[p] + deg[pb + p];
    const uint32_t nE = s.pose_ptr[nP];
    // request edges of this rank's poses, in request order.  Requests usually list the LM edges pose by pose (the reference's
    // simulator does): then the local edges are one contiguous request range already grouped by pose and the scatter
    // below is the identity
    bool grouped = true;
#pragma omp parallel for schedule(static) reduction(&& : grouped)
    for (long long e = 1; e < (long long)g.El; ++e) grouped = grouped && g.le_i[e - 1] <= g.le_i[e];
    s.le_local.resize(nE); s.lm_perm.resize(nE);
    if (grouped) {
        const uint32_t first = (uint32_t)(std::lower_bound(g.le_i.begin(), g.le_i.end(), pb) - g.le_i.begin());
#pragma omp parallel for schedule(static)
        for (long long k = 0; k < (long long)nE; ++k) { s.le_local[k] = first + (uint32_t)k; s.lm_perm[k] = (uint32_t)k; }
    } else {
        std::vector<uint32_t> cur(s.pose_ptr.begin(), s.pose_ptr.end() - 1);
        uint32_t nl = 0;
        for (uint32_t e = 0; e < g.El; ++e) {
            const uint32_t p = g.le_i[e];
            if (p < pb || p >= pe) continue;
            s.lm_perm[cur[p - pb]++] = nl;                               // stable: request order inside a pose
            s.le_local[nl++] = e;
        }
    }
#pragma omp parallel for schedule(dynamic, 1024)
    for (long long pp = 0; pp < (long long)nP; ++pp) {                   // then by landmark inside a pose
        const uint32_t p = (uint32_t)pp;
        auto b = s.lm_perm.begin() + s.pose_ptr[p], e = s.lm_perm.begin() + s.pose_ptr[p + 1];
        std::sort(b, e, [&](uint32_t x, uint32_t y) {
            const uint32_t lx = g.le_l[s.le_local[x]], ly = g.le_l[s.le_local[y]];
            return lx != ly ? lx < ly : x < y; });
    }
    tm.lap("partition + edge order");
    // ---- blocks and chunks ------------------------------------------------------------------------
    s.blk_chunk_ptr.push_back(0);
    for (uint32_t p = 0; p < nP;) {
        const uint32_t d = s.pose_ptr[p + 1] - s.pose_ptr[p];
        const uint32_t blk = (uint32_t)s.blk_chunk_ptr.size() - 1;
        if (d > kChunkEdges) {                                       // a single heavy pose: several chunks
            for (uint32_t e0 = s.pose_ptr[p]; e0 < s.pose_ptr[p + 1]; e0 += kChunkEdges)
                s.chunks.push_back({e0, std::min(e0 + kChunkEdges, s.pose_ptr[p + 1]), p, p + 1, 0, 0, blk, 0});
            ++p;
        } else {
            uint32_t q = p, cnt = 0;
            while (q < nP && q - p < kBlockPoses) {
                const uint32_t dq = s.pose_ptr[q + 1] - s.pose_ptr[q];
                if (cnt + dq > kChunkEdges) break;
                cnt += dq; ++q;
            }
            s.chunks.push_back({s.pose_ptr[p], s.pose_ptr[q], p, q, 0, 0, blk, 0});
            p = q;
        }
        s.blk_chunk_ptr.push_back((uint32_t)s.chunks.size());
    }
    tm.lap("chunks");
    // ---- slots: distinct landmarks per chunk ---------------------------------------------------------
    s.lm_packed.resize(nE); s.slot_edge.resize(nE);
    std::vector<uint32_t> elm(nE);                       // landmark of each LM edge in assembly order
#pragma omp parallel for schedule(static)
    for (long long k = 0; k < (long long)nE; ++k) elm[k] = g.le_l[s.le_local[s.lm_perm[k]]];
    std::vector<uint32_t> stamp(L, 0xFFFFFFFFu), lslot(L, 0), uniq, cnt;      // also used by the aggregate path below
    {
        // every thread takes a contiguous range of chunks (<= kChunkEdges edges each: balanced) and numbers its slots from
        // 0; the ranges are concatenated afterwards.  lm_packed / slot_edge / slot_ptr hold chunk-relative or edge-absolute
        // values, so only ChunkDesc::s0 / s1 need the thread's slot offset
        const uint32_t nCh = (uint32_t)s.chunks.size();
        int nt = 1;
#ifdef _OPENMP
        nt = omp_get_max_threads();
#endif
        nt = std::max(1, std::min<int>(nt, (int)(nCh / 256 + 1)));
        std::vector<std::vector<uint32_t>> t_slm(nt), t_sptr(nt);
#pragma omp parallel for num_threads(nt) schedule(static, 1)
        for (int t = 0; t < nt; ++t) {
            const uint32_t c_lo = (uint32_t)((uint64_t)nCh * t / nt), c_hi = (uint32_t)((uint64_t)nCh * (t + 1) / nt);
            std::vector<uint32_t> st_(L, 0xFFFFFFFFu), ls_(L, 0), uq, ucnt, order, srank, ct;
            auto& oslm = t_slm[t]; auto& osptr = t_sptr[t];
            for (uint32_t ci = c_lo; ci < c_hi; ++ci) {
                ChunkDesc& c = s.chunks[ci];
                const uint32_t n = c.e1 - c.e0;
                // distinct landmarks of the chunk (stamp table: O(n)), then sorted by landmark
                uq.clear();
                for (uint32_t k = 0; k < n; ++k) { const uint32_t l = elm[c.e0 + k]; if (st_[l] != ci) { st_[l] = ci; uq.push_back(l); } }
                std::sort(uq.begin(), uq.end());
                const uint32_t nu = (uint32_t)uq.size();
                for (uint32_t q = 0; q < nu; ++q) ls_[uq[q]] = q;          // landmark-sorted index
                // slots are numbered by descending edge count (ties: ascending landmark): the 32 lanes of a warp that sum
                // neighbouring slots then run the same number of iterations
                ucnt.assign(nu, 0); srank.resize(nu); order.resize(nu);
                for (uint32_t k = 0; k < n; ++k) ucnt[ls_[elm[c.e0 + k]]]++;
                std::iota(order.begin(), order.end(), 0u);
                std::stable_sort(order.begin(), order.end(), [&](uint32_t x, uint32_t y) { return ucnt[x] > ucnt[y]; });
                for (uint32_t q = 0; q < nu; ++q) srank[order[q]] = q;
                c.s0 = (uint32_t)oslm.size(); c.s1 = c.s0 + nu;            // thread-relative until the ranges are joined
                for (uint32_t q = 0; q < nu; ++q) oslm.push_back(uq[order[q]]);
                ct.assign(nu + 1, 0);
                uint32_t pose = c.p0;
                for (uint32_t k = 0; k < n; ++k) {
                    const uint32_t slot = srank[ls_[elm[c.e0 + k]]];
                    while (c.e0 + k >= s.pose_ptr[pose + 1]) ++pose;
                    s.lm_packed[c.e0 + k] = ((pose - c.p0) << 16) | slot;
                    ct[slot + 1]++;
                }
                for (size_t q = 0; q < nu; ++q) ct[q + 1] += ct[q];
                for (size_t q = 0; q < nu; ++q) osptr.push_back(c.e0 + ct[q + 1]);
                for (uint32_t k = 0; k < n; ++k) s.slot_edge[c.e0 + ct[s.lm_packed[c.e0 + k] & 0xffffu]++] = (uint16_t)k;
            }
        }
        std::vector<uint32_t> s_off(nt + 1, 0);
        for (int t = 0; t < nt; ++t) s_off[t + 1] = s_off[t] + (uint32_t)t_slm[t].size();
        s.slot_lm.resize(s_off[nt]); s.slot_ptr.resize((size_t)s_off[nt] + 1); s.slot_ptr[0] = 0;
#pragma omp parallel for num_threads(nt) schedule(static, 1)
        for (int t = 0; t < nt; ++t) {
            std::copy(t_slm[t].begin(), t_slm[t].end(), s.slot_lm.begin() + s_off[t]);
            std::copy(t_sptr[t].begin(), t_sptr[t].end(), s.slot_ptr.begin() + 1 + s_off[t]);
            const uint32_t c_lo = (uint32_t)((uint64_t)nCh * t / nt), c_hi = (uint32_t)((uint64_t)nCh * (t + 1) / nt);
            for (uint32_t ci = c_lo; ci < c_hi; ++ci) { s.chunks[ci].s0 += s_off[t]; s.chunks[ci].s1 += s_off[t]; }
        }
    }
    tm.lap("slots");
    // ---- landmark-major CSR over slots ---------------------------------------------------------------
    const uint32_t nS = (uint32_t)s.slot_lm.size();
    s.lm_part_ptr.assign(L + 1, 0);
    for (uint32_t q = 0; q < nS; ++q) s.lm_part_ptr[s.slot_lm[q] + 1]++;
    for (uint32_t l = 0; l < L; ++l) s.lm_part_ptr[l + 1] += s.lm_part_ptr[l];
    s.lm_part_idx.resize(nS);
    { std::vector<uint32_t> c2(s.lm_part_ptr.begin(), s.lm_part_ptr.end() - 1);
      for (uint32_t q = 0; q < nS; ++q) s.lm_part_idx[c2[s.slot_lm[q]]++] = q; }
    tm.lap("landmark CSR over slots");
    // ---- ODOM edges touching a local pose -----------------------------------------------------------
    auto local = [&](uint32_t p) { return p >= pb && p < pe; };
    std::vector<uint32_t> ghosts;
    for (uint32_t e = 0; e < g.Eo; ++e) {
        const bool a = local(g.od_i[e]), b = local(g.od_j[e]);
        if (a && !b) ghosts.push_back(g.od_j[e]);
        if (b && !a) ghosts.push_back(g.od_i[e]);
    }
    std::sort(ghosts.begin(), ghosts.end()); ghosts.erase(std::unique(ghosts.begin(), ghosts.end()), ghosts.end());
    s.ghost_pose = ghosts; s.nG = (uint32_t)ghosts.size();
    auto lidx = [&](uint32_t p) -> uint32_t {
        if (local(p)) return p - pb;
        return nP + (uint32_t)(std::lower_bound(ghosts.begin(), ghosts.end(), p) - ghosts.begin());
    };
    s.od_out_ptr.assign(nP + 1, 0);
    for (uint32_t e = 0; e < g.Eo; ++e) if (local(g.od_i[e])) s.od_out_ptr[g.od_i[e] - pb + 1]++;
    for (uint32_t p = 0; p < nP; ++p) s.od_out_ptr[p + 1] += s.od_out_ptr[p];
    s.n_od_owned = s.od_out_ptr[nP];
    s.od_local.assign(s.n_od_owned, 0);
    { std::vector<uint32_t> c2(s.od_out_ptr.begin(), s.od_out_ptr.end() - 1);
      for (uint32_t e = 0; e < g.Eo; ++e) if (local(g.od_i[e])) s.od_local[c2[g.od_i[e] - pb]++] = e; }
    for (uint32_t e = 0; e < g.Eo; ++e) if (!local(g.od_i[e]) && local(g.od_j[e])) s.od_local.push_back(e);
    const uint32_t nO = (uint32_t)s.od_local.size();
    s.od_a.resize(nO); s.od_b.resize(nO);
    for (uint32_t k = 0; k < nO; ++k) { s.od_a[k] = lidx(g.od_i[s.od_local[k]]); s.od_b[k] = lidx(g.od_j[s.od_local[k]]); }
    s.od_in_ptr.assign(nP + 1, 0);
    for (uint32_t k = 0; k < nO; ++k) if (s.od_b[k] < nP) s.od_in_ptr[s.od_b[k] + 1]++;
    for (uint32_t p = 0; p < nP; ++p) s.od_in_ptr[p + 1] += s.od_in_ptr[p];
    s.od_in_edge.resize(s.od_in_ptr[nP]);
    { std::vector<uint32_t> c2(s.od_in_ptr.begin(), s.od_in_ptr.end() - 1);
      for (uint32_t k = 0; k < nO; ++k) if (s.od_b[k] < nP) s.od_in_edge[c2[s.od_b[k]]++] = k; }
    tm.lap("odom lists");
    // ---- boundary poses (same list on every rank) ---------------------------------------------------
    if (world > 1) {
        auto owner = [&](uint32_t p) { return (uint32_t)(std::upper_bound(s.cuts.begin(), s.cuts.end(), p) - s.cuts.begin()) - 1; };
        for (uint32_t e = 0; e < g.Eo; ++e)
            if (owner(g.od_i[e]) != owner(g.od_j[e])) { s.boundary.push_back(g.od_i[e]); s.boundary.push_back(g.od_j[e]); }
        std::sort(s.boundary.begin(), s.boundary.end());
        s.boundary.erase(std::unique(s.boundary.begin(), s.boundary.end()), s.boundary.end());
        for (uint32_t k = 0; k < s.boundary.size(); ++k)
            if (local(s.boundary[k])) { s.own_bnd_slot.push_back(k); s.own_bnd_pose.push_back(s.boundary[k] - pb); }
        for (uint32_t gp : s.ghost_pose)
            s.ghost_bnd_slot.push_back((uint32_t)(std::lower_bound(s.boundary.begin(), s.boundary.end(), gp) - s.boundary.begin()));
        // peer exchange: every rank's ghost list (sorted, unique) -> where this rank's boundary poses go
        std::vector<std::vector<uint32_t>> gl(world);
        for (uint32_t e = 0; e < g.Eo; ++e) {
            const uint32_t oi = owner(g.od_i[e]), oj = owner(g.od_j[e]);
            if (oi != oj) { gl[oj].push_back(g.od_i[e]); gl[oi].push_back(g.od_j[e]); }
        }
        for (uint32_t r = 0; r < world; ++r) {
            std::sort(gl[r].begin(), gl[r].end()); gl[r].erase(std::unique(gl[r].begin(), gl[r].end()), gl[r].end());
            if (r == rank) continue;
            for (uint32_t k = 0; k < gl[r].size(); ++k)
                if (local(gl[r][k])) { s.x_exp.push_back(gl[r][k] - pb); s.x_exp.push_back(r); s.x_exp.push_back(k); }
        }
        s.lm_mask.assign(L, 0);
        for (uint32_t e = 0; e < g.El; ++e) s.lm_mask[g.le_l[e]] |= (uint8_t)(1u << owner(g.le_i[e]));
    }
    tm.lap("boundary / exchange lists");
    // ---- coarse space: aggregates of agg_K consecutive local poses ------------------------------------
    {
        // ---- resident plan: pose / landmark ranges per block, balanced by LM edges ----------------------------
        bool resident = false;
        if (rp) *rp = ResidentPlan();
        if (rp && resident_blocks > 0 && world <= 1 && nP > 0 && nE <= (uint64_t)resident_blocks * kResEdgeCap) {
            const uint32_t nb = resident_blocks;
            ResidentPlan& r = *rp;
            r.pose_cut.assign(nb + 1, nP); r.pose_cut[0] = 0;
            for (uint32_t b = 1; b < nb; ++b) {          // first pose whose edge offset reaches b/nb of the edges (ties: by pose count)
                const uint64_t target = ((uint64_t)nE + nP) * b / nb;
                uint32_t lo = r.pose_cut[b - 1], hi = nP;
                while (lo < hi) { const uint32_t mid = (lo + hi) / 2; if ((uint64_t)s.pose_ptr[mid] + mid < target) lo = mid + 1; else hi = mid; }
                r.pose_cut[b] = lo;
            }
            std::vector<uint32_t> ldeg(L + 1, 0);
            for (uint32_t k = 0; k < nE; ++k) ldeg[elm[k] + 1]++;
            r.lmo_ptr.assign(L + 1, 0);
            for (uint32_t l = 0; l < L; ++l) r.lmo_ptr[l + 1] = r.lmo_ptr[l] + ldeg[l + 1];
            r.lm_cut.assign(nb + 1, L); r.lm_cut[0] = 0;
            for (uint32_t b = 1; b < nb; ++b) {
                const uint64_t target = ((uint64_t)nE + L) * b / nb;
                uint32_t lo = r.lm_cut[b - 1], hi = L;
                while (lo < hi) { const uint32_t mid = (lo + hi) / 2; if ((uint64_t)r.lmo_ptr[mid] + mid < target) lo = mid + 1; else hi = mid; }
                r.lm_cut[b] = lo;
            }
            r.edge_lm.resize(nE); r.lmo_edge.resize(nE);
            { std::vector<uint32_t> c2(r.lmo_ptr.begin(), r.lmo_ptr.end() - 1);
              for (uint32_t k = 0; k < nE; ++k) { const uint32_t l = elm[k]; r.edge_lm[k] = l; r.lmo_edge[c2[l]++] = k; } }
            // ODOM neighbour entries per block (owned out-edges + in-edges of its poses)
            resident = true;
            r.agg_cut.assign(nb + 1, 0);
            for (uint32_t b = 0; b < nb && resident; ++b) {
                const uint32_t p0 = r.pose_cut[b], p1 = r.pose_cut[b + 1], l0 = r.lm_cut[b], l1 = r.lm_cut[b + 1];
                uint32_t nod = 0;
                for (uint32_t p = p0; p < p1; ++p) nod += (s.od_out_ptr[p + 1] - s.od_out_ptr[p]) + (s.od_in_ptr[p + 1] - s.od_in_ptr[p]);
                if (p1 - p0 > kResPoseCap || l1 - l0 > kResLmCap || s.pose_ptr[p1] - s.pose_ptr[p0] > kResEdgeCap ||
                    r.lmo_ptr[l1] - r.lmo_ptr[l0] > kResEdgeCap || nod > kResOdomCap) resident = false;
                const uint32_t m = p1 > p0 ? std::min<uint32_t>(kResAggCap, std::max<uint32_t>(1, (p1 - p0 + 16) / 32)) : 0;
                r.agg_cut[b + 1] = r.agg_cut[b] + m;
            }
            if (resident && r.agg_cut[nb] > kCoarseMaxAgg) resident = false;
            if (resident) {
                r.nb = nb; s.agg_K = 0; s.nA = r.agg_cut[nb];
                s.agg_ptr.assign(s.nA + 1, nP);
                for (uint32_t b = 0; b < nb; ++b) {
                    const uint32_t p0 = r.pose_cut[b], np = r.pose_cut[b + 1] - p0, m = r.agg_cut[b + 1] - r.agg_cut[b];
                    for (uint32_t q = 0; q < m; ++q) s.agg_ptr[r.agg_cut[b] + q] = p0 + (uint32_t)((uint64_t)np * q / m);
                }
            } else *rp = ResidentPlan();
        }
        // sharded: every rank builds the hierarchy of its own pose range; the top levels of all ranks together form the
        // exactly-solved global top level (ml_top nodes in total), so every rank needs every rank's node count -- and a
        // rank too small for a hierarchy switches it off everywhere (the ranks must take the same kernel phases)
        uint32_t top_rank = ml_top;
        if (ml_top > 0 && world > 1) {
            top_rank = std::max<uint32_t>(16, ml_top / world);
            s.top_nodes.assign(world, 0);
            for (uint32_t r = 0; r < world; ++r) { s.top_nodes[r] = ml_plan(s.cuts[r + 1] - s.cuts[r], top_rank); if (!s.top_nodes[r]) top_rank = 0; }
            if (!top_rank) s.top_nodes.clear();
        }
        if (!resident && top_rank > 0) {
            // streamed PCG: multilevel hierarchy instead of the uniform aggregates (no aslots, no block pairs)
            build_hierarchy(s, elm, top_rank);
            s.agg_K = 32; s.nA = 0; s.agg_ptr.assign(1, 0);
            s.aslot_ptr.push_back(0); s.cpair_ptr.push_back(0);
            s.edge_pose.resize(nE);
#pragma omp parallel for schedule(static)
            for (long long p = 0; p < (long long)nP; ++p) for (uint32_t k = s.pose_ptr[p]; k < s.pose_ptr[p + 1]; ++k) s.edge_pose[k] = (uint32_t)p;
        }
        const bool use_ml = !resident && top_rank > 0;
        if (!resident && !use_ml) {
            uint32_t K = 32;
            if (nP > 32ull * kCoarseMaxAggLarge) K = 32 * (uint32_t)((nP + 32ull * kCoarseMaxAggLarge - 1) / (32ull * kCoarseMaxAggLarge));
            s.agg_K = K; s.nA = (nP + K - 1) / K;
            s.agg_ptr.resize(s.nA + 1);
            for (uint32_t a = 0; a <= s.nA; ++a) s.agg_ptr[a] = std::min<uint64_t>(nP, (uint64_t)a * K);
        }
        if (!use_ml) {
        s.pose_agg.resize(nP);
        for (uint32_t a = 0; a < s.nA; ++a) for (uint32_t p = s.agg_ptr[a]; p < s.agg_ptr[a + 1]; ++p) s.pose_agg[p] = a;
        s.edge_pose.resize(nE);
        for (uint32_t p = 0; p < nP; ++p) for (uint32_t k = s.pose_ptr[p]; k < s.pose_ptr[p + 1]; ++k) s.edge_pose[k] = p;
        s.aslot_ptr.push_back(0); s.aslot_edge.resize(nE);
        std::fill(stamp.begin(), stamp.end(), 0xFFFFFFFFu);
        for (uint32_t a = 0; a < s.nA; ++a) {
            const uint32_t e0 = s.pose_ptr[s.agg_ptr[a]], e1 = s.pose_ptr[s.agg_ptr[a + 1]];
            uniq.clear();
            for (uint32_t k = e0; k < e1; ++k) { const uint32_t l = elm[k]; if (stamp[l] != a) { stamp[l] = a; uniq.push_back(l); } }
            std::sort(uniq.begin(), uniq.end());
            const uint32_t nu = (uint32_t)uniq.size(), base = (uint32_t)s.aslot_lm.size();
            cnt.assign(nu + 1, 0);
            for (uint32_t q = 0; q < nu; ++q) { lslot[uniq[q]] = q; s.aslot_lm.push_back(uniq[q]); s.aslot_agg.push_back(a); }
            for (uint32_t k = e0; k < e1; ++k) cnt[lslot[elm[k]] + 1]++;
            for (uint32_t q = 0; q < nu; ++q) { cnt[q + 1] += cnt[q]; s.aslot_ptr.push_back(e0 + cnt[q + 1]); }
            for (uint32_t k = e0; k < e1; ++k) s.aslot_edge[e0 + cnt[lslot[elm[k]]]++] = k;      // ascending edge index inside an aslot
            (void)base;
        }
        const uint32_t nAS = (uint32_t)s.aslot_lm.size();
        // landmark-major lists of aslots (ascending aslot id = ascending aggregate)
        std::vector<uint32_t> lptr(L + 1, 0), lidx(nAS);
        for (uint32_t q = 0; q < nAS; ++q) lptr[s.aslot_lm[q] + 1]++;
        for (uint32_t l = 0; l < L; ++l) lptr[l + 1] += lptr[l];
        { std::vector<uint32_t> c2(lptr.begin(), lptr.end() - 1);
          for (uint32_t q = 0; q < nAS; ++q) lidx[c2[s.aslot_lm[q]]++] = q; }
        // pairs (a, b) and their triples, grouped by a, then b, landmark ascending inside
        s.cpair_ptr.push_back(0);
        std::vector<uint32_t> tb, cntb(s.nA + 1), posb;        // per emitted triple: aggregate b; counting sort by b (stable)
        std::vector<std::pair<uint32_t, uint32_t>> ss;
        uint32_t q0 = 0;
        for (uint32_t a = 0; a < s.nA; ++a) {
            tb.clear(); ss.clear();
            uint32_t q1 = q0;
            while (q1 < nAS && s.aslot_agg[q1] == a) ++q1;
            for (uint32_t q = q0; q < q1; ++q) {
                const uint32_t l = s.aslot_lm[q];
                for (uint32_t j = lptr[l]; j < lptr[l + 1]; ++j) { tb.push_back(s.aslot_agg[lidx[j]]); ss.emplace_back(q, lidx[j]); }
            }
            // group by b, keeping the emission order (landmark ascending) inside each group
            std::fill(cntb.begin(), cntb.end(), 0u);
            for (uint32_t b : tb) cntb[b + 1]++;
            const size_t base = s.ctrip.size() / 2;
            for (uint32_t b = 0; b < s.nA; ++b) {
                if (cntb[b + 1]) { s.cpair_a.push_back(a); s.cpair_b.push_back(b); s.cpair_ptr.push_back((uint32_t)(base + cntb[b] + cntb[b + 1])); }
                cntb[b + 1] += cntb[b];
            }
            s.ctrip.resize(2 * (base + tb.size()));
            posb.assign(cntb.begin(), cntb.end() - 1);
            for (size_t t = 0; t < tb.size(); ++t) {
                const size_t at = base + posb[tb[t]]++;
                s.ctrip[2 * at] = ss[t].first; s.ctrip[2 * at + 1] = ss[t].second;
            }
            q0 = q1;
        }
        }
    }
    tm.lap("coarse space / hierarchy");
    // ---- fixed vertices ---------------------------------------------------------------------------
    s.fixed_pose_count.assign(nP, 0); s.fixed_lm_count.assign(L, 0);
    for (uint32_t v : g.fixed) {
        if (v < P) { if (local(v)) s.fixed_pose_count[v - pb]++; }
        else s.fixed_lm_count[v - P]++;
    }
}

}  // namespace tsb
