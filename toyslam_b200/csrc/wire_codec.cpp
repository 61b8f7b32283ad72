// wire_codec.cpp -- bulk wire parser / reply encoder (host).
// Request layout: remote/serialization/DeserializeGraph.h:9-174 (what python/remote/graph_to_bytes.py:32-67
// sends); reply layout: remote/serialization/SerializeGraph.h:8-72 with the CPU policy
// (SerializeGraphFuncCpu.h:10-65).  Unlike the reference parser this one is bounds-checked and
// never throws: a malformed payload is a returned error.
#include "host_graph.h"
#include <chrono>
#include <cstdio>
#include <cstdlib>

#include <algorithm>
#include <cmath>
#include <cstring>

namespace tsb {
namespace {

struct Reader {
    const uint8_t* p; const uint8_t* end;
    bool have(size_t n) const { return (size_t)(end - p) >= n; }
    uint32_t u32() { uint32_t v; std::memcpy(&v, p, 4); p += 4; return v; }
    float f32() { float v; std::memcpy(&v, p, 4); p += 4; return v; }
};

// general 3x3 inverse (adjugate / determinant); only the top two rows are kept
void zinv6(const double* m, double* o) {
    const double c00 = m[4] * m[8] - m[5] * m[7], c01 = m[5] * m[6] - m[3] * m[8], c02 = m[3] * m[7] - m[4] * m[6];
    const double id = 1.0 / (m[0] * c00 + m[1] * c01 + m[2] * c02);
    o[0] = c00 * id; o[1] = (m[2] * m[7] - m[1] * m[8]) * id; o[2] = (m[1] * m[5] - m[2] * m[4]) * id;
    o[3] = c01 * id; o[4] = (m[0] * m[8] - m[2] * m[6]) * id; o[5] = (m[2] * m[3] - m[0] * m[5]) * id;
}

// CreateTransform2d (remote/graph/Helper.h:6-18): rotation through a unit quaternion
void transform2d(double x, double y, double th, double* Z) {
    const double w = std::cos(0.5 * th), z = std::sin(0.5 * th), tz = 2.0 * z;
    const double c = 1.0 - tz * z, s = tz * w;
    Z[0] = c; Z[1] = -s; Z[2] = x; Z[3] = s; Z[4] = c; Z[5] = y; Z[6] = 0; Z[7] = 0; Z[8] = 1;
}

struct IdMap {
    bool direct = false;
    std::vector<uint32_t> table;              // direct: id -> code (0xFFFFFFFF = absent)
    std::vector<uint32_t> ids, codes;         // sorted
    uint32_t find(uint32_t id) const {
        if (direct) return id < table.size() ? table[id] : 0xFFFFFFFFu;
        auto it = std::lower_bound(ids.begin(), ids.end(), id);
        return (it == ids.end() || *it != id) ? 0xFFFFFFFFu : codes[it - ids.begin()];
    }
};

}  // namespace

bool parse_wire(const uint8_t* buf, uint32_t nbytes, HostGraph& g, std::string& err) {
    const bool tm_on = std::getenv("TSB_HOST_TIMING") != nullptr;
    auto tm_t = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) { if (!tm_on) return; const auto n = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[tsb] parse_wire %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(n - tm_t).count()); tm_t = n; };
    g.reset();
    Reader r{buf, buf + nbytes};
    auto fail = [&](const char* m) { err = m; return false; };
    if (!r.have(4)) return fail("truncated payload: vertex count");
    const uint32_t V = r.u32();
    if ((uint64_t)V * 16 > nbytes) return fail("vertex count exceeds payload size");
    // pass 1 over vertices: ids, types, raw values
    std::vector<uint32_t> vid(V), vtype(V); std::vector<float> vval(3 * (size_t)V);
    uint32_t maxid = 0;
    for (uint32_t i = 0; i < V; ++i) {
        if (!r.have(8)) return fail("truncated payload: vertex header");
        vid[i] = r.u32(); vtype[i] = r.u32();
        const uint32_t n = vtype[i] == 0 ? 3u : (vtype[i] == 1 ? 2u : 0u);
        if (n == 0) return fail("unknown vertex type");
        if (!r.have(4 * n)) return fail("truncated payload: vertex position");
        for (uint32_t k = 0; k < n; ++k) vval[3 * (size_t)i + k] = r.f32();
        maxid = std::max(maxid, vid[i]);
    }
    // id -> request index of the surviving (last) occurrence (AddVertex overwrites, GraphCpu.h:15-18)
    IdMap m;
    std::vector<uint32_t> order;              // request indices of survivors, ascending id
    m.direct = V > 0 && (uint64_t)maxid < 8ull * V + 1024;
    if (m.direct) {
        m.table.assign((size_t)maxid + 1, 0xFFFFFFFFu);
        for (uint32_t i = 0; i < V; ++i) m.table[vid[i]] = i;
        order.reserve(V);
        for (uint32_t id = 0; id <= maxid; ++id) if (m.table[id] != 0xFFFFFFFFu) order.push_back(m.table[id]);
    } else {
        std::vector<uint32_t> idx(V);
        for (uint32_t i = 0; i < V; ++i) idx[i] = i;
        std::stable_sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) { return vid[a] < vid[b]; });
        for (uint32_t k = 0; k < V; ++k) if (k + 1 == V || vid[idx[k + 1]] != vid[idx[k]]) order.push_back(idx[k]);
    }
    // canonical numbering; code = index | type << 31
    g.req_vertex.assign(V, 0xFFFFFFFFu);
    std::vector<uint32_t> code_of_survivor(order.size());
    for (size_t k = 0; k < order.size(); ++k) {
        const uint32_t i = order[k];
        if (vtype[i] == 0) {
            code_of_survivor[k] = g.P;
            g.pose_id.push_back(vid[i]);
            for (int q = 0; q < 3; ++q) g.pose_xyt.push_back((double)vval[3 * (size_t)i + q]);
            g.req_vertex[i] = g.P++;
        } else {
            code_of_survivor[k] = g.L | 0x80000000u;
            g.lm_id.push_back(vid[i]);
            for (int q = 0; q < 2; ++q) g.lm_xy.push_back((double)vval[3 * (size_t)i + q]);
            g.req_vertex[i] = (g.L++) | 0x80000000u;
        }
    }
    if (m.direct) { for (size_t k = 0; k < order.size(); ++k) m.table[vid[order[k]]] = code_of_survivor[k]; }
    else { m.ids.resize(order.size()); m.codes = code_of_survivor; for (size_t k = 0; k < order.size(); ++k) m.ids[k] = vid[order[k]]; }
    // a dropped duplicate must not reuse a survivor's slot in the reply
    for (uint32_t i = 0; i < V; ++i) if (g.req_vertex[i] != 0xFFFFFFFFu && m.find(vid[i]) != g.req_vertex[i]) g.req_vertex[i] = 0xFFFFFFFFu;

    lap("vertices");
    g.edges_offset = (uint32_t)(r.p - buf);
    if (!r.have(4)) return fail("truncated payload: edge count");
    const uint32_t E = r.u32();
    if ((uint64_t)E * 36 > nbytes) return fail("edge count exceeds payload size");
    // ---- pass 1 over edges (sequential, headers only): record offsets, per-type indices, shape checks --------------------
    // Records have no fixed stride in general (ODOM as (x, y, theta) or as a matrix, information diagonals of any length
    // up to the edge dimension), so the offsets come from one cheap scan; the values are then decoded in parallel.
    g.edge_off.resize(E); g.edge_slot.resize(E);
    uint32_t nO = 0, nL = 0;
    // Runs of records in the layout the reference client writes (graph_to_bytes.py:44-63: ODOM as a 3x3 matrix + 3 information
    // values = 76 B, LM as (range, bearing) + 2 = 44 B) are recognised at a fixed stride: the header words of a whole block of
    // records are compared in parallel (no pointer chasing from record to record, which is what bounds the scan below at ~10 ns
    // per record), and the block's offsets are written by the same loop.  Anything else -- and the first record that breaks a
    // run -- goes through the record-by-record scan, which also produces the error messages.
    auto std_size = [&](const uint8_t* q, size_t left, uint32_t& type) -> uint32_t {
        if (left < 76) return 0;
        uint32_t w[5]; std::memcpy(w, q, 4); std::memcpy(w + 1, q + 12, 8);
        type = w[0];
        if (w[0] == 0 && w[1] == 3 && w[2] == 3) { std::memcpy(w + 3, q + 56, 8); return (w[3] == 0 && w[4] == 3) ? 76u : 0u; }
        if (w[0] == 1 && w[1] == 0 && w[2] == 2) { std::memcpy(w + 3, q + 28, 8); return (w[3] == 0 && w[4] == 2) ? 44u : 0u; }
        return 0;
    };
    constexpr uint32_t kRunBlock = 1u << 16;
    for (uint32_t i = 0; i < E;) {
        {   // fast path: a block of same-layout records
            uint32_t type = 0;
            const size_t left = (size_t)(r.end - r.p);
            const uint32_t S = std_size(r.p, left, type);
            const uint32_t cnt = S ? (uint32_t)std::min<uint64_t>(std::min<uint64_t>(kRunBlock, E - i), (left - 76 + S) / S) : 0;   // every record of the block has 76 readable bytes
            if (cnt >= 1024) {
                const uint8_t* base = r.p;
                uint32_t first_bad = cnt;
#pragma omp parallel for schedule(static) reduction(min : first_bad)
                for (long long k = 0; k < (long long)cnt; ++k) {
                    uint32_t t2 = 0;
                    if (std_size(base + (size_t)k * S, 76, t2) != S || t2 != type) { first_bad = std::min(first_bad, (uint32_t)k); continue; }
                    g.edge_off[i + k] = (uint32_t)(base - buf) + (uint32_t)k * S;
                    g.edge_slot[i + k] = (type == 0 ? nO : nL) + (uint32_t)k;
                }
                if (first_bad > 0) {
                    (type == 0 ? nO : nL) += first_bad; i += first_bad; r.p += (size_t)first_bad * S;
                    continue;
                }
            }
        }
        if (!r.have(20)) return fail("truncated payload: edge header");
        g.edge_off[i] = (uint32_t)(r.p - buf);
        const uint32_t type = r.u32(); r.p += 8; const uint32_t rows = r.u32(), cols = r.u32();
        if (type > 1) return fail("NOT IMPLE INF edge type");
        uint32_t mbytes;
        if (rows == 0) mbytes = type == 0 ? 12u : 8u;      // DeserializeGraph.h:74-95
        else {                                             // DeserializeGraph.h:97-116: cols x cols values
            if (type != 0) return fail("NOT IMPLE INF edge type");
            if (cols > 3) return fail("ODOM measurement larger than 3x3");
            mbytes = 4u * cols * cols;
        }
        if (!r.have((size_t)mbytes + 8)) return fail("truncated payload: measurement");
        r.p += mbytes;
        const uint32_t ri = r.u32(), ci = r.u32();
        if (ri != 0) return fail("NOT IMPLE INF NxN");
        if (ci > (type == 0 ? 3u : 2u)) return fail("information diagonal longer than the edge dimension");
        if (!r.have(4ull * ci)) return fail("truncated payload: information");
        r.p += 4ull * ci;
        g.edge_slot[i] = type == 0 ? nO++ : nL++;
        ++i;
    }
    g.Eo = nO; g.El = nL;
    lap("edge offset scan");
    g.od_i.resize(nO); g.od_j.resize(nO); g.od_zinv.resize(6ull * nO); g.od_w.resize(3ull * nO);
    g.le_i.resize(nL); g.le_l.resize(nL); g.le_z.resize(2ull * nL); g.le_w.resize(2ull * nL);
    // ---- pass 2 (parallel over edges): ids -> canonical indices, values, topology hash ------------------------------------
    uint32_t bad = 0xFFFFFFFFu; int bad_kind = 0;           // lowest failing edge and why (same message as a sequential parse)
    uint64_t hsum = 0;
    auto mix = [](uint64_t x) { x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull; return x ^ (x >> 31); };
#pragma omp parallel for schedule(static) reduction(+ : hsum)
    for (long long ii = 0; ii < (long long)E; ++ii) {
        const uint32_t i = (uint32_t)ii;
        Reader q{buf + g.edge_off[i], buf + nbytes};
        const uint32_t type = q.u32(), id1 = q.u32(), id2 = q.u32(), rows = q.u32(), cols = q.u32();
        double Z[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}; double rb[2] = {0, 0};
        if (rows == 0) {
            if (type == 0) { const float x = q.f32(), y = q.f32(), t = q.f32(); transform2d(x, y, t, Z); }
            else { rb[0] = q.f32(); rb[1] = q.f32(); }
        } else for (uint32_t a = 0; a < cols; ++a) for (uint32_t b = 0; b < cols; ++b) Z[3 * a + b] = (double)q.f32();
        q.u32(); const uint32_t ci = q.u32();
        double w[3] = {1, 1, 1};                           // identity, diagonal overwritten (:124-142)
        for (uint32_t j = 0; j < ci; ++j) w[j] = (double)q.f32();
        hsum += mix(mix(mix((uint64_t)i << 32 | type << 16 | rows << 12 | cols << 8 | ci) ^ id1) ^ ((uint64_t)id2 << 1));
        const uint32_t a = m.find(id1), b = m.find(id2);
        int kind = 0;
        if (a == 0xFFFFFFFFu || b == 0xFFFFFFFFu) kind = 1;
        else if (type == 0 && ((a | b) & 0x80000000u)) kind = 2;
        else if (type == 1 && ((a & 0x80000000u) || !(b & 0x80000000u))) kind = 3;
        if (kind) {
#pragma omp critical(tsb_parse_error)
            if (i < bad) { bad = i; bad_kind = kind; }
            continue;
        }
        const size_t k = g.edge_slot[i];
        if (type == 0) {
            g.od_i[k] = a; g.od_j[k] = b;
            zinv6(Z, &g.od_zinv[6 * k]); g.od_w[3 * k] = w[0]; g.od_w[3 * k + 1] = w[1]; g.od_w[3 * k + 2] = w[2];
        } else {
            g.le_i[k] = a; g.le_l[k] = b & 0x7FFFFFFFu;
            double sb, cb; sincos(rb[1], &sb, &cb);          // same values as separate cos / sin calls (glibc)
            g.le_z[2 * k] = rb[0] * cb; g.le_z[2 * k + 1] = rb[0] * sb;
            g.le_w[2 * k] = w[0]; g.le_w[2 * k + 1] = w[1];
        }
    }
    if (bad != 0xFFFFFFFFu)
        return fail(bad_kind == 1 ? "edge references an unknown vertex id" : bad_kind == 2 ? "ODOM edge must connect two Se2 vertices" : "LM edge must be (Se2, Point2)");
    g.fixed_offset = (uint32_t)(r.p - buf);
    if (!r.have(4)) return fail("truncated payload: fixed count");
    const uint32_t F = r.u32();
    if (!r.have(4ull * F)) return fail("truncated payload: fixed ids");
    uint64_t h = mix(((uint64_t)V << 32) | E) ^ mix(F);
    for (uint32_t i = 0; i < F; ++i) {
        const uint32_t id = r.u32(), c = m.find(id);
        if (c == 0xFFFFFFFFu) return fail("fixed id is not a vertex");
        g.fixed.push_back((c & 0x80000000u) ? g.P + (c & 0x7FFFFFFFu) : c);   // FixVertex: push_back, repeats kept
        h = mix(h ^ id);
    }
    for (uint32_t i = 0; i < V; ++i) h += mix(((uint64_t)i << 33) ^ ((uint64_t)vtype[i] << 32) ^ vid[i]);
    g.topo_hash = h + hsum;
    g.request.resize(nbytes);
    {   // copy of the payload for the reply echo, in parallel slices
        const long long nsl = (long long)((nbytes + (8u << 20) - 1) / (8u << 20));
#pragma omp parallel for schedule(static)
        for (long long sl = 0; sl < nsl; ++sl) { const size_t o = (size_t)sl * (8u << 20); std::memcpy(&g.request[o], buf + o, std::min<size_t>(8u << 20, nbytes - o)); }
    }
    {   // the reply frame carries a signed 32-bit size: computed in 64 bits, saturated (tsb_load_wire rejects > INT32_MAX)
        const uint64_t rb = 4 + 20ull * g.P + 16ull * g.L + 4 + 76ull * nO + 44ull * nL + 4 + 4ull * F;
        g.reply_bytes = rb > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)rb;
    }
    lap("decode + copy + reply size");
    return true;
}

void encode_reply(const HostGraph& g, const double* pose_xyt, const double* lm_xy, uint8_t* out) {
    uint8_t* o = out;
    auto pu = [&](uint32_t v) { std::memcpy(o, &v, 4); o += 4; };
    auto pf = [&](float v) { std::memcpy(o, &v, 4); o += 4; };
    const uint8_t* in = g.request.data();
    Reader r{in, in + g.request.size()};
    const uint32_t V = r.u32();
    pu(g.P + g.L);
    for (uint32_t i = 0; i < V; ++i) {
        const uint32_t id = r.u32(), type = r.u32();
        r.p += type == 0 ? 12 : 8;
        const uint32_t c = g.req_vertex[i];
        if (c == 0xFFFFFFFFu) continue;
        pu(id);
        if (c & 0x80000000u) { const uint32_t l = c & 0x7FFFFFFFu; pu(1u); pf((float)lm_xy[2 * l]); pf((float)lm_xy[2 * l + 1]); }
        else {   // SerializeGraphFuncCpu.h:22-31: theta = atan2(R10, R00), i.e. wrapped to (-pi, pi]
            pu(0u); pf((float)pose_xyt[3 * c]); pf((float)pose_xyt[3 * c + 1]);
            const double t = pose_xyt[3 * c + 2]; pf((float)std::atan2(std::sin(t), std::cos(t)));
        }
    }
    const uint32_t E = r.u32(); pu(E);
    // edges are echoed in the reply shape (ODOM 76 B, LM 44 B): output offsets follow from the per-type indices
    uint8_t* const ebase = o;
#pragma omp parallel for schedule(static)
    for (long long ii = 0; ii < (long long)E; ++ii) {
        const uint32_t i = (uint32_t)ii;
        Reader q{in + g.edge_off[i], in + g.request.size()};
        const uint32_t type = q.u32(), id1 = q.u32(), id2 = q.u32(), rows = q.u32(), cols = q.u32();
        const size_t nod = type == 0 ? g.edge_slot[i] : i - g.edge_slot[i], nlm = i - nod;      // edges of each type before this one
        uint8_t* w = ebase + 76 * nod + 44 * nlm;
        auto wu = [&](uint32_t v) { std::memcpy(w, &v, 4); w += 4; };
        auto wf = [&](float v) { std::memcpy(w, &v, 4); w += 4; };
        wu(type); wu(id1); wu(id2);
        if (type == 0) {
            wu(3u); wu(3u);
            if (rows == 0) { const float x = q.f32(), y = q.f32(), t = q.f32(); double Z[9]; transform2d(x, y, t, Z);
                for (int k = 0; k < 9; ++k) wf((float)Z[k]); }
            else { float Z[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
                for (uint32_t a = 0; a < cols; ++a) for (uint32_t b = 0; b < cols; ++b) Z[3 * a + b] = q.f32();
                for (int k = 0; k < 9; ++k) wf(Z[k]); }
        } else { wu(2u); wu(1u); wf(q.f32()); wf(q.f32()); }   // Vector2 held as a 2 x 1 dynamic matrix
        q.u32(); const uint32_t ci = q.u32();
        const uint32_t nd = type == 0 ? 3u : 2u;
        wu(0u); wu(nd);
        for (uint32_t j = 0; j < nd; ++j) wf(j < ci ? q.f32() : 1.0f);
    }
    o = ebase + 76ull * g.Eo + 44ull * g.El;
    r.p = in + g.fixed_offset;
    const uint32_t F = r.u32(); pu(F);
    for (uint32_t i = 0; i < F; ++i) pu(r.u32());
}

}  // namespace tsb
