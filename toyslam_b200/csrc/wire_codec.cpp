// wire_codec.cpp -- bulk wire parser / reply encoder (host).
// Request layout: remote/serialization/DeserializeGraph.h:9-174 (what python/remote/graph_to_bytes.py:32-67
// sends); reply layout: remote/serialization/SerializeGraph.h:8-72 with the CPU policy
// (SerializeGraphFuncCpu.h:10-65).  Unlike the reference parser this one is bounds-checked and
// never throws: a malformed payload is a returned error.
#include "host_graph.h"

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
    g = HostGraph();
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

    g.edges_offset = (uint32_t)(r.p - buf);
    if (!r.have(4)) return fail("truncated payload: edge count");
    const uint32_t E = r.u32();
    if ((uint64_t)E * 44 > nbytes) return fail("edge count exceeds payload size");
    {   // no reallocation while the edges stream in (E bounds both kinds)
        const size_t cap = E;
        g.le_i.reserve(cap); g.le_l.reserve(cap); g.le_z.reserve(2 * cap); g.le_w.reserve(2 * cap);
        const size_t ocap = std::min<size_t>(cap, (size_t)g.P * 4 + 1024);
        g.od_i.reserve(ocap); g.od_j.reserve(ocap); g.od_zinv.reserve(6 * ocap); g.od_w.reserve(3 * ocap);
    }
    uint32_t reply = g.edges_offset + 4;     // bytes of the reply so far (vertex section is the same size
                                             // unless duplicates were dropped; fixed below)
    for (uint32_t i = 0; i < E; ++i) {
        if (!r.have(20)) return fail("truncated payload: edge header");
        const uint32_t type = r.u32(), id1 = r.u32(), id2 = r.u32(), rows = r.u32(), cols = r.u32();
        if (type > 1) return fail("NOT IMPLE INF edge type");
        double Z[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}; double rb[2] = {0, 0};
        if (rows == 0) {                                   // DeserializeGraph.h:74-95
            if (type == 0) { if (!r.have(12)) return fail("truncated payload: measurement");
                const float x = r.f32(), y = r.f32(), t = r.f32(); transform2d(x, y, t, Z); }
            else { if (!r.have(8)) return fail("truncated payload: measurement"); rb[0] = r.f32(); rb[1] = r.f32(); }
        } else {                                           // DeserializeGraph.h:97-116: cols x cols values
            if (type != 0) return fail("NOT IMPLE INF edge type");
            if (cols > 3) return fail("ODOM measurement larger than 3x3");
            if (!r.have(4ull * cols * cols)) return fail("truncated payload: measurement");
            for (uint32_t a = 0; a < cols; ++a) for (uint32_t b = 0; b < cols; ++b) Z[3 * a + b] = (double)r.f32();
        }
        if (!r.have(8)) return fail("truncated payload: information header");
        const uint32_t ri = r.u32(), ci = r.u32();
        if (ri != 0) return fail("NOT IMPLE INF NxN");
        const uint32_t nd = type == 0 ? 3u : 2u;
        if (ci > nd) return fail("information diagonal longer than the edge dimension");
        if (!r.have(4ull * ci)) return fail("truncated payload: information");
        double w[3] = {1, 1, 1};                           // identity, diagonal overwritten (:124-142)
        for (uint32_t j = 0; j < ci; ++j) w[j] = (double)r.f32();
        const uint32_t a = m.find(id1), b = m.find(id2);
        if (a == 0xFFFFFFFFu || b == 0xFFFFFFFFu) return fail("edge references an unknown vertex id");
        if (type == 0) {
            if ((a | b) & 0x80000000u) return fail("ODOM edge must connect two Se2 vertices");
            g.od_i.push_back(a); g.od_j.push_back(b);
            double zi[6]; zinv6(Z, zi);
            g.od_zinv.insert(g.od_zinv.end(), zi, zi + 6); g.od_w.insert(g.od_w.end(), w, w + 3);
            reply += 12 + 8 + 36 + 8 + 12;
        } else {
            if ((a & 0x80000000u) || !(b & 0x80000000u)) return fail("LM edge must be (Se2, Point2)");
            g.le_i.push_back(a); g.le_l.push_back(b & 0x7FFFFFFFu);
            double sb, cb; sincos(rb[1], &sb, &cb);          // same values as separate cos / sin calls (glibc)
            g.le_z.push_back(rb[0] * cb); g.le_z.push_back(rb[0] * sb);
            g.le_w.push_back(w[0]); g.le_w.push_back(w[1]);
            reply += 12 + 8 + 8 + 8 + 8;
        }
    }
    g.Eo = (uint32_t)g.od_i.size(); g.El = (uint32_t)g.le_i.size();
    if (!r.have(4)) return fail("truncated payload: fixed count");
    const uint32_t F = r.u32();
    if (!r.have(4ull * F)) return fail("truncated payload: fixed ids");
    for (uint32_t i = 0; i < F; ++i) {
        const uint32_t c = m.find(r.u32());
        if (c == 0xFFFFFFFFu) return fail("fixed id is not a vertex");
        g.fixed.push_back((c & 0x80000000u) ? g.P + (c & 0x7FFFFFFFu) : c);   // FixVertex: push_back, repeats kept
    }
    g.request.assign(buf, buf + nbytes);
    g.reply_bytes = 4 + 20 * g.P + 16 * g.L + (reply - g.edges_offset) + 4 + 4 * F;
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
    for (uint32_t i = 0; i < E; ++i) {
        const uint32_t type = r.u32(), id1 = r.u32(), id2 = r.u32(), rows = r.u32(), cols = r.u32();
        pu(type); pu(id1); pu(id2);
        if (type == 0) {
            pu(3u); pu(3u);
            if (rows == 0) { const float x = r.f32(), y = r.f32(), t = r.f32(); double Z[9]; transform2d(x, y, t, Z);
                for (int q = 0; q < 9; ++q) pf((float)Z[q]); }
            else { float Z[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
                for (uint32_t a = 0; a < cols; ++a) for (uint32_t b = 0; b < cols; ++b) Z[3 * a + b] = r.f32();
                for (int q = 0; q < 9; ++q) pf(Z[q]); }
        } else { pu(2u); pu(1u); pf(r.f32()); pf(r.f32()); }   // Vector2 held as a 2 x 1 dynamic matrix
        r.u32(); const uint32_t ci = r.u32();
        const uint32_t nd = type == 0 ? 3u : 2u;
        pu(0u); pu(nd);
        for (uint32_t j = 0; j < nd; ++j) pf(j < ci ? r.f32() : 1.0f);
    }
    const uint32_t F = r.u32(); pu(F);
    for (uint32_t i = 0; i < F; ++i) pu(r.u32());
}

}  // namespace tsb
