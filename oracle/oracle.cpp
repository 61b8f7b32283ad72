// oracle.cpp -- TEST INFRASTRUCTURE ONLY (not shipped, never on the product path).
//
// CPU restatement, in double precision, of ToySlam's `cpu eigen` Gauss-Newton pipeline.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may load this library.  The product (toyslam_b200/csrc) never links or calls it.
//
// Follows (paths relative to the reference checkout):
//   remote/optimizer/OptimizerCpu.h:25-183        GN loop, Huber, assembly, prior, exits
//   remote/graph/edge/EdgeSe2.h:24-38             ODOM residual, A=-I, B=+I
//   remote/graph/edge/EdgeSe2Point2d.h:28-70      LM residual and analytic Jacobians
//   remote/graph/Helper.h:6-31                    CreateTransform2d / ConvertMatToAngle / InverseTransform2d
//   remote/graph/vertex/VertexSe2.h:16-27, Vertex2d.h:16-19   updates
//   remote/graph/MatrixEigen.h:44-48,93-96        H += block, b -= segment
//   remote/solver/SolverEigen.h:20                colPivHouseholderQr().solve (Eigen 3.4.0, conanfile.txt:3)
//   remote/serialization/DeserializeGraph.h:9-174, SerializeGraph.h:8-72,
//   SerializeGraphFuncCpu.h:10-65, SerializeHelper.cpp:5-56    wire codec
//
// Parity status: the reference C++ cannot be compiled here (Eigen/Boost absent), and the
// reference has no tests or golden vectors.  This restatement is pinned against the
// reference's *Python* second implementation of the linearisation (python/optimizer/edges2d.py,
// graph_optimizer.py:94-156) and its wire encoder/decoder through tests/golden/ (see
// tools/make_golden.py).  The outer loop and the pivoted-QR solve are "parity unpinned"
// beyond that: no runnable reference artefact exists for them.
//
// Tiers:  T1 = dense H + column-pivoted Householder QR (literal algorithm);
//         T2 = same linearisation, block-sparse H, landmark Schur + block-Jacobi PCG
//              (timed CPU baseline where dense dim x dim is impossible).
// Vertex order (reference: unordered_map iteration, OptimizerCpu.h:62-68, a free
// permutation) is made canonical here: poses by ascending id, then landmarks by ascending id.

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

struct Pose { double c, s, x, y; };   // [[c,-s,x],[s,c,y],[0,0,1]]  (VertexSe2 position)

struct Graph {
    // canonical order
    std::vector<uint32_t> pose_id, lm_id;
    std::vector<Pose> pose;
    std::vector<double> lm;                 // 2 per landmark
    // edges in request order, split by type (GetEdges() is map<type, vector>)
    std::vector<uint32_t> od_i, od_j;       // canonical pose indices
    std::vector<double> od_Z;               // 9 per edge, row-major measurement
    std::vector<double> od_w;               // 3 per edge, information diagonal
    std::vector<uint32_t> lm_i, lm_l;       // pose index, landmark index
    std::vector<double> lm_meas;            // 2 per edge (range, bearing)
    std::vector<double> lm_w;               // 2 per edge
    std::vector<uint32_t> fixed_vertex;     // canonical vertex index: pose k -> k, landmark k -> P+k
    // request echo for the reply
    std::vector<uint8_t> request;
    std::vector<uint32_t> req_vertex_id, req_vertex_type;
    std::string error;
    // sparse-solver stats of the last run
    std::vector<int> pcg_iters;
    size_t P() const { return pose.size(); }
    size_t L() const { return lm_id.size(); }
    size_t dim() const { return 3 * P() + 2 * L(); }
};

// remote/graph/Helper.h:6-18  (AngleAxis(theta,Z)*AngleAxis(0,Y)*AngleAxis(0,X) -> quaternion -> matrix)
Pose create_transform2d(double x, double y, double theta) {
    const double w = std::cos(0.5 * theta), z = std::sin(0.5 * theta);
    const double tz = 2.0 * z, twz = tz * w, tzz = tz * z;
    Pose p;
    p.c = 1.0 - tzz;     // R(0,0) == R(1,1)
    p.s = twz;           // R(1,0);  R(0,1) = -twz
    p.x = x; p.y = y;
    return p;
}

// remote/graph/Helper.h:26-31  affine inverse: linear part by adj/det, translation -inv*t
void inverse_transform2d(const Pose& p, double out[6]) {
    const double a = p.c, b = -p.s, c = p.s, d = p.c;
    const double invdet = 1.0 / (a * d - b * c);
    const double i00 = d * invdet, i01 = -b * invdet, i10 = -c * invdet, i11 = a * invdet;
    out[0] = i00; out[1] = i01; out[2] = -(i00 * p.x + i01 * p.y);
    out[3] = i10; out[4] = i11; out[5] = -(i10 * p.x + i11 * p.y);
}

// general 3x3 inverse (EdgeSe2.h:32 `meas.inverse()`), adjugate / determinant
void inverse3x3(const double* m, double* o) {
    const double c00 = m[4] * m[8] - m[5] * m[7], c01 = m[5] * m[6] - m[3] * m[8], c02 = m[3] * m[7] - m[4] * m[6];
    const double det = m[0] * c00 + m[1] * c01 + m[2] * c02, id = 1.0 / det;
    o[0] = c00 * id; o[1] = (m[2] * m[7] - m[1] * m[8]) * id; o[2] = (m[1] * m[5] - m[2] * m[4]) * id;
    o[3] = c01 * id; o[4] = (m[0] * m[8] - m[2] * m[6]) * id; o[5] = (m[2] * m[3] - m[0] * m[5]) * id;
    o[6] = c02 * id; o[7] = (m[1] * m[6] - m[0] * m[7]) * id; o[8] = (m[0] * m[4] - m[1] * m[3]) * id;
}

// EdgeSe2.h:24-38.  e[3]; A = -I3, B = +I3 (constants)
void odom_error(const Graph& g, size_t k, double e[3]) {
    const Pose& p1 = g.pose[g.od_i[k]];
    const Pose& p2 = g.pose[g.od_j[k]];
    double inv[6];
    inverse_transform2d(p1, inv);
    // pp = posInv * pos_2  (3x3 affine product, last row 0 0 1)
    double pp[9];
    pp[0] = inv[0] * p2.c + inv[1] * p2.s;  pp[1] = inv[0] * (-p2.s) + inv[1] * p2.c;  pp[2] = inv[0] * p2.x + inv[1] * p2.y + inv[2];
    pp[3] = inv[3] * p2.c + inv[4] * p2.s;  pp[4] = inv[3] * (-p2.s) + inv[4] * p2.c;  pp[5] = inv[3] * p2.x + inv[4] * p2.y + inv[5];
    pp[6] = 0; pp[7] = 0; pp[8] = 1;
    double zi[9];
    inverse3x3(&g.od_Z[9 * k], zi);
    double d00 = zi[0] * pp[0] + zi[1] * pp[3] + zi[2] * pp[6];
    double d02 = zi[0] * pp[2] + zi[1] * pp[5] + zi[2] * pp[8];
    double d10 = zi[3] * pp[0] + zi[4] * pp[3] + zi[5] * pp[6];
    double d12 = zi[3] * pp[2] + zi[4] * pp[5] + zi[5] * pp[8];
    e[0] = d02; e[1] = d12; e[2] = std::atan2(d10, d00);
}

// EdgeSe2Point2d.h:28-70.  e[2], A[2x3] row-major, B[2x2] row-major
void lm_error(const Graph& g, size_t k, double e[2], double A[6], double B[4]) {
    const Pose& p = g.pose[g.lm_i[k]];
    const double lx = g.lm[2 * g.lm_l[k]], ly = g.lm[2 * g.lm_l[k] + 1];
    const double r = g.lm_meas[2 * k], beta = g.lm_meas[2 * k + 1];
    const double zx = r * std::cos(beta), zy = r * std::sin(beta);
    double inv[6];
    inverse_transform2d(p, inv);
    const double px = inv[0] * lx + inv[1] * ly + inv[2];
    const double py = inv[3] * lx + inv[4] * ly + inv[5];
    e[0] = px - zx; e[1] = py - zy;
    const double th = std::atan2(p.s, p.c);
    const double cA = std::cos(th), sA = std::sin(th);
    A[0] = -cA; A[1] = -sA; A[2] = cA * ly - sA * lx - cA * p.y + sA * p.x;
    A[3] = sA;  A[4] = -cA; A[5] = -sA * ly - cA * lx + sA * p.y + cA * p.x;
    B[0] = cA; B[1] = sA; B[2] = -sA; B[3] = cA;
}

// OptimizerCpu.h:36-46  (delta = 1.5f)
inline void robustify(double chi2, double& rho, double& w) {
    const double delta = 1.5, d2 = delta * delta;
    if (chi2 <= d2) { rho = chi2; w = 1.0; return; }
    const double sq = std::sqrt(chi2);
    rho = 2 * sq * delta - d2; w = delta / sq;
}

template <typename T> bool rd(const uint8_t*& p, const uint8_t* end, T& v) {
    if ((size_t)(end - p) < sizeof(T)) return false;
    std::memcpy(&v, p, sizeof(T)); p += sizeof(T); return true;
}

}  // namespace

// ------------------------------------------------------------------------------------------
// Wire parser: DeserializeGraph.h:9-174 with T=float on the wire, values widened to double.
// ------------------------------------------------------------------------------------------
namespace {

bool parse_wire(Graph& g, const uint8_t* buf, uint32_t nbytes) {
    const uint8_t* p = buf; const uint8_t* end = buf + nbytes;
    g.request.assign(buf, buf + nbytes);
    uint32_t V = 0;
    if (!rd(p, end, V)) { g.error = "truncated: vertex count"; return false; }
    struct RawV { uint32_t id, type; float v[3]; };
    std::vector<RawV> raw; raw.reserve(V);
    for (uint32_t i = 0; i < V; ++i) {
        RawV r{}; 
        if (!rd(p, end, r.id) || !rd(p, end, r.type)) { g.error = "truncated: vertex header"; return false; }
        const int n = r.type == 0 ? 3 : (r.type == 1 ? 2 : -1);
        if (n < 0) { g.error = "unknown vertex type"; return false; }
        for (int k = 0; k < n; ++k) if (!rd(p, end, r.v[k])) { g.error = "truncated: vertex"; return false; }
        raw.push_back(r);
        g.req_vertex_id.push_back(r.id); g.req_vertex_type.push_back(r.type);
    }
    // canonical order: poses by id, then landmarks by id.  A repeated id keeps the LAST
    // occurrence (AddVertex assigns vertices[id] = ..., GraphCpu.h).
    std::vector<uint32_t> order(raw.size());
    for (uint32_t i = 0; i < raw.size(); ++i) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return raw[a].id < raw[b].id; });
    std::vector<uint32_t> ids; std::vector<uint32_t> idx_of;  // sorted unique ids -> canonical vertex index
    std::vector<uint8_t> type_of;
    for (size_t k = 0; k < order.size(); ++k) {
        if (k + 1 < order.size() && raw[order[k + 1]].id == raw[order[k]].id) continue;
        const RawV& r = raw[order[k]];
        ids.push_back(r.id); type_of.push_back((uint8_t)r.type);
        if (r.type == 0) {
            idx_of.push_back((uint32_t)g.pose.size());
            g.pose_id.push_back(r.id);
            g.pose.push_back(create_transform2d((double)r.v[0], (double)r.v[1], (double)r.v[2]));
        } else {
            idx_of.push_back((uint32_t)g.lm_id.size());
            g.lm_id.push_back(r.id);
            g.lm.push_back((double)r.v[0]); g.lm.push_back((double)r.v[1]);
        }
    }
    auto find = [&](uint32_t id, uint8_t want, uint32_t& out) -> bool {
        auto it = std::lower_bound(ids.begin(), ids.end(), id);
        if (it == ids.end() || *it != id) return false;
        const size_t k = it - ids.begin();
        if (want != 255 && type_of[k] != want) return false;
        out = idx_of[k]; return true;
    };
    uint32_t E = 0;
    if (!rd(p, end, E)) { g.error = "truncated: edge count"; return false; }
    for (uint32_t i = 0; i < E; ++i) {
        uint32_t type, id1, id2, rows, cols;
        if (!rd(p, end, type) || !rd(p, end, id1) || !rd(p, end, id2) || !rd(p, end, rows) || !rd(p, end, cols)) {
            g.error = "truncated: edge header"; return false; }
        double Z[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}; double m2[2] = {0, 0};
        if (rows == 0) {                               // DeserializeGraph.h:74-95
            if (type == 0) {
                float x, y, t;
                if (!rd(p, end, x) || !rd(p, end, y) || !rd(p, end, t)) { g.error = "truncated: meas"; return false; }
                Pose z = create_transform2d(x, y, t);
                Z[0] = z.c; Z[1] = -z.s; Z[2] = z.x; Z[3] = z.s; Z[4] = z.c; Z[5] = z.y;
            } else if (type == 1) {
                float a, b;
                if (!rd(p, end, a) || !rd(p, end, b)) { g.error = "truncated: meas"; return false; }
                m2[0] = a; m2[1] = b;
            } else { g.error = "NOT IMPLE INF edge type"; return false; }
        } else {                                       // DeserializeGraph.h:97-116 (cols x cols values)
            if (type != 0) { g.error = "NOT IMPLE INF edge type"; return false; }
            if (cols > 3) { g.error = "measurement larger than 3x3"; return false; }
            for (uint32_t r = 0; r < cols; ++r) for (uint32_t c = 0; c < cols; ++c) {
                float v; if (!rd(p, end, v)) { g.error = "truncated: meas"; return false; }
                Z[3 * r + c] = v;
            }
        }
        uint32_t ri, ci;
        if (!rd(p, end, ri) || !rd(p, end, ci)) { g.error = "truncated: info header"; return false; }
        if (ri != 0) { g.error = "NOT IMPLE INF NxN"; return false; }
        const uint32_t nd = type == 0 ? 3u : 2u;
        if (type > 1) { g.error = "NOT IMPLE INF edge type"; return false; }
        if (ci > nd) { g.error = "information diagonal too long"; return false; }
        double w[3] = {1, 1, 1};                       // identity with the diagonal overwritten (:124-142)
        for (uint32_t j = 0; j < ci; ++j) { float v; if (!rd(p, end, v)) { g.error = "truncated: info"; return false; } w[j] = v; }
        uint32_t a, b;
        if (type == 0) {
            if (!find(id1, 0, a) || !find(id2, 0, b)) { g.error = "ODOM edge references a missing/non-Se2 vertex"; return false; }
            g.od_i.push_back(a); g.od_j.push_back(b);
            g.od_Z.insert(g.od_Z.end(), Z, Z + 9); g.od_w.insert(g.od_w.end(), w, w + 3);
        } else {
            if (!find(id1, 0, a) || !find(id2, 1, b)) { g.error = "LM edge must be (Se2, Point2)"; return false; }
            g.lm_i.push_back(a); g.lm_l.push_back(b);
            g.lm_meas.push_back(m2[0]); g.lm_meas.push_back(m2[1]);
            g.lm_w.push_back(w[0]); g.lm_w.push_back(w[1]);
        }
    }
    uint32_t F = 0;
    if (!rd(p, end, F)) { g.error = "truncated: fixed count"; return false; }
    for (uint32_t i = 0; i < F; ++i) {
        uint32_t id, k;
        if (!rd(p, end, id)) { g.error = "truncated: fixed id"; return false; }
        auto it = std::lower_bound(ids.begin(), ids.end(), id);
        if (it == ids.end() || *it != id) { g.error = "fixed id not a vertex"; return false; }
        k = (uint32_t)(it - ids.begin());
        const uint32_t vi = type_of[k] == 0 ? idx_of[k] : (uint32_t)g.P() + idx_of[k];
        // FixVertex is a vector push_back (GraphCpu.h:25-28): a repeated id adds the prior again
        g.fixed_vertex.push_back(vi);
    }
    return true;
}

}  // namespace

// ------------------------------------------------------------------------------------------
// T1: dense assembly (OptimizerCpu.h:80-138) and column-pivoted Householder QR (SolverEigen.h:20)
// ------------------------------------------------------------------------------------------
namespace {

// H row-major dim x dim, b dim.  Returns the robustified error sum (`err`, OptimizerCpu.h:117).
// Edge order: map<EdgeType, vector> iteration = all ODOM (type 0) then all LM (type 1), each in
// insertion order (OptimizerCpu.h:121-129).
double assemble_dense(const Graph& g, double* H, double* b) {
    const size_t n = g.dim(), P = g.P();
    std::fill(H, H + n * n, 0.0); std::fill(b, b + n, 0.0);
    double err = 0;
    for (size_t k = 0; k < g.od_i.size(); ++k) {
        double e[3]; odom_error(g, k, e);
        const double* w = &g.od_w[3 * k];
        const double chi2 = e[0] * (w[0] * e[0]) + e[1] * (w[1] * e[1]) + e[2] * (w[2] * e[2]);
        double rho, hw; robustify(chi2, rho, hw);
        const size_t i = 3 * g.od_i[k], j = 3 * g.od_j[k];
        for (int d = 0; d < 3; ++d) {
            const double inf = w[d] * hw, infw = inf * e[d];
            H[(i + d) * n + i + d] += inf;      // A^T INF A,  A = -I
            H[(j + d) * n + j + d] += inf;      // B^T INF B
            H[(i + d) * n + j + d] += -inf;     // A^T INF B
            H[(j + d) * n + i + d] += -inf;     // B^T INF A
            b[i + d] -= -infw;                  // b.segment -= A^T INF_W
            b[j + d] -= infw;
        }
        err += rho;
    }
    for (size_t k = 0; k < g.lm_i.size(); ++k) {
        double e[2], A[6], B[4]; lm_error(g, k, e, A, B);
        const double* w = &g.lm_w[2 * k];
        const double chi2 = e[0] * (w[0] * e[0]) + e[1] * (w[1] * e[1]);
        double rho, hw; robustify(chi2, rho, hw);
        const double inf[2] = {w[0] * hw, w[1] * hw};
        const double infw[2] = {inf[0] * e[0], inf[1] * e[1]};
        const size_t i = 3 * g.lm_i[k], j = 3 * P + 2 * g.lm_l[k];
        for (int r = 0; r < 3; ++r) {
            for (int c = 0; c < 3; ++c) H[(i + r) * n + i + c] += A[r] * inf[0] * A[c] + A[3 + r] * inf[1] * A[3 + c];
            for (int c = 0; c < 2; ++c) {
                const double v = A[r] * inf[0] * B[c] + A[3 + r] * inf[1] * B[2 + c];
                H[(i + r) * n + j + c] += v;
            }
            b[i + r] -= A[r] * infw[0] + A[3 + r] * infw[1];
        }
        for (int r = 0; r < 2; ++r) {
            for (int c = 0; c < 2; ++c) H[(j + r) * n + j + c] += B[r] * inf[0] * B[c] + B[2 + r] * inf[1] * B[2 + c];
            for (int c = 0; c < 3; ++c) H[(j + r) * n + i + c] += B[r] * inf[0] * A[c] + B[2 + r] * inf[1] * A[3 + c];
            b[j + r] -= B[r] * infw[0] + B[2 + r] * infw[1];
        }
        err += rho;
    }
    for (uint32_t v : g.fixed_vertex) {                 // OptimizerCpu.h:132-138
        const size_t off = v < P ? 3 * (size_t)v : 3 * P + 2 * (size_t)(v - P);
        const int sz = v < P ? 3 : 2;
        for (int d = 0; d < sz; ++d) H[(off + d) * n + off + d] += 1e6;
    }
    return err;
}

// Column-pivoted Householder QR solve, following Eigen 3.4.0 ColPivHouseholderQR
// (computeInPlace + _solve_impl): pivot = column of largest remaining squared norm with
// norm down-dating, Householder vectors with beta = -sign(c0)*norm, rank from
// |R_kk| > eps*n*max|R_ii|, solve on the leading rank x rank triangle, zero the rest.
// A is column-major n x n (H is symmetric so the transpose is immaterial); x overwrites b.
void colpiv_qr_solve(std::vector<double>& A, double* b, size_t n) {
    std::vector<size_t> perm(n); for (size_t i = 0; i < n; ++i) perm[i] = i;
    std::vector<double> cn_upd(n), cn_dir(n), hcoef(n, 0.0), tmp(n);
    for (size_t j = 0; j < n; ++j) {
        double s = 0; const double* col = &A[j * n];
        for (size_t i = 0; i < n; ++i) s += col[i] * col[i];
        cn_dir[j] = cn_upd[j] = std::sqrt(s);
    }
    const double eps = 2.220446049250313e-16;
    double maxnorm = 0; for (size_t j = 0; j < n; ++j) maxnorm = std::max(maxnorm, cn_upd[j]);
    const double thr_helper = (maxnorm * eps) * (maxnorm * eps) / (double)n;
    const double norm_downdate_thr = std::sqrt(eps);
    size_t nonzero_pivots = n; double maxpivot = 0;
    for (size_t k = 0; k < n; ++k) {
        size_t best = k; double bv = cn_upd[k];
        for (size_t j = k + 1; j < n; ++j) if (cn_upd[j] > bv) { bv = cn_upd[j]; best = j; }
        if (nonzero_pivots == n && bv * bv < thr_helper * (double)(n - k)) nonzero_pivots = k;
        if (best != k) {
            std::swap_ranges(&A[k * n], &A[k * n] + n, &A[best * n]);
            std::swap(cn_upd[k], cn_upd[best]); std::swap(cn_dir[k], cn_dir[best]); std::swap(perm[k], perm[best]);
        }
        // Householder on A[k:n, k]
        double* col = &A[k * n];
        double tail = 0; for (size_t i = k + 1; i < n; ++i) tail += col[i] * col[i];
        const double c0 = col[k];
        double beta, tau;
        if (tail <= std::numeric_limits<double>::min()) { tau = 0; beta = c0; for (size_t i = k + 1; i < n; ++i) col[i] = 0; }
        else {
            beta = std::sqrt(c0 * c0 + tail); if (c0 >= 0) beta = -beta;
            const double inv = 1.0 / (c0 - beta);
            for (size_t i = k + 1; i < n; ++i) col[i] *= inv;
            tau = (beta - c0) / beta;
        }
        col[k] = beta; hcoef[k] = tau;
        if (std::abs(beta) > maxpivot) maxpivot = std::abs(beta);
        if (tau != 0) {
            for (size_t j = k + 1; j < n; ++j) {           // apply H_k to the trailing columns
                double* cj = &A[j * n];
                double s = cj[k]; for (size_t i = k + 1; i < n; ++i) s += col[i] * cj[i];
                s *= tau; cj[k] -= s; for (size_t i = k + 1; i < n; ++i) cj[i] -= s * col[i];
            }
            double s = b[k]; for (size_t i = k + 1; i < n; ++i) s += col[i] * b[i];   // Q^T b on the fly
            s *= tau; b[k] -= s; for (size_t i = k + 1; i < n; ++i) b[i] -= s * col[i];
        }
        for (size_t j = k + 1; j < n; ++j) {               // norm down-dating (LAPACK WN 176)
            if (cn_upd[j] != 0) {
                double t = std::abs(A[j * n + k]) / cn_upd[j];
                t = (1 + t) * (1 - t); if (t < 0) t = 0;
                const double r = cn_upd[j] / cn_dir[j];
                const double t2 = t * r * r;
                if (t2 <= norm_downdate_thr) {
                    double s = 0; for (size_t i = k + 1; i < n; ++i) s += A[j * n + i] * A[j * n + i];
                    cn_dir[j] = cn_upd[j] = std::sqrt(s);
                } else cn_upd[j] *= std::sqrt(t);
            }
        }
    }
    // rank with Eigen's default threshold eps * diagonalSize
    size_t rank = 0; const double thr = maxpivot * eps * (double)n;
    for (size_t k = 0; k < nonzero_pivots; ++k) if (std::abs(A[k * n + k]) > thr) ++rank;
    for (size_t kk = rank; kk-- > 0;) {                     // back substitution on R[0:rank,0:rank]
        double s = b[kk];
        for (size_t j = kk + 1; j < rank; ++j) s -= A[j * n + kk] * tmp[j];
        tmp[kk] = s / A[kk * n + kk];
    }
    for (size_t k = rank; k < n; ++k) tmp[k] = 0;
    for (size_t k = 0; k < n; ++k) b[perm[k]] = tmp[k];
}

void update_vertices(Graph& g, const double* delta, double step) {    // OptimizerCpu.h:159-165
    const size_t P = g.P(), L = g.L();
    for (size_t i = 0; i < P; ++i) {                                   // VertexSe2.h:16-27
        Pose& p = g.pose[i];
        const double th = std::atan2(p.s, p.c) + delta[3 * i + 2] * step;
        p.c = std::cos(th); p.s = std::sin(th);
        p.x += delta[3 * i] * step; p.y += delta[3 * i + 1] * step;
    }
    for (size_t l = 0; l < L; ++l) {                                   // Vertex2d.h:16-19
        g.lm[2 * l] += delta[3 * P + 2 * l] * step; g.lm[2 * l + 1] += delta[3 * P + 2 * l + 1] * step;
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------
// T2: block-sparse restatement (same linearisation / Huber / prior / step / exits), landmark
// Schur complement + block-Jacobi PCG on the pose system.  Not in the reference (its H is
// dense); it exists to time a CPU baseline at sizes the dense form cannot reach and
// is itself checked against T1 in tests/.
// ------------------------------------------------------------------------------------------
namespace {

struct Sparse {
    size_t P = 0, L = 0, Eo = 0, El = 0;
    std::vector<uint32_t> pose_ptr, pose_edge;     // LM edges grouped by pose (edge order kept inside a pose)
    std::vector<uint32_t> lm_ptr, lm_edge;         // LM edges grouped by landmark
    std::vector<uint32_t> oo_ptr, oo_edge, oi_ptr, oi_edge;   // ODOM edges by id1 / by id2
    std::vector<double> Hpl;      // 6 per LM edge, 3x2 row-major  (A^T INF B)
    std::vector<double> eA;       // 9 per LM edge: A^T INF A (6 unique: 00 01 02 11 12 22) + A^T INF e (3)
    std::vector<double> eB;       // 5 per LM edge: B^T INF B (00 01 11) + B^T INF e (2)
    std::vector<double> rho_l, rho_o;
    std::vector<double> od_inf;   // 3 per ODOM edge: hw*w   (H_ii += inf, H_jj += inf, H_ij = -inf)
    std::vector<double> od_infw;  // 3 per ODOM edge: inf*e
    std::vector<double> Hpp, bp;  // 9 / 3 per pose
    std::vector<double> Hll, bl;  // 4 / 2 per landmark, then Hll holds the inverse
    std::vector<double> Minv;     // 9 per pose, inverse of diag block of S
    std::vector<double> rhs, x, r, z, p, Sp, w, dl;
};

void csr_group(const std::vector<uint32_t>& key, size_t nkeys, std::vector<uint32_t>& ptr, std::vector<uint32_t>& idx) {
    ptr.assign(nkeys + 1, 0);
    for (uint32_t k : key) ptr[k + 1]++;
    for (size_t i = 0; i < nkeys; ++i) ptr[i + 1] += ptr[i];
    idx.resize(key.size());
    std::vector<uint32_t> cur(ptr.begin(), ptr.end() - 1);
    for (uint32_t e = 0; e < key.size(); ++e) idx[cur[key[e]]++] = e;
}

void sparse_init(const Graph& g, Sparse& s) {
    s.P = g.P(); s.L = g.L(); s.Eo = g.od_i.size(); s.El = g.lm_i.size();
    csr_group(g.lm_i, s.P, s.pose_ptr, s.pose_edge);
    csr_group(g.lm_l, s.L, s.lm_ptr, s.lm_edge);
    csr_group(g.od_i, s.P, s.oo_ptr, s.oo_edge);
    csr_group(g.od_j, s.P, s.oi_ptr, s.oi_edge);
    s.Hpl.resize(6 * s.El); s.eA.resize(9 * s.El); s.eB.resize(5 * s.El); s.rho_l.resize(s.El); s.rho_o.resize(s.Eo);
    s.od_inf.resize(3 * s.Eo); s.od_infw.resize(3 * s.Eo);
    s.Hpp.resize(9 * s.P); s.bp.resize(3 * s.P); s.Hll.resize(4 * s.L); s.bl.resize(2 * s.L); s.Minv.resize(9 * s.P);
    for (auto* v : {&s.rhs, &s.x, &s.r, &s.z, &s.p, &s.Sp}) v->assign(3 * s.P, 0.0);
    s.w.assign(2 * s.L, 0.0); s.dl.assign(2 * s.L, 0.0);
}

bool inv3(const double* m, double* o) { inverse3x3(m, o); return std::isfinite(o[0]); }

double sparse_linearize(const Graph& g, Sparse& s) {
    const long El = (long)s.El, Eo = (long)s.Eo, P = (long)s.P, L = (long)s.L;
#pragma omp parallel for schedule(static)
    for (long k = 0; k < El; ++k) {
        double e[2], A[6], B[4]; lm_error(g, k, e, A, B);
        const double* w = &g.lm_w[2 * k];
        const double chi2 = e[0] * (w[0] * e[0]) + e[1] * (w[1] * e[1]);
        double rho, hw; robustify(chi2, rho, hw);
        const double i0 = w[0] * hw, i1 = w[1] * hw, f0 = i0 * e[0], f1 = i1 * e[1];
        double* h = &s.Hpl[6 * k];
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 2; ++c) h[2 * r + c] = A[r] * i0 * B[c] + A[3 + r] * i1 * B[2 + c];
        double* a = &s.eA[9 * k]; int q = 0;
        for (int r = 0; r < 3; ++r) for (int c = r; c < 3; ++c) a[q++] = A[r] * i0 * A[c] + A[3 + r] * i1 * A[3 + c];
        for (int r = 0; r < 3; ++r) a[6 + r] = A[r] * f0 + A[3 + r] * f1;
        double* bb = &s.eB[5 * k];
        bb[0] = B[0] * i0 * B[0] + B[2] * i1 * B[2]; bb[1] = B[0] * i0 * B[1] + B[2] * i1 * B[3]; bb[2] = B[1] * i0 * B[1] + B[3] * i1 * B[3];
        bb[3] = B[0] * f0 + B[2] * f1; bb[4] = B[1] * f0 + B[3] * f1;
        s.rho_l[k] = rho;
    }
#pragma omp parallel for schedule(static)
    for (long k = 0; k < Eo; ++k) {
        double e[3]; odom_error(g, k, e);
        const double* w = &g.od_w[3 * k];
        const double chi2 = e[0] * (w[0] * e[0]) + e[1] * (w[1] * e[1]) + e[2] * (w[2] * e[2]);
        double rho, hw; robustify(chi2, rho, hw);
        for (int d = 0; d < 3; ++d) { s.od_inf[3 * k + d] = w[d] * hw; s.od_infw[3 * k + d] = w[d] * hw * e[d]; }
        s.rho_o[k] = rho;
    }
    std::vector<char> fixed(P + L, 0); std::vector<int> nfix(P + L, 0);
    for (uint32_t v : g.fixed_vertex) nfix[v]++;
#pragma omp parallel for schedule(static)
    for (long i = 0; i < P; ++i) {
        double H[9] = {0}, b[3] = {0};
        for (uint32_t q = s.oo_ptr[i]; q < s.oo_ptr[i + 1]; ++q) { const uint32_t k = s.oo_edge[q];
            for (int d = 0; d < 3; ++d) { H[4 * d] += s.od_inf[3 * k + d]; b[d] += s.od_infw[3 * k + d]; } }
        for (uint32_t q = s.oi_ptr[i]; q < s.oi_ptr[i + 1]; ++q) { const uint32_t k = s.oi_edge[q];
            for (int d = 0; d < 3; ++d) { H[4 * d] += s.od_inf[3 * k + d]; b[d] -= s.od_infw[3 * k + d]; } }
        for (uint32_t q = s.pose_ptr[i]; q < s.pose_ptr[i + 1]; ++q) { const double* a = &s.eA[9 * (size_t)s.pose_edge[q]];
            H[0] += a[0]; H[1] += a[1]; H[2] += a[2]; H[4] += a[3]; H[5] += a[4]; H[8] += a[5];
            b[0] -= a[6]; b[1] -= a[7]; b[2] -= a[8]; }
        H[3] = H[1]; H[6] = H[2]; H[7] = H[5];
        for (int d = 0; d < 3; ++d) H[4 * d] += 1e6 * nfix[i];
        std::copy(H, H + 9, &s.Hpp[9 * i]); std::copy(b, b + 3, &s.bp[3 * i]);
    }
#pragma omp parallel for schedule(static)
    for (long l = 0; l < L; ++l) {
        double h0 = 0, h1 = 0, h2 = 0, b0 = 0, b1 = 0;
        for (uint32_t q = s.lm_ptr[l]; q < s.lm_ptr[l + 1]; ++q) { const double* bb = &s.eB[5 * (size_t)s.lm_edge[q]];
            h0 += bb[0]; h1 += bb[1]; h2 += bb[2]; b0 -= bb[3]; b1 -= bb[4]; }
        h0 += 1e6 * nfix[P + l]; h2 += 1e6 * nfix[P + l];
        s.Hll[4 * l] = h0; s.Hll[4 * l + 1] = h1; s.Hll[4 * l + 2] = h1; s.Hll[4 * l + 3] = h2;
        s.bl[2 * l] = b0; s.bl[2 * l + 1] = b1;
    }
    double err = 0;
    for (long k = 0; k < Eo; ++k) err += s.rho_o[k];
    for (long k = 0; k < El; ++k) err += s.rho_l[k];
    return err;
}

}  // namespace

namespace {

// S*v = Hpp*v - Hpl * Hll^-1 * Hpl^T * v   (matrix-free; Hll already inverted)
void schur_apply(const Graph& g, Sparse& s, const double* v, double* out) {
    const long P = (long)s.P, L = (long)s.L;
#pragma omp parallel for schedule(static)
    for (long l = 0; l < L; ++l) {
        double t0 = 0, t1 = 0;
        for (uint32_t q = s.lm_ptr[l]; q < s.lm_ptr[l + 1]; ++q) { const uint32_t k = s.lm_edge[q];
            const double* h = &s.Hpl[6 * (size_t)k]; const double* pv = &v[3 * (size_t)g.lm_i[k]];
            t0 += h[0] * pv[0] + h[2] * pv[1] + h[4] * pv[2]; t1 += h[1] * pv[0] + h[3] * pv[1] + h[5] * pv[2]; }
        const double* hi = &s.Hll[4 * l];
        s.w[2 * l] = hi[0] * t0 + hi[1] * t1; s.w[2 * l + 1] = hi[2] * t0 + hi[3] * t1;
    }
#pragma omp parallel for schedule(static)
    for (long i = 0; i < P; ++i) {
        const double* H = &s.Hpp[9 * i]; const double* pv = &v[3 * i];
        double y[3];
        for (int r = 0; r < 3; ++r) y[r] = H[3 * r] * pv[0] + H[3 * r + 1] * pv[1] + H[3 * r + 2] * pv[2];
        for (uint32_t q = s.oo_ptr[i]; q < s.oo_ptr[i + 1]; ++q) { const uint32_t k = s.oo_edge[q];
            const double* pj = &v[3 * (size_t)g.od_j[k]]; for (int d = 0; d < 3; ++d) y[d] -= s.od_inf[3 * k + d] * pj[d]; }
        for (uint32_t q = s.oi_ptr[i]; q < s.oi_ptr[i + 1]; ++q) { const uint32_t k = s.oi_edge[q];
            const double* pj = &v[3 * (size_t)g.od_i[k]]; for (int d = 0; d < 3; ++d) y[d] -= s.od_inf[3 * k + d] * pj[d]; }
        for (uint32_t q = s.pose_ptr[i]; q < s.pose_ptr[i + 1]; ++q) { const uint32_t k = s.pose_edge[q];
            const double* h = &s.Hpl[6 * (size_t)k]; const double* u = &s.w[2 * (size_t)g.lm_l[k]];
            y[0] -= h[0] * u[0] + h[1] * u[1]; y[1] -= h[2] * u[0] + h[3] * u[1]; y[2] -= h[4] * u[0] + h[5] * u[1]; }
        out[3 * i] = y[0]; out[3 * i + 1] = y[1]; out[3 * i + 2] = y[2];
    }
}

double dotv(const std::vector<double>& a, const std::vector<double>& b) {
    double s = 0; const long n = (long)a.size();
#pragma omp parallel for reduction(+ : s) schedule(static)
    for (long i = 0; i < n; ++i) s += a[i] * b[i];
    return s;
}

// Solves H*delta = b by landmark elimination + PCG.  delta: 3P pose entries then 2L landmark entries.
int sparse_solve(const Graph& g, Sparse& s, double tol, int max_it, double* delta) {
    const long P = (long)s.P, L = (long)s.L;
#pragma omp parallel for schedule(static)
    for (long l = 0; l < L; ++l) {                       // Hll <- Hll^-1 (2x2 closed form)
        double* h = &s.Hll[4 * l]; const double id = 1.0 / (h[0] * h[3] - h[1] * h[2]);
        const double a = h[3] * id, b = -h[1] * id, c = -h[2] * id, d = h[0] * id;
        h[0] = a; h[1] = b; h[2] = c; h[3] = d;
    }
#pragma omp parallel for schedule(static)
    for (long i = 0; i < P; ++i) {                       // reduced rhs and diag(S)^-1
        double D[9]; std::copy(&s.Hpp[9 * i], &s.Hpp[9 * i] + 9, D);
        double rb[3] = {s.bp[3 * i], s.bp[3 * i + 1], s.bp[3 * i + 2]};
        for (uint32_t q = s.pose_ptr[i]; q < s.pose_ptr[i + 1]; ++q) { const uint32_t k = s.pose_edge[q];
            const double* h = &s.Hpl[6 * (size_t)k]; const size_t l = g.lm_l[k]; const double* hi = &s.Hll[4 * l];
            double T[6];                                  // T = Hpl * Hll^-1   (3x2)
            for (int r = 0; r < 3; ++r) { T[2 * r] = h[2 * r] * hi[0] + h[2 * r + 1] * hi[2]; T[2 * r + 1] = h[2 * r] * hi[1] + h[2 * r + 1] * hi[3]; }
            for (int r = 0; r < 3; ++r) { rb[r] -= T[2 * r] * s.bl[2 * l] + T[2 * r + 1] * s.bl[2 * l + 1];
                for (int c = 0; c < 3; ++c) D[3 * r + c] -= T[2 * r] * h[2 * c] + T[2 * r + 1] * h[2 * c + 1]; }
        }
        inverse3x3(D, &s.Minv[9 * i]);
        s.rhs[3 * i] = rb[0]; s.rhs[3 * i + 1] = rb[1]; s.rhs[3 * i + 2] = rb[2];
    }
    auto precond = [&](const std::vector<double>& r, std::vector<double>& z) {
#pragma omp parallel for schedule(static)
        for (long i = 0; i < P; ++i) { const double* M = &s.Minv[9 * i];
            for (int a = 0; a < 3; ++a) z[3 * i + a] = M[3 * a] * r[3 * i] + M[3 * a + 1] * r[3 * i + 1] + M[3 * a + 2] * r[3 * i + 2]; }
    };
    std::fill(s.x.begin(), s.x.end(), 0.0); s.r = s.rhs;
    precond(s.r, s.z); s.p = s.z;
    double rz = dotv(s.r, s.z); const double rz0 = rz;
    int it = 0;
    while (it < max_it && rz > tol * tol * rz0 && rz0 > 0) {
        schur_apply(g, s, s.p.data(), s.Sp.data());
        const double alpha = rz / dotv(s.p, s.Sp);
#pragma omp parallel for schedule(static)
        for (long i = 0; i < 3 * P; ++i) { s.x[i] += alpha * s.p[i]; s.r[i] -= alpha * s.Sp[i]; }
        precond(s.r, s.z);
        const double rz_new = dotv(s.r, s.z), beta = rz_new / rz; rz = rz_new;
#pragma omp parallel for schedule(static)
        for (long i = 0; i < 3 * P; ++i) s.p[i] = s.z[i] + beta * s.p[i];
        ++it;
    }
    std::copy(s.x.begin(), s.x.end(), delta);
#pragma omp parallel for schedule(static)
    for (long l = 0; l < L; ++l) {                       // delta_l = Hll^-1 (b_l - Hpl^T delta_p)
        double t0 = s.bl[2 * l], t1 = s.bl[2 * l + 1];
        for (uint32_t q = s.lm_ptr[l]; q < s.lm_ptr[l + 1]; ++q) { const uint32_t k = s.lm_edge[q];
            const double* h = &s.Hpl[6 * (size_t)k]; const double* pv = &s.x[3 * (size_t)g.lm_i[k]];
            t0 -= h[0] * pv[0] + h[2] * pv[1] + h[4] * pv[2]; t1 -= h[1] * pv[0] + h[3] * pv[1] + h[5] * pv[2]; }
        const double* hi = &s.Hll[4 * l];
        delta[3 * P + 2 * l] = hi[0] * t0 + hi[1] * t1; delta[3 * P + 2 * l + 1] = hi[2] * t0 + hi[3] * t1;
    }
    return it;
}

// The GN loop, OptimizerCpu.h:70-182.  mode 0 = dense QR (T1), 1 = sparse Schur+PCG (T2).
// Returns the number of chi2 entries written (one per iteration entered).
int gn_loop(Graph& g, int mode, int iterations, int enable_exits, double tol, int max_pcg,
            double* chi2_traj, int* exit_code) {
    const size_t n = g.dim();
    std::vector<double> H, Hc, b(n);
    Sparse s;
    if (mode == 0) { H.resize(n * n); Hc.resize(n * n); } else sparse_init(g, s);
    g.pcg_iters.clear();
    double prevErr = -1, err = 0; int penalty = 0, done = 0; *exit_code = 0;
    for (int it = 0; it < iterations; ++it) {
        err = mode == 0 ? assemble_dense(g, H.data(), b.data()) : sparse_linearize(g, s);
        chi2_traj[done++] = err;
        if (prevErr > 0 && err > prevErr) {
            ++penalty;
            if (enable_exits && penalty > 2) { *exit_code = 1; break; }
        } else penalty = 0;
        if (mode == 0) {
            for (size_t r = 0; r < n; ++r) for (size_t c = 0; c < n; ++c) Hc[c * n + r] = H[r * n + c];
            colpiv_qr_solve(Hc, b.data(), n);
        } else g.pcg_iters.push_back(sparse_solve(g, s, tol, max_pcg, b.data()));
        update_vertices(g, b.data(), 0.2);
        if (enable_exits && std::abs(err - prevErr) < 0.001) { *exit_code = 2; break; }
        double nn = 0; for (size_t i = 0; i < n; ++i) nn += b[i] * b[i];
        if (enable_exits && std::sqrt(nn) < 0.001) { *exit_code = 3; break; }
        prevErr = err;
    }
    return done;
}

}  // namespace

// ------------------------------------------------------------------------------------------
// Reply encoder: SerializeGraph.h:8-72 + SerializeGraphFuncCpu.h:10-65 (payload without the
// leading size prefix).  Vertex order: first appearance in the request (the reference's
// unordered_map order is unspecified; the client re-keys by id).  Edges: ODOM then LM.
// ------------------------------------------------------------------------------------------
namespace {
template <typename T> void put(std::vector<uint8_t>& o, T v) { const uint8_t* p = (const uint8_t*)&v; o.insert(o.end(), p, p + sizeof(T)); }

std::vector<uint8_t> encode_reply(const Graph& g) {
    std::vector<uint8_t> o;
    put<uint32_t>(o, (uint32_t)(g.P() + g.L()));
    std::vector<char> seen_p(g.P(), 0), seen_l(g.L(), 0);
    for (size_t k = 0; k < g.req_vertex_id.size(); ++k) {
        const uint32_t id = g.req_vertex_id[k];
        // the surviving type for a repeated id is the last occurrence's; look the id up in both tables
        auto ip = std::lower_bound(g.pose_id.begin(), g.pose_id.end(), id);
        if (ip != g.pose_id.end() && *ip == id) {
            const size_t i = ip - g.pose_id.begin(); if (seen_p[i]) continue; seen_p[i] = 1;
            put<uint32_t>(o, id); put<uint32_t>(o, 0u);
            put<float>(o, (float)g.pose[i].x); put<float>(o, (float)g.pose[i].y);
            put<float>(o, (float)std::atan2(g.pose[i].s, g.pose[i].c));
            continue;
        }
        auto il = std::lower_bound(g.lm_id.begin(), g.lm_id.end(), id);
        const size_t l = il - g.lm_id.begin(); if (seen_l[l]) continue; seen_l[l] = 1;
        put<uint32_t>(o, id); put<uint32_t>(o, 1u);
        put<float>(o, (float)g.lm[2 * l]); put<float>(o, (float)g.lm[2 * l + 1]);
    }
    put<uint32_t>(o, (uint32_t)(g.od_i.size() + g.lm_i.size()));
    for (size_t k = 0; k < g.od_i.size(); ++k) {
        put<uint32_t>(o, 0u); put<uint32_t>(o, g.pose_id[g.od_i[k]]); put<uint32_t>(o, g.pose_id[g.od_j[k]]);
        put<uint32_t>(o, 3u); put<uint32_t>(o, 3u);
        for (int q = 0; q < 9; ++q) put<float>(o, (float)g.od_Z[9 * k + q]);
        put<uint32_t>(o, 0u); put<uint32_t>(o, 3u);
        for (int q = 0; q < 3; ++q) put<float>(o, (float)g.od_w[3 * k + q]);
    }
    for (size_t k = 0; k < g.lm_i.size(); ++k) {
        put<uint32_t>(o, 1u); put<uint32_t>(o, g.pose_id[g.lm_i[k]]); put<uint32_t>(o, g.lm_id[g.lm_l[k]]);
        put<uint32_t>(o, 2u); put<uint32_t>(o, 1u);          // Vector2 held in a dynamic matrix: 2 x 1
        put<float>(o, (float)g.lm_meas[2 * k]); put<float>(o, (float)g.lm_meas[2 * k + 1]);
        put<uint32_t>(o, 0u); put<uint32_t>(o, 2u);
        put<float>(o, (float)g.lm_w[2 * k]); put<float>(o, (float)g.lm_w[2 * k + 1]);
    }
    put<uint32_t>(o, (uint32_t)g.fixed_vertex.size());
    for (uint32_t v : g.fixed_vertex) put<uint32_t>(o, v < g.P() ? g.pose_id[v] : g.lm_id[v - g.P()]);
    return o;
}
}  // namespace

extern "C" {

void* orc_load_wire(const uint8_t* buf, uint32_t nbytes) { Graph* g = new Graph(); parse_wire(*g, buf, nbytes); return g; }
const char* orc_error(void* h) { return ((Graph*)h)->error.c_str(); }
void orc_free(void* h) { delete (Graph*)h; }
void orc_counts(void* h, uint32_t* out) {
    Graph& g = *(Graph*)h;
    out[0] = (uint32_t)g.P(); out[1] = (uint32_t)g.L(); out[2] = (uint32_t)g.od_i.size(); out[3] = (uint32_t)g.lm_i.size();
    out[4] = (uint32_t)g.fixed_vertex.size(); out[5] = (uint32_t)g.dim();
}
void orc_ids(void* h, uint32_t* pose_ids, uint32_t* lm_ids) {
    Graph& g = *(Graph*)h;
    std::copy(g.pose_id.begin(), g.pose_id.end(), pose_ids); std::copy(g.lm_id.begin(), g.lm_id.end(), lm_ids);
}
void orc_edge_index(void* h, uint32_t* od_i, uint32_t* od_j, uint32_t* lm_i, uint32_t* lm_l) {
    Graph& g = *(Graph*)h;
    std::copy(g.od_i.begin(), g.od_i.end(), od_i); std::copy(g.od_j.begin(), g.od_j.end(), od_j);
    std::copy(g.lm_i.begin(), g.lm_i.end(), lm_i); std::copy(g.lm_l.begin(), g.lm_l.end(), lm_l);
}
// pose_xyt: (x, y, atan2(R10,R00)) per pose (SerializeGraphFuncCpu.h:22-37 without the float cast)
void orc_get_state(void* h, double* pose_xyt, double* lm_xy) {
    Graph& g = *(Graph*)h;
    for (size_t i = 0; i < g.P(); ++i) { pose_xyt[3 * i] = g.pose[i].x; pose_xyt[3 * i + 1] = g.pose[i].y; pose_xyt[3 * i + 2] = std::atan2(g.pose[i].s, g.pose[i].c); }
    std::copy(g.lm.begin(), g.lm.end(), lm_xy);
}
void orc_set_state(void* h, const double* pose_xyt, const double* lm_xy) {
    Graph& g = *(Graph*)h;
    for (size_t i = 0; i < g.P(); ++i) { g.pose[i].x = pose_xyt[3 * i]; g.pose[i].y = pose_xyt[3 * i + 1]; g.pose[i].c = std::cos(pose_xyt[3 * i + 2]); g.pose[i].s = std::sin(pose_xyt[3 * i + 2]); }
    std::copy(lm_xy, lm_xy + 2 * g.L(), g.lm.begin());
}
void orc_edge_odom(void* h, uint32_t k, double* e) { odom_error(*(Graph*)h, k, e); }
void orc_edge_lm(void* h, uint32_t k, double* e, double* A, double* B) { lm_error(*(Graph*)h, k, e, A, B); }
double orc_assemble_dense(void* h, double* H, double* b) { return assemble_dense(*(Graph*)h, H, b); }
// solves the row-major n x n system in place (x overwrites b); H is consumed
void orc_solve_dense(const double* H, double* b, uint32_t n) {
    std::vector<double> Hc((size_t)n * n);
    for (size_t r = 0; r < n; ++r) for (size_t c = 0; c < n; ++c) Hc[c * n + r] = H[r * n + c];
    colpiv_qr_solve(Hc, b, n);
}
int orc_optimize_dense(void* h, int iterations, int enable_exits, double* chi2_traj, int* exit_code) {
    return gn_loop(*(Graph*)h, 0, iterations, enable_exits, 0, 0, chi2_traj, exit_code);
}
int orc_optimize_sparse(void* h, int iterations, int enable_exits, double tol, int max_pcg, int threads,
                        double* chi2_traj, int* pcg_iters, int* exit_code) {
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
    Graph& g = *(Graph*)h;
    const int n = gn_loop(g, 1, iterations, enable_exits, tol, max_pcg, chi2_traj, exit_code);
    if (pcg_iters) std::copy(g.pcg_iters.begin(), g.pcg_iters.end(), pcg_iters);
    return n;
}
// linearise only (sparse form), `reps` times: the CPU leg of the edges-linearised/s metric
double orc_linearize_sparse(void* h, int reps, int threads) {
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
    Graph& g = *(Graph*)h; Sparse s; sparse_init(g, s); double err = 0;
    for (int r = 0; r < reps; ++r) err = sparse_linearize(g, s);
    return err;
}
// True residual of a candidate step against the oracle's own linearisation at the oracle's current state:
//   || H delta - b ||_2 / || b ||_2  over the FULL system (poses and landmarks, nothing eliminated), one sparse mat-vec.
// Used by the parity tests at sizes where no dense solve exists (OptimizerCpu.h:82-156 assembles this H and b).
double orc_residual_sparse(void* h, const double* dp, const double* dl, int threads, double* chi2) {
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
    Graph& g = *(Graph*)h;
    Sparse s; sparse_init(g, s);
    const double err = sparse_linearize(g, s);
    if (chi2) *chi2 = err;
    const long P = (long)s.P, L = (long)s.L;
    double rr = 0, bb = 0;
#pragma omp parallel for reduction(+ : rr, bb) schedule(static)
    for (long i = 0; i < P; ++i) {
        const double* H = &s.Hpp[9 * i]; const double* pv = &dp[3 * i];
        double y[3];
        for (int r = 0; r < 3; ++r) y[r] = H[3 * r] * pv[0] + H[3 * r + 1] * pv[1] + H[3 * r + 2] * pv[2] - s.bp[3 * i + r];
        for (uint32_t q = s.oo_ptr[i]; q < s.oo_ptr[i + 1]; ++q) { const uint32_t k = s.oo_edge[q];
            const double* pj = &dp[3 * (size_t)g.od_j[k]]; for (int d = 0; d < 3; ++d) y[d] -= s.od_inf[3 * k + d] * pj[d]; }
        for (uint32_t q = s.oi_ptr[i]; q < s.oi_ptr[i + 1]; ++q) { const uint32_t k = s.oi_edge[q];
            const double* pj = &dp[3 * (size_t)g.od_i[k]]; for (int d = 0; d < 3; ++d) y[d] -= s.od_inf[3 * k + d] * pj[d]; }
        for (uint32_t q = s.pose_ptr[i]; q < s.pose_ptr[i + 1]; ++q) { const uint32_t k = s.pose_edge[q];
            const double* hh = &s.Hpl[6 * (size_t)k]; const double* u = &dl[2 * (size_t)g.lm_l[k]];
            y[0] += hh[0] * u[0] + hh[1] * u[1]; y[1] += hh[2] * u[0] + hh[3] * u[1]; y[2] += hh[4] * u[0] + hh[5] * u[1]; }
        for (int r = 0; r < 3; ++r) { rr += y[r] * y[r]; bb += s.bp[3 * i + r] * s.bp[3 * i + r]; }
    }
#pragma omp parallel for reduction(+ : rr, bb) schedule(static)
    for (long l = 0; l < L; ++l) {
        const double* hl = &s.Hll[4 * l]; const double* u = &dl[2 * l];
        double t0 = hl[0] * u[0] + hl[1] * u[1] - s.bl[2 * l], t1 = hl[2] * u[0] + hl[3] * u[1] - s.bl[2 * l + 1];
        for (uint32_t q = s.lm_ptr[l]; q < s.lm_ptr[l + 1]; ++q) { const uint32_t k = s.lm_edge[q];
            const double* hh = &s.Hpl[6 * (size_t)k]; const double* pv = &dp[3 * (size_t)g.lm_i[k]];
            t0 += hh[0] * pv[0] + hh[2] * pv[1] + hh[4] * pv[2]; t1 += hh[1] * pv[0] + hh[3] * pv[1] + hh[5] * pv[2]; }
        rr += t0 * t0 + t1 * t1; bb += s.bl[2 * l] * s.bl[2 * l] + s.bl[2 * l + 1] * s.bl[2 * l + 1];
    }
    return bb > 0 ? std::sqrt(rr / bb) : std::sqrt(rr);
}

int orc_max_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
uint32_t orc_reply_size(void* h) { return (uint32_t)encode_reply(*(Graph*)h).size(); }
void orc_write_reply(void* h, uint8_t* out) { auto v = encode_reply(*(Graph*)h); std::memcpy(out, v.data(), v.size()); }

}  // extern "C"
