// graph_gpu.h -- GraphGpu<T> / OptimizerGpu<T>: the class names remote/app/GraphManager.h:80,117 instantiate
// for PIPELINE=gpu, as thin C++ shims over the C ABI (include/toyslam_b200.h).
//
// GraphGpu keeps the Graph concept the codec templates use (remote/cuda/graph/GraphGpu.h:44-58,226-239):
// AddVertex / AddEdge / FixVertex fill flat fp32 arrays (no per-object allocation); the bulk path
// LoadFromWire hands the request bytes to the library untouched.  OptimizerGpu::Optimize uploads on first
// use, runs tsb_optimize and leaves the result in the context; Serialize() returns the reply payload
// (SerializeGraphStart, GraphManager.h:56-71).
#pragma once
#include <cstdint>
#include <cstdio>
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "../../include/toyslam_b200.h"
#include "interfaces.h"

enum class VertexType : unsigned { Se2 = 0, Point2 = 1 };       // remote/graph/vertex/VertexType.h:3-7
enum class EdgeType : unsigned { Se2 = 0, Se2Point2 = 1 };      // remote/graph/edge/EdgeType.h:3-7

// One process-wide set of contexts: the reference rebuilds solver/optimizer/graph per request
// (ConnectionHandler.h:18-20); device buffers, streams and the NCCL communicator persist here.
// TSB_GPUS=N (default 1) makes the ONE server process drive N GPUs (devices 0 .. N-1): one context per GPU, rank r on
// device r, a request is edge-sharded over them (contiguous pose ranges, NVLink peer exchange inside the kernels).  The
// library's calls are collective over the ranks, so every call runs on all contexts at once from N short-lived host
// threads (remote/app/main.cpp:42 has one thread for everything; the GPU work is what runs in parallel here).
class GpuContext {
public:
    static GpuContext& Instance(int enable_exits = 1, int verbose = 1) { static GpuContext inst(enable_exits, verbose); return inst; }
    static tsb_ctx* Get(int enable_exits = 1, int verbose = 1) { return Instance(enable_exits, verbose).ctx[0]; }
    int world() const { return (int)ctx.size(); }
    // fn(rank, ctx) on every rank concurrently; the first non-zero status wins (rank order)
    template <class F> int Each(F fn) {
        if (ctx.size() == 1) return fn(0, ctx[0]);
        std::vector<int> rc(ctx.size(), 0);
        std::vector<std::thread> th;
        for (size_t r = 0; r < ctx.size(); ++r) th.emplace_back([&, r] { rc[r] = fn((int)r, ctx[r]); });
        for (auto& t : th) t.join();
        for (int v : rc) if (v) return v;
        return 0;
    }
private:
    GpuContext(int enable_exits, int verbose) {
        int n = 1;
        if (const char* e = std::getenv("TSB_GPUS")) n = std::max(1, std::atoi(e));
        if (n > 8) throw std::runtime_error("TSB_GPUS: at most 8 GPUs");
        unsigned char id[128] = {};
        if (n > 1 && tsb_nccl_unique_id(id) != TSB_OK) throw std::runtime_error(std::string("tsb_nccl_unique_id: ") + tsb_last_error(nullptr));
        ctx.assign((size_t)n, nullptr);
        std::vector<int> rc((size_t)n, 0);
        std::vector<std::thread> th;
        for (int r = 0; r < n; ++r)            // ncclCommInitRank blocks until every rank has joined: create the contexts in parallel
            th.emplace_back([&, r] {
                tsb_config cfg{}; cfg.device = r; cfg.rank = r; cfg.world = n; cfg.nccl_id = n > 1 ? id : nullptr; cfg.pcg_tol = 1e-12;
                cfg.pcg_max_iters = 20000; cfg.enable_exits = enable_exits; cfg.verbose = r == 0 ? verbose : 0;
                rc[(size_t)r] = tsb_create(&cfg, &ctx[(size_t)r]);
            });
        for (auto& t : th) t.join();
        for (int v : rc) if (v != TSB_OK) throw std::runtime_error(std::string("tsb_create: ") + tsb_last_error(nullptr));
    }
    ~GpuContext() { for (tsb_ctx* c : ctx) tsb_destroy(c); }
    std::vector<tsb_ctx*> ctx;
};

template <typename T>
class GraphGpu : public IGraph {
public:
    // --- Graph concept (per-object path, small graphs) ---
    void AddVertexSe2(unsigned id, T x, T y, T theta) { pose_id.push_back(id); pose_xyt.insert(pose_xyt.end(), {(float)x, (float)y, (float)theta}); }
    void AddVertexPoint2(unsigned id, T x, T y) { lm_id.push_back(id); lm_xy.insert(lm_xy.end(), {(float)x, (float)y}); }
    void AddEdgeSe2(unsigned id1, unsigned id2, const T* meas3x3, const T* info_diag3) {
        od_ij.insert(od_ij.end(), {id1, id2});
        for (int k = 0; k < 9; ++k) od_Z.push_back((float)meas3x3[k]);
        for (int k = 0; k < 3; ++k) od_w.push_back((float)info_diag3[k]);
    }
    void AddEdgeSe2Point2(unsigned id1, unsigned id2, T range, T bearing, const T* info_diag2) {
        le_ij.insert(le_ij.end(), {id1, id2}); le_meas.insert(le_meas.end(), {(float)range, (float)bearing});
        le_w.insert(le_w.end(), {(float)info_diag2[0], (float)info_diag2[1]});
    }
    void FixVertex(unsigned id) { fixed.push_back(id); }
    // --- bulk path: the request payload as received (DeserializeGraphStart, GraphManager.h:35-54) ---
    void LoadFromWire(const char* bytes, uint32_t nbytes) { wire.assign(bytes, bytes + nbytes); use_wire = true; }

    // ToDevice (GraphGpu.h:80-187): parse / upload on every rank; throws with the library's message on a malformed payload
    void ToDevice(GpuContext& gc) {
        tsb_graph_soa s{};
        if (!use_wire) {
            s.n_pose = (uint32_t)pose_id.size(); s.n_lm = (uint32_t)lm_id.size(); s.n_odom = (uint32_t)od_ij.size() / 2;
            s.n_lmedge = (uint32_t)le_ij.size() / 2; s.n_fixed = (uint32_t)fixed.size();
            s.pose_id = pose_id.data(); s.pose_xyt = pose_xyt.data(); s.lm_id = lm_id.data(); s.lm_xy = lm_xy.data();
            s.odom_ij = od_ij.data(); s.odom_Z = od_Z.data(); s.odom_w = od_w.data();
            s.lmedge_ij = le_ij.data(); s.lmedge_meas = le_meas.data(); s.lmedge_w = le_w.data(); s.fixed_id = fixed.data();
        }
        std::string msg;
        const int rc = gc.Each([&](int r, tsb_ctx* ctx) {
            const int v = use_wire ? tsb_load_wire(ctx, wire.data(), (uint32_t)wire.size()) : tsb_load_soa(ctx, &s);
            if (v != TSB_OK && r == 0) msg = tsb_last_error(ctx);
            return v;
        });
        if (rc != TSB_OK) throw std::runtime_error(msg.empty() ? "tsb_load failed on a rank" : msg);
        on_device = true;
    }
    void ToDevice(tsb_ctx*) { ToDevice(GpuContext::Instance()); }
    // SerializeGraph (SerializeGraph.h:8-72): reply payload with the 4-byte size prefix, as the reference returns it
    // (every rank takes part in the state gather; rank 0's bytes are the reply)
    std::vector<uint8_t> Serialize(GpuContext& gc) const {
        uint32_t n = 0;
        if (!on_device || tsb_reply_size(GpuContext::Get(), &n) != TSB_OK) return {};
        std::vector<std::vector<uint8_t>> out((size_t)gc.world());
        const int rc = gc.Each([&](int r, tsb_ctx* ctx) {
            out[(size_t)r].resize(4 + (size_t)n);
            std::memcpy(out[(size_t)r].data(), &n, 4);
            return tsb_write_reply(ctx, out[(size_t)r].data() + 4, n);
        });
        if (rc != TSB_OK) return {};
        return std::move(out[0]);
    }
    std::vector<uint8_t> Serialize(tsb_ctx*) const { return Serialize(GpuContext::Instance()); }
    bool on_device = false;
private:
    bool use_wire = false;
    std::vector<char> wire;
    std::vector<uint32_t> pose_id, lm_id, od_ij, le_ij, fixed;
    std::vector<float> pose_xyt, lm_xy, od_Z, od_w, le_meas, le_w;
};

template <typename T>
class OptimizerGpu : public IOptimizer<T> {
public:
    OptimizerGpu(const unsigned iterations, std::unique_ptr<ISolver<T>> solver) : IOptimizer<T>(iterations, std::move(solver)) {}
    // OptimizerGpu.cu:16-126 in the reference; semantics here are OptimizerCpu<T>::Optimize's (the parity target)
    void Optimize(IGraph* graphIn) override {
        if (graphIn == nullptr) return;                                    // OptimizerCpu.h:27-30
        auto* graph = dynamic_cast<GraphGpu<T>*>(graphIn);
        if (graph == nullptr) { std::printf("OptimizerGpu: not a GraphGpu\n"); return; }   // OptimizerGpu.cu:22-27
        GpuContext& gc = GpuContext::Instance();
        if (!graph->on_device) graph->ToDevice(gc);
        std::vector<tsb_report> rep((size_t)gc.world());
        std::string msg;
        const int rc = gc.Each([&](int r, tsb_ctx* ctx) {
            const int v = tsb_optimize(ctx, this->iterations, &rep[(size_t)r], nullptr);
            if (v != TSB_OK && r == 0) msg = tsb_last_error(ctx);
            return v;
        });
        if (rc != TSB_OK) throw std::runtime_error(msg.empty() ? "tsb_optimize failed on a rank" : msg);
        report = rep[0];
    }
    tsb_report report{};
};
