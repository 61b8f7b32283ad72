// host_timing.cpp -- times the host side of a request (parse_wire, build_structure, encode_reply) on a payload file.
//   g++ -O3 -std=c++17 -fopenmp -Itoyslam_b200/csrc tools/host_timing.cpp toyslam_b200/csrc/build/wire_codec.o toyslam_b200/csrc/build/structure.o -o /tmp/host_timing
//   python -c "from toyslam_b200 import synth; open('/tmp/c4.bin','wb').write(synth.to_payload(synth.make_config('c4')))"
//   TSB_HOST_TIMING=1 /tmp/host_timing /tmp/c4.bin [world] [ml_top]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iterator>
#include <string>
#include <vector>
#include "host_graph.h"

int main(int argc, char** argv) {
    if (argc < 2) return 2;
    std::ifstream f(argv[1], std::ios::binary);
    std::vector<uint8_t> buf((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    const uint32_t world = argc > 2 ? (uint32_t)std::atoi(argv[2]) : 1, ml_top = argc > 3 ? (uint32_t)std::atoi(argv[3]) : 1536;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto ms = [](auto a, auto b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
    tsb::HostGraph g; tsb::Structure st; std::string err;
    for (int rep = 0; rep < 2; ++rep) {
        g.reset();
        auto t0 = now();
        if (!tsb::parse_wire(buf.data(), (uint32_t)buf.size(), g, err)) { std::printf("parse: %s\n", err.c_str()); return 1; }
        auto t1 = now();
        tsb::build_structure(g, 0, world, st, 0, nullptr, ml_top);
        auto t2 = now();
        std::printf("rep %d: parse %.1f ms, build_structure %.1f ms (P=%u El=%u chunks=%zu slots=%zu levels=%zu)\n", rep, ms(t0, t1), ms(t1, t2),
                    g.P, g.El, st.chunks.size(), st.slot_lm.size(), st.ml.size());
    }
    // checksum of the integer structure: a refactoring must not change it
    uint64_t h = 1469598103934665603ull;
    auto mix = [&](const void* p, size_t n) { const uint8_t* b = (const uint8_t*)p; for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 1099511628211ull; } };
    auto mv = [&](const auto& v) { if (!v.empty()) mix(v.data(), v.size() * sizeof(v[0])); };
    mv(st.le_local); mv(st.lm_perm); mv(st.lm_packed); mv(st.pose_ptr); mv(st.chunks); mv(st.blk_chunk_ptr); mv(st.slot_lm); mv(st.slot_ptr); mv(st.slot_edge);
    mv(st.lm_part_ptr); mv(st.lm_part_idx); mv(st.od_local); mv(st.od_a); mv(st.od_b); mv(st.od_out_ptr); mv(st.od_in_ptr); mv(st.od_in_edge);
    mv(st.fixed_pose_count); mv(st.fixed_lm_count); mv(st.edge_pose); mv(st.ghost_pose); mv(st.boundary); mv(st.x_exp); mv(st.lm_mask); mv(st.top_nodes);
    for (const auto& lv : st.ml) { mv(lv.node_ptr); mv(lv.edge_lm); mv(lv.ref_ptr); mv(lv.ref); }
    std::printf("structure checksum %016llx\n", (unsigned long long)h);
    // ... and of the parsed graph (offsets, per-type indices, decoded values, topology hash)
    h = 1469598103934665603ull;
    mv(g.edge_off); mv(g.edge_slot); mv(g.od_i); mv(g.od_j); mv(g.od_zinv); mv(g.od_w); mv(g.le_i); mv(g.le_l); mv(g.le_z); mv(g.le_w); mv(g.fixed); mv(g.req_vertex);
    mix(&g.topo_hash, 8); mix(&g.reply_bytes, 4);
    std::printf("parse checksum %016llx (Eo %u El %u hash %016llx)\n", (unsigned long long)h, g.Eo, g.El, (unsigned long long)g.topo_hash);
    return 0;
}
