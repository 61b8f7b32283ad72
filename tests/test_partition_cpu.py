"""CPU tests of the multi-GPU host logic: pose partition / edge ownership / ghosts (tsb_host_plan), and a
world_size-2 gloo run of the rendezvous helper that ships the NCCL unique id."""
import os
import subprocess
import sys
import numpy as np
import pytest
from conftest import ROOT
from toyslam_b200 import synth, capi


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_partition_tiles_the_graph(world):
    g = synth.make_config("c2", scale=0.2, seed=4)
    payload = synth.to_payload(g)
    c0 = synth.counts(g)
    plans = [capi.host_plan(payload, r, world) for r in range(world)]
    assert plans[0][0].pose_begin == 0 and plans[-1][0].pose_end == c0["P"]
    for a, b in zip(plans[:-1], plans[1:]):
        assert a[0].pose_end == b[0].pose_begin
    assert sum(p[0].n_lmedge_local for p in plans) == c0["El"]      # every LM edge linearised exactly once
    assert sum(p[0].n_odom_local for p in plans) == c0["Eo"]        # every ODOM edge owned exactly once
    assert len({p[3] for p in plans}) == 1                           # cuts + boundary list agree on all ranks
    assert len({p[2] for p in plans}) == 1
    if world == 1:
        assert plans[0][1] == 0 and plans[0][2] == 0
    else:
        # a chain cut at world-1 places: 2 boundary poses per cut, each interior rank has 2 ghosts
        assert plans[0][2] == 2 * (world - 1)
        assert [p[1] for p in plans] == [1] + [2] * (world - 2) + [1]
        work = np.array([p[0].n_lmedge_local for p in plans], float)
        assert work.max() / work.mean() < 1.15                       # balanced by edge count


def test_gloo_world2_unique_id_and_plan(tmp_path):
    script = tmp_path / "w2.py"
    script.write_text(f'''
import os, sys
sys.path.insert(0, {ROOT!r})
import torch.distributed as dist
from toyslam_b200 import synth, capi, dist as tdist
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
uid = tdist.broadcast_unique_id(rank)
payload = synth.to_payload(synth.make_config("c2", scale=0.05, seed=1))
c, ng, nb, h = capi.host_plan(payload, rank, world)
out = [None] * world
dist.all_gather_object(out, (uid.hex(), c.pose_begin, c.pose_end, c.n_lmedge_local, ng, nb, h))
if rank == 0:
    assert out[0][0] == out[1][0] and len(uid) == 128 and any(uid)
    assert out[0][2] == out[1][1] and out[0][1] == 0 and out[1][2] == c.n_pose
    assert out[0][3] + out[1][3] == c.n_lmedge and out[0][5] == out[1][5] == 2 and out[0][6] == out[1][6]
    print("W2 OK")
dist.destroy_process_group()
''')
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29731", str(script)],
                       capture_output=True, text=True, timeout=240, env=env)
    assert r.returncode == 0 and "W2 OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_multilevel_hierarchy_of_large_graphs():
    """Graphs beyond the resident kernel get the multilevel hierarchy (structure.cpp build_hierarchy): ratio 4 per
    level, a final ratio of 2 where 4 would leave the dense top level under half of kMlTopDefault = 1024 nodes."""
    from toyslam_b200 import synth, capi
    for scale in (0.02, 0.05, 0.2):
        payload = synth.to_payload(synth.make_config("c5", scale=scale, seed=1))
        c, _, _, _ = capi.host_plan(payload, 0, 1)
        n, levels = c.n_pose, 0
        while levels == 0 or n > 1024:
            ratio = 2 if (levels > 0 and (n + 1) // 2 <= 1024) else 3 if (levels > 0 and (n + 2) // 3 <= 1024) else 4
            n = -(-n // ratio); levels += 1
        assert (c.n_ml_levels, c.n_ml_top) == (levels, n), (c.n_pose, c.n_ml_levels, c.n_ml_top, levels, n)
        assert 341 < c.n_ml_top <= 1024 or levels == 1


def test_per_rank_structure_is_the_restriction_of_the_one_rank_structure():
    """The edge -> block index map of rank r is the 1-rank map restricted to r's pose range, bit for bit: the same LM
    edges in the same assembly order (sorted by pose, landmark, request index), the same row pointers up to the
    offset, chunks that tile the rank's edges and poses, and per chunk the slots an independent numpy derivation gives."""
    from toyslam_b200 import synth, capi
    payload = synth.to_payload(synth.make_config("c3", scale=0.03, seed=4))
    c1, s1 = capi.host_structure(payload, 0, 1)
    g = synth.make_config("c3", scale=0.03, seed=4)
    lm_i = g["lm_ij"][:, 0].astype(np.int64); lm_l = g["lm_ij"][:, 1].astype(np.int64) - c1.n_pose
    assert np.array_equal(s1["lm_perm"], np.lexsort((np.arange(len(lm_i)), lm_l, lm_i)).astype(np.uint32))
    for world in (2, 4, 8):
        seen = 0
        for rank in range(world):
            c, s = capi.host_structure(payload, rank, world)
            pb, pe = c.pose_begin, c.pose_end
            e0, e1 = int(s1["pose_ptr"][pb]), int(s1["pose_ptr"][pe])
            assert np.array_equal(s["lm_perm"], s1["lm_perm"][e0:e1])
            assert np.array_equal(s["pose_ptr"], s1["pose_ptr"][pb:pe + 1] - e0)
            cd = s["chunk_desc"]
            assert cd[0, 0] == 0 and cd[-1, 1] == e1 - e0 and np.array_equal(cd[1:, 0], cd[:-1, 1])
            assert cd[0, 2] == 0 and cd[-1, 3] == pe - pb
            for (a, b, p0, p1, s0, sl1, blk, _) in cd[:: max(1, len(cd) // 40)]:          # a sample of chunks, slot by slot
                lms = lm_l[s["lm_perm"][a:b]]; uniq, cnt = np.unique(lms, return_counts=True)
                by_len = np.lexsort((uniq, -cnt))
                assert np.array_equal(s["slot_lm"][s0:sl1], uniq[by_len])
                rank_of = np.empty(len(uniq), np.uint32); rank_of[by_len] = np.arange(len(uniq), dtype=np.uint32)
                packed = ((lm_i[s["lm_perm"][a:b]] - pb - p0).astype(np.uint32) << 16) | rank_of[np.searchsorted(uniq, lms)]
                assert np.array_equal(s["lm_packed"][a:b], packed)
            seen += e1 - e0
        assert seen == c1.n_lmedge


def test_bulk_copies_of_the_streamed_kernels_stay_inside_their_arrays():
    """The TMA producers (ls_issue in linearize_stream.cuh, ps_issue in pcg_stream.cuh) copy from sources aligned DOWN to
    16 B with sizes rounded UP to 16 B; every device array carries 64 B of slack (dalloc in api.cu).  compute-sanitizer
    is closed on the GPU pool, so the address arithmetic is replayed here for every chunk of a graph whose LM edge count
    is not a multiple of 4: each copy must start inside its array, end within array + 64 B, and fit its ring-stage slot."""
    from toyslam_b200 import synth, capi
    g = synth.make_graph(1301, 300, 19, "manhattan", seed=3)          # 24 719 LM edges
    payload = synth.to_payload(g)
    c, s = capi.host_structure(payload, 0, 1)
    nE, nS, nP = c.n_lmedge_local, c.n_slots, c.pose_end - c.pose_begin
    assert nE % 4 != 0
    up = lambda v, m: (v + m - 1) // m * m
    slots = {"lm_packed": 2080, "slot_edge": 1056, "slot_lm": 2080, "slot_ptr": 2080, "pose_ptr": 544, "pose": 4096, "hpl": 16 * 512}
    for (e0, e1, p0, p1, s0, s1, blk, _) in s["chunk_desc"].astype(np.int64):
        n, ns, npz = e1 - e0, s1 - s0, p1 - p0
        assert n <= 512 and npz <= 128 and ns <= n
        e4, e8, s4, p4 = e0 & ~3, e0 & ~7, s0 & ~3, p0 & ~3
        copies = [("lm_packed", 4 * e4, 4 * up(e1 - e4, 4), 4 * nE), ("slot_edge", 2 * e8, 2 * up(e1 - e8, 8), 2 * nE),
                  ("slot_lm", 4 * s4, 4 * up(s1 - s4, 4), 4 * nS), ("slot_ptr", 4 * s4, 4 * up(s1 + 1 - s4, 4), 4 * (nS + 1)),
                  ("pose_ptr", 4 * p4, 4 * up(p1 + 1 - p4, 4), 4 * (nP + 1)), ("pose", 32 * p0, 32 * npz, 32 * nP),
                  ("hpl", 16 * e0, 16 * n, 16 * nE)]
        for name, off, nbytes, size in copies:
            assert off % 16 == 0 and nbytes % 16 == 0, (name, off, nbytes)
            assert 0 <= off < size + 16 and off + nbytes <= size + 64, (name, off, nbytes, size)
            assert nbytes <= slots[name], (name, nbytes)

def test_chunk_descriptors_fit_the_packed_form_of_the_chunk_loops():
    """The streamed kernels carry chunk descriptors packed into four words (PDesc, linearize_stream.cuh): n and ns in 10 bits, np
    in 8, and ONE bit of the unit index to tell consecutive chunks of different units apart.  That needs n, ns <= 512, np <= 128,
    and consecutive chunks in the same or in consecutive units -- checked here on graphs with ordinary, sparse and heavy poses."""
    from toyslam_b200 import synth, capi
    cases = [synth.make_graph(1301, 300, 19, "manhattan", seed=3), synth.make_graph(4000, 40, 1, "chain", seed=4),
             synth.make_graph(60, 700, 650, "manhattan", seed=5)]            # the last one: poses with more LM edges than a chunk holds
    for g in cases:
        for world, rank in ((1, 0), (3, 1)):
            c, s = capi.host_structure(synth.to_payload(g), rank, world)
            d = s["chunk_desc"].astype(np.int64)
            if not len(d):
                continue
            n, npz, ns, unit = d[:, 1] - d[:, 0], d[:, 3] - d[:, 2], d[:, 5] - d[:, 4], d[:, 6]
            assert n.min() >= 0 and n.max() <= 512 and npz.min() >= 1 and npz.max() <= 128 and ns.max() <= 512 and (ns <= np.maximum(n, 0)).all()
            step = np.diff(unit)
            assert ((step == 0) | (step == 1)).all() and unit[0] == 0
            assert (d[1:, 0] == d[:-1, 1]).all() and (d[1:, 4] == d[:-1, 5]).all()      # edges and slots of consecutive chunks are contiguous

def test_parse_of_interleaved_record_layouts_matches_the_standard_layout():
    """parse_wire recognises runs of records in the reference client's layout (ODOM 76 B, LM 44 B) at a fixed stride and checks
    them in parallel; everything else goes record by record.  A payload that interleaves blocks of LM records with ODOM records
    in three layouts (3x3 matrix, (x, y, theta) with rows = 0, a short information diagonal) must give the same LM-side structure
    as the plain payload of the same graph (same LM edges in the same relative order), and a payload cut in the middle of a run
    must be rejected with the scan's message."""
    import struct
    from toyslam_b200 import synth, wire, capi
    g = synth.make_config("c3", scale=0.03, seed=7)
    plain = wire.encode_request(g["pose_id"], g["pose_xyt"], g["lm_id"], g["lm_xy"], g["od_ij"], g["od_Z"], g["od_w"], g["lm_ij"], g["lm_meas"],
                                g["lm_w"], g["fixed_ids"])
    P, L = len(g["pose_id"]), len(g["lm_id"])
    head = [struct.pack("<I", P + L)]
    head += [struct.pack("<IIfff", int(g["pose_id"][i]), 0, *map(float, g["pose_xyt"][i])) for i in range(P)]
    head += [struct.pack("<IIff", int(g["lm_id"][i]), 1, *map(float, g["lm_xy"][i])) for i in range(L)]
    od = []
    for k, (a, b) in enumerate(g["od_ij"].tolist()):
        Z, w = np.asarray(g["od_Z"][k], "<f4").tobytes(), [float(v) for v in g["od_w"][k]]
        if k % 7 == 3:
            od.append(struct.pack("<IIIII", 0, a, b, 0, 3) + struct.pack("<fff", 0.5, 0.1, 0.01) + struct.pack("<II", 0, 3) + struct.pack("<fff", *w))
        elif k % 11 == 5:
            od.append(struct.pack("<IIIII", 0, a, b, 3, 3) + Z + struct.pack("<II", 0, 1) + struct.pack("<f", w[0]))
        else:
            od.append(struct.pack("<IIIII", 0, a, b, 3, 3) + Z + struct.pack("<II", 0, 3) + struct.pack("<fff", *w))
    lm = [struct.pack("<IIIII", 1, a, b, 0, 2) + struct.pack("<ff", *map(float, g["lm_meas"][k])) + struct.pack("<II", 0, 2)
          + struct.pack("<ff", *map(float, g["lm_w"][k])) for k, (a, b) in enumerate(g["lm_ij"].tolist())]
    recs, i, j = [], 0, 0
    while i < len(lm) or j < len(od):
        recs += lm[i:i + 3000]; i += 3000
        recs += od[j:j + 500]; j += 500
    fx = np.asarray(g["fixed_ids"], "<u4")
    mixed = b"".join(head + [struct.pack("<I", len(recs))] + recs + [struct.pack("<I", len(fx)), fx.tobytes()])
    c0, s0 = capi.host_structure(plain, 0, 1)
    c1, s1 = capi.host_structure(mixed, 0, 1)
    assert (c0.n_pose, c0.n_lm, c0.n_odom, c0.n_lmedge, c0.n_chunks, c0.n_slots) == (c1.n_pose, c1.n_lm, c1.n_odom, c1.n_lmedge, c1.n_chunks, c1.n_slots)
    # lm_perm holds request-order LM indices: the same in both payloads because LM edges keep their relative order
    for k in ("lm_perm", "pose_ptr", "chunk_desc", "slot_lm", "lm_packed"):
        assert np.array_equal(s0[k], s1[k]), k
    for cut in (len(mixed) - 30001, len(mixed) - 4 * len(fx) - 4 - 13):      # inside a run near the end / inside the last record
        with pytest.raises(capi.TsbError) as ei:
            capi.host_plan(mixed[:cut], 0, 1)
        assert ei.value.code == 2 and "truncated payload" in str(ei.value), str(ei.value)
