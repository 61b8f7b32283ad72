"""GPU test of the allocation-free entry point tsb_odom_linearize_noalloc
(it has no counterpart in the reference).  Collected last on purpose (file name):
a failure here must not hide the parity tests.

The new entry point is called through the C ABI with caller-owned device buffers (torch tensors, allocated BEFORE the
calls) and compared with the existing entry points on the same ODOM-only graph at the same perturbed state:
tsb_load_wire + tsb_set_state + tsb_linearize (chi2) and tsb_read_system (hpp_off, b_pose).  Tolerances are the ones
tests/test_gpu_parity.py uses for the assembled system: chi2 1e-12 relative, blocks 1e-12 / 1e-11 of the largest entry.
Two calls in a row on different graphs (the second smaller, so stale scratch of the first lies beyond what it may read)
and a third that must reproduce the first bit for bit."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def odom_only_case(P, seed):
    """ODOM-only chain of P poses as a request payload + the arrays the new entry point takes (host copies)."""
    from toyslam_b200 import synth, wire
    g = synth.make_graph(P, 16, 2, "chain", seed)
    z = lambda *shape, dt=np.float32: np.zeros(shape, dt)
    payload = wire.encode_request(g["pose_id"], g["pose_xyt"], z(0, dt=np.uint32), z(0, 2), g["od_ij"], g["od_Z"], g["od_w"],
                                  z(0, 2, dt=np.uint32), z(0, 2), z(0, 2), g["fixed_ids"])
    rng = np.random.default_rng(seed + 100)
    pose = g["pose_xyt"].astype(np.float64)
    pose[:, :2] += rng.normal(0.0, 0.3, (P, 2)); pose[:, 2] += rng.normal(0.0, 0.05, P)      # residuals on both sides of the Huber threshold
    Z = g["od_Z"].astype(np.float64).reshape(-1, 3, 3)
    zinv = np.ascontiguousarray(np.linalg.inv(Z)[:, :2, :].reshape(-1, 6))
    return payload, pose, g["od_ij"].astype(np.int32), zinv, g["od_w"].astype(np.float64)


def existing_entry_points(payload, pose):
    from toyslam_b200.capi import GraphOptimizerGpu
    g = GraphOptimizerGpu()
    c = g.load_wire(payload)
    assert c.n_lm == 0 and c.n_lmedge == 0
    g.set_state(pose, np.zeros((0, 2)))
    st = g.linearize()
    s = g.system()
    g.close()
    return st.chi2, s["hpp_off"], s["b_pose"]


def test_noalloc_entry_point_matches_existing_entry_points():
    import torch
    from toyslam_b200 import capi
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    dev = torch.device("cuda", 0)
    cases = [odom_only_case(70_000, 5), odom_only_case(3_000, 6)]
    ref = [existing_entry_points(c[0], c[1]) for c in cases]
    # caller-owned device buffers, all allocated before the first call
    bufs = []
    for payload, pose, ij, zinv, w in cases:
        E = len(ij)
        bufs.append(dict(pose=torch.from_numpy(pose).to(dev), ij=torch.from_numpy(ij).to(dev), zinv=torch.from_numpy(zinv).to(dev),
                         w=torch.from_numpy(w).to(dev), hoff=torch.full((E, 3), np.nan, dtype=torch.float64, device=dev),
                         we=torch.full((E, 3), np.nan, dtype=torch.float64, device=dev),
                         chi2=torch.full((1,), np.nan, dtype=torch.float64, device=dev)))
    again = dict(hoff=torch.empty_like(bufs[0]["hoff"]), we=torch.empty_like(bufs[0]["we"]), chi2=torch.empty_like(bufs[0]["chi2"]))
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream(dev).cuda_stream

    def call(b, out):
        capi.odom_linearize_noalloc(0, stream, b["pose"].shape[0], b["ij"].shape[0], b["pose"].data_ptr(), b["ij"].data_ptr(),
                                    b["zinv"].data_ptr(), b["w"].data_ptr(), out["hoff"].data_ptr(), out["we"].data_ptr(), out["chi2"].data_ptr())

    mem0 = torch.cuda.memory_allocated(dev)
    call(bufs[0], bufs[0]); call(bufs[1], bufs[1]); call(bufs[0], again)          # back to back on one stream, no sync in between
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated(dev) == mem0
    for (payload, pose, ij, zinv, w), b, (chi2_ref, hoff_ref, b_ref) in zip(cases, bufs, ref):
        chi2 = float(b["chi2"].item()); hoff = b["hoff"].cpu().numpy(); we = b["we"].cpu().numpy()
        assert chi2_ref > 0 and abs(chi2 - chi2_ref) <= 1e-12 * chi2_ref, (chi2, chi2_ref)
        assert hoff.shape == hoff_ref.shape and np.max(np.abs(hoff - hoff_ref)) <= 1e-12 * np.max(np.abs(hoff_ref))
        # b of an ODOM-only graph from W e: + on the edge's first pose, - on its second (A = -I, B = +I)
        bp = np.zeros_like(b_ref)
        np.add.at(bp, ij[:, 0], we); np.add.at(bp, ij[:, 1], -we)
        assert np.max(np.abs(bp - b_ref)) <= 1e-11 * np.max(np.abs(b_ref))
        assert np.any(-hoff[:, 0] < w[:, 0] * (1 - 1e-9)) and np.any(-hoff[:, 0] == w[:, 0])      # both Huber branches were taken
    # the third call (graph A again, after B overwrote the head of the scratch) reproduces the first bit for bit
    assert torch.equal(again["chi2"], bufs[0]["chi2"]) and torch.equal(again["hoff"], bufs[0]["hoff"]) and torch.equal(again["we"], bufs[0]["we"])


def test_noalloc_entry_point_rejects_oversize_and_null():
    import torch
    from toyslam_b200 import capi
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    one = torch.zeros(8, dtype=torch.float64, device="cuda")
    p = one.data_ptr()
    with pytest.raises(capi.TsbError) as ei:
        capi.odom_linearize_noalloc(0, 0, 4, capi.NOALLOC_MAX_ODOM_EDGES + 1, p, p, p, p, p, p, p)
    assert ei.value.code == 1 and "too many ODOM edges" in str(ei.value) and str(capi.NOALLOC_MAX_ODOM_EDGES) in str(ei.value)
    with pytest.raises(capi.TsbError) as ei:
        capi.odom_linearize_noalloc(0, 0, 4, 2, p, p, p, p, p, p, 0)
    assert ei.value.code == 1 and "null pointer" in str(ei.value)
    # an empty edge list is fine: chi2 = 0
    capi.odom_linearize_noalloc(0, 0, 0, 0, 0, 0, 0, 0, 0, 0, p)
    torch.cuda.synchronize()
    assert float(one[0].item()) == 0.0
