"""rmb_pointcloud_from_depth_static: the allocation-free variant of rmb_pointcloud_from_depth (scratch in a file-scope
static __device__ array, include/rmb_b200.h) against the existing entry point with caller-provided scratch, through the
C ABI, on two calls in a row with different inputs (the shared array keeps the first call's contents). The arithmetic
is fp64 / integer, so the tolerance is the one tests/test_pointcloud_gpu.py uses: bit-exact.

Collected last on purpose (a failure here must not hide other tests)."""
import ctypes
import os
import sys

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from cases import POINTCLOUD_CASES, pointcloud_inputs  # noqa: E402

pytestmark = pytest.mark.gpu

STATIC_MAX = 32 << 20     # RMB_POINTCLOUD_STATIC_SCRATCH_BYTES


def _params(c, K):
    from robomanipbaselines_b200._lib import PointcloudParams

    p = PointcloudParams()
    p.H, p.W, p.num_points = c["H"], c["W"], K
    p.focal_scaling = float((1.0 / np.tan(np.deg2rad(c["fovy"]) / 2.0)) * c["H"] / 2.0)
    p.near_clip, p.far_clip = 0.0, float("inf")
    p.has_min_bound = p.has_max_bound = 1
    for i in range(3):
        p.min_bound[i], p.max_bound[i] = float(c["min_bound"][i]), float(c["max_bound"][i])
    p.norm_type = 0
    for i in range(6):
        p.norm_p0[i], p.norm_p1[i] = 0.1 * i, 1.0 + 0.25 * i
    p.norm_out_min = 0.0
    return p


def test_static_scratch_entry_point_matches_the_allocating_one(cuda):
    from robomanipbaselines_b200._lib import check, load

    lib = load()
    st = torch.cuda.current_stream().cuda_stream
    for call, (name, K, B) in enumerate([("img84", 256, 3), ("small", 64, 2)]):
        c = POINTCLOUD_CASES[name]
        depth_h, rgb_h = pointcloud_inputs(c, batch=B)
        depth = torch.from_numpy(depth_h).to(cuda).contiguous()
        rgb = torch.from_numpy(rgb_h).to(cuda).contiguous()
        p = _params(c, K)
        nbytes = lib.rmb_pointcloud_scratch_bytes(B, c["H"], c["W"], K)
        assert 0 < nbytes <= STATIC_MAX
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=cuda)
        want = torch.full((B, K, 6), -7.0, device=cuda)
        want_cnt = torch.full((B,), -1, dtype=torch.int32, device=cuda)
        want_idx = torch.full((B, K), -5, dtype=torch.int32, device=cuda)
        check(lib.rmb_pointcloud_from_depth(depth.data_ptr(), rgb.data_ptr(), ctypes.byref(p), scratch.data_ptr(),
                                            want.data_ptr(), want_cnt.data_ptr(), want_idx.data_ptr(), B, st), "pointcloud")
        got = torch.full((B, K, 6), 9.0, device=cuda)
        got_cnt = torch.full((B,), -2, dtype=torch.int32, device=cuda)
        got_idx = torch.full((B, K), -6, dtype=torch.int32, device=cuda)
        check(lib.rmb_pointcloud_from_depth_static(depth.data_ptr(), rgb.data_ptr(), ctypes.byref(p), got.data_ptr(),
                                                   got_cnt.data_ptr(), got_idx.data_ptr(), B, st), "pointcloud static")
        # without the optional outputs the index list lives in the static array too
        got2 = torch.full((B, K, 6), 3.0, device=cuda)
        check(lib.rmb_pointcloud_from_depth_static(depth.data_ptr(), rgb.data_ptr(), ctypes.byref(p), got2.data_ptr(),
                                                   None, None, B, st), "pointcloud static (no optional outputs)")
        torch.cuda.synchronize()
        assert int(want_cnt.min()) > 0
        assert torch.equal(got_cnt, want_cnt), (call, got_cnt.tolist(), want_cnt.tolist())
        assert torch.equal(got_idx, want_idx), call
        assert torch.equal(got, want) and torch.equal(got2, want), call
        assert torch.isfinite(got).all()
    assert lib.rmb_device_fault(0) == 0


def test_static_scratch_entry_point_rejects_oversized_requests(cuda):
    from robomanipbaselines_b200._lib import load

    lib = load()
    c = dict(POINTCLOUD_CASES["img84"])
    c["H"], c["W"] = 480, 640
    p = _params(c, 1024)
    B = 2                                          # 2 x 480 x 640 x 56 B = 34.4 MB > 32 MiB
    dummy = torch.zeros(16, device=cuda)
    rc = lib.rmb_pointcloud_from_depth_static(dummy.data_ptr(), dummy.data_ptr(), ctypes.byref(p), dummy.data_ptr(), None,
                                              None, B, torch.cuda.current_stream().cuda_stream)
    assert rc == -1
    msg = lib.rmb_last_error().decode()
    assert "static maximum" in msg and str(STATIC_MAX) in msg, msg
    torch.cuda.synchronize()
