"""GPU parity of the device-side DP3 observation pipeline (rmb_pointcloud_from_depth) against oracle/pointcloud_oracle.py,
whose back-projection and crop are pinned on the reference's own functions (tests/golden/pointcloud_golden.npz):
valid-point counts, FPS indices and the float32 output are bit-exact."""
import os
import sys

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from cases import POINTCLOUD_CASES, pointcloud_inputs  # noqa: E402
from oracle import pointcloud_oracle as po  # noqa: E402
from robomanipbaselines_b200.common.synthetic import synthetic_norm_stats  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,K,norm", [("small", 32, None), ("small", 128, "gaussian"), ("img84", 256, "limits"),
                                         ("img84", 1024, "gaussian")])
def test_pointcloud_pipeline_bit_exact(cuda, name, K, norm):
    from robomanipbaselines_b200.common.pointcloud import DepthToPointcloud

    c = POINTCLOUD_CASES[name]
    B = 3
    depth, rgb = pointcloud_inputs(c, batch=B)
    stats = synthetic_norm_stats(6, 9, norm) if norm else None
    pipe = DepthToPointcloud(B, (c["W"], c["H"]), c["fovy"], K, c["min_bound"], c["max_bound"], stats, device=str(cuda))
    got = pipe(torch.from_numpy(depth).to(cuda), torch.from_numpy(rgb).to(cuda)).cpu().numpy()
    cnt, idx = pipe.count.cpu().numpy(), pipe.indices.cpu().numpy()
    for b in range(B):
        pc = np.concat(po.convert_depth_image_to_pointcloud(depth[b], c["fovy"], rgb[b]), axis=1)
        pc = po.crop_pointcloud_bb(pc, c["min_bound"], c["max_bound"])
        want_idx = po.farthest_point_sampling(pc, K)
        want, n = po.get_pointcloud(depth[b], rgb[b], c["fovy"], c["min_bound"], c["max_bound"], K, stats)
        assert cnt[b] == n == pc.shape[0]
        assert np.array_equal(idx[b], want_idx), (name, b, np.flatnonzero(idx[b] != want_idx)[:5])
        assert got[b].dtype == want.dtype and np.array_equal(got[b], want), (name, b)
    print(f"[pointcloud {name} K={K} norm={norm}] valid points {cnt.tolist()}, output bit-exact")


def test_pointcloud_large_cloud_takes_the_global_memory_sweep(cuda):
    """More kept points than the resident FPS loop holds on chip (7168): the sweep over the scratch buffer, and a cloud
    just under the limit beside it in the same launch; both bit-exact against the oracle."""
    from robomanipbaselines_b200.common.pointcloud import DepthToPointcloud

    K = 48
    big = dict(H=96, W=128, fovy=50.0, seed=77, min_bound=[-5.0, -5.0, 0.1], max_bound=[5.0, 5.0, 5.0], holes=False)
    depth, rgb = pointcloud_inputs(big, batch=2)
    depth[1, 55:] = 0.0                                        # frame 1: 55 x 128 = 7040 valid pixels (resident path)
    pipe = DepthToPointcloud(2, (big["W"], big["H"]), big["fovy"], K, big["min_bound"], big["max_bound"], None, device=str(cuda))
    got = pipe(torch.from_numpy(depth).to(cuda), torch.from_numpy(rgb).to(cuda)).cpu().numpy()
    cnt, idx = pipe.count.cpu().numpy(), pipe.indices.cpu().numpy()
    assert cnt.tolist() == [96 * 128, 55 * 128]
    for b in range(2):
        pc = np.concat(po.convert_depth_image_to_pointcloud(depth[b], big["fovy"], rgb[b]), axis=1)
        pc = po.crop_pointcloud_bb(pc, big["min_bound"], big["max_bound"])
        assert np.array_equal(idx[b], po.farthest_point_sampling(pc, K)), b
        want, _ = po.get_pointcloud(depth[b], rgb[b], big["fovy"], big["min_bound"], big["max_bound"], K, None)
        assert np.array_equal(got[b], want), b


def test_pointcloud_matches_reference_golden_directly(cuda):
    """The first frame of each case is the very input the reference functions were run on: device output rows must be
    rows of the stored reference crop (float64 -> float32)."""
    from robomanipbaselines_b200.common.pointcloud import DepthToPointcloud

    g = np.load(os.path.join(HERE, "golden", "pointcloud_golden.npz"))
    for name, c in POINTCLOUD_CASES.items():
        depth, rgb = pointcloud_inputs(c)
        pipe = DepthToPointcloud(1, (c["W"], c["H"]), c["fovy"], 64, c["min_bound"], c["max_bound"], None, device=str(cuda))
        got = pipe(torch.from_numpy(depth[None]).to(cuda), torch.from_numpy(rgb[None]).to(cuda)).cpu().numpy()[0]
        ref = g[name + "_cropped"]
        assert int(pipe.count[0]) == ref.shape[0]
        idx = pipe.indices.cpu().numpy()[0]
        assert np.array_equal(got, ref[idx].astype(np.float32))


def test_dp3_observation_pipeline_with_resize(cuda):
    """Full-resolution frames -> cv2.resize -> point cloud: the device pipeline against cv2 + the pinned oracle. The
    uint8 resize is bit-exact; the float32 depth resize differs from cv2 in the last bits, so points are compared
    where both pipelines kept the same pixels (identical valid-point count) with a tolerance on xyz."""
    cv2 = pytest.importorskip("cv2")
    from robomanipbaselines_b200.common.pointcloud import Dp3ObservationPipeline

    rng = np.random.default_rng(51)
    B, H, W = 2, 120, 160
    meta = {"data": {"image_size": (84, 84), "num_points": 128, "min_bound": [-2.0, -2.0, 0.0], "max_bound": [2.0, 2.0, 3.0]},
            "pointcloud": synthetic_norm_stats(6, 10, "gaussian")}
    # smooth depth so that bilinear last-bit differences cannot move a point across the (wide) bounding box
    yy, xx = np.meshgrid(np.linspace(0, 1, H), np.linspace(0, 1, W), indexing="ij")
    depth = np.stack([(1.0 + 0.3 * np.sin(3 * xx + b) + 0.2 * yy).astype(np.float32) for b in range(B)])
    rgb = rng.integers(0, 256, (B, H, W, 3), dtype=np.uint8)
    pipe = Dp3ObservationPipeline(B, meta, 58.0, device=str(cuda))
    got = pipe(torch.from_numpy(rgb).to(cuda), torch.from_numpy(depth).to(cuda)).cpu().numpy()
    for b in range(B):
        d2, r2 = cv2.resize(depth[b], (84, 84)), cv2.resize(rgb[b], (84, 84))
        want, n = po.get_pointcloud(d2, r2, 58.0, meta["data"]["min_bound"], meta["data"]["max_bound"], 128, meta["pointcloud"])
        assert int(pipe.to_cloud.count[b]) == n == 84 * 84
        # FPS order can differ after a last-bit change of the depth: compare as point sets through the first point and
        # through the exactly-resized colours
        assert np.allclose(got[b, 0], want[0], atol=1e-4)
        dmin = np.abs(got[b][:, None, :] - want[None, :, :]).max(axis=2).min(axis=1)
        assert (dmin <= 1e-4).mean() >= 0.9, (dmin <= 1e-4).mean()
