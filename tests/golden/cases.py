"""Definition of the golden ACT cases (shared by make_golden.py and the tests): configs, seeds, input generation."""
import numpy as np

ACT_CASES = {
    # small: every structural feature (2 cameras, >1 decoder layer, Ds != Da, batch 3) at a size whose taps fit in git
    "small": dict(img_h=64, img_w=96, num_cams=2, hidden_dim=512, dim_feedforward=128, enc_layers=2, dec_layers=2,
                  nheads=8, chunk_size=10, state_dim=5, action_dim=6, batch=3, weight_seed=11, input_seed=21,
                  store_taps=True),
    # BASELINE.json configs[0]: ACT (ResNet-18, 4 enc / 7 dec, hidden 512, ff 3200, chunk 100), MujocoUR5eCable shapes
    "c1_1cam": dict(img_h=480, img_w=640, num_cams=1, hidden_dim=512, dim_feedforward=3200, enc_layers=4, dec_layers=7,
                    nheads=8, chunk_size=100, state_dim=7, action_dim=7, batch=1, weight_seed=12, input_seed=22),
    "c1_2cam": dict(img_h=480, img_w=640, num_cams=2, hidden_dim=512, dim_feedforward=3200, enc_layers=4, dec_layers=7,
                    nheads=8, chunk_size=100, state_dim=7, action_dim=7, batch=2, weight_seed=13, input_seed=23),
}


def policy_args(case):
    """model_meta_info["policy"]["args"] as policy/act/TrainAct.py:45-57 writes it."""
    return {
        "lr": 1e-5, "num_queries": case["chunk_size"], "kl_weight": 10, "hidden_dim": case["hidden_dim"],
        "dim_feedforward": case["dim_feedforward"], "lr_backbone": 1e-5, "backbone": "resnet18",
        "enc_layers": case["enc_layers"], "dec_layers": case["dec_layers"], "nheads": case["nheads"],
        "camera_names": [f"cam{i}" for i in range(case["num_cams"])],
    }


def act_inputs(case):
    """(qpos float32 (B, Ds) already normalised, images uint8 (B, Ncam, H, W, 3))."""
    rng = np.random.default_rng(case["input_seed"])
    qpos = rng.standard_normal(size=(case["batch"], case["state_dim"])).astype(np.float32)
    # smooth-ish images (low-frequency blobs + noise) so the backbone sees structure, not only white noise
    B, N, H, W = case["batch"], case["num_cams"], case["img_h"], case["img_w"]
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    img = np.zeros((B, N, H, W, 3), dtype=np.float32)
    for b in range(B):
        for n in range(N):
            for c in range(3):
                fx, fy, ph = rng.uniform(0.005, 0.05), rng.uniform(0.005, 0.05), rng.uniform(0, 6.28)
                img[b, n, :, :, c] = 127 + 90 * np.sin(fx * xx + fy * yy + ph)
    img += rng.normal(0, 20, size=img.shape).astype(np.float32)
    return qpos, np.clip(img, 0, 255).astype(np.uint8)


# ---- DP3 observation pipeline (depth image -> point cloud), see make_pointcloud_golden.py
POINTCLOUD_CASES = {
    "small": dict(H=12, W=16, fovy=45.0, seed=31, min_bound=[-0.5, -0.5, 0.4], max_bound=[0.5, 0.5, 1.5], holes=True),
    "img84": dict(H=84, W=84, fovy=58.0, seed=32, min_bound=[-0.6, -0.6, 0.3], max_bound=[0.6, 0.6, 1.6], holes=False),
}


def pointcloud_inputs(case, batch=None):
    """float32 depth (H, W) in metres (zeros = invalid pixels when `holes`) and uint8 rgb (H, W, 3); with `batch` a
    leading batch dimension of independent draws."""
    import numpy as np

    rng = np.random.default_rng(case["seed"])
    shape = (case["H"], case["W"]) if batch is None else (batch, case["H"], case["W"])
    depth = rng.uniform(0.3, 2.0, shape).astype(np.float32)
    if case.get("holes"):
        depth[rng.uniform(size=shape) < 0.1] = 0.0
    rgb = rng.integers(0, 256, shape + (3,), dtype=np.uint8)
    return depth, rgb
