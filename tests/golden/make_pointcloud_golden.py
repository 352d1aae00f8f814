#!/usr/bin/env python
"""Golden vectors for the DP3 observation pipeline, made by executing the reference's OWN functions
(robo_manip_baselines/common/utils/VisionUtils.py:55-87, Vision3dUtils.py:6-14) from /root/reference (build container
only). Inputs are regenerated from the seeds in cases.POINTCLOUD_CASES; only outputs are stored.

    python tests/golden/make_pointcloud_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import ref_loader  # noqa: E402
from cases import POINTCLOUD_CASES, pointcloud_inputs  # noqa: E402


def main():
    ref_loader.install()
    from robo_manip_baselines.common.utils import Vision3dUtils, VisionUtils

    out = {}
    for name, case in POINTCLOUD_CASES.items():
        depth, rgb = pointcloud_inputs(case)
        xyz, col = VisionUtils.convert_depth_image_to_pointcloud(depth, case["fovy"], rgb)
        pc = np.concat((xyz, col), axis=1)
        cropped = Vision3dUtils.crop_pointcloud_bb(pc, case["min_bound"], case["max_bound"])
        out[f"{name}_pointcloud"] = pc
        out[f"{name}_cropped"] = cropped
        print(name, pc.dtype, pc.shape, cropped.shape)
    np.savez_compressed(os.path.join(HERE, "pointcloud_golden.npz"), **out)


if __name__ == "__main__":
    main()
