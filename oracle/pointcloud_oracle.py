"""TEST INFRASTRUCTURE ONLY (parity checker; never on the product path).

numpy restatement of the DP3 observation pipeline that precedes the policy
(robo_manip_baselines/policy/diffusion_policy_3d/RolloutDiffusionPolicy3d.py:125-171):
  * convert_depth_image_to_pointcloud   -- robo_manip_baselines/common/utils/VisionUtils.py:55-87
  * crop_pointcloud_bb                  -- robo_manip_baselines/common/utils/Vision3dUtils.py:6-14
  * downsample_pointcloud_fps           -- robo_manip_baselines/common/utils/Vision3dUtils.py:17-25
  * normalize_data + float32 cast       -- RolloutDiffusionPolicy3d.py:152-153

Parity: back-projection and bounding-box crop are PINNED -- tests/golden/make_pointcloud_golden.py executes the
reference's own VisionUtils / Vision3dUtils functions (imported from /root/reference) and tests/test_oracle.py compares
this restatement with the stored outputs bit for bit. Farthest-point sampling is UNPINNED: the reference calls
pytorch3d.ops.sample_farthest_points (not installed here); `farthest_point_sampling` restates its published CPU
algorithm (start at index 0, squared distances accumulated dimension by dimension in the tensor's dtype, running
minimum per point, next = FIRST arg-max, indices beyond the cloud's length are -1).

dtype flow of the reference under NumPy >= 2 (NEP 50), kept here: pixel grid float32, minus 0.5*shape (float32),
divided by the np.float64 scalar focal_scaling -> float64, times depth -> float64; rgb uint8 -> float32 / 255.0 ->
float32; np.concat -> float64 (N, 6); FPS runs on the 6-D float64 points (the reference passes xyz AND rgb columns).
"""
import numpy as np

from .rollout_oracle import normalize_data


def focal_scaling(fovy, height):
    """VisionUtils.py:59 (np.float64 scalar)."""
    return (1.0 / np.tan(np.deg2rad(fovy) / 2.0)) * height / 2.0


def convert_depth_image_to_pointcloud(depth_image, fovy, rgb_image=None, near_clip=0.0, far_clip=np.inf):
    """VisionUtils.py:55-87, vectorised (the reference builds the pixel grid with a Python double loop)."""
    h, w = depth_image.shape[:2]
    fs = focal_scaling(fovy, h)
    ii, jj = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
    xyz_array = np.stack([ii.ravel(), jj.ravel()], axis=1).astype(np.float32)          # (i, j) per pixel, row-major
    xyz_array = (xyz_array - 0.5 * np.array(depth_image.shape[:2], dtype=np.float32)) / fs
    xyz_array *= depth_image.flatten()[:, np.newaxis]
    xyz_array = np.hstack((xyz_array[:, [1, 0]], depth_image.flatten()[:, np.newaxis]))
    clip = np.argwhere((near_clip < depth_image.flatten()) & (depth_image.flatten() < far_clip))[:, 0]
    xyz_array = xyz_array[clip]
    if rgb_image is None:
        return xyz_array
    rgb_array = rgb_image.reshape(-1, 3).astype(np.float32) / 255.0
    return xyz_array, rgb_array[clip]


def crop_pointcloud_bb(pointcloud, min_bound=None, max_bound=None):
    """Vision3dUtils.py:6-14."""
    if min_bound is not None:
        pointcloud = pointcloud[np.all(pointcloud[:, :3] > min_bound, axis=1)]
    if max_bound is not None:
        pointcloud = pointcloud[np.all(pointcloud[:, :3] < max_bound, axis=1)]
    return pointcloud


def farthest_point_sampling(points, num_points):
    """pytorch3d.ops.sample_farthest_points (CPU path, random_start_point=False) for one cloud -> indices (K,), -1 padded."""
    n, dim = points.shape
    idx = np.full(num_points, -1, dtype=np.int64)
    if n == 0:
        return idx
    dists = np.full(n, np.inf, dtype=points.dtype)
    sel = 0
    idx[0] = 0
    for k in range(1, min(num_points, n)):
        d2 = np.zeros(n, dtype=points.dtype)
        for d in range(dim):                                   # accumulated dimension by dimension, no FMA
            diff = points[:, d] - points[sel, d]
            d2 = d2 + diff * diff
        dists = np.minimum(dists, d2)
        sel = int(np.argmax(dists))                            # first arg-max
        idx[k] = sel
    return idx


def downsample_pointcloud_fps(pointcloud, num_points=512):
    """Vision3dUtils.py:17-25: pointcloud[sampled_indices] (a -1 index picks the last point, as numpy indexing does)."""
    return pointcloud[farthest_point_sampling(pointcloud, num_points)]


def get_pointcloud(depth_image, rgb_image, fovy, min_bound, max_bound, num_points, pointcloud_stats=None):
    """RolloutDiffusionPolicy3d.get_pointcloud (:125-153) after the cv2.resize calls -> float32 (num_points, 6)."""
    pc = np.concat(convert_depth_image_to_pointcloud(depth_image, fovy, rgb_image), axis=1)
    pc = crop_pointcloud_bb(pc, min_bound, max_bound)
    n_valid = pc.shape[0]
    pc = downsample_pointcloud_fps(pc, num_points)
    if pointcloud_stats is not None:
        pc = normalize_data(pc, pointcloud_stats)
    return pc.astype(np.float32), n_valid
