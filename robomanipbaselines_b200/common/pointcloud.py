"""Device-side DP3 observation pipeline: RolloutDiffusionPolicy3d.get_pointcloud
(policy/diffusion_policy_3d/RolloutDiffusionPolicy3d.py:125-171) after its two cv2.resize calls, for B camera frames at
once: convert_depth_image_to_pointcloud (common/utils/VisionUtils.py:55-87) -> crop_pointcloud_bb ->
downsample_pointcloud_fps (common/utils/Vision3dUtils.py:6-25) -> normalize_data -> float32, all in ONE kernel
(rmb_pointcloud_from_depth), bit-exact with the numpy path.
"""
import numpy as np
import torch

from .. import _lib
from .._lib import check, cur_stream, ptr
from .batched_rollout import _norm_params


class DepthToPointcloud:
    def __init__(self, batch, image_size, fovy, num_points, min_bound=None, max_bound=None, pointcloud_stats=None,
                 near_clip=0.0, far_clip=np.inf, device="cuda:0"):
        """image_size = (W, H) as model_meta_info["data"]["image_size"] (cv2 convention); fovy in degrees
        (env.unwrapped.get_camera_fovy); pointcloud_stats = model_meta_info["pointcloud"] or None."""
        if not torch.cuda.is_available():
            raise _lib.RmbError("DepthToPointcloud needs a CUDA device (sm_100a); there is no CPU fallback")
        self.lib = _lib.load()
        self.B, self.device = int(batch), torch.device(device)
        self.W, self.H = int(image_size[0]), int(image_size[1])
        self.K = int(num_points)
        p = _lib.PointcloudParams()
        p.H, p.W, p.num_points = self.H, self.W, self.K
        # the reference's own expression (VisionUtils.py:59), evaluated with numpy so the device divides by the same double
        p.focal_scaling = float((1.0 / np.tan(np.deg2rad(fovy) / 2.0)) * self.H / 2.0)
        p.near_clip, p.far_clip = float(near_clip), float(far_clip)
        p.has_min_bound, p.has_max_bound = int(min_bound is not None), int(max_bound is not None)
        for i in range(3):
            p.min_bound[i] = float(min_bound[i]) if min_bound is not None else 0.0
            p.max_bound[i] = float(max_bound[i]) if max_bound is not None else 0.0
        p.norm_type = -1
        if pointcloud_stats is not None:
            t, p0, p1, om = _norm_params(pointcloud_stats, "normalize")
            if p0.shape != (6,) or p1.shape != (6,):
                raise ValueError("pointcloud statistics must have 6 entries (xyz + rgb)")
            p.norm_type, p.norm_out_min = t, float(om)
            for i in range(6):
                p.norm_p0[i], p.norm_p1[i] = float(p0[i]), float(p1[i])
        self.params = p
        nbytes = check(self.lib.rmb_pointcloud_scratch_bytes(self.B, self.H, self.W, self.K), "rmb_pointcloud_scratch_bytes")
        self.scratch = torch.empty(nbytes, device=self.device, dtype=torch.uint8)
        self.count = torch.zeros(self.B, device=self.device, dtype=torch.int32)
        self.indices = torch.zeros(self.B, self.K, device=self.device, dtype=torch.int32)

    def __call__(self, depth, rgb, out=None):
        """depth float32 (B, H, W) metres, rgb uint8 (B, H, W, 3), on device -> float32 (B, num_points, 6).
        self.count holds the number of points that survived clip + crop, self.indices the FPS indices (-1 padded)."""
        import ctypes

        if tuple(depth.shape) != (self.B, self.H, self.W) or depth.dtype != torch.float32 or not depth.is_cuda:
            raise ValueError(f"depth must be a float32 CUDA tensor ({self.B}, {self.H}, {self.W})")
        if tuple(rgb.shape) != (self.B, self.H, self.W, 3) or rgb.dtype != torch.uint8 or not rgb.is_cuda:
            raise ValueError(f"rgb must be a uint8 CUDA tensor ({self.B}, {self.H}, {self.W}, 3)")
        if out is None:
            out = torch.empty(self.B, self.K, 6, device=self.device, dtype=torch.float32)
        check(self.lib.rmb_pointcloud_from_depth(ptr(depth.contiguous()), ptr(rgb.contiguous()), ctypes.byref(self.params),
                                                 ptr(self.scratch), ptr(out), ptr(self.count), ptr(self.indices), self.B,
                                                 cur_stream()), "rmb_pointcloud_from_depth")
        return out


class Dp3ObservationPipeline:
    """RolloutDiffusionPolicy3d.get_pointcloud (:125-171) for B frames, entirely on the device: cv2.resize of the rgb
    and depth frames to data["image_size"], back-projection, bounding-box crop, farthest point sampling and
    normalisation. Output float32 (B, num_points, 6) is what the rollout pushes into its point-cloud ring buffer."""

    def __init__(self, batch, model_meta_info, fovy, device="cuda:0"):
        data = model_meta_info["data"]
        self.image_size = tuple(int(v) for v in data["image_size"])        # (W, H), cv2 convention
        self.to_cloud = DepthToPointcloud(batch, self.image_size, fovy, data["num_points"], data.get("min_bound"),
                                          data.get("max_bound"), model_meta_info.get("pointcloud"), device=device)

    def __call__(self, rgb_u8, depth_f32):
        """rgb_u8 uint8 (B, H, W, 3), depth_f32 float32 (B, H, W) as the env renders them."""
        from .image_ops import resize_depth, resize_u8

        w, h = self.image_size
        if rgb_u8.shape[1:3] != (h, w):
            rgb_u8 = resize_u8(rgb_u8, self.image_size)
        if depth_f32.shape[1:3] != (h, w):
            depth_f32 = resize_depth(depth_f32, self.image_size)
        return self.to_cloud(depth_f32, rgb_u8)
