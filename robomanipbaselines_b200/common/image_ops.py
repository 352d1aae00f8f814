"""Host handles of the device image kernels that replace the reference's per-frame OpenCV / torchvision calls:
cv2.resize (policy/diffusion_policy/RolloutDiffusionPolicy.py:117, policy/diffusion_policy_3d/RolloutDiffusionPolicy3d.py:134-137)
and the uint8 -> float32 scaling of RolloutBase.image_transforms (common/base/RolloutBase.py:353)."""
import torch

from .. import _lib
from .._lib import check, cur_stream, ptr


def _check_u8(images):
    if images.dtype != torch.uint8 or images.dim() != 4 or images.shape[-1] != 3 or not images.is_cuda:
        raise ValueError("images must be a uint8 CUDA tensor (n, H, W, 3)")
    return images.contiguous()


def resize_u8(images, size, out=None):
    """cv2.resize(img, size) per frame: uint8 (n, H, W, 3) -> uint8 (n, size[1], size[0], 3); size = (W2, H2) as cv2."""
    images = _check_u8(images)
    n, H, W, _ = images.shape
    W2, H2 = int(size[0]), int(size[1])
    if out is None:
        out = torch.empty(n, H2, W2, 3, device=images.device, dtype=torch.uint8)
    check(_lib.load().rmb_resize_u8_bilinear(ptr(images), n, H, W, H2, W2, 0, 1.0, 0.0, ptr(out), cur_stream()),
          "rmb_resize_u8_bilinear")
    return out


def resize_u8_to_policy_input(images, size, scale=2.0, offset=-1.0, out=None):
    """RolloutDiffusionPolicy.get_images (:113-125) per frame in one kernel: cv2.resize -> HWC to CHW -> uint8 to float32
    with 1/255 scaling -> `image * 2.0 - 1.0`. uint8 (n, H, W, 3) -> float32 (n, 3, H2, W2)."""
    images = _check_u8(images)
    n, H, W, _ = images.shape
    W2, H2 = int(size[0]), int(size[1])
    if out is None:
        out = torch.empty(n, 3, H2, W2, device=images.device, dtype=torch.float32)
    check(_lib.load().rmb_resize_u8_bilinear(ptr(images), n, H, W, H2, W2, 1, float(scale), float(offset), ptr(out),
                                             cur_stream()), "rmb_resize_u8_bilinear")
    return out


def resize_depth(depth, size, out=None):
    """cv2.resize(depth_image, size) per frame: float32 (n, H, W) -> float32 (n, size[1], size[0])."""
    if depth.dtype != torch.float32 or depth.dim() != 3 or not depth.is_cuda:
        raise ValueError("depth must be a float32 CUDA tensor (n, H, W)")
    depth = depth.contiguous()
    n, H, W = depth.shape
    W2, H2 = int(size[0]), int(size[1])
    if out is None:
        out = torch.empty(n, H2, W2, device=depth.device, dtype=torch.float32)
    check(_lib.load().rmb_resize_f32_bilinear(ptr(depth), n, H, W, H2, W2, ptr(out), cur_stream()), "rmb_resize_f32_bilinear")
    return out
