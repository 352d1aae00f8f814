"""ctypes binding of librmb_b200.so (the C ABI in include/rmb_b200.h).

There is no CPU fallback: if the shared library is missing, loading raises, and every entry point that needs a GPU
returns an error code that is turned into an exception here.
"""
import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_int32, c_int64, c_uint8, c_void_p

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
# Product default: the fp16-operand build. RMB_LIB_DTYPE=bf16 (read once, at the first load) selects the bf16-operand
# build of the same sources instead (librmb_b200_bf16.so); one process uses exactly one of them.
LIB_DTYPE = os.environ.get("RMB_LIB_DTYPE", "fp16")
if LIB_DTYPE not in ("fp16", "bf16"):
    raise ValueError(f"RMB_LIB_DTYPE must be fp16 or bf16, got {LIB_DTYPE!r}")
LIB_PATH = os.path.join(_PKG_DIR, "librmb_b200.so" if LIB_DTYPE == "fp16" else "librmb_b200_bf16.so")


class RmbError(RuntimeError):
    pass


class ActConfig(ctypes.Structure):
    """Mirror of rmb_act_config_t."""

    _fields_ = [
        ("max_batch", c_int32), ("num_cams", c_int32), ("img_h", c_int32), ("img_w", c_int32),
        ("state_dim", c_int32), ("action_dim", c_int32), ("hidden_dim", c_int32), ("dim_feedforward", c_int32),
        ("nheads", c_int32), ("enc_layers", c_int32), ("dec_layers", c_int32), ("chunk_size", c_int32),
        ("device", c_int32), ("multi_task", c_int32), ("task_emb_dim", c_int32), ("task_token", c_int32),
    ]


class LaunchProf(ctypes.Structure):
    """Mirror of rmb_launch_prof_t."""

    _fields_ = [("name", ctypes.c_char * 48), ("kernel", ctypes.c_char * 40), ("ms", c_float), ("flops", c_double),
                ("bytes", c_double)]


class UnetConfig(ctypes.Structure):
    """Mirror of rmb_unet_config_t."""

    _fields_ = [
        ("max_batch", c_int32), ("horizon", c_int32), ("input_dim", c_int32), ("global_cond_dim", c_int32),
        ("dsed", c_int32), ("n_levels", c_int32), ("down_dims", c_int32 * 4), ("kernel_size", c_int32),
        ("n_groups", c_int32), ("num_train_timesteps", c_int32), ("device", c_int32),
    ]


class SchedStep(ctypes.Structure):
    """Mirror of rmb_sched_step_t."""

    _fields_ = [
        ("t", c_int32), ("kind", c_int32), ("pred_sample", c_int32), ("clip", c_int32), ("derive_eps", c_int32),
        ("s_beta", c_float), ("s_alpha", c_float), ("c_x0", c_float), ("c_xt", c_float), ("sigma", c_float),
    ]


class PointNetConfig(ctypes.Structure):
    """Mirror of rmb_pointnet_config_t."""

    _fields_ = [
        ("max_clouds", c_int32), ("num_points", c_int32), ("in_channels", c_int32), ("out_channels", c_int32),
        ("state_dim", c_int32), ("device", c_int32),
    ]


class ImgEncConfig(ctypes.Structure):
    """Mirror of rmb_imgenc_config_t."""

    _fields_ = [
        ("max_images", c_int32), ("in_h", c_int32), ("in_w", c_int32), ("crop_h", c_int32), ("crop_w", c_int32),
        ("num_kp", c_int32), ("feat_dim", c_int32), ("device", c_int32),
    ]


class PointcloudParams(ctypes.Structure):
    """Mirror of rmb_pointcloud_params_t."""

    _fields_ = [
        ("H", c_int32), ("W", c_int32), ("num_points", c_int32), ("focal_scaling", c_double), ("near_clip", c_double),
        ("far_clip", c_double), ("has_min_bound", c_int32), ("has_max_bound", c_int32), ("min_bound", c_double * 3),
        ("max_bound", c_double * 3), ("norm_type", c_int32), ("norm_p0", c_double * 6), ("norm_p1", c_double * 6),
        ("norm_out_min", c_double),
    ]


ARM_MAX_JOINTS, ARM_MAX_GRIPPER = 8, 4


class ArmModel(ctypes.Structure):
    """Mirror of rmb_arm_model_t."""

    _fields_ = [("n_joints", c_int32), ("ik_joint_id", c_int32), ("placement_R", (c_double * 9) * ARM_MAX_JOINTS),
                ("placement_p", (c_double * 3) * ARM_MAX_JOINTS), ("axis", (c_double * 3) * ARM_MAX_JOINTS)]


class ArmLayout(ctypes.Structure):
    """Mirror of rmb_arm_layout_t."""

    _fields_ = [("n_gripper", c_int32), ("eef_idx", c_int32), ("arm_idx", c_int32 * ARM_MAX_JOINTS),
                ("gripper_idx", c_int32 * ARM_MAX_GRIPPER), ("gripper_idx_in_gripper", c_int32 * ARM_MAX_GRIPPER),
                ("init_arm", c_double * ARM_MAX_JOINTS), ("init_gripper", c_double * ARM_MAX_GRIPPER),
                ("gripper_low", c_double * ARM_MAX_GRIPPER), ("gripper_high", c_double * ARM_MAX_GRIPPER)]


# name -> (restype, argtypes); must list every symbol include/rmb_b200.h declares (tests check this)
SIGNATURES = {
    "rmb_last_error": (c_char_p, []),
    "rmb_version": (c_char_p, []),
    "rmb_act_dtype": (c_char_p, []),
    "rmb_check_device": (c_int, [c_int]),
    "rmb_device_fault": (c_int, [c_int]),
    "rmb_act_set_attention_debug": (c_int, [c_void_p, c_int, c_int]),
    "rmb_act_read_attention_maps": (c_int64, [c_void_p, c_void_p, c_void_p]),
    "rmb_act_set_check_saturation": (c_int, [c_void_p, c_int]),
    "rmb_saturation_count": (c_int64, [c_int]),
    "rmb_act_engine_create": (c_int, [POINTER(ActConfig), POINTER(c_void_p)]),
    "rmb_act_engine_destroy": (None, [c_void_p]),
    "rmb_act_load_tensor": (c_int, [c_void_p, c_char_p, c_void_p, c_int64]),
    "rmb_act_finalize": (c_int, [c_void_p]),
    "rmb_act_set_task_embedding": (c_int, [c_void_p, c_void_p, c_int, c_void_p]),
    "rmb_act_forward": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p]),
    "rmb_act_forward_profile": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p,
                                        POINTER(LaunchProf), c_int]),
    "rmb_act_forward_num_launches": (c_int, [c_void_p]),
    "rmb_act_forward_flops": (c_double, [c_void_p, c_int]),
    "rmb_act_debug_read": (c_int64, [c_void_p, c_int, c_void_p, c_int, c_void_p]),
    "rmb_ensemble_step": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p,
                                  c_double, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "rmb_normalize": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_double, c_void_p, c_int64, c_int, c_void_p]),
    "rmb_denormalize": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_double, c_void_p, c_int64, c_int, c_void_p]),
    "rmb_preprocess_images_u8": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "rmb_preprocess_images_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "rmb_gemm": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int,
                              c_void_p]),
    "rmb_gemm_residual_layernorm": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                            c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "rmb_conv2d_nhwc": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                                     c_int, c_int, c_int, c_int, c_void_p]),
    "rmb_stem_pool": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "rmb_conv1_stem": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "rmb_maxpool3x3s2_nhwc": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "rmb_attention": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_int, c_int, c_void_p, c_int, c_int,
                                   c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "rmb_set_attention_variant": (c_int, [c_int]),
    "rmb_layernorm512": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p,
                                 c_int64, c_int, c_void_p]),
    "rmb_layernorm512_act": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p,
                                 c_int64, c_int, c_void_p]),
    "rmb_scheduler_step": (c_int, [c_void_p, c_void_p, c_void_p, POINTER(SchedStep), c_void_p, c_int64, c_void_p]),
    "rmb_unet_engine_create": (c_int, [POINTER(UnetConfig), POINTER(c_void_p)]),
    "rmb_unet_engine_destroy": (None, [c_void_p]),
    "rmb_unet_load_tensor": (c_int, [c_void_p, c_char_p, c_void_p, c_int64]),
    "rmb_unet_finalize": (c_int, [c_void_p]),
    "rmb_unet1d_denoise": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, POINTER(SchedStep), c_int, c_void_p, c_int,
                                   c_void_p]),
    "rmb_unet1d_forward": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p]),
    "rmb_unet_launches_per_step": (c_int, [c_void_p]),
    "rmb_unet_step_flops": (c_double, [c_void_p, c_int]),
    "rmb_unet1d_forward_profile": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p,
                                           POINTER(LaunchProf), c_int]),
    "rmb_pointnet_engine_create": (c_int, [POINTER(PointNetConfig), POINTER(c_void_p)]),
    "rmb_pointnet_engine_destroy": (None, [c_void_p]),
    "rmb_pointnet_load_tensor": (c_int, [c_void_p, c_char_p, c_void_p, c_int64]),
    "rmb_pointnet_finalize": (c_int, [c_void_p]),
    "rmb_pointnet_encode": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p]),
    "rmb_imgenc_engine_create": (c_int, [POINTER(ImgEncConfig), POINTER(c_void_p)]),
    "rmb_imgenc_engine_destroy": (None, [c_void_p]),
    "rmb_imgenc_load_tensor": (c_int, [c_void_p, c_char_p, c_void_p, c_int64]),
    "rmb_imgenc_finalize": (c_int, [c_void_p]),
    "rmb_imgenc_encode": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "rmb_imgenc_debug_read": (c_int64, [c_void_p, c_void_p, c_int, c_void_p]),
    "rmb_imgenc_flops": (c_double, [c_void_p, c_int]),
    "rmb_resize_u8_bilinear": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_float, c_float, c_void_p,
                                       c_void_p]),
    "rmb_resize_f32_bilinear": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "rmb_pointcloud_scratch_bytes": (c_int64, [c_int, c_int, c_int, c_int]),
    "rmb_pointcloud_from_depth": (c_int, [c_void_p, c_void_p, POINTER(PointcloudParams), c_void_p, c_void_p, c_void_p,
                                          c_void_p, c_int, c_void_p]),
    "rmb_pointcloud_from_depth_static": (c_int, [c_void_p, c_void_p, POINTER(PointcloudParams), c_void_p, c_void_p, c_void_p,
                                                 c_int, c_void_p]),
    "rmb_arm_command": (c_int, [POINTER(ArmModel), POINTER(ArmLayout), c_int, c_void_p, c_int, c_int, c_void_p, c_void_p,
                                c_void_p, c_void_p, c_int, c_void_p]),
    "rmb_arm_get_command": (c_int, [POINTER(ArmModel), POINTER(ArmLayout), c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                    c_void_p, c_int, c_void_p]),
    "rmb_obs_ring_push": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int64, c_void_p]),
    "rmb_gather_rows": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int64, c_void_p]),
    "rmb_action_fifo_step": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int, c_void_p, c_void_p,
                                     c_double, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
}

_lib = None


def load():
    """dlopen librmb_b200.so and attach signatures. Raises RmbError when the library was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RmbError(
            f"{LIB_PATH} not found: build it with `python -m robomanipbaselines_b200.build` "
            "(there is no CPU fallback for the CUDA path)"
        )
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def check(code, what=""):
    """Turn a negative return code into an exception carrying rmb_last_error()."""
    if code is not None and code < 0:
        msg = load().rmb_last_error()
        raise RmbError(f"{what} failed ({code}): {msg.decode() if msg else '?'}")
    return code


def ptr(t):
    """Device/host pointer of a torch tensor (None -> NULL)."""
    if t is None:
        return None
    return ctypes.c_void_p(t.data_ptr())


def act_torch_dtype():
    """torch dtype of the library's 16-bit operand type (rmb_act_dtype())."""
    import torch

    return {"fp16": torch.float16, "bf16": torch.bfloat16}[load().rmb_act_dtype().decode()]


def cur_stream():
    import torch

    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
