"""Golden-generation helper (runs only in the build container, where /root/reference is mounted).

Imports the UNMODIFIED reference sources from /root/reference. Packages the reference needs but this container lacks
(GUI / robotics / dataset stacks that the inference path never executes) are replaced by inert stub modules; the
reference files themselves are not patched or copied.
"""
import importlib.abc
import importlib.machinery
import os
import sys
import types
from unittest import mock

REF_ROOT = os.environ.get("RMB_REFERENCE_ROOT", "/root/reference")

# top-level packages that are absent here and irrelevant to the arithmetic under test
STUB_TOPLEVEL = (
    "IPython", "pinocchio", "matplotlib", "gymnasium", "mujoco", "h5py", "videoio", "torchcodec", "pyspacemouse",
    "pynput", "tensorboard", "tqdm_stub_never", "schedule", "pytorch3d", "sentence_transformers", "diffusers",
    "robomimic", "zarr", "hydra", "omegaconf", "gello", "isaacgym", "pyrealsense2", "serial", "rtde_control",
    "rtde_receive", "xarm", "dm_control", "einops_stub_never", "numcodecs", "av", "imageio",
    "diffusion_policy", "diffusion_policy_3d",
)


class _StubLoader(importlib.abc.Loader):
    def create_module(self, spec):
        m = mock.MagicMock(name=spec.name)
        m.__name__ = spec.name
        m.__path__ = []          # behave like a package so sub-imports resolve through the finder too
        m.__spec__ = spec
        m.__loader__ = self
        return m

    def exec_module(self, module):
        pass


class _StubFinder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path, target=None):
        if fullname.split(".")[0] in STUB_TOPLEVEL:
            return importlib.machinery.ModuleSpec(fullname, _StubLoader(), is_package=True)
        return None


_installed = False


def install():
    """Make `import robo_manip_baselines...`, `import detr...`, `import policy` resolve to /root/reference."""
    global _installed
    if _installed:
        return
    if not os.path.isdir(REF_ROOT):
        raise RuntimeError(f"reference tree not found at {REF_ROOT}")
    missing = []
    for name in STUB_TOPLEVEL:
        try:
            __import__(name)
        except Exception:
            missing.append(name)
    # only stub what is really absent
    global STUB_TOPLEVEL_ACTIVE
    STUB_TOPLEVEL_ACTIVE = tuple(missing)
    finder = _StubFinder()
    finder.find_spec = lambda fullname, path, target=None: (
        importlib.machinery.ModuleSpec(fullname, _StubLoader(), is_package=True)
        if fullname.split(".")[0] in STUB_TOPLEVEL_ACTIVE else None)
    sys.meta_path.append(finder)
    act = os.path.join(REF_ROOT, "third_party", "act")
    for p in (REF_ROOT, act, os.path.join(act, "detr")):
        if p not in sys.path:
            sys.path.insert(0, p)
    _installed = True


def build_reference_detrvae(policy_args, state_dim, action_dim):
    """The reference's DETRVAE built on CPU (third_party/act/detr/models/detr_vae.py:234-259) with the argparse
    defaults of third_party/act/detr/main.py:12-66 overridden exactly as build_ACT_model_and_optimizer does, minus its
    hard-wired .cuda() (main.py:78) and the torchvision weight download (backbone.py:92-94, is_main_process())."""
    install()
    import detr.models.backbone as ref_backbone
    from detr.main import get_args_parser
    from detr.models.detr_vae import DETRVAE, build

    ref_backbone.is_main_process = lambda: False     # random-init ResNet-18 instead of downloading ImageNet weights
    args = get_args_parser().parse_args([])
    for k, v in policy_args.items():
        setattr(args, k, v)
    DETRVAE.set_state_dim(state_dim)
    DETRVAE.set_action_dim(action_dim)
    model = build(args)
    model.eval()
    return model


class ReferenceACTPolicy:
    """ACTPolicy.__call__'s inference branch (third_party/act/policy.py:18-38) around the reference DETRVAE, on CPU:
    the ImageNet Normalize then model(qpos, image, None)[0]. State-dict keys carry the same `model.` prefix."""

    def __init__(self, policy_args, state_dim, action_dim):
        import torch

        self.torch = torch
        self.model = build_reference_detrvae(policy_args, state_dim, action_dim)

    def state_dict(self):
        return {"model." + k: v for k, v in self.model.state_dict().items()}

    def load_state_dict(self, sd):
        self.model.load_state_dict({k[len("model."):]: v for k, v in sd.items()}, strict=True)

    def __call__(self, qpos, image):
        import torchvision.transforms as transforms

        normalize = transforms.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])
        with self.torch.no_grad():
            a_hat, _, _ = self.model(qpos, normalize(image), None)
        return a_hat
