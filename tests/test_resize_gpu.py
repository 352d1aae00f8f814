"""Device cv2.resize against cv2 itself (the very dependency the reference calls: RolloutDiffusionPolicy.py:117,
RolloutDiffusionPolicy3d.py:134-137): uint8 bit-exact for the reference's down-scaling, float32 depth within 1e-4."""
import numpy as np
import pytest
import torch

cv2 = pytest.importorskip("cv2")
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("hw,size", [((480, 640), (320, 240)), ((480, 640), (84, 84)), ((480, 640), (224, 224)),
                                     ((123, 77), (50, 31)), ((240, 320), (320, 240))])
def test_resize_u8_matches_cv2(cuda, hw, size):
    from robomanipbaselines_b200.common.image_ops import resize_u8, resize_u8_to_policy_input

    rng = np.random.default_rng(41)
    imgs = rng.integers(0, 256, (3,) + hw + (3,), dtype=np.uint8)
    want = np.stack([cv2.resize(im, size) for im in imgs])
    got = resize_u8(torch.from_numpy(imgs).to(cuda), size).cpu().numpy()
    assert got.shape == want.shape and np.array_equal(got, want)
    # the fused policy-input form: (cv2.resize -> CHW -> float32 * (1/255)) * 2 - 1, as the reference computes it
    t = torch.from_numpy(np.moveaxis(want, -1, -3)).to(torch.float32) * (1.0 / 255.0)
    want_f = t * 2.0 - 1.0
    got_f = resize_u8_to_policy_input(torch.from_numpy(imgs).to(cuda), size).cpu()
    assert torch.equal(got_f, want_f)


@pytest.mark.parametrize("hw,size", [((480, 640), (84, 84)), ((480, 640), (320, 240)), ((123, 77), (50, 31))])
def test_resize_depth_matches_cv2(cuda, hw, size):
    from robomanipbaselines_b200.common.image_ops import resize_depth

    rng = np.random.default_rng(42)
    d = rng.uniform(0.3, 2.0, (2,) + hw).astype(np.float32)
    want = np.stack([cv2.resize(x, size) for x in d])
    got = resize_depth(torch.from_numpy(d).to(cuda), size).cpu().numpy()
    err = np.abs(got - want).max()
    print(f"[resize depth {hw}->{size}] max-abs err {err:.3e}")
    assert got.shape == want.shape and err <= 1e-4
