"""Host side of the pyngp shell that needs no GPU: the module imports under the
reference's name, its enums carry the reference's values (python_api.cu:268-352, common.h:79-160),
the image-mode file loader follows Testbed::load_image (src/testbed_image.cu:362-428), and a Testbed
cannot be made without a device (no CPU fallback)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def pyngp():
    sys.path.insert(0, os.path.join(ROOT, "instant-ngp-rendering_b200"))
    import pyngp as mod
    return mod


def test_enums(pyngp):
    want = {"TestbedMode": {"Nerf": 0, "Sdf": 1, "Image": 2, "Volume": 3, "None": 5},
            "LossType": {"L2": 0, "L1": 1, "Mape": 2, "Smape": 3, "Huber": 4, "LogL1": 5, "RelativeL2": 6},
            "NerfActivation": {"None": 0, "ReLU": 1, "Logistic": 2, "Exponential": 3},
            "ColorSpace": {"Linear": 0, "SRGB": 1},
            "TonemapCurve": {"Identity": 0, "ACES": 1, "Hable": 2, "Reinhard": 3},
            "LensMode": {"Perspective": 0, "OpenCV": 1, "FTheta": 2, "LatLong": 3, "OpenCVFisheye": 4, "Equirectangular": 5}}
    for enum_name, members in want.items():
        e = getattr(pyngp, enum_name)
        assert {m.name: int(m) for m in e} == members, enum_name
    assert pyngp.LossType.SmoothL1 is pyngp.LossType.Huber


def test_testbed_needs_a_device(pyngp):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    with pytest.raises(pyngp.NgpbError, match="no CUDA device"):
        pyngp.Testbed()


def test_image_file_loader(tmp_path):
    import importlib
    loader = importlib.import_module("instant-ngp-rendering_b200.loader")
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(0)
    rgba = rng.integers(0, 256, size=(5, 7, 4), dtype=np.uint8)
    p = str(tmp_path / "img.png")
    assert cv2.imwrite(p, rgba[:, :, [2, 1, 0, 3]])
    got = loader.load_image_file(p)
    assert got.dtype == np.float32 and got.shape == (5, 7, 4)
    s = rgba[..., :3].astype(np.float32) / np.float32(255.0)
    lin = np.where(s <= 0.04045, s / 12.92, ((s + 0.055) / 1.055) ** 2.4)
    a = rgba[..., 3:4].astype(np.float32) / np.float32(255.0)
    np.testing.assert_allclose(got[..., :3], lin * a, rtol=1e-5, atol=1e-7)   # from_rgba32: linear, premultiplied
    np.testing.assert_allclose(got[..., 3:], a, atol=1e-7)
    # opaque 3-channel file
    p3 = str(tmp_path / "img.jpg")
    assert cv2.imwrite(p3, rgba[:, :, :3])
    assert (loader.load_image_file(p3)[..., 3] == 1.0).all()
    # *.bin: int32 height, int32 width, fp16 RGBA texels (load_binary_image)
    px = rng.random((3, 4, 4)).astype(np.float16)
    pb = str(tmp_path / "img.bin")
    with open(pb, "wb") as f:
        f.write(np.array([3, 4], dtype=np.int32).tobytes())
        f.write(px.tobytes())
    got = loader.load_image_file(pb)
    assert got.dtype == np.float16 and np.array_equal(got, px)
    with open(pb, "wb") as f:
        f.write(np.array([3, 4], dtype=np.int32).tobytes())
        f.write(px.tobytes()[:40])
    with pytest.raises(ValueError, match="truncated"):
        loader.load_image_file(pb)
    with pytest.raises(ValueError, match="does not exist"):
        loader.load_image_file(str(tmp_path / "missing.png"))
