"""transforms.json loader (instant-ngp-rendering_b200/loader.py) against the rules of the reference's
load_nerf (src/nerf_loader.cu:175-271, 297-720): frame order and culling, focal-length keys, lens
keys, scale / offset / aabb, the byte-image fix-ups. Host logic: runs without a GPU."""
import importlib
import json
import os

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")
L = importlib.import_module("instant-ngp-rendering_b200.loader")


def _write_png(path, rgba):
    os.makedirs(os.path.dirname(path), exist_ok=True)
    a = np.asarray(rgba, dtype=np.uint8)
    assert cv2.imwrite(path, a[:, :, [2, 1, 0, 3]] if a.shape[2] == 4 else a[:, :, ::-1])


def _pose(tx, ty, tz):
    m = np.eye(4)
    m[:3, 3] = (tx, ty, tz)
    return m.tolist()


def _scene(tmp_path, extra=None, frames=None, n=4):
    rng = np.random.default_rng(0)
    root = str(tmp_path)
    names = ["images/img_10", "images/img_2", "images/img_1", "images\\\\img_3"][:n]
    out = []
    for k, name in enumerate(names):
        px = rng.integers(1, 255, size=(6, 8, 4), dtype=np.uint8)
        _write_png(os.path.join(root, name.replace("\\\\", "/") + ".png"), px)
        out.append({"file_path": name.replace("\\\\", "\\"), "transform_matrix": _pose(0.1 * k, 0.2, 1.0 + k)})
    j = {"camera_angle_x": 0.6911112070083618, "w": 8, "h": 6, "aabb_scale": 4, "frames": frames if frames is not None else out}
    j.update(extra or {})
    text = "// a comment, as nlohmann's ignore_comments allows\n" + json.dumps(j, indent=1)
    with open(os.path.join(root, "transforms.json"), "w") as f:
        f.write(text)
    return root, j


def test_focal_length_keys():
    # the classic lego value: camera_angle_x 0.6911 rad at 800 px -> 1111.11 px
    fx, fy = L.read_focal_length({"camera_angle_x": 0.6911112070083618}, (800, 800))
    assert abs(fx - 1111.111) < 0.01 and fy == fx
    assert L.read_focal_length({"x_fov": 90.0}, (100, 50)) == (pytest.approx(50.0, rel=1e-6),) * 2          # degrees
    assert L.read_focal_length({"fl_x": 10.0, "fl_y": 12.0}, (1, 1)) == (10.0, 12.0)
    assert L.read_focal_length({"camera_angle_y": np.pi / 2}, (100, 50)) == (pytest.approx(25.0, rel=1e-6),) * 2   # y alone sets both
    assert L.read_focal_length({"x_fov": 90.0, "fl_x": 7.0}, (100, 50))[0] == pytest.approx(50.0, rel=1e-6)        # _fov wins over fl_
    assert L.read_focal_length({}, (10, 10)) is None


def test_lens_keys_and_unsupported_modes():
    lens = {"mode": 0, "params": [0.0] * 7, "principal": [0.5, 0.5]}
    L.read_lens({"k1": 0.0, "k2": 0.0, "cx": 4.0, "cy": 1.5, "w": 8, "h": 6}, lens)
    assert lens["mode"] == 0 and lens["principal"] == [0.5, 0.25]        # all-zero distortion stays Perspective
    L.read_lens({"k1": 0.05, "k2": -0.02, "p1": 0.001, "p2": -0.002}, lens)
    assert lens["mode"] == 1 and np.allclose(lens["params"][:4], [0.05, -0.02, 0.001, -0.002])
    for bad in ({"is_fisheye": True, "k1": 0.1}, {"ftheta_p0": 1.0}, {"latlong": True}, {"equirectangular": True}, {"rolling_shutter": [0, 0, 1, 0]}):
        with pytest.raises(ValueError):
            L.read_lens(bad, {"mode": 0, "params": [0.0] * 7, "principal": [0.5, 0.5]})


def test_frame_order_defaults_and_matrix_convention(tmp_path):
    root, j = _scene(tmp_path)
    ds = L.describe_dataset(root)                                      # a directory is searched for *.json
    assert ds["n_images"] == 4 and ds["aabb_scale"] == 4 and ds["scale"] == pytest.approx(0.33) and ds["offset"] == [0.5] * 3
    # natural order (2 before 10); the reference sorts BEFORE it turns '\\' into '/' (:347-357), so the Windows path sorts last
    assert [os.path.basename(f["path"]) for f in ds["frames"]] == ["img_1.png", "img_2.png", "img_10.png", "img_3.png"]
    # nerf_matrix_to_ngp: flip y / z columns, scale + offset the origin, cycle the rows xyz <- yzx
    f = ds["frames"][1]                                                # img_2: translation (0.1, 0.2, 2.0)
    t = np.array([0.1, 0.2, 2.0], dtype=np.float32) * np.float32(0.33) + np.float32(0.5)
    want = np.array([[0, -1, 0, t[1]], [0, 0, -1, t[2]], [1, 0, 0, t[0]]], dtype=np.float32)
    assert np.array_equal(f["xform"], want)
    assert L.finish_frame(f, 8, 6)[0] == pytest.approx(0.5 * 8 / np.tan(0.5 * 0.6911112070083618), rel=1e-6)
    px = L.load_frame_pixels(f, ds)
    assert px.dtype == np.uint8 and px.shape == (6, 8, 4)


def test_culling_scale_offset_aabb_and_fixups(tmp_path):
    rng = np.random.default_rng(1)
    root = str(tmp_path)
    frames = []
    for k in range(6):
        px = rng.integers(1, 255, size=(4, 4, 4), dtype=np.uint8)
        px[0, 0, :3] = 255                                            # a pure white pixel
        _write_png(os.path.join(root, "f%d.png" % k), px)
        frames.append({"file_path": "f%d" % k, "transform_matrix": _pose(k, 0, 0), "sharpness": [10, 10, 1, 10, 10, 10][k]})   # no extension: resolved
    frames[3]["fl_x"] = 123.0                                          # per-frame override
    _write_png(os.path.join(root, "f1.alpha.png"), np.full((4, 4, 4), 128, dtype=np.uint8))
    mask = np.zeros((4, 4, 4), dtype=np.uint8)
    mask[1, 2, 0] = 9
    _write_png(os.path.join(root, "dynamic_mask_f4.png"), mask)
    j = {"fl_x": 50.0, "fl_y": 60.0, "cx": 1.0, "cy": 3.0, "w": 4, "h": 4, "k1": 0.01, "frames": frames, "n_frames": 5,
         "sharpness_discard_threshold": 0.5, "aabb": [[-1, -2, -3], [3, 2, 5]], "white_transparent": True, "up": [0, 0, 1]}
    with open(os.path.join(root, "t.json"), "w") as f:
        json.dump(j, f)
    ds = L.describe_dataset(os.path.join(root, "t.json"))
    # n_frames keeps 5, the blurry one (sharpness 1 against a neighbourhood mean of ~8) goes
    assert [os.path.basename(f["path"]) for f in ds["frames"]] == ["f0.png", "f1.png", "f3.png", "f4.png"]
    # aabb -> isotropic scale 1 / longest edge, centre mapped to 0.5
    assert ds["scale"] == pytest.approx(1 / 8) and np.allclose(ds["offset"], [0.5 - 1 / 8, 0.5, 0.5 - 1 / 8])
    assert ds["up"] == [0.0, 1.0, 0.0]
    f3 = next(f for f in ds["frames"] if f["path"].endswith("f3.png"))
    # a per-frame fl_x alone overrides BOTH components (read_focal_length sets vec2(x_fl), :260-264)
    assert L.finish_frame(f3, 4, 4) == (123.0, 123.0) and L.finish_frame(ds["frames"][0], 4, 4) == (50.0, 60.0)
    assert f3["lens_mode"] == 1 and f3["principal"] == [0.25, 0.75]
    p1 = L.load_frame_pixels(ds["frames"][1], ds)                     # alpha file: red channel, sRGB -> linear
    assert (p1[1:, :, 3] == int(255 * (((128 / 255 + 0.055) / 1.055) ** 2.4))).all()
    assert p1[0, 0, 3] == 0 and (p1[0, 0, :3] == 255).all()           # white_transparent
    p4 = L.load_frame_pixels(next(f for f in ds["frames"] if f["path"].endswith("f4.png")), ds)
    assert p4[1, 2].tolist() == [255, 0, 255, 0]                       # the mask colour the trainer skips (0x00FF00FF)


def test_errors(tmp_path):
    root, _ = _scene(tmp_path, frames=[])
    with pytest.raises(ValueError, match="No training images"):
        L.describe_dataset(root)
    root, _ = _scene(tmp_path, frames=[{"file_path": "images/nope", "transform_matrix": _pose(0, 0, 0)}])
    with pytest.raises(FileNotFoundError, match="Could not find image file"):
        L.describe_dataset(root)
    root, j = _scene(tmp_path, n=2)
    del j["camera_angle_x"]
    with open(os.path.join(root, "transforms.json"), "w") as f:
        json.dump(j, f)
    ds = L.describe_dataset(root)
    with pytest.raises(ValueError, match="Couldn't read fov"):
        L.finish_frame(ds["frames"][0], 8, 6)


@pytest.mark.gpu
def test_load_training_data_trains(tmp_path):
    ngp = importlib.import_module("instant-ngp-rendering_b200")
    syn = importlib.import_module("instant-ngp-rendering_b200.synthetic")
    import torch
    src = ngp.Testbed()
    scene, poses, focal = syn.make_dataset(src, n_images=6, width=64, height=48, aabb_scale=1)
    frames = []
    for i, pose in enumerate(poses):
        img = syn.linear_premul_to_srgb8(scene.render(pose, 64, 48, focal)).cpu().numpy()
        _write_png(os.path.join(str(tmp_path), "images", "%04d.png" % i), img)
        # back to the NeRF convention: the inverse of nerf_matrix_to_ngp
        m = np.asarray(pose, dtype=np.float64)[[2, 0, 1], :]
        m[:, 3] = (m[:, 3] - 0.5) / 0.33
        m[:, 1] *= -1
        m[:, 2] *= -1
        frames.append({"file_path": "images/%04d.png" % i, "transform_matrix": np.vstack([m, [0, 0, 0, 1]]).tolist()})
    with open(os.path.join(str(tmp_path), "transforms.json"), "w") as f:
        json.dump({"fl_x": focal, "fl_y": focal, "w": 64, "h": 48, "aabb_scale": 1, "frames": frames}, f)
    src.close()
    tb = ngp.Testbed()
    ds = tb.load_training_data(str(tmp_path))
    assert ds["n_images"] == 6
    for fr, pose in zip(ds["frames"], poses):
        np.testing.assert_allclose(fr["xform"], np.asarray(pose, dtype=np.float32)[:3, :4], atol=2e-6)
    tb.reload_network_from_json(dict(syn.NERF_BASE_CONFIG))
    first = None
    for step in range(97):
        tb.train(1 << 14)
        if step == 16:
            first = tb.loss
    assert tb.loss < 0.6 * first, (first, tb.loss)
    tb.close()
