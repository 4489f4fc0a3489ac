"""The C++ loader's decisions (csrc/host_dataset.cu, through ngpb_host_describe_dataset: no device, no pixels) against
loader.describe_dataset - the restatement of src/nerf_loader.cu:273-740 that is itself held to the reference's own loader output
(tests/test_ref_fixture_cpu.py) - on datasets that exercise every rule: natural frame order, n_frames, sharpness culling,
extension search, alpha / mask side files, per-frame lens override, scale / offset (scalar and vector) / aabb / up, Mitsuba
axes, several transforms files in one directory, the reference's own fox scene."""
import ctypes
import importlib
import json
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _describe(lib, path):
    n = lib.ngpb_host_describe_dataset(str(path).encode(), None, 0)
    assert n > 0, lib.ngpb_last_error().decode()
    buf = ctypes.create_string_buffer(n)
    lib.ngpb_host_describe_dataset(str(path).encode(), buf, n)
    return json.loads(buf.value.decode())


def _compare(got, want, loader, w=None, h=None):
    assert got["n_images"] == want["n_images"] == len(want["frames"])
    assert got["aabb_scale"] == want["aabb_scale"]
    for key in ("from_mitsuba", "is_hdr", "white_transparent", "black_transparent", "fix_premult"):
        assert got[key] == bool(want[key]), key
    assert np.float32(got["scale"]) == np.float32(want["scale"])
    assert np.array_equal(np.float32(got["offset"]), np.float32(want["offset"]))
    assert np.array_equal(np.float32(got["up"]), np.float32(want["up"]))
    if want["render_aabb"] is None:
        assert "render_aabb" not in got
    else:
        assert np.array_equal(np.float32(got["render_aabb"]), np.float32(want["render_aabb"]).reshape(-1))
    for i, (g, f) in enumerate(zip(got["frames"], want["frames"])):
        assert g["path"] == f["path"], i
        assert g["alpha_path"] == (f["alpha_path"] or "") and g["mask_path"] == (f["mask_path"] or ""), i
        assert np.array_equal(np.float32(g["xform"]).reshape(3, 4).view(np.uint32), np.float32(f["xform"]).view(np.uint32)), (i, g["xform"], f["xform"])
        assert g["lens_mode"] == f["lens_mode"]
        assert np.array_equal(np.float32(g["lens_params"]), np.float32(f["lens_params"])), i
        assert np.array_equal(np.float32(g["principal"]), np.float32(f["principal"])), i
        if w and h:
            fl = loader.finish_frame(f, w, h)
            assert np.array_equal(np.float32(g["focal_at_json_resolution"]).view(np.uint32), np.float32(fl).view(np.uint32)), (i, g["focal_at_json_resolution"], fl)


def _touch(path):
    os.makedirs(os.path.dirname(path), exist_ok=True)
    open(path, "wb").close()


def _pose(rng):
    m = np.eye(4)
    q, _ = np.linalg.qr(rng.standard_normal((3, 3)))
    m[:3, :3] = q
    m[:3, 3] = rng.standard_normal(3) * 2
    return m.tolist()


@pytest.fixture()
def loader():
    return importlib.import_module("instant-ngp-rendering_b200.loader")


def test_describe_matches_python_loader_on_a_rule_heavy_scene(ngp, loader, tmp_path):
    rng = np.random.default_rng(11)
    frames = []
    for i in [10, 2, 1, 33, 9, 100, 20, 3]:
        name = "img/frame_%d" % i                      # no extension: searched; natural order 1 2 3 9 10 20 33 100
        _touch(str(tmp_path / (name + (".png" if i % 2 else ".jpg"))))
        fr = {"file_path": name.replace("/", "\\") if i == 9 else name, "sharpness": float(5 + (i * 7) % 13), "transform_matrix": _pose(rng)}
        if i == 3:
            fr.update({"k1": -0.1, "p2": 0.002, "cx": 301.5, "cy": 202.25, "w": 640, "h": 400, "fl_y": 512.5})
        if i == 20:
            fr.update({"camera_angle_x": 0.9})
        frames.append(fr)
    _touch(str(tmp_path / "img/frame_2.jpg.alpha.jpg"))          # file_path + ".alpha." + ext
    _touch(str(tmp_path / "img/dynamic_mask_frame_33.png"))
    _touch(str(tmp_path / "img/frame_999.png"))                  # not referenced
    meta = {"x_fov": 52.0, "k1": 0.03, "k2": 0.004, "cx": 322.0, "cy": 198.0, "w": 640, "h": 400, "aabb_scale": 8, "scale": 0.21, "offset": 0.4,
            "up": [0.1, -0.2, 0.97], "render_aabb": [[-0.25, 0.0, 0.125], [1.0, 0.75, 0.5]], "white_transparent": True, "n_frames": 7,
            "sharpness_discard_threshold": 0.6, "fix_premult": True, "frames": frames}
    path = tmp_path / "transforms.json"
    path.write_text("/* scene */\n" + json.dumps(meta, indent=1))
    got, want = _describe(ngp._lib.lib, path), loader.describe_dataset(str(path))
    assert 0 < want["n_images"] < 7, "the culling rule must drop something"
    _compare(got, want, loader, 640, 400)


def test_describe_matches_python_loader_on_mitsuba_aabb_and_several_files(ngp, loader, tmp_path):
    rng = np.random.default_rng(5)
    for part, n in (("train", 3), ("val", 2)):
        frames = []
        for i in range(n):
            name = "./%s/r_%d.png" % (part, i)
            _touch(str(tmp_path / name))
            frames.append({"file_path": name, "transform_matrix": _pose(rng)})
        meta = {"camera_angle_x": 0.6911, "frames": frames}
        if part == "train":
            meta.update({"normal_mts_args": "x", "aabb": [[-1.0, -2.0, 0.5], [3.0, 1.0, 2.5]], "offset": [0.1, 0.2, 0.3], "black_transparent": True})
        (tmp_path / ("transforms_%s.json" % part)).write_text(json.dumps(meta))
    got, want = _describe(ngp._lib.lib, tmp_path), loader.describe_dataset(str(tmp_path))
    assert want["from_mitsuba"] and want["n_images"] == 5
    _compare(got, want, loader)


def test_describe_matches_python_loader_on_the_reference_fox_scene(ngp, loader):
    fox = os.path.join(ROOT, "data_stage", "fox")
    if not os.path.isdir(fox):
        pytest.skip("data_stage/fox is not staged")
    got, want = _describe(ngp._lib.lib, fox), loader.describe_dataset(fox)
    assert want["n_images"] == 50
    _compare(got, want, loader, 1080, 1920)


def test_describe_reports_the_reference_errors(ngp, tmp_path):
    lib = ngp._lib.lib
    (tmp_path / "empty.json").write_text('{"camera_angle_x": 0.5, "frames": []}')
    assert lib.ngpb_host_describe_dataset(str(tmp_path / "empty.json").encode(), None, 0) == 0
    assert "No training images were found" in lib.ngpb_last_error().decode()
    (tmp_path / "missing.json").write_text(json.dumps({"camera_angle_x": 0.5, "frames": [{"file_path": "nope.png", "transform_matrix": np.eye(4).tolist()}]}))
    assert lib.ngpb_host_describe_dataset(str(tmp_path / "missing.json").encode(), None, 0) == 0
    assert "Could not find image file" in lib.ngpb_last_error().decode()
    (tmp_path / "fisheye.json").write_text(json.dumps({"is_fisheye": True, "frames": [{"file_path": "nope.png", "transform_matrix": np.eye(4).tolist()}]}))
    assert lib.ngpb_host_describe_dataset(str(tmp_path / "fisheye.json").encode(), None, 0) == 0
    assert "Perspective or OpenCV" in lib.ngpb_last_error().decode()


def test_describe_resolves_non_ascii_file_names(ngp, loader, tmp_path):
    """transforms.json written by json.dumps escapes non-ASCII names (\\u00e9, surrogate pairs): the C++ parser turns them into the
    UTF-8 bytes the files carry on disk, so the frames are found - by both loaders, at the same paths."""
    rng = np.random.default_rng(12)
    names = ["vues/été_01", "vues/写真_02", "vues/view_\U0001F600_03"]
    for n in names:
        _touch(str(tmp_path / (n + ".png")))
    meta = {"camera_angle_x": 0.7, "w": 64, "h": 48, "frames": [{"file_path": n, "transform_matrix": _pose(rng)} for n in names]}
    path = tmp_path / "transforms.json"
    text = json.dumps(meta)
    assert "\\u00e9" in text and "\\ud83d\\ude00" in text
    path.write_text(text)
    got, want = _describe(ngp._lib.lib, path), loader.describe_dataset(str(path))
    assert got["n_images"] == 3
    _compare(got, want, loader, 64, 48)
    assert all(os.path.exists(f["path"]) for f in got["frames"])


def test_frame_order_is_the_references_natural_compare(ngp, loader):
    """tests/golden/ref_natural_sort.json.gz: SI::natural::compare of the reference's vendored NaturalSort header on 9 591 pairs of
    names (digit runs, fractions after a dot, leading zeros, repeated spaces, case, non-ASCII bytes, its bool-from-int quirks).
    The C++ loader's order and loader.py's must answer every pair the same way; the live header is asked too where it is built."""
    import gzip
    with gzip.open(os.path.join(ROOT, "tests", "golden", "ref_natural_sort.json.gz"), "rb") as f:
        pairs = json.loads(f.read().decode())["pairs"]
    assert len(pairs) > 9000 and 0.3 < sum(p[2] for p in pairs) / len(pairs) < 0.7
    lib = ngp._lib.lib
    for a, b, want in pairs:
        assert lib.ngpb_host_natural_less(a.encode(), b.encode()) == want, (a, b, want)
        assert int(loader.natural_less(a, b)) == want, (a, b, want)
    ref_path = os.path.join(ROOT, "oracle", "_ref", "libref_json.so")
    if os.path.exists(ref_path):
        ref = ctypes.CDLL(ref_path)
        if hasattr(ref, "ref_natural_less"):
            ref.ref_natural_less.argtypes = [ctypes.c_char_p, ctypes.c_char_p]
            rng = np.random.default_rng(3)
            alphabet = list("aAzZ0159 ._-/") + ["é"]
            for _ in range(3000):
                a = "".join(rng.choice(alphabet, size=rng.integers(1, 8)))
                b = "".join(rng.choice(alphabet, size=rng.integers(1, 8)))
                if a.rstrip(" ") != a and b.rstrip(" ") != b:
                    continue   # the reference reads past both ends there (see make_ref_natural_sort.py)
                assert lib.ngpb_host_natural_less(a.encode(), b.encode()) == ref.ref_natural_less(a.encode(), b.encode()), (a, b)
