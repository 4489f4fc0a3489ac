"""The file-level entry points in the library's host C++ (ngpb_testbed_load_training_data / reload_network_from_file /
save_snapshot / load_snapshot) against the Python mirror (loader.py, snapshot.py, testbed.load_network_config), which is
itself pinned against the reference's files (tests/test_ref_fixture_cpu.py, profiles/r02_*_snapshot.json).

Bars: metadata (focal, principal, lens, camera-to-world, resolution) bit-exact; PNG pixels bit-exact; JPEG pixels within
the decoders' IDCT / chroma-upsampling difference (nvjpeg on the GPU vs libjpeg in OpenCV): mean |d| <= 1.5 levels and
99 % of the samples within 8 levels; parameters, occupancy grid and optimizer moments through a snapshot: the same bytes.
"""
import ctypes
import json
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CONFIG = {
    "loss": {"otype": "Huber"},
    "optimizer": {"otype": "Ema", "decay": 0.95, "nested": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000, "decay_base": 0.33,
                                                           "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}}},
    "encoding": {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 4, "log2_hashmap_size": 15, "base_resolution": 16},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 1},
    "rgb_network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}


def _device_bytes(ptr, n):
    out = torch.empty(n, dtype=torch.uint8, device="cuda")

    class _Iface:
        __cuda_array_interface__ = {"shape": (n,), "typestr": "|u1", "data": (ptr, False), "version": 2}
    out.copy_(torch.as_tensor(_Iface(), device="cuda"))
    return out.cpu().numpy()


def _write_scene(tmp_path, ext, n=6, w=96, h=64):
    """A small scene on disk: OpenCV lens at file level, one per-frame override, sharpness culling, natural-order names."""
    import cv2
    rng = np.random.default_rng(5)
    os.makedirs(tmp_path / "images", exist_ok=True)
    frames = []
    yy, xx = np.mgrid[0:h, 0:w]
    for i in range(n):
        img = np.stack([(xx * 2 + 13 * i) % 256, (yy * 3 + 7 * i) % 256, ((xx + yy) * 2 + 29 * i) % 256], axis=-1).astype(np.uint8)
        img = cv2.GaussianBlur(img, (5, 5), 0)
        name = "images/%d.%s" % (i + 1, ext)   # 1, 2, ... 10 must sort by value
        if ext == "png":
            rgba = np.dstack([img, np.where((xx + yy + i) % 11 == 0, 0, 255).astype(np.uint8)])
            cv2.imwrite(str(tmp_path / name), rgba[:, :, [2, 1, 0, 3]])
        else:
            cv2.imwrite(str(tmp_path / name), img[:, :, ::-1], [cv2.IMWRITE_JPEG_QUALITY, 92])
        ang = 2 * np.pi * i / n
        c2w = np.eye(4)
        c2w[:3, :3] = [[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]]
        c2w[:3, 3] = [2.5 * np.sin(ang), 0.3 * i, 2.5 * np.cos(ang)]
        fr = {"file_path": name if i % 2 else name.rsplit(".", 1)[0], "sharpness": 50.0 + i, "transform_matrix": c2w.tolist()}
        if i == 3:
            fr.update({"k1": 0.01, "fl_x": 111.0})
        frames.append(fr)
    rng.shuffle(frames)
    meta = {"camera_angle_x": 0.7, "k1": 0.02, "k2": -0.003, "p1": 1e-4, "p2": -2e-4, "cx": w / 2 + 1.5, "cy": h / 2 - 0.5, "w": w, "h": h,
            "aabb_scale": 2, "scale": 0.3, "offset": [0.5, 0.45, 0.5], "up": [0, 0, 1], "sharpness_discard_threshold": 0.1, "frames": frames}
    with open(tmp_path / "transforms.json", "w") as f:
        f.write("// written by the test\n" + json.dumps(meta, indent=1))
    return str(tmp_path / "transforms.json")


@pytest.mark.parametrize("ext", ["png", "jpg"])
def test_cpp_loader_matches_python_loader(ngp, tmp_path, ext):
    scene = _write_scene(tmp_path, ext)
    a, b = ngp.Testbed(), ngp.Testbed()
    ds_py = a.load_training_data(scene)
    ds_cc = b.load_training_data_native(scene)
    n = int(a._get("n_images"))
    assert n == int(b._get("n_images")) == 6
    assert int(a._get("aabb_scale")) == int(b._get("aabb_scale")) == 2
    assert [os.path.basename(fr["path"]) for fr in ds_py["frames"]] == [os.path.basename(fr["path"]) for fr in ds_cc["frames"]] == ["%d.%s" % (i + 1, ext) for i in range(6)]
    for key in ("scale", "offset", "up"):
        assert np.array_equal(np.float32(ds_py[key]), np.float32(ds_cc[key])), key
    for i in range(n):
        ma, mb = a.image_metadata_native(i), b.image_metadata_native(i)
        for key in ("image_data_type", "lens_mode", "resolution"):
            assert ma[key] == mb[key], (i, key)
        for key in ("lens_params", "principal_point", "focal_length", "xform"):
            assert np.array_equal(np.float32(ma[key]).view(np.uint32), np.float32(mb[key]).view(np.uint32)), (i, key, ma[key], mb[key])
        w, h = ma["resolution"]
        pa, pb = _device_bytes(ma["pixels"], w * h * 4), _device_bytes(mb["pixels"], w * h * 4)
        if ext == "png":
            assert np.array_equal(pa, pb), i
        else:
            d = np.abs(pa.astype(np.int32) - pb.astype(np.int32)).reshape(-1, 4)
            assert (d[:, 3] == 0).all()
            # (this test image wraps modulo 256: hard chroma edges, where the two decoders' upsampling filters differ most; the fox test holds the docstring's bar)
            assert d[:, :3].mean() <= 2.0 and (d[:, :3] <= 8).mean() >= 0.97, (i, float(d[:, :3].mean()), int(d.max()))
    # both train
    for tb in (a, b):
        tb.reload_network_from_json(CONFIG)
        for _ in range(20):
            tb.train(1 << 14)
        assert np.isfinite(tb.loss)
    # same pixels, same cameras, same seed: the same training up to the order of the fp16 gradient atomics
    assert abs(a.loss - b.loss) <= 0.5 * max(a.loss, b.loss)


def test_cpp_loader_on_the_reference_fox_scene(ngp):
    fox = os.path.join(ROOT, "data_stage", "fox")
    if not os.path.isdir(fox):
        pytest.skip("data_stage/fox is not staged")
    a, b = ngp.Testbed(), ngp.Testbed()
    a.load_training_data(fox, max_images=8)
    b.load_training_data_native(fox, max_images=8)
    for i in range(8):
        ma, mb = a.image_metadata_native(i), b.image_metadata_native(i)
        for key in ("lens_params", "principal_point", "focal_length", "xform"):
            assert np.array_equal(np.float32(ma[key]).view(np.uint32), np.float32(mb[key]).view(np.uint32)), (i, key)
        assert ma["resolution"] == mb["resolution"] == [1080, 1920]
        w, h = ma["resolution"]
        d = np.abs(_device_bytes(ma["pixels"], w * h * 4).astype(np.int32) - _device_bytes(mb["pixels"], w * h * 4).astype(np.int32)).reshape(-1, 4)[:, :3]
        assert d.mean() <= 1.5 and (d <= 8).mean() >= 0.99, (i, float(d.mean()), int(d.max()))


def test_config_file_with_parent_through_the_c_entry_point(ngp, tmp_path):
    with open(tmp_path / "base.json", "w") as f:
        f.write("{ // base\n" + json.dumps(CONFIG)[1:])
    with open(tmp_path / "child.json", "w") as f:
        json.dump({"parent": "base.json", "encoding": {"log2_hashmap_size": 14}, "optimizer": {"nested": {"nested": {"learning_rate": 5e-3}}}}, f)
    a, b = ngp.Testbed(), ngp.Testbed()
    for tb in (a, b):
        tb.create_empty_nerf_dataset(1, aabb_scale=1)
    a.reload_network_from_file(str(tmp_path / "child.json"))
    b.reload_network_from_file_native(str(tmp_path / "child.json"))
    assert a.network_config == b.network_config
    assert b.network_config["encoding"]["log2_hashmap_size"] == 14 and b.network_config["optimizer"]["nested"]["nested"]["learning_rate"] == 5e-3
    assert a.n_params == b.n_params
    with pytest.raises(ngp.NgpbError, match="does not exist"):
        b.reload_network_from_file_native(str(tmp_path / "missing.json"))


def _state(tb):
    names = [("params", torch.float16), ("params_inference", torch.float16), ("density_grid", torch.float32), ("optimizer_first_moments", torch.float32),
             ("optimizer_second_moments", torch.float32), ("optimizer_param_steps", torch.int32)]
    return {n: tb.device_buffer(n, t).cpu().numpy().copy() for n, t in names}


@pytest.mark.parametrize("writer,reader", [("native", "python"), ("python", "native"), ("native", "native")])
def test_snapshots_interchange_between_the_two_hosts(ngp, tmp_path, writer, reader):
    scene = _write_scene(tmp_path, "png")
    src = ngp.Testbed()
    src.load_training_data_native(scene) if writer == "native" else src.load_training_data(scene)
    src.reload_network_from_json(CONFIG)
    for _ in range(40):
        src.train(1 << 14)
    src.set_camera_to_training_view(2)
    src.exposure = 0.5
    path = str(tmp_path / "snap.ingp")
    (src.save_snapshot_native if writer == "native" else src.save_snapshot)(path, include_optimizer_state=True)
    dst = ngp.Testbed()
    (dst.load_snapshot_native if reader == "native" else dst.load_snapshot)(path)
    want, got = _state(src), _state(dst)
    n = src.n_params
    assert np.array_equal(got["params_inference"][:n].view(np.uint16), want["params_inference"][:n].view(np.uint16))
    assert np.array_equal(got["optimizer_first_moments"][:n].view(np.uint32), want["optimizer_first_moments"][:n].view(np.uint32))
    assert np.array_equal(got["optimizer_second_moments"][:n].view(np.uint32), want["optimizer_second_moments"][:n].view(np.uint32))
    assert np.array_equal(got["optimizer_param_steps"][:n], want["optimizer_param_steps"][:n])
    assert np.array_equal(got["density_grid"].astype(np.float16).view(np.uint16), want["density_grid"].astype(np.float16).view(np.uint16))
    assert dst.training_step == src.training_step == 40 and dst.loss == pytest.approx(src.loss, rel=1e-6)
    for key in ("optimizer_step", "optimizer_base_learning_rate", "optimizer_learning_rate_factor", "rays_per_batch", "measured_batch_size", "aabb_scale", "n_images", "exposure"):
        assert dst._get(key) == src._get(key), key
    for i in range(int(src._get("n_images"))):
        ma, mb = src.image_metadata_native(i), dst.image_metadata_native(i)
        assert ma["lens_mode"] == mb["lens_mode"]
        # (snapshot.py keeps the resolutions of an image-less dataset in the Python mirror only)
        assert (mb["resolution"] if reader == "native" else list(dst._resolutions[i])) == ma["resolution"]
        for key in ("lens_params", "principal_point", "focal_length", "xform"):
            assert np.array_equal(np.float32(ma[key]), np.float32(mb[key])), (i, key)
    # the restored model renders the frame the trained one renders
    cam = np.ascontiguousarray(src._extrinsics[1], dtype=np.float32)
    ra, _ = src.render_nerf(64, 48, cam, 70.0)
    rb, _ = dst.render_nerf(64, 48, cam, 70.0)
    # (the restored occupancy grid went through fp16, so a few cells on the threshold flip and the frames differ in the third digit)
    assert float((ra - rb).abs().mean()) <= 5e-3
    # and both containers hold the same tree (keys and scalar values; binaries by content)
    snap = __import__("importlib").import_module("instant-ngp-rendering_b200.snapshot")
    other = str(tmp_path / "other.ingp")
    (src.save_snapshot if writer == "native" else src.save_snapshot_native)(other, include_optimizer_state=True)
    ta, tb_ = snap.read_container(path), snap.read_container(other)

    def same(x, y, where=""):
        if isinstance(x, dict):
            assert isinstance(y, dict) and sorted(x) == sorted(y), (where, sorted(x), sorted(y) if isinstance(y, dict) else y)
            for k in x:
                same(x[k], y[k], where + "/" + k)
        elif isinstance(x, (list, tuple)):
            assert len(x) == len(y), where
            for i, (p, q) in enumerate(zip(x, y)):
                same(p, q, where + "[%d]" % i)
        elif isinstance(x, float) or isinstance(y, float):
            assert np.float32(x) == np.float32(y), (where, x, y)
        else:
            assert x == y, (where, x if not isinstance(x, bytes) else len(x))
    same(ta, tb_)


def test_cpp_host_program_trains_snapshots_and_renders(ngp, tmp_path):
    """examples/ngp_host: a C++ process with no Python in it does what `instant-ngp scene --network config.json` does headless."""
    import subprocess
    exe = os.path.join(ROOT, "examples", "ngp_host")
    if not os.path.exists(exe):
        pytest.skip("examples/ngp_host is not built")
    scene = _write_scene(tmp_path, "png")
    with open(tmp_path / "base.json", "w") as f:
        json.dump(CONFIG, f)
    with open(tmp_path / "small.json", "w") as f:
        f.write('{"parent": "base.json", /* smaller table */ "encoding": {"log2_hashmap_size": 14}}')
    snap, frame = str(tmp_path / "out.ingp"), str(tmp_path / "frame.ppm")
    res = subprocess.run([exe, scene, str(tmp_path / "small.json"), "60", snap, frame], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr
    assert "6 views" in res.stdout and "wrote " + snap in res.stdout
    # the snapshot loads in the Python mirror and carries the trained model
    tb = ngp.Testbed()
    tb.load_snapshot(snap)
    assert tb.training_step == 60 and tb.network_config["encoding"]["log2_hashmap_size"] == 14
    assert int(tb._get("n_images")) == 6 and np.isfinite(tb.loss)
    with open(frame, "rb") as f:
        head = f.readline() + f.readline() + f.readline()
        px = np.frombuffer(f.read(), dtype=np.uint8)
    assert head == b"P6\n24 16\n255\n" and px.size == 24 * 16 * 3 and px.std() > 1.0


def test_cpp_loader_image_mode(ngp, tmp_path):
    """An image file selects the neural-image mode: albert.exr (when staged) and an 8-bit PNG through from_rgba32 give the
    training targets the Python loader's image gives (same seed, same stratified positions): bit for bit for the EXR, to the last bits of powf for the PNG."""
    import cv2
    cfgs = __import__("importlib").import_module("baseline.configs")
    files = []
    albert = os.path.join(ROOT, "data_stage", "albert.exr")
    if os.path.exists(albert):
        files.append(albert)
    rng = np.random.default_rng(3)
    png = str(tmp_path / "img.png")
    cv2.imwrite(png, rng.integers(0, 256, (200, 300, 4), dtype=np.uint8))
    files.append(png)
    for path in files:
        a, b = ngp.Testbed(), ngp.Testbed()
        a.load_training_data(path)
        b.load_training_data_native(path)
        for tb in (a, b):
            tb.reload_network_from_json(cfgs.IMAGE_BASE_CONFIG)
            tb.train(1 << 16)
        for name, dt in (("image_positions", torch.float32), ("image_targets", torch.float32)):
            x, y = a.device_buffer(name, dt).cpu().numpy(), b.device_buffer(name, dt).cpu().numpy()
            assert x.size > 0
            if path.endswith(".exr") or name == "image_positions":
                assert np.array_equal(x.view(np.uint32), y.view(np.uint32)), (path, name)
            else:   # sRGB -> linear: powf on the device here, numpy's float32 pow in the Python loader (last-bit differences)
                np.testing.assert_allclose(x, y, rtol=4e-6, atol=1e-7)
        assert np.isfinite(b.loss)


def test_load_file_dispatches_by_content(ngp, tmp_path):
    """ngpb_testbed_load_file = Testbed::load_file: a scene description, a network config and a snapshot through one entry point."""
    lib, check = ngp._lib.lib, ngp._lib.check
    scene = _write_scene(tmp_path, "png")
    with open(tmp_path / "net.json", "w") as f:
        json.dump(CONFIG, f)
    tb = ngp.Testbed()
    check(lib.ngpb_testbed_load_file(tb._h, scene.encode()))
    assert int(tb._get("n_images")) == 6
    check(lib.ngpb_testbed_load_file(tb._h, str(tmp_path / "net.json").encode()))
    assert tb.n_params > 0
    for _ in range(10):
        tb.train(1 << 14)
    snap = str(tmp_path / "s.msgpack")
    tb.save_snapshot_native(snap)
    other = ngp.Testbed()
    check(lib.ngpb_testbed_load_file(other._h, snap.encode()))
    assert int(other._get("training_step")) == 10 and int(other._get("n_images")) == 6
    assert lib.ngpb_testbed_load_file(other._h, str(tmp_path / "missing.ingp").encode()) != 0
    assert "does not exist" in lib.ngpb_last_error().decode()
