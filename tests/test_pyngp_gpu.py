"""The pyngp shell: `import pyngp as ngp` and the calls of scripts/run.py's
evaluation protocol (scripts/run.py:216-268) against the C ABI -
  * render_ground_truth + set_camera_to_training_view + render(w, h, 1, True) returns the training
    image through overlay_image_kernel: compared with the numpy restatement, float rounding;
  * a view with an OpenCV lens is rendered with that lens: per-pixel sample counts BIT-EXACT with the
    oracle's march through the same lens;
  * the protocol itself: PSNR of rendered against ground-truth frames rises with training.
"""
import importlib
import os
import sys

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def pyngp():
    sys.path.insert(0, os.path.join(ROOT, "instant-ngp-rendering_b200"))
    import pyngp as mod
    return mod


@pytest.fixture(scope="module")
def scene(pyngp):
    syn = importlib.import_module("instant-ngp-rendering_b200.synthetic")
    tb = pyngp.Testbed()
    assert tb.mode == getattr(pyngp.TestbedMode, "None")
    _, poses, focal = syn.make_dataset(tb, n_images=6, width=96, height=72, aabb_scale=2, opaque_background=False)
    assert tb.mode == pyngp.TestbedMode.Nerf
    tb.reload_network_from_json(dict(syn.NERF_BASE_CONFIG))
    yield tb, poses, focal, syn
    tb.close()


def _srgb8_of_view(tb, syn, scene_obj, pose, w, h, focal):
    return syn.linear_premul_to_srgb8(scene_obj.render(pose, w, h, focal)).cpu().numpy()


def test_enums_are_the_reference_values(pyngp):
    assert int(pyngp.TestbedMode.Nerf) == 0 and int(pyngp.TestbedMode.Image) == 2 and int(getattr(pyngp.TestbedMode, "None")) == 5
    assert [int(v) for v in pyngp.TonemapCurve] == [0, 1, 2, 3]
    assert int(pyngp.ColorSpace.SRGB) == 1 and int(pyngp.LossType.Huber) == 4 and pyngp.LossType.SmoothL1 is pyngp.LossType.Huber
    assert int(pyngp.NerfActivation.Exponential) == 3 and int(pyngp.LensMode.OpenCV) == 1


def test_ground_truth_render_matches_restatement(pyngp, oracle, scene):
    tb, poses, focal, syn = scene
    blob = syn.BlobScene(n_blobs=6, seed=7, device="cuda")
    ds = tb.nerf.training.dataset
    assert ds.n_images == 6 and ds.metadata[3].resolution == [96, 72]
    img = _srgb8_of_view(tb, syn, blob, poses[3], 96, 72, focal)
    try:
        tb.render_ground_truth = True
        tb.set_camera_to_training_view(3)
        for (w, h, linear, curve, exposure, bg, zoom, axis) in ((96, 72, True, 0, 0.0, (0, 0, 0, 1), 1.0, 1), (64, 48, False, 0, 0.0, (0.2, 0.5, 0.7, 1.0), 1.0, 1),
                                                                (80, 40, True, 1, 0.5, (1.0, 1.0, 1.0, 0.5), 1.5, 0), (50, 70, False, 2, -0.3, (0, 0, 0, 0), 0.7, 1)):
            tb.tonemap_curve, tb.exposure, tb.background_color, tb.zoom, tb.fov_axis = curve, exposure, bg, zoom, axis
            got = tb.render(w, h, 1, linear).reshape(-1, 4)
            want = oracle.overlay_image(img, w, h, fov_axis=axis, zoom=zoom, exposure=exposure, background=bg, output_color_space=0 if linear else 1, curve=curve)
            np.testing.assert_allclose(got, want, rtol=2e-5, atol=2e-6, err_msg=str((w, h, linear, curve)))
            assert got[:, :3].max() > 0.05
        # at the native resolution and linear output the frame IS the image: premultiplied linear colours of the texels
        tb.tonemap_curve, tb.exposure, tb.background_color, tb.zoom, tb.fov_axis = 0, 0.0, (0, 0, 0, 0), 1.0, 1
        got = tb.render(96, 72, 1, True)
        a = img[..., 3].astype(np.float32) / 255.0
        np.testing.assert_allclose(got[..., 3], a, atol=1e-6)
        with pytest.raises(pyngp.NgpbError):
            tb.set_camera_to_training_view(17)
    finally:
        tb.render_ground_truth = False
        tb.tonemap_curve, tb.exposure, tb.background_color, tb.zoom, tb.fov_axis = 0, 0.0, (0.0, 0.0, 0.0, 1.0), 1.0, 1


def test_training_view_with_opencv_lens_counts_bit_exact(pyngp, oracle, scene):
    tb, poses, focal, syn = scene
    for _ in range(60):
        tb.train(1 << 14)
    lens = (0.08, -0.02, 0.003, -0.004)
    w, h = 96, 72
    tb.nerf.training.set_camera_intrinsics(2, fx=focal, fy=focal * 1.05, cx=0.45 * w, cy=0.55 * h, k1=lens[0], k2=lens[1], p1=lens[2], p2=lens[3])
    try:
        tb.set_camera_to_training_view(2)
        assert tb.nerf.render_with_lens_distortion
        md = tb.nerf.training.dataset.metadata[2]
        assert md.lens_mode == int(pyngp.LensMode.OpenCV)
        np.testing.assert_allclose(tb.screen_center, (1 - 0.45, 1 - 0.55), atol=1e-7)
        # what Testbed.render hands the tracer: focal = relative focal x res[fov_axis], centre flipped back
        rel = np.asarray(tb._relative_focal_length)
        fx, fy = float(rel[0] * h), float(rel[1] * h)
        cx, cy = 0.5 - tb.screen_center[0] + 0.5, 0.5 - tb.screen_center[1] + 0.5
        rgba, depth, steps = tb.render_nerf(w, h, poses[2], (fx, fy), principal=(cx, cy), return_steps=True)
        bitfield = tb.device_buffer("density_grid_bitfield", torch.uint8).cpu().numpy()
        got = {}
        for name, lens_arg in (("lens", lens), ("pinhole", None)):
            rc = oracle.render_config([-0.5] * 3, [1.5] * 3, poses[2], w, h, (fx, fy), 1, 1.0 / 256.0, principal=(cx, cy), lens=lens_arg)
            numsteps, coords = oracle.nerf_render_march(rc, bitfield)
            got[name] = (numsteps, coords)
        assert not np.array_equal(got["lens"][0][:, 0], got["pinhole"][0][:, 0]), "the lens must move the rays"
        # every sample the device composited is one of the oracle's: count <= marched count, equal where the ray did not terminate early
        numsteps = got["lens"][0]
        steps = steps.cpu().numpy().reshape(-1).astype(np.uint32)
        assert (steps <= numsteps[:, 0]).all()
        a = rgba.cpu().numpy().reshape(-1, 4)[:, 3]
        open_rays = a < 0.9
        assert open_rays.sum() > 100 and np.array_equal(steps[open_rays], numsteps[open_rays, 0])
        assert (steps[numsteps[:, 0] == 0] == 0).all()
        # and render() itself goes through the same lens: identical to the epilogue of that frame
        tb.snap_to_pixel_centers = False
        frame = tb.render(w, h, 1, True).reshape(-1, 4)
        want = oracle.render_buffer_epilogue(rgba.cpu().numpy().reshape(1, -1, 4), 0.0, (0, 0, 0, 1), 0, 0, 0)
        np.testing.assert_allclose(frame, want, rtol=2e-5, atol=2e-6)
    finally:
        tb.nerf.training.set_camera_intrinsics(2, fx=focal, fy=focal, cx=w / 2, cy=h / 2)
        tb.nerf.render_with_lens_distortion = False


def test_run_py_evaluation_protocol(pyngp, scene, tmp_path):
    """scripts/run.py:216-268: background black, snap to pixel centres, min transmittance 1e-4, then for
    every training view the ground-truth frame and the rendered frame, PSNR of the two."""
    tb, poses, focal, syn = scene

    def evaluate():
        tb.background_color = [0.0, 0.0, 0.0, 1.0]
        tb.snap_to_pixel_centers = True
        tb.nerf.render_min_transmittance = 1e-4
        tb.shall_train = False
        total = 0.0
        n = tb.nerf.training.dataset.n_images
        for i in range(n):
            resolution = tb.nerf.training.dataset.metadata[i].resolution
            tb.render_ground_truth = True
            tb.set_camera_to_training_view(i)
            ref_image = tb.render(resolution[0], resolution[1], 1, True)
            tb.render_ground_truth = False
            image = tb.render(resolution[0], resolution[1], 4, True)
            assert image.shape == ref_image.shape == (resolution[1], resolution[0], 4)
            mse = float(np.mean((image[..., :3] - ref_image[..., :3]) ** 2))
            total += -10.0 * np.log10(max(mse, 1e-12))
        tb.shall_train = True
        return total / n

    before = evaluate()
    tb.training_batch_size = 1 << 15
    n0 = tb.training_step
    while tb.frame():
        if tb.want_repl():
            break
        if tb.training_step >= n0 + 400:
            break
    after = evaluate()
    assert after > before + 3.0 and after > 20.0, (before, after)
    # the setters run.py --nerf_compatibility uses
    tb.nerf.cone_angle_constant = 0
    assert tb.nerf.cone_angle_constant == 0.0
    tb.nerf.training.random_bg_color = False
    tb.color_space = pyngp.ColorSpace.SRGB
    assert tb.color_space == pyngp.ColorSpace.SRGB
    tb.train(1 << 14)
    tb.color_space = pyngp.ColorSpace.Linear
    tb.nerf.cone_angle_constant = 1.0 / 256.0
    with pytest.raises(pyngp.NgpbError):
        tb.init_window(1920, 1080)
    with pytest.raises(pyngp.NgpbError):
        tb.compute_and_save_marching_cubes_mesh("mesh.obj", [256, 256, 256])
    with pytest.raises(pyngp.NgpbError):
        tb.nerf.sharpen = 1.0
    tb.nerf.sharpen = 0.0
    # a snapshot taken with the camera on a training view (vec2 relative focal length, flipped screen centre) loads back
    tb.set_camera_to_training_view(1)
    path = str(tmp_path / "view.ingp")
    tb.save_snapshot(path, False)
    other = pyngp.Testbed()
    try:
        other.load_file(path)
        assert other.training_step == tb.training_step
        np.testing.assert_allclose(np.broadcast_to(other._relative_focal_length, (2,)), np.broadcast_to(tb._relative_focal_length, (2,)), rtol=1e-6)
        np.testing.assert_allclose(other.screen_center, tb.screen_center, atol=1e-7)
    finally:
        other.close()
