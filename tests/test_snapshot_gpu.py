"""Snapshot round trips on the device: save_snapshot -> load_snapshot into a fresh testbed gives the
same inference parameters, occupancy grid, step, loss, camera and pixels; with the optimizer state a
run resumes where it stopped (src/testbed.cu:4775-4972, tcnn trainer.h:275-314)."""
import importlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ngp = importlib.import_module("instant-ngp-rendering_b200")
syn = importlib.import_module("instant-ngp-rendering_b200.synthetic")
S = importlib.import_module("instant-ngp-rendering_b200.snapshot")


def _trained(steps=80, aabb_scale=1):
    tb = ngp.Testbed()
    scene, poses, focal = syn.make_dataset(tb, n_images=8, width=96, height=72, aabb_scale=aabb_scale)
    tb.reload_network_from_json(dict(syn.NERF_BASE_CONFIG))
    for _ in range(steps):
        tb.train(1 << 16)
    return tb, poses, focal


def _render(tb, pose, focal):
    rgba, depth = tb.render_nerf(96, 72, pose, focal)
    torch.cuda.synchronize()
    return rgba.cpu().numpy(), depth.cpu().numpy()


@pytest.mark.parametrize("name,compress", [("s.ingp", True), ("s.ingp", False), ("s.msgpack", True)])
def test_snapshot_round_trip_into_a_fresh_testbed(tmp_path, name, compress):
    tb, poses, focal = _trained()
    tb.exposure = 0.75
    tb.background_color = [0.25, 0.5, 0.75, 1.0]
    tb.set_ngp_camera_matrix(poses[2])
    tb.fov = 40.0
    path = str(tmp_path / name)
    tb.save_snapshot(path, include_optimizer_state=False, compress=compress)

    cfg = S.read_container(path)
    snap = cfg["snapshot"]
    assert snap["version"] == 1 and snap["mode"] == "nerf" and snap["density_grid_size"] == 128
    assert snap["n_params"] == tb.n_params and snap["params_type"] == "__half" and len(snap["params_binary"]) == 2 * tb.n_params
    assert len(snap["density_grid_binary"]) == 2 * 128 ** 3 and "optimizer" not in snap
    assert snap["training_step"] == 80 and snap["nerf"]["dataset"]["n_images"] == 8 and len(snap["nerf"]["dataset"]["xforms"]) == 8
    assert {"encoding", "network", "optimizer", "loss"} <= set(cfg)                 # the network config travels with it

    fresh = ngp.Testbed()
    fresh.load_snapshot(path)
    assert fresh.training_step == tb.training_step and fresh.loss == pytest.approx(tb.loss, rel=1e-6)
    assert fresh.rays_per_batch == tb.rays_per_batch and fresh.measured_batch_size == tb.measured_batch_size
    assert fresh.exposure == pytest.approx(0.75) and fresh.background_color == pytest.approx([0.25, 0.5, 0.75, 1.0])
    assert fresh.fov == pytest.approx(40.0, rel=1e-5) and np.array_equal(fresh._camera, np.asarray(poses[2], dtype=np.float32)[:3, :4])
    want = tb.device_buffer("params_inference", torch.float16)[:tb.n_params]
    for buf in ("params_inference", "params"):
        assert torch.equal(fresh.device_buffer(buf, torch.float16)[:tb.n_params], want), buf
    assert torch.equal(fresh.device_buffer("params_fp32", torch.float32)[:tb.n_params], want.float())
    grid = tb.device_buffer("density_grid", torch.float32)
    assert torch.equal(fresh.device_buffer("density_grid", torch.float32), grid.half().float())
    # the bitfield is rebuilt from the fp16 grid: the same but for cells the rounding moves across the threshold
    # (the threshold itself is min(0.01, mean), and the mean moves with the rounding too)
    b0, b1 = tb.device_buffer("density_grid_bitfield", torch.uint8), fresh.device_buffer("density_grid_bitfield", torch.uint8)
    flipped = int((b0 ^ b1).to(torch.int32).bitwise_count().sum()) if hasattr(torch.Tensor, "bitwise_count") else int((b0 != b1).sum())
    assert flipped <= 2048, flipped   # 0.1 % of the 2^21 cells
    a, _ = _render(tb, poses[5], focal)
    b, _ = _render(fresh, poses[5], focal)
    assert np.abs(a - b).mean() < 1e-4 and a[..., 3].max() > 0.5
    # and a second load of the same file is the same state, bit for bit
    again = ngp.Testbed()
    again.load_snapshot(path)
    c, _ = _render(again, poses[5], focal)
    assert np.array_equal(b, c)
    for t in (tb, fresh, again):
        t.close()


def test_resume_with_optimizer_state(tmp_path):
    tb, poses, focal = _trained(steps=96)
    path = str(tmp_path / "resume.ingp")
    tb.save_snapshot(path, include_optimizer_state=True)
    snap = S.read_container(path)["snapshot"]
    opt = snap["optimizer"]                                    # Ema{ExponentialDecay{Adam}}: the chain of configs/nerf/base.json
    adam = opt["nested"]["nested"]
    assert opt["weights_ema_binary"] == snap["params_binary"] and adam["current_step"] == 96
    assert len(adam["first_moments_binary"]) == 4 * tb.n_params == len(adam["second_moments_binary"]) == len(adam["param_steps_binary"])
    assert opt["nested"]["learning_rate_factor"] == 1.0 and opt["nested"]["learning_rate"] == pytest.approx(adam["base_learning_rate"])

    # the same scene loaded first, then the snapshot: training carries on from step 96
    cont = ngp.Testbed()
    syn.make_dataset(cont, n_images=8, width=96, height=72, aabb_scale=1)
    cont.load_snapshot(path)
    assert cont.training_step == 96 and int(cont._get("optimizer_step")) == 96
    for name, dt in (("optimizer_first_moments", torch.float32), ("optimizer_second_moments", torch.float32), ("optimizer_param_steps", torch.int32)):
        assert torch.equal(cont.device_buffer(name, dt), tb.device_buffer(name, dt)), name
    loss_at_save = tb.loss
    for _ in range(32):
        cont.train(1 << 16)
    assert cont.training_step == 128 and int(cont._get("optimizer_step")) == 128
    assert cont.loss < 1.5 * loss_at_save, (cont.loss, loss_at_save)
    # without the optimizer state the moments start from zero (Trainer::deserialize only restores what the file holds)
    tb.save_snapshot(str(tmp_path / "weights_only.msgpack"))
    bare = ngp.Testbed()
    bare.load_snapshot(str(tmp_path / "weights_only.msgpack"))
    assert int(bare._get("optimizer_step")) == 0 and float(bare.device_buffer("optimizer_second_moments", torch.float32).abs().max()) == 0.0
    for t in (tb, cont, bare):
        t.close()


def test_snapshot_errors(tmp_path):
    tb, _, _ = _trained(steps=4)
    cfg = S.build_snapshot(tb)
    bad = dict(cfg, snapshot=dict(cfg["snapshot"], params_binary=cfg["snapshot"]["params_binary"][:-2], n_params=tb.n_params - 1))
    S.write_container(str(tmp_path / "short.msgpack"), bad)
    with pytest.raises(ngp.NgpbError, match="Can't set params because buffer has the wrong size."):
        ngp.Testbed().load_snapshot(str(tmp_path / "short.msgpack"))
    bad = dict(cfg, snapshot=dict(cfg["snapshot"], density_grid_binary=cfg["snapshot"]["density_grid_binary"] * 2))
    S.write_container(str(tmp_path / "cascades.msgpack"), bad)
    with pytest.raises(ngp.NgpbError, match="Incompatible number of grid cascades."):
        ngp.Testbed().load_snapshot(str(tmp_path / "cascades.msgpack"))
    # a model that never trained: an empty grid is a valid state, and fp32 parameters are accepted
    p32 = tb.device_buffer("params_fp32", torch.float32)[:tb.n_params].cpu().numpy()
    ok = dict(cfg, snapshot=dict(cfg["snapshot"], density_grid_binary=b"", params_type="float", params_binary=p32.tobytes()))
    S.write_container(str(tmp_path / "fp32.msgpack"), ok)
    t2 = ngp.Testbed()
    t2.load_snapshot(str(tmp_path / "fp32.msgpack"))
    assert torch.equal(t2.device_buffer("params_fp32", torch.float32)[:tb.n_params].cpu(), torch.from_numpy(p32))
    assert torch.equal(t2.device_buffer("params", torch.float16)[:tb.n_params].cpu(), torch.from_numpy(p32).half())
    assert int(t2.device_buffer("density_grid_bitfield", torch.uint8).sum()) == 0
    tb.close()
    t2.close()
