"""Snapshot files of the reference: Testbed::save_snapshot / load_snapshot (src/testbed.cu:4775-4972).

The container is host work: nlohmann::json::to_msgpack of the network config with a "snapshot"
object inside, wrapped in a zlib stream when the file is called *.ingp (zstr::ostream), plain
msgpack otherwise (src/testbed.cu:4826-4836; load_network_config, :99-132, also takes *.json).
Binary members (params_binary, density_grid_binary, the optimizer moments) are msgpack `bin` values,
which is what nlohmann writes for json::binary_t without a subtype. The device side of a restore is
in csrc/snapshot.cu. Key names and the row-major matrix layout follow include/neural-graphics-
primitives/json_binding.h and tiny-cuda-nn's vec_json.h (a tmat is a list of ROWS).
"""
import ctypes
import json
import os
import zlib

import numpy as np

from ._lib import lib, check

SNAPSHOT_FORMAT_VERSION = 1   # src/testbed.cu:4773
NERF_GRIDSIZE = 128
MODES = ("nerf", "sdf", "image", "volume", "none")


def _is_ingp(path):
    return str(path).lower().endswith(".ingp")     # equals_case_insensitive(extension, "ingp")


def write_container(path, config, compress=True):
    """json::to_msgpack into the file; *.ingp gets the zlib stream (level 0 when not `compress`)."""
    import msgpack
    blob = msgpack.packb(config, use_bin_type=True, use_single_float=False)
    if _is_ingp(path):
        blob = zlib.compress(blob, zlib.Z_DEFAULT_COMPRESSION if compress else 0)
    with open(path, "wb") as f:
        f.write(blob)


def read_container(path):
    """load_network_config (src/testbed.cu:246-272): *.ingp = zlib + msgpack, *.msgpack = msgpack,
    *.json = text with // comments and "parent" inheritance; any other extension reads as an empty
    config. zstr passes a stream without a zlib / gzip header through unchanged."""
    import msgpack
    p = str(path)
    if not p or not os.path.exists(p):
        raise FileNotFoundError("Network %s '%s' does not exist." % ("snapshot" if p.lower().endswith((".msgpack", ".ingp")) else "config", p))
    with open(p, "rb") as f:
        blob = f.read()
    low = p.lower()
    if low.endswith(".ingp"):
        if blob[:1] == b"\x78":
            blob = zlib.decompress(blob)
        elif blob[:2] == b"\x1f\x8b":
            blob = zlib.decompress(blob, 16 + zlib.MAX_WBITS)
        return msgpack.unpackb(blob, raw=False, strict_map_key=False)
    if low.endswith(".msgpack"):
        return msgpack.unpackb(blob, raw=False, strict_map_key=False)
    if low.endswith(".json"):
        from .testbed import load_network_config
        return load_network_config(p)
    return {}


def _binary(j):
    """from_json(GPUMemory<T>) (tcnn gpu_memory_json.h:52-71): msgpack bin, or the {"bytes": [...]}
    object nlohmann prints binaries as in text JSON."""
    if isinstance(j, (bytes, bytearray, memoryview)):
        return bytes(j)
    if isinstance(j, dict):
        return bytes(bytearray(int(v) & 0xFF for v in j["bytes"]))
    raise ValueError("Invalid json type: must be either binary or object")


def _mat(m):
    return [[float(v) for v in row] for row in np.asarray(m, dtype=np.float32)]


def _vec(v):
    return [float(x) for x in np.asarray(v, dtype=np.float32).reshape(-1)]


def lens_to_json(mode, params):
    """to_json(Lens), json_binding.h:37-63. Perspective leaves the value null."""
    if mode == 1:
        return {"is_fisheye": False, "k1": float(params[0]), "k2": float(params[1]), "p1": float(params[2]), "p2": float(params[3])}
    if mode == 0:
        return None
    raise ValueError("lens mode %d is not part of the path" % mode)


def lens_from_json(j):
    """from_json(Lens), json_binding.h:65-96 -> (mode, [k1, k2, p1, p2])."""
    if isinstance(j, dict) and "k1" in j:
        if j.get("is_fisheye", False):
            raise ValueError("OpenCV fisheye lenses are not part of the path")
        return 1, [float(j["k1"]), float(j["k2"]), float(j["p1"]), float(j["p2"])]
    if isinstance(j, dict) and any(k in j for k in ("ftheta_p0", "latlong", "equirectangular")):
        raise ValueError("F-theta / lat-long / equirectangular lenses are not part of the path")
    return 0, [0.0, 0.0, 0.0, 0.0]


def _aabb(aabb_scale):
    """m_aabb of a NeRF scene: the unit cube's centre inflated by half the scale (load_nerf_post, src/testbed_nerf.cu:2700-2712)."""
    h = 0.5 * float(aabb_scale)
    return {"min": [0.5 - h] * 3, "max": [0.5 + h] * 3}


def dataset_to_json(tb):
    """to_json(NerfDataset), json_binding.h:105-132, from what the mirror knows about the frames."""
    ds = getattr(tb, "dataset", None) or {}
    n = int(tb._get("n_images"))
    paths = [fr.get("path", "") for fr in ds.get("frames", [])][:n]
    paths += [""] * (n - len(paths))
    meta, xforms = [], []
    for i in range(n):
        fx, fy, cx, cy, k1, k2, p1, p2 = tb._intrinsics.get(i, (0.0, 0.0, -0.5, -0.5, 0.0, 0.0, 0.0, 0.0))
        w, h = tb._resolutions.get(i, (0, 0))
        mode = 1 if any(v != 0.0 for v in (k1, k2, p1, p2)) else 0
        meta.append({"focal_length": [float(fx), float(fy if fy > 0 else fx)], "lens": lens_to_json(mode, (k1, k2, p1, p2)),
                     "principal_point": [float(cx / w) if cx >= 0 and w else 0.5, float(cy / h) if cy >= 0 and h else 0.5],
                     "rolling_shutter": [0.0, 0.0, 0.0, 0.0], "resolution": [int(w), int(h)]})
        x = _mat(tb._extrinsics.get(i, np.eye(4, dtype=np.float32)[:3]))
        xforms.append({"start": x, "end": x})
    aabb_scale = int(tb._get("aabb_scale"))
    return {"n_images": n, "paths": paths, "metadata": meta, "xforms": xforms,
            "render_aabb": ds.get("render_aabb") or _aabb(aabb_scale), "render_aabb_to_local": _mat(np.eye(3)),
            "up": _vec(ds.get("up", (0.0, 1.0, 0.0))), "offset": _vec(ds.get("offset", (0.5, 0.5, 0.5))), "envmap_resolution": [0, 0],
            "scale": float(ds.get("scale", 0.33)), "aabb_scale": aabb_scale, "from_mitsuba": bool(ds.get("from_mitsuba", False)),
            "is_hdr": bool(ds.get("is_hdr", False)), "wants_importance_sampling": True, "n_extra_learnable_dims": 0}


def optimizer_to_json(tb):
    """Ema{ExponentialDecay{Adam}}::serialize (tcnn ema.h:190-195, exponential_decay.h:136-142,
    adam.h:278-286), following the chain the network config names."""
    import torch

    def host(name, dtype):
        return tb.device_buffer(name, dtype).cpu().numpy().tobytes()

    def chain(cfg):
        otype = str(cfg.get("otype", "Adam")).lower()
        if otype == "ema":
            return {"nested": chain(cfg.get("nested", {})), "weights_ema_binary": host("params_inference", torch.float16)}
        if otype == "exponentialdecay":
            return {"nested": chain(cfg.get("nested", {})), "learning_rate": tb._get("optimizer_base_learning_rate"),
                    "learning_rate_factor": tb._get("optimizer_learning_rate_factor")}
        if otype == "adam":
            return {"current_step": int(tb._get("optimizer_step")),
                    "base_learning_rate": tb._get("optimizer_base_learning_rate") * tb._get("optimizer_learning_rate_factor"),
                    "first_moments_binary": host("optimizer_first_moments", torch.float32),
                    "second_moments_binary": host("optimizer_second_moments", torch.float32),
                    "param_steps_binary": host("optimizer_param_steps", torch.int32)}
        raise ValueError("optimizer '%s' is not part of the path" % cfg.get("otype"))
    return chain(tb.network_config.get("optimizer", {}))


def optimizer_from_json(tb, data):
    """The deserialize() counterparts (ema.h:197-206, exponential_decay.h:144-148, adam.h:288-299).
    The EMA weights need no copy of their own: Trainer::set_params has already written the same
    bytes (params_binary IS the EMA copy at save time, trainer.h:281)."""
    import torch

    def put(name, dtype, blob):
        dst = tb.device_buffer(name, dtype)
        src = torch.from_numpy(np.frombuffer(_binary(blob), dtype={torch.float32: np.float32, torch.int32: np.int32}[dtype]).copy())
        if src.numel() != dst.numel():
            raise ValueError("optimizer state '%s' has %d elements, the network has %d" % (name, src.numel(), dst.numel()))
        dst.copy_(src)

    def chain(cfg, d):
        otype = str(cfg.get("otype", "Adam")).lower()
        if otype == "ema":
            chain(cfg.get("nested", {}), d["nested"])
        elif otype == "exponentialdecay":
            tb._set("optimizer_base_learning_rate", d["learning_rate"])
            tb._set("optimizer_learning_rate_factor", d.get("learning_rate_factor", 1.0))
            chain(cfg.get("nested", {}), d["nested"])
        elif otype == "adam":
            put("optimizer_first_moments", torch.float32, d["first_moments_binary"])
            put("optimizer_second_moments", torch.float32, d["second_moments_binary"])
            if "param_steps_binary" in d:
                put("optimizer_param_steps", torch.int32, d["param_steps_binary"])
            else:
                tb.device_buffer("optimizer_param_steps", torch.int32).zero_()
            tb._set("optimizer_step", d["current_step"])
            if "exponentialdecay" not in json.dumps(tb.network_config.get("optimizer", {})).lower():
                tb._set("optimizer_base_learning_rate", d["base_learning_rate"])
        else:
            raise ValueError("optimizer '%s' is not part of the path" % cfg.get("otype"))
    chain(tb.network_config.get("optimizer", {}), data)
    torch.cuda.synchronize()


def build_snapshot(tb, include_optimizer_state=False):
    """The config object save_snapshot writes: m_network_config with ["snapshot"] filled in
    (src/testbed.cu:4776-4824; Trainer::serialize, tcnn trainer.h:275-288)."""
    import torch
    if tb.n_params == 0:
        raise RuntimeError("save_snapshot: there is no network to save (reload_network_from_json first)")
    config = dict(tb.network_config)
    n_params = tb.n_params
    snap = {"n_params": n_params, "params_type": "__half",
            "params_binary": tb.device_buffer("params_inference", torch.float16)[:n_params].cpu().numpy().tobytes()}
    if include_optimizer_state:
        snap["optimizer"] = optimizer_to_json(tb)
    snap["version"] = SNAPSHOT_FORMAT_VERSION
    snap["mode"] = "nerf"
    snap["density_grid_size"] = NERF_GRIDSIZE
    snap["density_grid_binary"] = tb.device_buffer("density_grid", torch.float32).to(torch.float16).cpu().numpy().tobytes()
    aabb_scale = int(tb._get("aabb_scale"))
    # per-image pose / latent optimizers of the reference (AdamOptimizer<vec3>, adam_optimizer.h; to_json in
    # json_binding.h): never stepped on this path, so written in their initial state, one per image, exactly as
    # the reference writes them for an untouched camera (tests/golden/ref_fox_step5000.snapshot.json)
    n_images = int(tb._get("n_images"))
    f32 = lambda v: float(np.float32(v))
    def idle_adam(dims):
        return {"iter": 0, "beta1": f32(0.9), "beta2": f32(0.99), "epsilon": f32(1e-8), "learning_rate": f32(1e-4),
                "first_moment": [0.0] * dims, "second_moment": [0.0] * dims, "variable": [0.0] * dims}
    snap["nerf"] = {"aabb_scale": aabb_scale, "cam_pos_offset": [idle_adam(3) for _ in range(n_images)],
                    "cam_rot_offset": [idle_adam(3) for _ in range(n_images)], "extra_dims_opt": [idle_adam(0) for _ in range(n_images)],
                    "rgb": {"rays_per_batch": tb.rays_per_batch, "measured_batch_size": tb.measured_batch_size,
                            "measured_batch_size_before_compaction": tb.measured_batch_size_before_compaction},
                    "dataset": dataset_to_json(tb)}
    snap["training_step"] = tb.training_step
    snap["loss"] = float(tb.loss)
    snap["aabb"] = _aabb(aabb_scale)
    snap["bounding_radius"] = 1.0
    snap["render_aabb_to_local"] = _mat(np.eye(3))
    snap["render_aabb"] = snap["nerf"]["dataset"]["render_aabb"]
    snap["up_dir"] = _vec(getattr(tb, "up_dir", (0.0, 1.0, 0.0)))
    snap["sun_dir"] = _vec(getattr(tb, "sun_dir", np.float32([1.0, 1.0, 1.0]) / np.sqrt(np.float32(3.0))))
    snap["exposure"] = float(tb.exposure)
    snap["background_color"] = [float(v) for v in tb.background_color]
    cam = tb._camera if tb._camera is not None else np.float32([[1, 0, 0, 0.5], [0, -1, 0, 0.5], [0, 0, -1, 0.5]])
    fxy = [float(v) for v in np.broadcast_to(np.asarray(tb._relative_focal_length, dtype=np.float64), (2,))]   # m_relative_focal_length is a vec2
    snap["camera"] = {"matrix": _mat(cam), "fov_axis": int(tb.fov_axis), "relative_focal_length": fxy,
                      "screen_center": [float(v) for v in tb.screen_center], "zoom": float(getattr(tb, "zoom", 1.0)), "scale": 1.5,
                      "aperture_size": 0.0, "autofocus": False, "autofocus_target": [0.5, 0.5, 0.5], "autofocus_depth": 0.5}
    config["snapshot"] = snap
    return config


def save(tb, path, include_optimizer_state=False, compress=True):
    write_container(path, build_snapshot(tb, include_optimizer_state), compress)


def restore(tb, config):
    """Testbed::load_snapshot(nlohmann::json config), src/testbed.cu:4841-4955."""
    if "snapshot" not in config:
        raise RuntimeError("Given stream does not contain a snapshot.")
    snap = config["snapshot"]
    if snap.get("version", 0) < SNAPSHOT_FORMAT_VERSION:
        raise RuntimeError("Snapshot uses an old format and can not be loaded.")
    if "mode" in snap:
        if snap["mode"] not in MODES or snap["mode"] == "none":
            raise RuntimeError("Unknown snapshot mode. Snapshot must be regenerated with a new version of instant-ngp.")
        if snap["mode"] != "nerf":
            raise RuntimeError("snapshot mode '%s' is not part of the path (NeRF snapshots only)" % snap["mode"])
    if snap["density_grid_size"] != NERF_GRIDSIZE:
        raise RuntimeError("Incompatible grid size.")
    nerf = snap["nerf"]
    ds = nerf.get("dataset")
    if int(tb._get("n_images")) == 0 and ds is not None:
        # no scene loaded: the dataset's metadata comes from the snapshot, enough to render (:4867-4870)
        n = int(ds["n_images"])
        tb.create_empty_nerf_dataset(n, aabb_scale=int(ds["aabb_scale"]), is_hdr=bool(ds.get("is_hdr", False)))
        tb._extrinsics, tb._intrinsics, tb._resolutions = {}, {}, {}
        for i in range(n):
            m = ds["metadata"][i]
            w, h = (int(v) for v in m["resolution"])
            mode, lens = lens_from_json(m.get("lens", m.get("camera_distortion")))
            tb._resolutions[i] = (w, h)
            tb._intrinsics[i] = (float(m["focal_length"][0]), float(m["focal_length"][1]), float(m["principal_point"][0]) * w,
                                 float(m["principal_point"][1]) * h, *lens)
            tb._extrinsics[i] = np.ascontiguousarray(np.asarray(ds["xforms"][i]["start"], dtype=np.float32)[:3, :4])
            if w > 0 and h > 0:
                tb._push_camera(i)
        tb.dataset = {"n_images": n, "aabb_scale": int(ds["aabb_scale"]), "scale": float(ds["scale"]), "offset": _vec(ds["offset"]),
                      "up": _vec(ds["up"]), "render_aabb": ds["render_aabb"], "is_hdr": bool(ds.get("is_hdr", False)),
                      "from_mitsuba": bool(ds.get("from_mitsuba", False)),
                      "frames": [{"path": p} for p in (ds.get("paths") or [""] * n)]}
    elif "aabb_scale" in nerf and int(nerf["aabb_scale"]) != int(tb._get("aabb_scale")):
        # the reference re-derives the scene box from the snapshot's scale here; the images on the device were laid out for the other one
        raise RuntimeError("ngp-b200: the snapshot's aabb_scale (%d) differs from the loaded scene's (%d)" % (int(nerf["aabb_scale"]), int(tb._get("aabb_scale"))))
    tb.exposure = snap.get("exposure", tb.exposure)
    tb.background_color = snap.get("background_color", tb.background_color)
    cam = snap.get("camera")
    if cam is not None:
        if "matrix" in cam:
            tb.set_ngp_camera_matrix(cam["matrix"])
        tb.fov_axis = int(cam.get("fov_axis", tb.fov_axis))
        if "relative_focal_length" in cam:
            rf = [float(v) for v in cam["relative_focal_length"]]
            tb._relative_focal_length = rf[0] if rf[0] == rf[1] else np.array(rf, dtype=np.float64)
        if "screen_center" in cam:
            tb.screen_center = tuple(float(v) for v in cam["screen_center"])
    if "up_dir" in snap:
        tb.up_dir = _vec(snap["up_dir"])
    # m_network_config = config; reset_network(false) rebuilds the trainer; the occupancy grid is the snapshot's
    # (a grid of 0 cells is a model that never trained: it stays cleared)
    tb.reload_network_from_json({k: v for k, v in config.items() if k != "snapshot"})
    grid = np.frombuffer(_binary(snap["density_grid_binary"]), dtype=np.float16)
    check(lib.ngpb_testbed_set_density_grid_fp16(tb._h, grid.ctypes.data_as(ctypes.c_void_p), grid.size))
    check(lib.ngpb_testbed_restore_training_state(tb._h, int(snap["training_step"]), float(snap["loss"])))
    # Trainer::deserialize (tcnn trainer.h:290-314)
    kind = snap.get("params_type", "__half")
    if kind not in ("float", "__half"):
        raise RuntimeError("Trainer: snapshot parameters must be of type float of __half")
    params = np.frombuffer(_binary(snap["params_binary"]), dtype=np.float32 if kind == "float" else np.float16)
    check(lib.ngpb_testbed_set_params(tb._h, params.ctypes.data_as(ctypes.c_void_p), int(kind == "float"), params.size))
    if "optimizer" in snap:
        optimizer_from_json(tb, snap["optimizer"])
    tb._set("rays_per_batch", int(nerf["rgb"]["rays_per_batch"]))
    tb._set("measured_batch_size", int(nerf["rgb"]["measured_batch_size"]))
    tb._set("measured_batch_size_before_compaction", int(nerf["rgb"]["measured_batch_size_before_compaction"]))


def load(tb, path):
    config = read_container(path)
    if "snapshot" not in config:
        raise RuntimeError("File '%s' does not contain a snapshot." % path)
    restore(tb, config)
