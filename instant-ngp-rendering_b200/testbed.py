"""Python mirror of pyngp.Testbed (src/python_api.cu:383-567) for the NeRF mode, over the C ABI.

Names and argument meaning follow the reference binding: create_empty_nerf_dataset,
nerf.training.set_image / set_camera_extrinsics / set_camera_intrinsics, reload_network_from_json /
reload_network_from_file (with "parent" inheritance and // comments, src/testbed.cu:86-97,246-272),
train(batch_size), frame(), render(width, height, spp, linear), training_step, loss, shall_train.
torch owns device buffers only; all compute is in libngp_b200.so.
"""
import ctypes
import json
import os
import struct
import re

import numpy as np
import torch

from ._lib import lib, check, NgpbError

IMAGE_BYTE, IMAGE_HALF, IMAGE_FLOAT = 1, 2, 3
LENS_PERSPECTIVE, LENS_OPENCV = 0, 1
MODE_NERF, MODE_SDF, MODE_IMAGE, MODE_VOLUME, MODE_GEOMETRY, MODE_NONE = 0, 1, 2, 3, 4, 5   # ETestbedMode (common.h:132-139)
LOSS_TYPES = {"L2": 0, "L1": 1, "Mape": 2, "Smape": 3, "Huber": 4, "LogL1": 5, "RelativeL2": 6}
ACTIVATIONS = {"None": 0, "ReLU": 1, "Logistic": 2, "Exponential": 3}

NERF_SCALE = 0.33               # NerfDataset::scale default (nerf_loader.h:66)
NERF_OFFSET = (0.5, 0.5, 0.5)   # NerfDataset::offset default (nerf_loader.h:67)


def nerf_matrix_to_ngp(m, scale=NERF_SCALE, offset=NERF_OFFSET):
    """NerfDataset::nerf_matrix_to_ngp (nerf_loader.h:101-120): 3x4 (or 4x4) camera-to-world in the
    NeRF / blender convention -> NGP convention. Returns float32 [3][4]."""
    r = np.array(m, dtype=np.float32)[:3, :4].copy()
    r[:, 1] *= -1
    r[:, 2] *= -1
    r[:, 3] = r[:, 3] * np.float32(scale) + np.asarray(offset, dtype=np.float32)
    return np.ascontiguousarray(r[[1, 2, 0], :])  # cycle axes xyz <- yzx


def _strip_json_comments(text):
    # nlohmann::json::parse(..., ignore_comments = true): /* */ and // comments
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return re.sub(r"(^|[^:\"])//[^\n]*", r"\1", text)


def merge_patch(target, patch):
    """RFC 7386, as nlohmann::json::merge_patch."""
    if not isinstance(patch, dict):
        return patch
    if not isinstance(target, dict):
        target = {}
    out = dict(target)
    for k, v in patch.items():
        if v is None:
            out.pop(k, None)
        else:
            out[k] = merge_patch(out.get(k), v)
    return out


def load_network_config(path):
    """load_network_config + merge_parent_network_config (src/testbed.cu:86-97, 246-272)."""
    with open(path) as f:
        cfg = json.loads(_strip_json_comments(f.read()))
    if "parent" in cfg:
        parent = load_network_config(os.path.join(os.path.dirname(path), cfg["parent"]))
        cfg = merge_patch(parent, {k: v for k, v in cfg.items() if k != "parent"})
    return cfg


class _ImageMetadata:
    """One entry of testbed.nerf.training.dataset.metadata (TrainingImageMetadata, python_api.cu:578-590)."""

    def __init__(self, tb, idx):
        w, h = tb._resolutions[idx]
        fx, fy, cx, cy, k1, k2, p1, p2 = tb._intrinsics.get(idx, (0.0, 0.0, -0.5, -0.5, 0.0, 0.0, 0.0, 0.0))
        self.resolution = [w, h]
        self.focal_length = [fx, fy if fy > 0 else fx]
        self.principal_point = [cx / w if cx >= 0 else 0.5, cy / h if cy >= 0 else 0.5]
        self.lens_params = [k1, k2, p1, p2]
        self.lens_mode = LENS_OPENCV if any(v != 0.0 for v in self.lens_params) else LENS_PERSPECTIVE


class _Dataset:
    """testbed.nerf.training.dataset (NerfDataset, python_api.cu:592-612): what scripts/run.py reads."""

    def __init__(self, tb):
        self._tb = tb

    @property
    def n_images(self):
        return int(self._tb._get("n_images"))

    @property
    def metadata(self):
        return [_ImageMetadata(self._tb, i) for i in range(self.n_images)]

    @property
    def aabb_scale(self):
        return int(self._tb._get("aabb_scale"))

    def _record(self, key, default):
        ds = getattr(self._tb, "dataset", None)
        return ds[key] if ds is not None and key in ds else default

    scale = property(lambda self: self._record("scale", NERF_SCALE))
    offset = property(lambda self: self._record("offset", list(NERF_OFFSET)))
    is_hdr = property(lambda self: bool(self._record("is_hdr", False)))
    paths = property(lambda self: [fr.get("path", "") for fr in self._record("frames", [])])


class _Training:
    """testbed.nerf.training (src/python_api.cu:645-697)."""

    def __init__(self, tb):
        self._tb = tb
        self._dataset = _Dataset(tb)

    dataset = property(lambda self: self._dataset)
    n_images_for_training = property(lambda self: self._dataset.n_images)

    def set_image(self, frame_idx, img, depth_img=None, depth_scale=1.0):
        self._tb.set_nerf_image(frame_idx, img)

    def set_camera_extrinsics(self, frame_idx, camera_to_world, convert_to_ngp=True):
        self._tb._extrinsics[frame_idx] = nerf_matrix_to_ngp(camera_to_world) if convert_to_ngp else np.ascontiguousarray(np.array(camera_to_world, dtype=np.float32)[:3, :4])
        self._tb._push_camera(frame_idx)

    def set_camera_intrinsics(self, frame_idx, fx=0.0, fy=0.0, cx=-0.5, cy=-0.5, k1=0.0, k2=0.0, p1=0.0, p2=0.0):
        self._tb._intrinsics[frame_idx] = (fx, fy, cx, cy, k1, k2, p1, p2)
        self._tb._push_camera(frame_idx)

    def __getattr__(self, name):
        if name in ("random_bg_color", "linear_colors", "snap_to_pixel_centers", "near_distance", "density_grid_decay", "loss_type", "train_network", "train_encoding"):
            return self._tb._get(name)
        raise AttributeError(name)

    def __setattr__(self, name, value):
        if name.startswith("_"):
            object.__setattr__(self, name, value)
        else:
            self._tb._set(name, float(value))


class _Nerf:
    def __init__(self, tb):
        self.training = _Training(tb)
        self._tb = tb

    @property
    def render_min_transmittance(self):
        return self._tb._get("render_min_transmittance")

    @render_min_transmittance.setter
    def render_min_transmittance(self, v):
        self._tb._set("render_min_transmittance", float(v))

    # python_api.cu:622-643
    cone_angle_constant = property(lambda self: self._tb._get("cone_angle_constant"), lambda self, v: self._tb._set("cone_angle_constant", float(v)))
    render_with_lens_distortion = property(lambda self: bool(self._tb._get("render_with_lens_distortion")),
                                           lambda self, v: self._tb._set("render_with_lens_distortion", float(bool(v))))
    rgb_activation = property(lambda self: int(self._tb._get("rgb_activation")), lambda self, v: self._tb._set("rgb_activation", int(v)))
    density_activation = property(lambda self: int(self._tb._get("density_activation")), lambda self, v: self._tb._set("density_activation", int(v)))

    @property
    def sharpen(self):
        return self._sharpen

    @sharpen.setter
    def sharpen(self, v):
        """Dataset sharpening at load time (nerf_loader.cu:762-770). Only 0 (off, the default) is built."""
        if float(v) != 0.0:
            raise NgpbError("ngp-b200: nerf.sharpen != 0 is not implemented (out of the path, DESIGN.md section 7)")
        self._sharpen = 0.0

    _sharpen = 0.0


class Testbed:
    def __init__(self, device="cuda"):
        h = ctypes.c_void_p()
        check(lib.ngpb_testbed_create(ctypes.byref(h)))
        self._h = h
        self._device = device
        self._extrinsics = {}
        self._intrinsics = {}
        self._resolutions = {}
        self.shall_train = True
        self.nerf = _Nerf(self)
        self.mode = MODE_NONE
        self.root_dir = ""
        self.training_batch_size = 1 << 18   # Testbed::m_training_batch_size (testbed.h)
        self.render_ground_truth = False
        self.ground_truth_alpha = 1.0
        self.zoom = 1.0

    # -- plumbing --------------------------------------------------------------------------------
    def _get(self, name):
        v = ctypes.c_double()
        check(lib.ngpb_testbed_get(self._h, name.encode(), ctypes.byref(v)))
        return v.value

    def _set(self, name, value):
        check(lib.ngpb_testbed_set(self._h, name.encode(), float(value)))

    def device_buffer(self, name, dtype):
        """torch view (no copy) of a named device buffer of the testbed."""
        p, n = ctypes.c_void_p(), ctypes.c_size_t()
        check(lib.ngpb_testbed_device_ptr(self._h, name.encode(), ctypes.byref(p), ctypes.byref(n)))
        itemsize = torch.empty((), dtype=dtype).element_size()
        count = n.value // itemsize

        class _Iface:
            __cuda_array_interface__ = {"shape": (count,), "typestr": {torch.float16: "<f2", torch.float32: "<f4", torch.uint8: "|u1", torch.int32: "<i4"}[dtype],
                                        "data": (p.value, False), "version": 2}
        return torch.as_tensor(_Iface(), device=self._device)

    def close(self):
        if self._h:
            lib.ngpb_testbed_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def use_current_stream(self):
        """Run every call on torch's current stream (so that torch CUDA events bracket the work and
        torch.distributed collectives are ordered with it). torch's default stream has handle 0,
        which the C ABI reads as "the testbed's own stream": pass cudaStreamLegacy (0x1) instead."""
        handle = torch.cuda.current_stream().cuda_stream or 1
        check(lib.ngpb_testbed_set_stream(self._h, ctypes.c_void_p(handle)))

    # -- dataset ---------------------------------------------------------------------------------
    def create_empty_nerf_dataset(self, n_images, aabb_scale=1, is_hdr=False):
        check(lib.ngpb_testbed_create_empty_nerf_dataset(self._h, n_images, int(aabb_scale), int(is_hdr)))
        self.mode = MODE_NERF
        self._extrinsics, self._intrinsics, self._resolutions = {}, {}, {}

    def set_nerf_image(self, idx, img):
        """img: [h][w][4] uint8 (sRGB, straight alpha), float16 or float32 (linear, premultiplied);
        numpy array or torch tensor (host or device)."""
        if isinstance(img, np.ndarray):
            img = torch.from_numpy(np.ascontiguousarray(img))
        img = img.contiguous()
        assert img.dim() == 3 and img.shape[2] == 4
        kind = {torch.uint8: IMAGE_BYTE, torch.float16: IMAGE_HALF, torch.float32: IMAGE_FLOAT}[img.dtype]
        h, w = int(img.shape[0]), int(img.shape[1])
        if img.is_cuda:
            torch.cuda.current_stream().synchronize()  # the library copies on the testbed's own stream
        check(lib.ngpb_testbed_set_nerf_image(self._h, idx, ctypes.c_void_p(img.data_ptr()), kind, w, h))
        self._resolutions[idx] = (w, h)
        self._push_camera(idx)

    def _push_camera(self, idx):
        if idx not in self._extrinsics or idx not in self._intrinsics or idx not in self._resolutions:
            return
        fx, fy, cx, cy, k1, k2, p1, p2 = self._intrinsics[idx]
        w, h = self._resolutions[idx]
        if fy <= 0:
            fy = fx
        px = cx / w if cx >= 0 else 0.5
        py = cy / h if cy >= 0 else 0.5
        lens = (ctypes.c_float * 7)(k1, k2, p1, p2, 0, 0, 0)
        mode = LENS_OPENCV if any(v != 0.0 for v in (k1, k2, p1, p2)) else LENS_PERSPECTIVE
        x = np.ascontiguousarray(self._extrinsics[idx], dtype=np.float32)
        check(lib.ngpb_testbed_set_nerf_camera(self._h, idx, x.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), fx, fy, px, py, mode, lens))

    # -- neural image mode (pyngp: load_training_data of an image file, src/python_api.cu:415) -------
    def set_image(self, img):
        """img: [h][w][4] float32 or float16, linear colours; numpy array or torch tensor."""
        if isinstance(img, np.ndarray):
            img = torch.from_numpy(np.ascontiguousarray(img))
        img = img.contiguous()
        assert img.dim() == 3 and img.shape[2] == 4
        kind = {torch.float16: IMAGE_HALF, torch.float32: IMAGE_FLOAT}[img.dtype]
        if img.is_cuda:
            torch.cuda.current_stream().synchronize()
        check(lib.ngpb_testbed_set_image(self._h, ctypes.c_void_p(img.data_ptr()), kind, int(img.shape[1]), int(img.shape[0])))

    def render_image(self, width, height):
        """Linear RGBA float32 [h][w][4] of the fitted image at the pixel centres."""
        rgba = torch.empty((height, width, 4), dtype=torch.float32, device=self._device)
        torch.cuda.current_stream().synchronize()
        check(lib.ngpb_testbed_render_image(self._h, width, height, ctypes.c_void_p(rgba.data_ptr())))
        return rgba

    # -- network ---------------------------------------------------------------------------------
    def reload_network_from_json(self, config):
        check(lib.ngpb_testbed_reload_network_from_json(self._h, json.dumps(config).encode()))

    def reload_network_from_file(self, path):
        self.reload_network_from_json(load_network_config(path))

    def reset_network(self, clear_density_grid=True):
        check(lib.ngpb_testbed_reset_network(self._h, int(clear_density_grid)))

    @property
    def network_config(self):
        n = lib.ngpb_testbed_network_config(self._h, None, 0)
        buf = ctypes.create_string_buffer(n)
        lib.ngpb_testbed_network_config(self._h, buf, n)
        return json.loads(buf.value.decode())

    # -- training --------------------------------------------------------------------------------
    def train(self, batch_size=1 << 18):
        check(lib.ngpb_testbed_train(self._h, batch_size))

    # -- data-parallel training (one process per GPU; no reference counterpart) --------------------
    def set_shard(self, rank, world):
        """This testbed computes rays [rank R/P, (rank+1) R/P) of every step's R rays."""
        self._set("shard_world", 1)
        self._set("shard_rank", 0)
        self._set("shard_world", world)
        self._set("shard_rank", rank)
        self._dp = None

    def init_data_parallel(self, group=None):
        """Sharded training inside the library (csrc/dist.cu): rank 0 makes two NCCL unique ids, torch.distributed carries
        them to the other ranks (bootstrap only), every rank builds its communicators, and from then on train() /
        train_distributed() run the data-parallel step - reduce-scatter, sharded Adam, all-gather - without Python in
        the data path. A C++ host does the same through ngpb_nccl_get_unique_id / ngpb_testbed_init_data_parallel."""
        import torch.distributed as dist
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        ids = torch.zeros(256, dtype=torch.uint8)
        if rank == 0:
            buf = (ctypes.c_uint8 * 256)()
            check(lib.ngpb_nccl_get_unique_id(ctypes.byref(buf)))
            check(lib.ngpb_nccl_get_unique_id(ctypes.byref(buf, 128)))
            ids = torch.frombuffer(bytearray(buf), dtype=torch.uint8).clone()
        dev = ids.to(self._device) if dist.get_backend(group) == "nccl" else ids
        dist.broadcast(dev, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        raw = bytes(dev.cpu().numpy().tobytes())
        a, b = ctypes.create_string_buffer(raw[:128], 128), ctypes.create_string_buffer(raw[128:], 128)
        torch.cuda.synchronize()
        check(lib.ngpb_testbed_init_data_parallel(self._h, a, b, rank, world))
        self._dp = None
        self._dp_native = world > 1

    def train_distributed(self, batch_size=1 << 18, group=None):
        """Testbed.train for a sharded run. After init_data_parallel() this is train(): the library runs the whole step.
        Without it (the CPU tests of the host logic use gloo, which NCCL cannot replace) the protocol is driven from
        here: local gradients, all-reduce (sum) of the counters and the gradients over `group` with torch.distributed,
        the full optimizer step on every rank. Needs use_current_stream() so that the collectives are ordered with the
        kernels."""
        if getattr(self, "_dp_native", False) or self._get("shard_world") <= 1:
            check(lib.ngpb_testbed_train(self._h, batch_size))
            return
        # buffers are looked up every step: reload_network_from_json / load_snapshot / mode changes reallocate them
        c = self.device_buffer("counters", torch.int32)
        gradients, counts, loss_sum = self.device_buffer("gradients", torch.float16), c[:3], c[3:4].view(torch.float32)
        from .sharding import allreduce_counters, allreduce_gradients
        with_loss = self.training_step % 16 == 0   # the loss sum is only filled on these steps
        check(lib.ngpb_testbed_train_forward(self._h, batch_size))
        allreduce_counters(counts, loss_sum if with_loss else None, group=group)
        check(lib.ngpb_testbed_train_backward(self._h, batch_size))
        allreduce_gradients(gradients, group=group)
        check(lib.ngpb_testbed_train_end(self._h, batch_size))

    def frame(self):
        """Headless Testbed::frame: one training step when shall_train (src/testbed.cu:3380-3479)."""
        if self.shall_train:
            self.train(self.training_batch_size)
        return True

    def want_repl(self):
        return False

    def init_window(self, width, height, hidden=False, second_window=False):
        raise NgpbError("init_window: no GUI in ngp-b200 (GLFW / OpenGL are unavailable on this box); train() / frame() / render() run headless")

    def init_vr(self):
        raise NgpbError("init_vr: no VR in ngp-b200 (unavailable on this box)")

    def load_file(self, path):
        """Testbed::load_file (src/testbed.cu:318-391): snapshot, network config, or training data by what the file is."""
        path = str(path)
        ext = os.path.splitext(path)[1].lower()
        if not os.path.exists(path):
            if ext == ".json":
                for cand in (path, os.path.join(self.root_dir, "configs", {MODE_NERF: "nerf", MODE_IMAGE: "image"}.get(self.mode, "nerf"), os.path.basename(path))):
                    if os.path.exists(cand):
                        return self.reload_network_from_file(cand)
            raise NgpbError(f"File '{path}' does not exist.")
        if ext in (".ingp", ".msgpack"):
            return self.load_snapshot(path)
        if ext == ".json":
            with open(path) as f:
                doc = json.loads(_strip_json_comments(f.read()))
            if "snapshot" in doc:
                raise NgpbError("ngp-b200: snapshots in JSON form are not supported; use .ingp / .msgpack")
            if "path" in doc and "frames" not in doc:
                raise NgpbError("ngp-b200: camera paths are not part of the path (DESIGN.md section 7)")
            if "parent" in doc or "network" in doc or "encoding" in doc or "loss" in doc or "optimizer" in doc:
                return self.reload_network_from_file(path)
        return self.load_training_data(path)

    training_step = property(lambda self: int(self._get("training_step")))
    loss = property(lambda self: self._get("loss"))
    n_params = property(lambda self: int(self._get("n_params")))
    rays_per_batch = property(lambda self: int(self._get("rays_per_batch")))
    measured_batch_size = property(lambda self: int(self._get("measured_batch_size")))
    measured_batch_size_before_compaction = property(lambda self: int(self._get("measured_batch_size_before_compaction")))

    def update_density_grid(self, decay, n_uniform, n_nonuniform):
        check(lib.ngpb_testbed_update_density_grid(self._h, decay, n_uniform, n_nonuniform))

    # -- rendering -------------------------------------------------------------------------------
    def render_nerf(self, width, height, camera_3x4, focal, principal=(0.5, 0.5), spp_index=0, pixel_range=None, return_steps=False, shard=None):
        """Linear premultiplied RGBA float32 [h][w][4] and depth [h][w] on the device. camera_3x4 in
        NGP convention; focal in pixels. pixel_range=(begin, end) renders a slice of the row-major
        pixel list; shard=(rank, world) traces only this rank's 256-pixel tiles of that range (tile
        sharding across GPUs, no communication); the rest stays zero. return_steps adds the per-pixel count of
        composited samples (int32 [h][w])."""
        rgba = torch.zeros((height, width, 4), dtype=torch.float32, device=self._device)
        depth = torch.zeros((height, width), dtype=torch.float32, device=self._device)
        steps = torch.zeros((height, width), dtype=torch.int32, device=self._device) if return_steps else None
        cam = np.ascontiguousarray(camera_3x4, dtype=np.float32)
        fx, fy = (focal, focal) if np.isscalar(focal) else focal
        b, e = pixel_range if pixel_range is not None else (0, width * height)
        rank, world = shard if shard is not None else (0, 1)
        self._set("render_shard_rank", 0)
        self._set("render_shard_world", world)
        self._set("render_shard_rank", rank)
        torch.cuda.current_stream().synchronize()  # the testbed runs on its own stream unless use_current_stream()
        check(lib.ngpb_testbed_set_render_steps_out(self._h, ctypes.c_void_p(steps.data_ptr()) if return_steps else None))
        try:
            check(lib.ngpb_testbed_render_nerf(self._h, width, height, cam.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), fx, fy,
                                               principal[0], principal[1], spp_index, b, e, ctypes.c_void_p(rgba.data_ptr()), ctypes.c_void_p(depth.data_ptr())))
        finally:
            lib.ngpb_testbed_set_render_steps_out(self._h, None)
        return (rgba, depth, steps) if return_steps else (rgba, depth)

    last_render_samples = property(lambda self: int(self._get("last_render_samples")))

    # -- dataset files (Testbed::load_training_data, src/testbed.cu; load_nerf, src/nerf_loader.cu:273-740)
    def load_training_data(self, path, max_images=None):
        """transforms.json (or a directory holding some) -> images, cameras and lens models on the
        device, as pyngp's load_training_data does for NeRF scenes. Returns the dataset record of
        loader.describe_dataset (scale, offset, up, render_aabb, frames ...). An image file selects
        the neural-image mode (mode_from_scene, src/testbed.cu:180-200); meshes and volumes are not
        part of the path."""
        from . import loader
        path = str(path)
        ext = os.path.splitext(path)[1].lower()
        if ext in (".obj", ".stl", ".nvdb"):
            raise NgpbError("ngp-b200: SDF and volume modes are out of scope (DESIGN.md section 7)")
        if os.path.isfile(path) and ext not in (".json", ""):
            img = loader.load_image_file(path)
            self.set_image(img)
            self.mode = MODE_IMAGE
            return None
        if self.mode == MODE_IMAGE:
            raise NgpbError("ngp-b200: this testbed is in image mode; create a new Testbed for a NeRF scene")
        ds = loader.describe_dataset(path)
        frames = ds["frames"][:max_images] if max_images else ds["frames"]
        self.create_empty_nerf_dataset(len(frames), aabb_scale=ds["aabb_scale"], is_hdr=ds["is_hdr"])
        for i, fr in enumerate(frames):
            px = loader.load_frame_pixels(fr, ds)
            h, w = px.shape[:2]
            fx, fy = loader.finish_frame(fr, w, h)
            self.nerf.training.set_image(i, px)
            cx, cy = fr["principal"]
            k1, k2, p1, p2 = fr["lens_params"][:4] if fr["lens_mode"] == 1 else (0.0, 0.0, 0.0, 0.0)
            self.nerf.training.set_camera_intrinsics(i, fx=fx, fy=fy, cx=cx * w, cy=cy * h, k1=k1, k2=k2, p1=p1, p2=p2)
            self.nerf.training.set_camera_extrinsics(i, fr["xform"], convert_to_ngp=False)
        self.dataset = ds
        self.up_dir = ds["up"]
        self.mode = MODE_NERF
        return ds

    def set_camera_to_training_view(self, trainview):
        """Testbed::set_camera_to_training_view (src/testbed.cu:472-487): camera, focal length, lens and
        screen centre of a training view; it also becomes the view render_ground_truth shows."""
        trainview = int(trainview)
        if trainview not in self._extrinsics or trainview not in self._resolutions:
            raise NgpbError("Invalid frame index")
        md = _ImageMetadata(self, trainview)
        self._camera = np.ascontiguousarray(self._extrinsics[trainview], dtype=np.float32)
        self._relative_focal_length = np.array(md.focal_length, dtype=np.float64) / float(md.resolution[self.fov_axis])
        self.nerf.render_with_lens_distortion = True
        lens = (ctypes.c_float * 7)(*md.lens_params, 0, 0, 0)
        check(lib.ngpb_testbed_set_render_lens(self._h, md.lens_mode, lens))
        self.screen_center = (1.0 - md.principal_point[0], 1.0 - md.principal_point[1])
        self.training_view = trainview

    training_view = 0
    snap_to_pixel_centers = property(lambda self: bool(self._get("render_snap_to_pixel_centers")), lambda self, v: self._set("render_snap_to_pixel_centers", float(bool(v))))

    @property
    def color_space(self):
        return int(self._get("render_color_space"))

    @color_space.setter
    def color_space(self, v):
        """Testbed::m_color_space: the space colours are trained and accumulated in (python_api.cu:503)."""
        self._set("color_space", int(v))
        self._set("render_color_space", int(v))

    def reset(self, reset_density_grid=True):
        self.reset_network(reset_density_grid)

    def n_encoding_params(self):
        return int(self.n_params) - sum(int(np.prod(sh)) for sh in getattr(self, "_mlp_shapes", []))

    def compute_and_save_marching_cubes_mesh(self, *a, **k):
        raise NgpbError("ngp-b200: marching-cubes mesh extraction is out of scope (DESIGN.md section 7)")

    compute_marching_cubes_mesh = compute_and_save_marching_cubes_mesh

    def load_camera_path(self, path):
        raise NgpbError("ngp-b200: camera paths are out of scope (DESIGN.md section 7)")

    # -- snapshots (pyngp save_snapshot / load_snapshot, src/python_api.cu:432-434; src/testbed.cu:4775-4972)
    def save_snapshot(self, path, include_optimizer_state=False, compress=True):
        """The reference's snapshot file: msgpack of the network config with the "snapshot" object
        (fp16 inference parameters, fp16 occupancy grid, step, loss, camera, dataset metadata);
        *.ingp is zlib-compressed, any other name is plain msgpack."""
        from . import snapshot
        snapshot.save(self, path, include_optimizer_state, compress)

    def load_snapshot(self, path):
        from . import snapshot
        snapshot.load(self, path)

    # -- the same file-level entry points in the library's host C++ (csrc/host_io.cu, host_snapshot.cu, host_dataset.cu) --
    def image_metadata_native(self, idx):
        """ngpb_nerf_image_t of view idx as the C++ testbed holds it."""
        buf = (ctypes.c_uint8 * 128)()
        check(lib.ngpb_testbed_image_metadata(self._h, int(idx), ctypes.cast(buf, ctypes.c_void_p)))
        raw = bytes(buf)
        pixels, = struct.unpack_from("<Q", raw, 0)
        image_type, lens_mode = struct.unpack_from("<ii", raw, 8)
        lens = struct.unpack_from("<7f", raw, 16)
        res = struct.unpack_from("<ii", raw, 44)
        principal = struct.unpack_from("<ff", raw, 52)
        focal = struct.unpack_from("<ff", raw, 60)
        cols = np.array(struct.unpack_from("<12f", raw, 68), dtype=np.float32).reshape(4, 3)   # Mat4x3: four columns
        return {"pixels": pixels, "image_data_type": image_type, "lens_mode": lens_mode, "lens_params": list(lens), "resolution": list(res),
                "principal_point": list(principal), "focal_length": list(focal), "xform": np.ascontiguousarray(cols.T)}

    def _sync_from_native(self):
        """The Python mirror's per-view records and dataset dict, read back from the C++ testbed."""
        n = int(self._get("n_images"))
        self._extrinsics, self._intrinsics, self._resolutions = {}, {}, {}
        for i in range(n):
            m = self.image_metadata_native(i)
            w, h = m["resolution"]
            self._resolutions[i] = (w, h)
            k = m["lens_params"][:4] if m["lens_mode"] == LENS_OPENCV else [0.0] * 4
            self._intrinsics[i] = (m["focal_length"][0], m["focal_length"][1], m["principal_point"][0] * w, m["principal_point"][1] * h, *k)
            self._extrinsics[i] = m["xform"]
        size = lib.ngpb_testbed_dataset_info(self._h, None, 0)
        buf = ctypes.create_string_buffer(size)
        lib.ngpb_testbed_dataset_info(self._h, buf, size)
        info = json.loads(buf.value.decode())
        ra = info.get("render_aabb")
        self.dataset = {"n_images": n, "aabb_scale": int(self._get("aabb_scale")), "scale": info["scale"], "offset": info["offset"], "up": info["up"],
                        "render_aabb": {"min": ra[:3], "max": ra[3:]} if ra else None, "is_hdr": info["is_hdr"], "from_mitsuba": info["from_mitsuba"],
                        "frames": [{"path": p} for p in info["paths"]]}
        self.up_dir = info["up"]
        view = (ctypes.c_float * 24)()
        check(lib.ngpb_testbed_view_state(self._h, view, 0))
        v = list(view)
        self._camera = np.array(v[:12], dtype=np.float32).reshape(3, 4)
        self._relative_focal_length = v[12] if v[12] == v[13] else np.array(v[12:14], dtype=np.float64)
        self.screen_center = (v[14], v[15])
        self.zoom = v[16]
        self.fov_axis = int(v[23])

    def _push_dataset_to_native(self):
        """NerfDataset's scalar members of the Python mirror -> the C++ testbed (what its save_snapshot writes)."""
        ds = getattr(self, "dataset", None)
        if not ds:
            return
        info = {"scale": float(ds.get("scale", 0.33)), "offset": [float(v) for v in ds.get("offset", (0.5, 0.5, 0.5))], "up": [float(v) for v in ds.get("up", (0.0, 1.0, 0.0))],
                "paths": [str(fr.get("path", "")) for fr in ds.get("frames", [])][: int(self._get("n_images"))],
                "from_mitsuba": bool(ds.get("from_mitsuba", False)), "is_hdr": bool(ds.get("is_hdr", False))}
        ra = ds.get("render_aabb")
        if isinstance(ra, dict):
            info["render_aabb"] = [float(v) for v in ra["min"]] + [float(v) for v in ra["max"]]
        elif ra:
            info["render_aabb"] = [float(v) for v in ra[0]] + [float(v) for v in ra[1]]
        check(lib.ngpb_testbed_set_dataset_info(self._h, json.dumps(info).encode()))

    def _push_view_to_native(self):
        cam = self._camera if self._camera is not None else np.float32([[1, 0, 0, 0.5], [0, -1, 0, 0.5], [0, 0, -1, 0.5]])
        rf = np.broadcast_to(np.asarray(self._relative_focal_length, dtype=np.float64), (2,))
        up = getattr(self, "up_dir", (0.0, 1.0, 0.0))
        sun = getattr(self, "sun_dir", np.float32([1.0, 1.0, 1.0]) / np.sqrt(np.float32(3.0)))
        vals = [float(x) for x in np.asarray(cam, dtype=np.float32).reshape(-1)] + [float(rf[0]), float(rf[1])] + [float(x) for x in self.screen_center] + \
               [float(self.zoom)] + [float(x) for x in up] + [float(x) for x in sun] + [float(self.fov_axis)]
        check(lib.ngpb_testbed_view_state(self._h, (ctypes.c_float * 24)(*vals), 1))

    def load_training_data_native(self, path, max_images=0):
        """ngpb_testbed_load_training_data: the C++ loader (JPEG on the GPU through nvjpeg, PNG on the host)."""
        check(lib.ngpb_testbed_load_training_data(self._h, str(path).encode(), int(max_images)))
        if os.path.isfile(str(path)) and os.path.splitext(str(path))[1].lower() not in (".json", ""):
            self.mode = MODE_IMAGE   # an image file selects the neural-image mode (mode_from_scene, src/testbed.cu:180-200)
            return None
        self.mode = MODE_NERF
        self._sync_from_native()
        return self.dataset

    def reload_network_from_file_native(self, path):
        check(lib.ngpb_testbed_reload_network_from_file(self._h, str(path).encode()))

    def save_snapshot_native(self, path, include_optimizer_state=False, compress=True):
        self._push_dataset_to_native()
        self._push_view_to_native()
        check(lib.ngpb_testbed_save_snapshot(self._h, str(path).encode(), int(include_optimizer_state), int(compress)))

    def load_snapshot_native(self, path):
        check(lib.ngpb_testbed_load_snapshot(self._h, str(path).encode()))
        self.mode = MODE_NERF
        self._sync_from_native()

    # -- pyngp's camera and render() (src/python_api.cu:124-221, 416-425, 481-566) ---------------
    def set_nerf_camera_matrix(self, camera_to_world):
        """Testbed::set_nerf_camera_matrix (src/testbed.cu:420-422): NeRF-convention 3x4 -> the view camera."""
        self._camera = nerf_matrix_to_ngp(camera_to_world)

    def set_ngp_camera_matrix(self, camera_3x4):
        self._camera = np.ascontiguousarray(np.array(camera_3x4, dtype=np.float32)[:3, :4])

    @property
    def fov(self):
        """Field of view in degrees along fov_axis (Testbed::fov / set_fov, src/testbed.cu:3526-3536)."""
        rel = float(np.broadcast_to(np.asarray(self._relative_focal_length, dtype=np.float64), (2,))[self.fov_axis])
        return float(np.degrees(2.0 * np.arctan(0.5 / rel)))

    @fov.setter
    def fov(self, degrees):
        self._relative_focal_length = 0.5 / np.tan(0.5 * np.radians(float(degrees)))

    @property
    def background_color(self):
        return [self._get("background_color_" + c) for c in "rgba"]

    @background_color.setter
    def background_color(self, rgba):
        for c, v in zip("rgba", rgba):
            self._set("background_color_" + c, v)

    exposure = property(lambda self: self._get("exposure"), lambda self, v: self._set("exposure", v))
    tonemap_curve = property(lambda self: int(self._get("tonemap_curve")), lambda self, v: self._set("tonemap_curve", int(v)))

    def render(self, width=1920, height=1080, spp=1, linear=True, start_t=-1.0, end_t=-1.0, fps=30.0, shutter_fraction=1.0):
        """pyngp's Testbed.render: float32 [h][w][4] on the host - `spp` samples per pixel averaged, the
        background blended in, exposure, tonemap curve, sRGB transfer unless `linear`. The view is the
        one set by set_nerf_camera_matrix / fov / screen_center (focal length = relative focal length
        x resolution[fov_axis], Testbed::calc_focal_length)."""
        if start_t >= 0.0 or end_t >= 0.0:
            raise NgpbError("render: camera-path rendering (start_t / end_t) is out of scope (DESIGN.md section 7)")
        width, height, spp = int(width), int(height), int(spp)
        rgba = torch.zeros((height, width, 4), dtype=torch.float32, device=self._device)
        torch.cuda.current_stream().synchronize()
        gt = bool(self.render_ground_truth) and self.mode != MODE_IMAGE
        if not gt or self.ground_truth_alpha < 1.0:   # the reference skips the trace under an opaque overlay (src/testbed.cu:4383)
            if getattr(self, "_camera", None) is None:
                raise NgpbError("render: set the view with set_nerf_camera_matrix first")
            # calc_focal_length and render_screen_center (src/testbed.cu:4111-4118)
            rel = np.broadcast_to(np.asarray(self._relative_focal_length, dtype=np.float64), (2,))
            fx, fy = (rel * (height if self.fov_axis == 1 else width) * self.zoom).tolist()
            cx, cy = ((0.5 - np.asarray(self.screen_center, dtype=np.float64)) * self.zoom + 0.5).tolist()
            cam = np.ascontiguousarray(self._camera, dtype=np.float32)
            self._set("render_shard_rank", 0)
            self._set("render_shard_world", 1)
            check(lib.ngpb_testbed_render(self._h, width, height, spp, int(bool(linear)), cam.ctypes.data_as(ctypes.POINTER(ctypes.c_float)),
                                          fx, fy, cx, cy, ctypes.c_void_p(rgba.data_ptr())))
        if gt:
            check(lib.ngpb_testbed_overlay_ground_truth(self._h, width, height, int(bool(linear)), int(self.training_view), int(self.fov_axis),
                                                        float(self.zoom), float(self.ground_truth_alpha), ctypes.c_void_p(rgba.data_ptr())))
        self.device_buffer("counters", torch.int32)   # orders the read below behind the testbed's stream
        return rgba.cpu().numpy()

    fov_axis = 1
    screen_center = (0.5, 0.5)
    _relative_focal_length = 1.0
    _camera = None


def profile_enable(on=True):
    lib.ngpb_profile_enable(int(on))


def profile_collect(reset=True):
    """{name: (total_ms, count)} of the library's phase timer."""
    n = lib.ngpb_profile_collect(None, 0, 0)
    buf = ctypes.create_string_buffer(n + 1)
    lib.ngpb_profile_collect(buf, n + 1, int(reset))
    out = {}
    for item in buf.value.decode().split(";"):
        if "=" in item:
            k, v = item.split("=")
            ms, cnt = v.split(":")
            out[k] = (float(ms), int(cnt))
    return out
