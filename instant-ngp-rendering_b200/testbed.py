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
import re

import numpy as np
import torch

from ._lib import lib, check, NgpbError

IMAGE_BYTE, IMAGE_HALF, IMAGE_FLOAT = 1, 2, 3
LENS_PERSPECTIVE, LENS_OPENCV = 0, 1
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
    return re.sub(r"//[^\n]*", "", text)


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


class _Training:
    """testbed.nerf.training (src/python_api.cu:645-697)."""

    def __init__(self, tb):
        self._tb = tb

    def set_image(self, frame_idx, img, depth_img=None, depth_scale=1.0):
        self._tb.set_nerf_image(frame_idx, img)

    def set_camera_extrinsics(self, frame_idx, camera_to_world, convert_to_ngp=True):
        self._tb._extrinsics[frame_idx] = nerf_matrix_to_ngp(camera_to_world) if convert_to_ngp else np.ascontiguousarray(np.array(camera_to_world, dtype=np.float32)[:3, :4])
        self._tb._push_camera(frame_idx)

    def set_camera_intrinsics(self, frame_idx, fx=0.0, fy=0.0, cx=-0.5, cy=-0.5, k1=0.0, k2=0.0, p1=0.0, p2=0.0):
        self._tb._intrinsics[frame_idx] = (fx, fy, cx, cy, k1, k2, p1, p2)
        self._tb._push_camera(frame_idx)

    def __getattr__(self, name):
        if name in ("random_bg_color", "linear_colors", "snap_to_pixel_centers", "near_distance", "density_grid_decay"):
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

    def train_distributed(self, batch_size=1 << 18, group=None):
        """Testbed.train for a sharded run: local gradients, all-reduce (sum) of the gradients and
        the step counters over `group` with torch.distributed (NCCL over NVLink on GPUs, gloo in
        the CPU tests of the host logic), then the optimizer step on every rank. Needs
        use_current_stream() so that the collectives are ordered with the kernels."""
        if self._get("shard_world") <= 1:
            check(lib.ngpb_testbed_train(self._h, batch_size))
            return
        if getattr(self, "_dp", None) is None:
            c = self.device_buffer("counters", torch.int32)   # fixed address while the testbed is sharded
            self._dp = (self.device_buffer("gradients", torch.float16), c[:3], c[3:4].view(torch.float32))
        from .sharding import allreduce_counters, allreduce_gradients
        gradients, counts, loss_sum = self._dp
        with_loss = self.training_step % 16 == 0   # the loss sum is only filled on these steps
        # counters first: with the global sample count on its way to the host, the next step's rays
        # are marched under this step's backward pass, gradient all-reduce and optimizer
        check(lib.ngpb_testbed_train_forward(self._h, batch_size))
        allreduce_counters(counts, loss_sum if with_loss else None, group=group)
        check(lib.ngpb_testbed_train_backward(self._h, batch_size))
        allreduce_gradients(gradients, group=group)
        check(lib.ngpb_testbed_train_end(self._h, batch_size))

    def frame(self):
        """Headless Testbed::frame: one training step when shall_train (src/testbed.cu:3380-3479)."""
        if self.shall_train:
            self.train(1 << 18)
        return True

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
        loader.describe_dataset (scale, offset, up, render_aabb, frames ...)."""
        from . import loader
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
        return ds

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

    # -- pyngp's camera and render() (src/python_api.cu:124-221, 416-425, 481-566) ---------------
    def set_nerf_camera_matrix(self, camera_to_world):
        """Testbed::set_nerf_camera_matrix (src/testbed.cu:420-422): NeRF-convention 3x4 -> the view camera."""
        self._camera = nerf_matrix_to_ngp(camera_to_world)

    def set_ngp_camera_matrix(self, camera_3x4):
        self._camera = np.ascontiguousarray(np.array(camera_3x4, dtype=np.float32)[:3, :4])

    @property
    def fov(self):
        """Field of view in degrees along fov_axis (Testbed::fov / set_fov, src/testbed.cu:3526-3536)."""
        return float(np.degrees(2.0 * np.arctan(0.5 / self._relative_focal_length)))

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

    def render(self, width=1920, height=1080, spp=1, linear=True):
        """pyngp's Testbed.render: float32 [h][w][4] on the host - `spp` samples per pixel averaged, the
        background blended in, exposure, tonemap curve, sRGB transfer unless `linear`. The view is the
        one set by set_nerf_camera_matrix / fov / screen_center (focal length = relative focal length
        x resolution[fov_axis], Testbed::calc_focal_length)."""
        if getattr(self, "_camera", None) is None:
            raise NgpbError("render: set the view with set_nerf_camera_matrix first")
        focal = float(getattr(self, "_relative_focal_length", 1.0)) * (height if self.fov_axis == 1 else width)
        cx, cy = self.screen_center
        rgba = torch.empty((height, width, 4), dtype=torch.float32, device=self._device)
        cam = np.ascontiguousarray(self._camera, dtype=np.float32)
        self._set("render_shard_rank", 0)
        self._set("render_shard_world", 1)
        torch.cuda.current_stream().synchronize()
        check(lib.ngpb_testbed_render(self._h, width, height, spp, int(bool(linear)), cam.ctypes.data_as(ctypes.POINTER(ctypes.c_float)),
                                      focal, focal, cx, cy, ctypes.c_void_p(rgba.data_ptr())))
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
