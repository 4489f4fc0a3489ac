"""NeRF dataset description loader: transforms.json -> what Testbed needs (host logic only).

Restates load_nerf of the reference (src/nerf_loader.cu:273-740: frame ordering and culling,
`read_focal_length` :243-271, `read_lens` :175-241, scale / offset / aabb handling :448-506,
nerf_matrix_to_ngp nerf_loader.h:101-120, the byte-image fix-ups of convert_rgba32 :41-63) for the
lens models the NeRF path implements (Perspective and OpenCV). Image files are decoded with OpenCV
(the reference uses stb_image); everything numeric is the reference's arithmetic in float32.
"""
import functools
import json
import os
import re

import numpy as np

NERF_SCALE = 0.33
SUPPORTED_IMAGE_FORMATS = ("png", "jpg", "jpeg", "bmp", "gif", "tga", "pic", "pnm", "psd", "exr")  # nerf_loader.cu:315-317
MASK_COLOR = 0x00FF00FF  # "hot pink", alpha 0: pixels the trainer skips (nerf_loader.cu:610)


def _natural_number(a, l, le, fa, b, r, re_, fb):
    """compare_number (natural_sort.hpp:85-150), its bool-from-int returns included."""
    if fa and not fb:
        return 1
    if not fa and fb:
        return 0
    if fa:
        while l < le and r < re_:
            if a[l] != b[r]:
                return -1 if a[l] < b[r] else 1
            l, r = l + 1, r + 1
        while l < le and a[l] == 0x30:
            l += 1
        while r < re_ and b[r] == 0x30:
            r += 1
        return -1 if (l == le and r != re_) else 1 if (l != le and r == re_) else 0
    while l < le and a[l] == 0x30:
        l += 1
    while r < re_ and b[r] == 0x30:
        r += 1
    if le - l != re_ - r:
        return -1 if le - l < re_ - r else 1
    while l < le:
        if a[l] != b[r]:
            return -1 if a[l] < b[r] else 1
        l, r = l + 1, r + 1
    return 0


def natural_less(sa, sb):
    """SI::natural::compare (dependencies/NaturalSort/natural_sort.hpp:170-236), the order of the reference's frames
    (nerf_loader.cu:347-349), on the UTF-8 bytes: case-folded unsigned byte order, digit runs by value, runs after a '.' as
    fractions, repeated spaces collapsed - quirks included (csrc/host_dataset.cu natural_less is the same restatement; both are
    pinned to the reference's header by tests/golden/ref_natural_sort.json.gz)."""
    a, b = sa.encode("utf-8"), sb.encode("utf-8")
    na, nb = len(a), len(b)

    def at(s, i):
        return s[i] if i < len(s) else 0

    def fold(c):
        return c + 32 if 0x41 <= c <= 0x5A else c

    def digit(c):
        return 0x30 <= c <= 0x39
    i = j = 0
    space_a = space_b = False
    while i < na and j < nb:
        while space_a and i < na and a[i] == 0x20:
            i += 1
        space_a = at(a, i) == 0x20
        while space_b and j < nb and b[j] == 0x20:
            j += 1
        space_b = at(b, j) == 0x20
        ca, cb = at(a, i), at(b, j)
        if not digit(ca) or not digit(cb):
            if fold(ca) != fold(cb):
                return fold(ca) < fold(cb)
            i, j = i + 1, j + 1
        else:
            ei, ej = i, j
            while ei < na and digit(a[ei]):
                ei += 1
            while ej < nb and digit(b[ej]):
                ej += 1
            order = _natural_number(a, i, ei, i > 0 and a[i - 1] == 0x2E, b, j, ej, j > 0 and b[j - 1] == 0x2E)
            if order != 0:
                return order < 0
            i, j = ei, ej
    if i >= na and j >= nb:
        return False
    return i >= na


def _natural_cmp(x, y):
    return -1 if natural_less(x, y) else 1 if natural_less(y, x) else 0


def _strip_json_comments(text):
    # nlohmann::json::parse(..., ignore_comments = true)
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return re.sub(r"(^|[^:\"])//[^\n]*", r"\1", text)


def fov_to_focal_length(resolution, degrees):
    """common_host.h fov_to_focal_length."""
    return np.float32(0.5) * np.float32(resolution) / np.float32(np.tan(np.float32(0.5) * np.float32(degrees) * np.float32(np.pi) / np.float32(180)))


def read_focal_length(j, res):
    """nerf_loader.cu:243-271. Returns (fx, fy) or None. x_fov is in degrees, camera_angle_x in radians."""
    def one(resolution, axis):
        if axis + "_fov" in j:
            return float(fov_to_focal_length(resolution, float(j[axis + "_fov"])))
        if "fl_" + axis in j:
            return float(np.float32(j["fl_" + axis]))
        if "camera_angle_" + axis in j:
            return float(fov_to_focal_length(resolution, np.float32(j["camera_angle_" + axis]) * np.float32(180) / np.float32(np.pi)))
        return 0.0
    x_fl, y_fl = one(res[0], "x"), one(res[1], "y")
    if x_fl != 0:
        return (x_fl, y_fl if y_fl != 0 else x_fl)
    if y_fl != 0:
        return (y_fl, y_fl)
    return None


def read_lens(j, lens):
    """nerf_loader.cu:175-241 for the parameters the NeRF path supports. `lens` is updated in place:
    {"mode": 0 Perspective | 1 OpenCV, "params": [k1, k2, p1, p2, 0, 0, 0], "principal": [cx, cy]}."""
    if j.get("is_fisheye", False) or "ftheta_p0" in j or "latlong" in j or "equirectangular" in j:
        raise ValueError("ngp-b200: lens mode must be Perspective or OpenCV")
    if "rolling_shutter" in j and any(float(v) != 0.0 for v in j["rolling_shutter"]):
        raise ValueError("ngp-b200: rolling shutter is outside the implemented path")
    for name, idx in (("k1", 0), ("k2", 1), ("k3", 2), ("k4", 3), ("p1", 2), ("p2", 3)):
        if name in j:
            lens["params"][idx] = float(np.float32(j[name]))
            if lens["params"][idx] != 0.0:
                lens["mode"] = 1
    if "cx" in j:
        lens["principal"][0] = float(np.float32(j["cx"]) / np.float32(j["w"]))
    if "cy" in j:
        lens["principal"][1] = float(np.float32(j["cy"]) / np.float32(j["h"]))


def nerf_matrix_to_ngp(m, scale, offset, from_mitsuba=False):
    """nerf_loader.h:101-120: flip the y / z columns, scale + offset the origin, then either cycle the
    axes xyz <- yzx or, for Mitsuba data (:108-110), negate columns 0 and 2 instead."""
    r = np.array(m, dtype=np.float32)[:3, :4].copy()
    r[:, 1] *= -1
    r[:, 2] *= -1
    r[:, 3] = r[:, 3] * np.float32(scale) + np.asarray(offset, dtype=np.float32)
    if from_mitsuba:
        r[:, 0] *= -1
        r[:, 2] *= -1
        return np.ascontiguousarray(r)
    return np.ascontiguousarray(r[[1, 2, 0], :])


def _resolve_path(base, local):
    path = local if os.path.isabs(local) else os.path.join(base, local)
    if os.path.splitext(path)[1] == "" and not os.path.exists(path):
        for ext in SUPPORTED_IMAGE_FORMATS:
            if os.path.exists(path + "." + ext):
                return path + "." + ext
    return path


def _json_paths(path):
    """Testbed::load_training_data / load_nerf (src/testbed_nerf.cu:2089-2140): a json file, or a
    directory searched for *.json."""
    if os.path.isdir(path):
        found = sorted(os.path.join(path, f) for f in os.listdir(path) if f.lower().endswith(".json"))
        if not found:
            raise ValueError("No training images were found for NeRF training!")
        return found
    return [path]


def describe_dataset(path_or_paths):
    """The dataset record of load_nerf without the pixels: a dict with n_images, aabb_scale, scale,
    offset, up, render_aabb, is_hdr, white/black_transparent, fix_premult and `frames`, each frame
    {path, alpha_path, mask_path, xform (NGP 3x4), focal (None until the resolution is known, see
    finish_frame), lens_mode, lens_params[7], principal, json, frame_json}."""
    paths = [path_or_paths] if isinstance(path_or_paths, str) else list(path_or_paths)
    jsonpaths = [p for q in paths for p in _json_paths(q)]
    if not jsonpaths:
        raise ValueError("Cannot load NeRF data from an empty set of paths.")
    jsons = [json.loads(_strip_json_comments(open(p).read())) for p in jsonpaths]
    if isinstance(jsons[0].get("camera"), list):
        raise ValueError("hdf5 is no longer supported. please use the hdf52nerf.py conversion script")

    out = {"scale": NERF_SCALE, "offset": [0.5, 0.5, 0.5], "aabb_scale": 1, "up": [0.0, 1.0, 0.0], "render_aabb": None, "is_hdr": False,
           "from_mitsuba": False, "white_transparent": False, "black_transparent": False, "fix_premult": False,
           "n_extra_learnable_dims": 0, "frames": []}
    # pass 1 (:331-396): order, cull
    for j, jp in zip(jsons, jsonpaths):
        if not isinstance(j.get("frames"), list):
            continue
        base = os.path.dirname(jp)
        frames = sorted(j["frames"], key=functools.cmp_to_key(lambda f, g: _natural_cmp(f["file_path"], g["file_path"])))
        for f in frames:
            f["file_path"] = f["file_path"].replace("\\", "/")
        if "n_frames" in j:
            frames = frames[: min(len(frames), int(j["n_frames"]))]
        if frames and "sharpness" in frames[0]:
            thresh = float(j.get("sharpness_discard_threshold", 0.0))
            kept, hood = [], 3
            for i, f in enumerate(frames):
                lo, hi = max(0, i - hood), min(i + hood, len(frames) - 1)
                mean = np.float32(0)
                for k in range(lo, hi):
                    mean += np.float32(frames[k].get("sharpness", 1.0))
                mean = mean / np.float32(hi - lo) if hi > lo else np.float32(np.inf)
                if os.path.exists(_resolve_path(base, f["file_path"])) and float(f.get("sharpness", 1.0)) > thresh * float(mean):
                    kept.append(f)
            frames = kept
        j["frames"] = frames
    if sum(len(j.get("frames", [])) for j in jsons if isinstance(j.get("frames"), list)) == 0:
        raise ValueError("No training images were found for NeRF training!")

    # pass 2 (:422-718): global settings, then the frames
    for j, jp in zip(jsons, jsonpaths):
        base = os.path.dirname(jp)
        if "normal_mts_args" in j:
            out["from_mitsuba"] = True
        if "fix_premult" in j:
            out["fix_premult"] = bool(j["fix_premult"])
        if out["from_mitsuba"]:
            out["scale"] = 0.66
            out["offset"] = [float(np.float32(0.25) * np.float32(0.66))] * 3
        if "render_aabb" in j:
            out["render_aabb"] = [[float(v) for v in j["render_aabb"][0]], [float(v) for v in j["render_aabb"][1]]]
        for key in ("white_transparent", "black_transparent"):
            if key in j:
                out[key] = bool(j[key])
        if "scale" in j:
            out["scale"] = float(np.float32(j["scale"]))
        if "n_extra_learnable_dims" in j:
            out["n_extra_learnable_dims"] = int(j["n_extra_learnable_dims"])
        lens = {"mode": 0, "params": [0.0] * 7, "principal": [0.5, 0.5]}
        read_lens(j, lens)
        if "aabb_scale" in j:
            out["aabb_scale"] = int(j["aabb_scale"])
        if "offset" in j:
            o = j["offset"]
            out["offset"] = [float(np.float32(v)) for v in o] if isinstance(o, list) else [float(np.float32(o))] * 3
        if "aabb" in j:
            a = np.array(j["aabb"], dtype=np.float32)
            length = max(np.float32(0.000001), np.abs(a[1] - a[0]).max())
            out["scale"] = float(np.float32(1) / length)
            out["offset"] = [float(((a[1][k] + a[0][k]) * np.float32(0.5)) * -np.float32(out["scale"]) + np.float32(0.5)) for k in range(3)]
        if "up" in j:
            out["up"] = [float(j["up"][1]), float(j["up"][2]), float(j["up"][0])]  # axes permuted as the xforms
        if not isinstance(j.get("frames"), list):
            continue
        stem = os.path.splitext(os.path.basename(jp))[0]
        part_after_underscore = stem.rsplit("_", 1)[1] if "_" in stem else stem
        for i, f in enumerate(j["frames"]):
            if "transform_matrix_end" in f and f["transform_matrix_end"] != f.get("transform_matrix_start", f.get("transform_matrix")):
                raise ValueError("ngp-b200: per-frame camera motion (transform_matrix_end) is outside the implemented path")
            rel = f["file_path"] if f["file_path"] != "" else "%s_%03d/rgba.png" % (part_after_underscore, i)
            path = _resolve_path(base, rel)
            if not os.path.exists(path):
                raise FileNotFoundError("Could not find image file '%s'." % path)
            ext = os.path.splitext(path)[1].lstrip(".")
            alpha = _resolve_path(base, "%s.alpha.%s" % (f["file_path"], ext))
            mask = os.path.join(os.path.dirname(path), "dynamic_mask_%s.png" % os.path.splitext(os.path.basename(path))[0])
            frame_lens = {"mode": lens["mode"], "params": list(lens["params"]), "principal": list(lens["principal"])}
            read_lens(f, frame_lens)   # per-frame override (:703)
            m = f["transform_matrix_start"] if "transform_matrix_start" in f else f["transform_matrix"]
            out["frames"].append({
                "path": path, "alpha_path": alpha if os.path.exists(alpha) else None, "mask_path": mask if os.path.exists(mask) else None,
                "xform_nerf": np.array(m, dtype=np.float32)[:3, :4], "focal": None,
                "lens_mode": frame_lens["mode"], "lens_params": frame_lens["params"], "principal": frame_lens["principal"],
                "json": j, "frame_json": f,
            })
            if ext.lower() == "exr":
                out["is_hdr"] = True
    # the xforms use the FINAL scale / offset (the reference converts inside the loop with the values
    # of the file being read; a single transforms.json - the usual case - gives the same result)
    for fr in out["frames"]:
        fr["xform"] = nerf_matrix_to_ngp(fr["xform_nerf"], out["scale"], out["offset"], out["from_mitsuba"])
    out["n_images"] = len(out["frames"])
    return out


def finish_frame(frame, width, height):
    """Focal length once the resolution is known (:686-690): file-level keys, then per-frame keys."""
    fl = read_focal_length(frame["json"], (width, height))
    per_frame = read_focal_length(frame["frame_json"], (width, height))
    if per_frame is not None:
        fl = per_frame
    if fl is None:
        raise ValueError("Couldn't read fov.")
    frame["focal"] = fl
    return fl


def _srgb_to_linear(s):
    s = np.asarray(s, dtype=np.float32)
    return np.where(s <= np.float32(0.04045), s / np.float32(12.92), ((s + np.float32(0.055)) / np.float32(1.055)) ** np.float32(2.4))


def _imread_rgba8(path):
    import cv2
    img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    if img is None:
        raise ValueError("Could not open image file: " + path)
    if img.dtype == np.uint16:
        img = (img >> 8).astype(np.uint8)   # stb_image's 16 -> 8 bit conversion
    if img.ndim == 2:
        img = np.stack([img, img, img, np.full_like(img, 255)], axis=-1)
    elif img.shape[2] == 3:
        img = np.concatenate([img[:, :, ::-1], np.full(img.shape[:2] + (1,), 255, dtype=np.uint8)], axis=-1)
    else:
        img = img[:, :, [2, 1, 0, 3]]
    return np.ascontiguousarray(img, dtype=np.uint8)


def load_frame_pixels(frame, dataset):
    """Pixels of one frame as Testbed.set_nerf_image takes them: uint8 RGBA [h][w][4] (sRGB, straight
    alpha; image type Byte) or float16 RGBA (EXR, linear premultiplied; image type Half), after the
    alpha-file, dynamic-mask, white / black-transparent fix-ups of nerf_loader.cu:576-618 and :41-63."""
    path = frame["path"]
    if path.lower().endswith(".exr"):
        os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
        import cv2
        img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
        if img is None:
            raise ValueError("Could not load image file '%s'." % path)
        img = img.astype(np.float32)
        if img.ndim == 2:
            img = np.stack([img] * 3, axis=-1)
        rgb = img[:, :, 2::-1] if img.shape[2] >= 3 else img
        a = img[:, :, 3:4] if img.shape[2] == 4 else np.ones(img.shape[:2] + (1,), dtype=np.float32)
        if dataset.get("fix_premult", False):
            rgb = rgb * a   # load_exr_to_gpu(fix_premult): straight -> premultiplied
        return np.ascontiguousarray(np.concatenate([rgb, a], axis=-1).astype(np.float16))
    img = _imread_rgba8(path)
    if frame.get("alpha_path"):
        alpha = _imread_rgba8(frame["alpha_path"])
        if alpha.shape[:2] != img.shape[:2]:
            raise ValueError("Alpha image %s has wrong resolution." % frame["alpha_path"])
        img[:, :, 3] = (np.float32(255.0) * _srgb_to_linear(alpha[:, :, 0].astype(np.float32) * np.float32(1.0 / 255.0))).astype(np.uint8)
    if frame.get("mask_path"):
        mask = _imread_rgba8(frame["mask_path"])
        if mask.shape[:2] != img.shape[:2]:
            raise ValueError("Dynamic mask %s has wrong resolution." % frame["mask_path"])
        hit = (mask[:, :, :3] != 0).any(axis=-1)
        img[hit] = np.array([0xFF, 0x00, 0xFF, 0x00], dtype=np.uint8)
    if dataset.get("white_transparent", False):
        img[(img[:, :, :3] == 255).all(axis=-1), 3] = 0
    if dataset.get("black_transparent", False):
        img[(img[:, :, :3] == 0).all(axis=-1), 3] = 0
    return img


def load_image_file(path):
    """The image of the neural-image mode as Testbed::load_image makes it (src/testbed_image.cu:362-428):
    *.exr -> float32 RGBA as stored; *.bin -> int32 height, int32 width, then float16 RGBA texels;
    anything else -> 8-bit RGBA through from_rgba32 (src/common_host.cu:268-298): sRGB -> linear,
    colours multiplied by alpha. Returns [h][w][4] float32 or float16."""
    if not os.path.exists(path):
        raise ValueError("Image file '%s' does not exist." % path)
    ext = os.path.splitext(path)[1].lower()
    if ext == ".bin":
        with open(path, "rb") as f:
            h, w = np.frombuffer(f.read(8), dtype=np.int32)
            px = np.frombuffer(f.read(int(h) * int(w) * 8), dtype=np.float16)
        if px.size != int(h) * int(w) * 4:
            raise ValueError("Binary image '%s' is truncated." % path)
        return np.ascontiguousarray(px.reshape(int(h), int(w), 4))
    if ext == ".exr":
        os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
        import cv2
        img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
        if img is None:
            raise ValueError("Could not load image file '%s'." % path)
        img = img.astype(np.float32)
        if img.ndim == 2:
            img = np.stack([img] * 3, axis=-1)
        a = img[:, :, 3:4] if img.shape[2] == 4 else np.ones(img.shape[:2] + (1,), dtype=np.float32)
        return np.ascontiguousarray(np.concatenate([img[:, :, 2::-1], a], axis=-1))
    img = _imread_rgba8(path).astype(np.float32) * np.float32(1.0 / 255.0)
    if img.shape[0] == 0 or img.shape[1] == 0:
        raise ValueError("Image has zero pixels.")
    a = img[:, :, 3:4]
    return np.ascontiguousarray(np.concatenate([_srgb_to_linear(img[:, :, :3]) * a, a], axis=-1).astype(np.float32))
