"""Writes tests/golden/exr/*.exr with OpenCV's bundled OpenEXR (an encoder that is not ours) and records, per file, the
sha256 of the float32 RGBA pixels OpenCV's own OpenEXR reader returns for it (a one-channel file spread over R, G, B and A,
which is what the reference's reader - tinyexr's LoadEXR - does with it) and, where oracle/_ref/libref_images.so is built,
the sha256 of what the reference's own tinyexr returns (`sha256_tinyexr`; the two agree). tests/test_host_io_cpu.py decodes the files
with csrc/host_exr.cu and compares the digests, so the PIZ / RLE / ZIP paths stay pinned where cv2 has no EXR support.

    python tests/golden/make_exr_fixtures.py
"""
import hashlib
import json
import os

os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
import cv2  # noqa: E402
import numpy as np  # noqa: E402

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "exr")


def rgba_of(path):
    img = cv2.imread(path, cv2.IMREAD_UNCHANGED).astype(np.float32)
    if img.ndim == 2:
        img = np.stack([img] * 4, -1)   # LoadEXR: one channel fills all four
    a = img[:, :, 3:4] if img.shape[2] == 4 else np.ones(img.shape[:2] + (1,), dtype=np.float32)
    return np.ascontiguousarray(np.concatenate([img[:, :, 2::-1], a], axis=-1))


def tinyexr_rgba(path):
    lib_path = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "oracle", "_ref", "libref_images.so")
    if not os.path.exists(lib_path):
        return None
    import ctypes
    lib = ctypes.CDLL(lib_path)
    w, h, why = ctypes.c_int(), ctypes.c_int(), ctypes.create_string_buffer(200)
    assert lib.ref_tinyexr_load(path.encode(), ctypes.byref(w), ctypes.byref(h), None, ctypes.c_size_t(0), why, ctypes.c_size_t(200)) == 0, why.value
    out = np.zeros((h.value, w.value, 4), np.float32)
    lib.ref_tinyexr_load(path.encode(), ctypes.byref(w), ctypes.byref(h), out.ctypes.data_as(ctypes.c_void_p), ctypes.c_size_t(out.size), why, ctypes.c_size_t(200))
    return out


def main():
    rng = np.random.default_rng(20)
    yy, xx = np.mgrid[0:37, 0:53]
    smooth = np.stack([np.sin(xx * 0.11 + k) * np.cos(yy * 0.07 * (k + 1)) for k in range(3)], -1).astype(np.float32)
    cases = {
        # below 2^14 distinct 16-bit values: the exact 14-bit wavelet
        "piz_half_rgb": (smooth + 0.01 * rng.standard_normal(smooth.shape).astype(np.float32), cv2.IMWRITE_EXR_TYPE_HALF, cv2.IMWRITE_EXR_COMPRESSION_PIZ),
        # more than 2^14 distinct values and Huffman codes longer than 14 bits: the modulo-2^16 wavelet, the long-code search
        "piz_float_rgba": (rng.standard_normal((80, 80, 4)).astype(np.float32), cv2.IMWRITE_EXR_TYPE_FLOAT, cv2.IMWRITE_EXR_COMPRESSION_PIZ),
        # odd sizes below one wavelet level, and runs (the run-length symbol)
        "piz_half_y_flat": (np.where(np.mgrid[0:5, 0:131][1] > 60, 1.0, 0.25).astype(np.float32), cv2.IMWRITE_EXR_TYPE_HALF, cv2.IMWRITE_EXR_COMPRESSION_PIZ),
        "rle_half_rgba": (np.concatenate([smooth, np.ones((37, 53, 1), np.float32)], -1), cv2.IMWRITE_EXR_TYPE_HALF, cv2.IMWRITE_EXR_COMPRESSION_RLE),
        "zip_float_rgb": (smooth, cv2.IMWRITE_EXR_TYPE_FLOAT, cv2.IMWRITE_EXR_COMPRESSION_ZIP),
    }
    os.makedirs(HERE, exist_ok=True)
    digests = {}
    for name, (img, pixel_type, compression) in cases.items():
        path = os.path.join(HERE, name + ".exr")
        assert cv2.imwrite(path, img, [cv2.IMWRITE_EXR_COMPRESSION, compression, cv2.IMWRITE_EXR_TYPE, pixel_type])
        rgba = rgba_of(path)
        digests[name] = {"height": int(rgba.shape[0]), "width": int(rgba.shape[1]), "sha256": hashlib.sha256(rgba.tobytes()).hexdigest(),
                         "bytes": os.path.getsize(path)}
        ref = tinyexr_rgba(path)
        if ref is not None:
            digests[name]["sha256_tinyexr"] = hashlib.sha256(ref.tobytes()).hexdigest()
    with open(os.path.join(HERE, "digests.json"), "w") as f:
        json.dump({"made_with": "opencv-python " + cv2.__version__, "files": digests}, f, indent=1, sort_keys=True)
    print(json.dumps(digests, indent=1))


if __name__ == "__main__":
    main()
