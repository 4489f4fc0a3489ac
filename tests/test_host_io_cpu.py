"""Host C++ of the file-level entry points (csrc/host_io.cu, msgpack.h, host_png.cu), through the C ABI, without a GPU:
config files with "parent" inheritance against the Python mirror, the msgpack reader / writer against the msgpack
package (what snapshot.py uses, and nlohmann's encoding choices), the PNG decoder against OpenCV.
"""
import ctypes
import json
import os
import struct
import zlib

import numpy as np
import pytest


def _config_text(lib, path):
    n = lib.ngpb_host_load_network_config(str(path).encode(), None, 0)
    assert n > 0, lib.ngpb_last_error().decode()
    buf = ctypes.create_string_buffer(n)
    lib.ngpb_host_load_network_config(str(path).encode(), buf, n)
    return json.loads(buf.value.decode())


def _reencode(lib, blob):
    n = lib.ngpb_host_msgpack_reencode(blob, len(blob), None, 0)
    assert n > 0, lib.ngpb_last_error().decode()
    out = ctypes.create_string_buffer(n)
    assert lib.ngpb_host_msgpack_reencode(blob, len(blob), out, n) == n
    return out.raw


def test_config_file_with_parent_and_comments(ngp, tmp_path):
    lib = ngp._lib.lib
    (tmp_path / "base.json").write_text("""{
        // the reference's configs carry comments (configs/image/base.json:6-9)
        "loss": {"otype": "L2"},
        "optimizer": {"otype": "Ema", "decay": 0.95, "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9}},
        "encoding": {"otype": "HashGrid", "n_levels": 16, "log2_hashmap_size": 19},
        "network": {"otype": "FullyFusedMLP", "n_neurons": 64, "n_hidden_layers": 1}
    }""")
    os.makedirs(tmp_path / "sub")
    (tmp_path / "sub" / "big.json").write_text('{"parent": "../base.json", "encoding": {"log2_hashmap_size": 21}, /* drop */ "loss": null}')
    (tmp_path / "sub" / "bigger.json").write_text('{"parent": "big.json", "encoding": {"n_levels": 8}, "network": {"n_hidden_layers": 2}}')
    got = _config_text(lib, tmp_path / "sub" / "bigger.json")
    from importlib import import_module
    want = import_module("instant-ngp-rendering_b200.testbed").load_network_config(str(tmp_path / "sub" / "bigger.json"))
    assert got == want
    assert got["encoding"] == {"otype": "HashGrid", "n_levels": 8, "log2_hashmap_size": 21}
    assert "loss" not in got and "parent" not in got and got["network"]["n_hidden_layers"] == 2
    # missing file: the reference's text
    assert lib.ngpb_host_load_network_config(str(tmp_path / "nope.json").encode(), None, 0) == 0
    assert "does not exist" in lib.ngpb_last_error().decode()


def test_msgpack_round_trip_against_the_msgpack_package(ngp):
    msgpack = pytest.importorskip("msgpack")
    lib = ngp._lib.lib
    rng = np.random.default_rng(1)
    doc = {
        "ints": [0, 1, 127, 128, 255, 256, 65535, 65536, 2 ** 32 - 1, 2 ** 32, 2 ** 53, -1, -32, -33, -128, -129, -32768, -32769, -2 ** 31, -2 ** 31 - 1],
        "floats": [0.5, 1.0 / 3.0, 1e-8, -2.25, 3.4e38, 1e300, float(np.float32(0.95)), 0.1],
        "strings": ["", "a", "x" * 31, "y" * 32, "z" * 255, "w" * 256, "v" * 70000],
        "nested": {"k%d" % i: {"list": list(range(i)), "flag": bool(i % 2), "none": None} for i in range(20)},
        "small_bin": b"\x00\x01\x02", "bin16": rng.bytes(300), "bin32": rng.bytes(70000),
        "long_list": list(range(70000)),
    }
    blob = msgpack.packb(doc, use_bin_type=True, use_single_float=False)
    back = msgpack.unpackb(_reencode(lib, blob), raw=False, strict_map_key=False)
    assert back == doc
    # nlohmann's choices: float32 when exact, the smallest integer type, keys in map order (std::map: sorted)
    assert _reencode(lib, msgpack.packb({"b": 0.5, "a": 0.1})) == b"\x82\xa1a\xcb" + struct.pack(">d", 0.1) + b"\xa1b\xca" + struct.pack(">f", 0.5)
    assert _reencode(lib, msgpack.packb([300, -200])) == b"\x92\xcd\x01\x2c\xd1\xff\x38"
    # truncated input and ext types fail loudly
    assert lib.ngpb_host_msgpack_reencode(blob[:-5], len(blob) - 5, None, 0) == 0
    assert "unexpected end" in lib.ngpb_last_error().decode()
    ext = msgpack.packb(msgpack.ExtType(5, b"abcd"))
    assert lib.ngpb_host_msgpack_reencode(ext, len(ext), None, 0) == 0


def test_snapshot_container_of_the_python_mirror_reads_here(ngp, tmp_path):
    """A *.ingp written by snapshot.py's container (zlib + msgpack) parses in C++ to the same tree; gzip framing too."""
    msgpack = pytest.importorskip("msgpack")
    lib = ngp._lib.lib
    snap = importlib_snapshot()
    cfg = {"encoding": {"otype": "HashGrid", "n_levels": 8}, "snapshot": {"version": 1, "n_params": 3, "params_binary": b"\x01\x02\x03\x04\x05\x06",
                                                                          "nerf": {"rgb": {"rays_per_batch": 4096}}, "loss": 0.125}}
    for name, compress in (("a.ingp", True), ("b.ingp", False), ("c.msgpack", True)):
        snap.write_container(str(tmp_path / name), cfg, compress)
        got = _config_text(lib, tmp_path / name)
        assert got["snapshot"]["params_binary"] == {"bytes": [1, 2, 3, 4, 5, 6], "subtype": None}
        assert got["snapshot"]["nerf"]["rgb"]["rays_per_batch"] == 4096 and got["snapshot"]["loss"] == 0.125
        assert got["encoding"] == cfg["encoding"]
    import gzip
    with open(tmp_path / "g.ingp", "wb") as f:
        f.write(gzip.compress(msgpack.packb(cfg, use_bin_type=True)))
    assert _config_text(lib, tmp_path / "g.ingp")["snapshot"]["n_params"] == 3
    # a *.ingp without a zlib header passes through (zstr)
    with open(tmp_path / "raw.ingp", "wb") as f:
        f.write(msgpack.packb(cfg, use_bin_type=True))
    assert _config_text(lib, tmp_path / "raw.ingp")["snapshot"]["version"] == 1


def importlib_snapshot():
    from importlib import import_module
    return import_module("instant-ngp-rendering_b200.snapshot")


@pytest.mark.parametrize("kind", ["rgb8", "rgba8", "grey8", "grey_alpha8", "rgb16", "palette", "grey1"])
def test_png_decoder_matches_opencv(ngp, tmp_path, kind):
    cv2 = pytest.importorskip("cv2")
    lib = ngp._lib.lib
    rng = np.random.default_rng(len(kind))
    h, w = 37, 53
    path = str(tmp_path / (kind + ".png"))
    if kind == "rgb8":
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8); cv2.imwrite(path, img)
        want = np.dstack([img[:, :, ::-1], np.full((h, w), 255, np.uint8)])
    elif kind == "rgba8":
        img = rng.integers(0, 256, (h, w, 4), dtype=np.uint8); cv2.imwrite(path, img)
        want = img[:, :, [2, 1, 0, 3]]
    elif kind == "grey8":
        img = rng.integers(0, 256, (h, w), dtype=np.uint8); cv2.imwrite(path, img)
        want = np.dstack([img, img, img, np.full((h, w), 255, np.uint8)])
    elif kind == "rgb16":
        img = rng.integers(0, 65536, (h, w, 3), dtype=np.uint16); cv2.imwrite(path, img)
        hi = (img >> 8).astype(np.uint8)
        want = np.dstack([hi[:, :, ::-1], np.full((h, w), 255, np.uint8)])
    else:
        want = _write_raw_png(path, kind, h, w, rng)
    cw, ch = ctypes.c_int(), ctypes.c_int()
    out = np.zeros((h, w, 4), dtype=np.uint8)
    rc = lib.ngpb_host_decode_png(path.encode(), out.ctypes.data_as(ctypes.c_void_p), out.nbytes, ctypes.byref(cw), ctypes.byref(ch))
    assert rc == 0, lib.ngpb_last_error().decode()
    assert (cw.value, ch.value) == (w, h)
    assert np.array_equal(out, want)


def _write_raw_png(path, kind, h, w, rng):
    """PNG kinds OpenCV does not write: built by hand (filter type 0..4 cycled over the rows)."""
    import struct

    def chunk(t, d):
        return struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d) & 0xFFFFFFFF)
    extra = b""
    if kind == "grey_alpha8":
        px = rng.integers(0, 256, (h, w, 2), dtype=np.uint8)
        rows, depth, color, bpp = px.reshape(h, -1), 8, 4, 2
        want = np.dstack([px[:, :, 0]] * 3 + [px[:, :, 1]])
    elif kind == "palette":
        pal = rng.integers(0, 256, (16, 3), dtype=np.uint8)
        alpha = rng.integers(0, 256, 10, dtype=np.uint8)
        idx = rng.integers(0, 16, (h, w), dtype=np.uint8)
        packed = np.zeros((h, (w + 1) // 2), dtype=np.uint8)
        for x in range(w):
            packed[:, x // 2] |= idx[:, x] << (4 if x % 2 == 0 else 0)
        rows, depth, color, bpp = packed, 4, 3, 1
        extra = chunk(b"PLTE", pal.tobytes()) + chunk(b"tRNS", alpha.tobytes())
        a = np.where(idx < 10, alpha[np.minimum(idx, 9)], 255).astype(np.uint8)
        want = np.dstack([pal[idx][:, :, 0], pal[idx][:, :, 1], pal[idx][:, :, 2], a])
    else:  # grey1
        bits = rng.integers(0, 2, (h, w), dtype=np.uint8)
        packed = np.packbits(bits, axis=1)
        rows, depth, color, bpp = packed, 1, 0, 1
        want = np.dstack([bits * 255] * 3 + [np.full((h, w), 255, np.uint8)])
    raw = bytearray()
    prev = np.zeros(rows.shape[1], dtype=np.int32)
    for y in range(h):
        cur = rows[y].astype(np.int32)
        ft = y % 5
        left = np.concatenate([np.zeros(bpp, np.int32), cur[:-bpp]])
        upleft = np.concatenate([np.zeros(bpp, np.int32), prev[:-bpp]])
        if ft == 0: pred = 0
        elif ft == 1: pred = left
        elif ft == 2: pred = prev
        elif ft == 3: pred = (left + prev) // 2
        else:
            p = left + prev - upleft
            pa, pb, pc = abs(p - left), abs(p - prev), abs(p - upleft)
            pred = np.where((pa <= pb) & (pa <= pc), left, np.where(pb <= pc, prev, upleft))
        raw += bytes([ft]) + ((cur - pred) & 0xFF).astype(np.uint8).tobytes()
        prev = cur
    data = zlib.compress(bytes(raw))
    ihdr = struct.pack(">IIBBBBB", w, h, depth, color, 0, 0, 0)
    with open(path, "wb") as f:   # two IDAT chunks: the decoder must concatenate them
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", ihdr) + extra + chunk(b"IDAT", data[:len(data) // 2]) + chunk(b"IDAT", data[len(data) // 2:]) + chunk(b"IEND", b""))
    return want.astype(np.uint8)


def _decode_exr(lib, path):
    cw, ch = ctypes.c_int(), ctypes.c_int()
    assert lib.ngpb_host_decode_exr(str(path).encode(), None, 0, ctypes.byref(cw), ctypes.byref(ch)) == 0, lib.ngpb_last_error().decode()
    out = np.zeros((ch.value, cw.value, 4), dtype=np.float32)
    assert lib.ngpb_host_decode_exr(str(path).encode(), out.ctypes.data_as(ctypes.c_void_p), out.size, ctypes.byref(cw), ctypes.byref(ch)) == 0
    return out


def _cv2_exr(path):
    os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
    import cv2
    img = cv2.imread(str(path), cv2.IMREAD_UNCHANGED)
    if img is None:
        pytest.skip("this OpenCV build cannot read EXR")
    img = img.astype(np.float32)
    if img.ndim == 2:
        img = np.stack([img] * 4, -1)   # the reference's reader (tinyexr LoadEXR) spreads a one-channel file over R, G, B and A
    a = img[:, :, 3:4] if img.shape[2] == 4 else np.ones(img.shape[:2] + (1,), dtype=np.float32)
    return np.concatenate([img[:, :, 2::-1], a], axis=-1)


def test_exr_decoder_on_the_reference_image(ngp):
    """data/image/albert.exr (FLOAT RGBA, ZIP, decreasing line order): the same floats OpenCV's OpenEXR reads."""
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "data_stage", "albert.exr")
    if not os.path.exists(path):
        pytest.skip("data_stage/albert.exr is not staged")
    got = _decode_exr(ngp._lib.lib, path)
    want = _cv2_exr(path)
    assert got.shape == want.shape == (1024, 1024, 4)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("kind", ["half_rgb_zip", "float_rgba_zips", "half_y_none"])
def test_exr_decoder_on_hand_written_files(ngp, tmp_path, kind):
    """HALF / FLOAT channels, NONE / ZIPS / ZIP, missing alpha, luminance only: files written here chunk by chunk."""
    rng = np.random.default_rng(len(kind))
    h, w = 37, 29
    if kind == "half_rgb_zip":
        names, dtype, comp, lines = ["B", "G", "R"], np.float16, 3, 16
    elif kind == "float_rgba_zips":
        names, dtype, comp, lines = ["A", "B", "G", "R"], np.float32, 2, 1
    else:
        names, dtype, comp, lines = ["Y"], np.float16, 0, 1
    data = {n: rng.standard_normal((h, w)).astype(dtype) for n in names}
    path = tmp_path / (kind + ".exr")
    chlist = b"".join(n.encode() + b"\0" + struct.pack("<iBBBBii", 1 if dtype == np.float16 else 2, 0, 0, 0, 0, 1, 1) for n in names) + b"\0"

    def attr(name, typ, val):
        return name.encode() + b"\0" + typ.encode() + b"\0" + struct.pack("<I", len(val)) + val
    header = struct.pack("<II", 20000630, 2) + attr("channels", "chlist", chlist) + attr("compression", "compression", bytes([comp])) + \
        attr("dataWindow", "box2i", struct.pack("<4i", 0, 0, w - 1, h - 1)) + attr("displayWindow", "box2i", struct.pack("<4i", 0, 0, w - 1, h - 1)) + \
        attr("lineOrder", "lineOrder", b"\0") + attr("pixelAspectRatio", "float", struct.pack("<f", 1.0)) + \
        attr("screenWindowCenter", "v2f", struct.pack("<2f", 0, 0)) + attr("screenWindowWidth", "float", struct.pack("<f", 1.0)) + b"\0"
    chunks = []
    for y0 in range(0, h, lines):
        raw = b"".join(data[n][y].tobytes() for y in range(y0, min(y0 + lines, h)) for n in names)
        if comp == 0:
            packed = raw
        else:
            b = np.frombuffer(raw, dtype=np.uint8)
            inter = np.concatenate([b[0::2], b[1::2]]).astype(np.int32)
            pred = inter.copy()
            pred[1:] = (inter[1:] - inter[:-1] + 128 + 256) & 0xFF
            packed = zlib.compress(pred.astype(np.uint8).tobytes())
            if len(packed) >= len(raw):
                packed = raw
        chunks.append(struct.pack("<iI", y0, len(packed)) + packed)
    table_at = len(header)
    offs, pos = [], table_at + 8 * len(chunks)
    for c in chunks:
        offs.append(pos)
        pos += len(c)
    path.write_bytes(header + struct.pack("<%dQ" % len(offs), *offs) + b"".join(chunks))
    got = _decode_exr(ngp._lib.lib, path)
    want = np.zeros((h, w, 4), dtype=np.float32)
    want[..., 3] = data["A"].astype(np.float32) if "A" in data else 1.0
    if "Y" in data:
        want[...] = data["Y"].astype(np.float32)[..., None]   # image mode: LoadEXR fills all four with a lone channel
        # the NeRF-frame rule (load_exr_to_gpu) takes channels by name only: nothing named R / G / B / A -> colour 0, alpha 1
        frame = np.zeros_like(want)
        cw, ch = ctypes.c_int(), ctypes.c_int()
        assert ngp._lib.lib.ngpb_host_decode_exr_frame(str(path).encode(), frame.ctypes.data_as(ctypes.c_void_p), frame.size, ctypes.byref(cw), ctypes.byref(ch)) == 0
        assert not frame[..., :3].any() and (frame[..., 3] == 1.0).all()
    else:
        for k, n in enumerate("RGB"):
            want[..., k] = data[n].astype(np.float32)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    # unsupported compression is refused, not misread
    bad = header.replace(attr("compression", "compression", bytes([comp])), attr("compression", "compression", bytes([5])))
    (tmp_path / "pxr24.exr").write_bytes(bad + b"\0" * 64)
    cw, ch = ctypes.c_int(), ctypes.c_int()
    assert ngp._lib.lib.ngpb_host_decode_exr(str(tmp_path / "pxr24.exr").encode(), None, 0, ctypes.byref(cw), ctypes.byref(ch)) != 0
    assert "not supported" in ngp._lib.lib.ngpb_last_error().decode()


def test_exr_decoder_on_openexr_written_fixtures(ngp):
    """tests/golden/exr/*.exr were written by OpenCV's bundled OpenEXR (make_exr_fixtures.py): PIZ with the 14-bit and the
    modulo-2^16 wavelet, long Huffman codes, the run-length symbol, RLE, ZIP. The decoded float RGBA must hash to what
    OpenEXR's own reader returned for the same file."""
    import hashlib
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "exr")
    with open(os.path.join(here, "digests.json")) as f:
        files = json.load(f)["files"]
    assert len(files) >= 5
    for name, want in files.items():
        got = _decode_exr(ngp._lib.lib, os.path.join(here, name + ".exr"))
        assert got.shape == (want["height"], want["width"], 4), name
        assert hashlib.sha256(got.tobytes()).hexdigest() == want["sha256"] == want["sha256_tinyexr"], name   # OpenEXR's reader and the reference's tinyexr


@pytest.mark.parametrize("compression", ["NO", "RLE", "ZIPS", "ZIP", "PIZ"])
def test_exr_decoder_against_opencv_round_trips(ngp, tmp_path, compression):
    """Every compression tinyexr reads (the reference's loader, src/tinyexr_wrapper.cu:153-260), HALF and FLOAT, 1 / 3 / 4
    channels, sizes around the chunk heights and below one wavelet level: written and read back by OpenCV's OpenEXR, bit for bit."""
    os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(7)
    flag = getattr(cv2, "IMWRITE_EXR_COMPRESSION_" + compression)
    for (h, w, c) in [(37, 53, 3), (1, 1, 3), (33, 1, 4), (2, 77, 1), (70, 90, 4)]:
        yy, xx = np.mgrid[0:h, 0:w]
        for kind in ("noise", "smooth", "flat"):
            if kind == "noise":
                img = rng.standard_normal((h, w, c)).astype(np.float32)
            elif kind == "smooth":
                img = np.stack([np.sin(xx * 0.05 + k) + np.cos(yy * 0.03 * k) for k in range(c)], -1).astype(np.float32)
            else:
                img = np.full((h, w, c), 0.25, np.float32)
                img[h // 2:, :, 0] = 1.0
            if c == 1:
                img = img[..., 0]
            for pixel_type in (cv2.IMWRITE_EXR_TYPE_HALF, cv2.IMWRITE_EXR_TYPE_FLOAT):
                path = tmp_path / "t.exr"
                try:
                    ok = cv2.imwrite(str(path), img, [cv2.IMWRITE_EXR_COMPRESSION, flag, cv2.IMWRITE_EXR_TYPE, pixel_type])
                except cv2.error:
                    ok = False
                if not ok:
                    pytest.skip("this OpenCV build cannot write EXR")
                want = _cv2_exr(path)
                got = _decode_exr(ngp._lib.lib, path)
                assert got.shape == want.shape, (h, w, c, kind)
                assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (h, w, c, kind, pixel_type)


def test_exr_decoder_refuses_corrupt_piz_chunks(ngp, tmp_path):
    """Truncated or bit-flipped PIZ data ends in an error (or other pixels), never in a crash or an out-of-bounds write."""
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "exr")
    blob = open(os.path.join(here, "piz_half_rgb.exr"), "rb").read()
    rng = np.random.default_rng(3)
    cw, ch = ctypes.c_int(), ctypes.c_int()
    out = np.zeros((37, 53, 4), dtype=np.float32)
    failures = 0
    for trial in range(40):
        bad = bytearray(blob)
        if trial % 2:
            bad = bad[:int(rng.integers(400, len(bad)))]
        else:
            for k in rng.integers(700, len(bad), size=4):
                bad[int(k)] ^= 1 << int(rng.integers(0, 8))
        (tmp_path / "bad.exr").write_bytes(bytes(bad))
        failures += ngp._lib.lib.ngpb_host_decode_exr(str(tmp_path / "bad.exr").encode(), out.ctypes.data_as(ctypes.c_void_p), out.size, ctypes.byref(cw), ctypes.byref(ch)) != 0
    assert failures >= 20


def _ref_json_vectors():
    import gzip
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_json_vectors.json.gz")
    with gzip.open(path, "rb") as f:
        return json.loads(f.read().decode())


def test_msgpack_writer_against_the_references_json_library(ngp):
    """tests/golden/ref_json_vectors.json.gz holds json::to_msgpack of the reference's own nlohmann json (oracle/ref_json.cc):
    reading those bytes and writing them again must reproduce them exactly - integer widths, float32-when-exact, size classes of
    str / bin / array / map, key order - which is what makes a snapshot written here byte-compatible with one written there."""
    msgpack = pytest.importorskip("msgpack")
    lib = ngp._lib.lib
    vec = _ref_json_vectors()["to_msgpack"]
    assert set(vec) >= {"integers", "floats", "strings", "containers", "binaries", "key_order", "snapshot_shaped"}
    for name, v in vec.items():
        want = bytes.fromhex(v["msgpack_hex"])
        assert _reencode(lib, want) == want, name
        # and the bytes mean what the JSON says (binaries are written as "bin:<hex>" in the JSON text)
        def expect(x):
            if isinstance(x, str) and x.startswith("bin:"):
                return bytes.fromhex(x[4:])
            if isinstance(x, list):
                return [expect(e) for e in x]
            if isinstance(x, dict):
                return {k: expect(e) for k, e in x.items()}
            return x
        got = msgpack.unpackb(want, raw=False, strict_map_key=False)
        doc = expect(json.loads(v["json"]))
        assert got == doc, name   # float32-encoded values are exactly representable, so they decode to the same doubles


def test_merge_patch_against_the_references_json_library(ngp, tmp_path):
    """parent.merge_patch(child) of the reference's nlohmann json (merge_parent_network_config) = load_network_config here."""
    lib = ngp._lib.lib
    for name, v in _ref_json_vectors()["merge_patch"].items():
        (tmp_path / "parent.json").write_text(v["parent"])
        child = json.loads(__import__("re").sub(r"/\*.*?\*/|//[^\n]*", "", v["child"]))
        child["parent"] = "parent.json"
        (tmp_path / "child.json").write_text(json.dumps(child))
        assert _config_text(lib, tmp_path / "child.json") == json.loads(v["merged"]), name


def _decode_image(lib, path):
    cw, ch = ctypes.c_int(), ctypes.c_int()
    assert lib.ngpb_host_decode_image(str(path).encode(), None, 0, ctypes.byref(cw), ctypes.byref(ch)) == 0, lib.ngpb_last_error().decode()
    out = np.zeros((ch.value, cw.value, 4), dtype=np.uint8)
    assert lib.ngpb_host_decode_image(str(path).encode(), out.ctypes.data_as(ctypes.c_void_p), out.size, ctypes.byref(cw), ctypes.byref(ch)) == 0
    return out


@pytest.mark.parametrize("shape", [(13, 17), (1, 1), (8, 3), (30, 64)])
def test_stb_formats_against_pillow(ngp, tmp_path, shape):
    """BMP (24 / 32 bit, 8 / 4 / 1 bit palette), TGA (true colour, grey, grey + alpha, colour-mapped, run-length coded, both row
    orders), binary PGM / PPM and PNG through the content-sniffing decoder the loader uses for every 8-bit file but JPEG
    (the reference: stb_image behind load_stbi, src/common_host.cu:214-217). Files written by Pillow, texels bit for bit."""
    Image = pytest.importorskip("PIL.Image")
    lib = ngp._lib.lib
    h, w = shape
    rng = np.random.default_rng(h * 100 + w)
    rgba = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    rgba[..., 3] |= 1   # a 32-bit BMP whose alpha is zero everywhere reads as opaque (stb_image's rule), tested below
    rgb, grey = np.ascontiguousarray(rgba[..., :3]), np.ascontiguousarray(rgba[..., 0])
    opaque = np.concatenate([rgb, np.full((h, w, 1), 255, np.uint8)], -1)
    grey4 = np.stack([grey] * 3 + [np.full((h, w), 255, np.uint8)], -1)
    pal = Image.fromarray(rgb).quantize(16)
    bw = Image.fromarray(grey).convert("1")
    cases = [
        ("bmp", Image.fromarray(rgb), {}, opaque), ("bmp", Image.fromarray(rgba), {}, rgba), ("bmp", Image.fromarray(grey), {}, grey4),
        ("bmp", pal, {}, np.array(pal.convert("RGBA"))), ("bmp", bw, {}, np.array(bw.convert("RGBA"))),
        ("tga", Image.fromarray(rgb), {}, opaque), ("tga", Image.fromarray(rgba), {"compression": "tga_rle"}, rgba),
        ("tga", Image.fromarray(grey), {"compression": "tga_rle"}, grey4), ("tga", pal, {}, np.array(pal.convert("RGBA"))),
        ("tga", Image.fromarray(rgb), {"orientation": 1}, opaque),
        ("tga", Image.fromarray(np.ascontiguousarray(rgba[..., [0, 3]]), "LA"), {}, np.concatenate([grey4[..., :3], rgba[..., 3:]], -1)),
        ("ppm", Image.fromarray(rgb), {}, opaque), ("pgm", Image.fromarray(grey), {}, grey4), ("png", Image.fromarray(rgba), {}, rgba),
    ]
    for k, (ext, im, kw, want) in enumerate(cases):
        # the name does not decide the decoder: every file is also read under a wrong extension
        for name in ("c%d.%s" % (k, ext), "c%d.dat" % k):
            path = tmp_path / name
            im.save(str(path), format={"bmp": "BMP", "tga": "TGA", "ppm": "PPM", "pgm": "PPM", "png": "PNG"}[ext], **kw)
            got = _decode_image(lib, path)
            assert got.shape == want.shape and np.array_equal(got, want), (k, ext, kw)


def test_bmp_bit_fields_and_alpha_rules(ngp, tmp_path):
    """Hand-written BMPs Pillow does not write: 16-bit 5-6-5 bit fields, 16-bit BI_RGB (5-5-5), top-down rows, a 32-bit file whose
    alpha bytes are all zero (opaque by stb_image's rule), and the refusals (run-length coding, JPEG handed to the host decoder)."""
    lib = ngp._lib.lib
    rng = np.random.default_rng(4)
    h, w = 5, 7

    def widen(v, bits):
        v = v.astype(np.uint32) << (8 - bits)
        return (v | (v >> bits) | (v >> (2 * bits) if bits < 4 else 0)).astype(np.uint8)

    def bmp(bpp, rows, compression=0, masks=b"", height=None):
        stride = (w * bpp + 31) // 32 * 4
        data = b"".join(r.tobytes() + b"\0" * (stride - len(r.tobytes())) for r in rows)
        info = struct.pack("<IiiHHIIiiII", 40, w, h if height is None else height, 1, bpp, compression, len(data), 2835, 2835, 0, 0) + masks
        return b"BM" + struct.pack("<IHHI", 14 + len(info) + len(data), 0, 0, 14 + len(info)) + info + data

    v565 = rng.integers(0, 1 << 16, (h, w), dtype=np.uint16)
    (tmp_path / "565.bmp").write_bytes(bmp(16, v565[::-1], 3, struct.pack("<III", 0xF800, 0x07E0, 0x001F)))
    got = _decode_image(lib, tmp_path / "565.bmp")
    assert np.array_equal(got[..., 0], widen(v565 >> 11, 5)) and np.array_equal(got[..., 1], widen((v565 >> 5) & 63, 6))
    assert np.array_equal(got[..., 2], widen(v565 & 31, 5)) and (got[..., 3] == 255).all()

    (tmp_path / "555.bmp").write_bytes(bmp(16, v565, 0, height=-h))   # top-down
    got = _decode_image(lib, tmp_path / "555.bmp")
    assert np.array_equal(got[..., 0], widen((v565 >> 10) & 31, 5)) and np.array_equal(got[..., 2], widen(v565 & 31, 5))

    bgr0 = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    bgr0[..., 3] = 0
    (tmp_path / "x888.bmp").write_bytes(bmp(32, bgr0[::-1]))
    got = _decode_image(lib, tmp_path / "x888.bmp")
    assert np.array_equal(got[..., :3], bgr0[..., 2::-1]) and (got[..., 3] == 255).all()

    cw, ch = ctypes.c_int(), ctypes.c_int()
    (tmp_path / "rle.bmp").write_bytes(bmp(8, np.zeros((h, w), np.uint8), 1))
    assert lib.ngpb_host_decode_image(str(tmp_path / "rle.bmp").encode(), None, 0, ctypes.byref(cw), ctypes.byref(ch)) != 0
    assert "run-length" in lib.ngpb_last_error().decode()
    (tmp_path / "a.jpg").write_bytes(b"\xff\xd8\xff\xe0" + b"\0" * 64)
    assert lib.ngpb_host_decode_image(str(tmp_path / "a.jpg").encode(), None, 0, ctypes.byref(cw), ctypes.byref(ch)) != 0
    assert "a.jpg" in lib.ngpb_last_error().decode()   # a header without a picture: the host JPEG decoder refuses it
    (tmp_path / "noise.bin").write_bytes(bytes(rng.integers(128, 256, 64, dtype=np.uint8)))
    assert lib.ngpb_host_decode_image(str(tmp_path / "noise.bin").encode(), None, 0, ctypes.byref(cw), ctypes.byref(ch)) != 0
    assert "unknown image type" in lib.ngpb_last_error().decode()


def test_parsers_refuse_pathological_nesting(ngp, tmp_path):
    """A config or snapshot made of 100 000 opening brackets ends in an error message, not in a stack overflow; ordinary depth passes."""
    lib = ngp._lib.lib
    (tmp_path / "deep.json").write_text("[" * 100000)
    assert lib.ngpb_host_load_network_config(str(tmp_path / "deep.json").encode(), None, 0) == 0
    assert "nesting" in lib.ngpb_last_error().decode()
    blob = bytes([0x91]) * 100000
    assert lib.ngpb_host_msgpack_reencode(blob, len(blob), None, 0) == 0
    assert "nesting" in lib.ngpb_last_error().decode()
    ok = {"a": 1}
    for _ in range(100):
        ok = {"n": [ok]}
    (tmp_path / "ok.json").write_text(json.dumps(ok))
    assert lib.ngpb_host_load_network_config(str(tmp_path / "ok.json").encode(), None, 0) > 0


def test_json_text_keeps_every_digit_of_a_double(ngp, tmp_path):
    """Json::dump prints the shortest text that reads back as the same double (nlohmann's rule): a learning rate keeps its digits,
    0.1 stays 0.1, a float widened to double shows as what it is, non-finite values become null."""
    lib = ngp._lib.lib
    vals = {"lr": 0.0012345678901234567, "tenth": 0.1, "third": 1.0 / 3.0, "widened": float(np.float32(1.3819)), "big": 1.5e300, "tiny": 4.9e-324, "int": 7, "neg": -2.5}
    (tmp_path / "v.json").write_text(json.dumps(vals))
    n = lib.ngpb_host_load_network_config(str(tmp_path / "v.json").encode(), None, 0)
    buf = ctypes.create_string_buffer(n)
    lib.ngpb_host_load_network_config(str(tmp_path / "v.json").encode(), buf, n)
    text = buf.value.decode()
    assert json.loads(text) == vals
    assert '"tenth":0.1,' in text and '"widened":1.3818999528884888' in text and '"int":7' in text


def _ref_images():
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libref_images.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libref_images.so (the reference's stb_image and tinyexr, oracle/build_ref.sh) is not built")
    return ctypes.CDLL(path)


def _ref_stbi(ref, path):
    w, h, c, why = ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.create_string_buffer(200)
    if ref.ref_stbi_load(str(path).encode(), ctypes.byref(w), ctypes.byref(h), ctypes.byref(c), None, ctypes.c_size_t(0), why, ctypes.c_size_t(200)):
        return why.value.decode()
    out = np.zeros((h.value, w.value, 4), np.uint8)
    ref.ref_stbi_load(str(path).encode(), ctypes.byref(w), ctypes.byref(h), ctypes.byref(c), out.ctypes.data_as(ctypes.c_void_p), ctypes.c_size_t(out.size), why, ctypes.c_size_t(200))
    return out


def _ref_tinyexr(ref, path):
    w, h, why = ctypes.c_int(), ctypes.c_int(), ctypes.create_string_buffer(200)
    if ref.ref_tinyexr_load(str(path).encode(), ctypes.byref(w), ctypes.byref(h), None, ctypes.c_size_t(0), why, ctypes.c_size_t(200)):
        return why.value.decode()
    out = np.zeros((h.value, w.value, 4), np.float32)
    ref.ref_tinyexr_load(str(path).encode(), ctypes.byref(w), ctypes.byref(h), out.ctypes.data_as(ctypes.c_void_p), ctypes.c_size_t(out.size), why, ctypes.c_size_t(200))
    return out


def test_8bit_decoders_against_the_references_stb_image(ngp, tmp_path):
    """The reference decodes every non-EXR image with its vendored stb_image (stbi_load(..., 4), src/common_host.cu:214-217). That
    header, compiled as it is (oracle/ref_images.cc), and this library's host decoders must return the same RGBA8 texels for PNG
    (8 / 16 bit, grey, grey + alpha, palette, tRNS keys, 1-bit), BMP, TGA and binary PNM files."""
    ref = _ref_images()
    Image = pytest.importorskip("PIL.Image")
    cv2 = pytest.importorskip("cv2")
    lib = ngp._lib.lib
    rng = np.random.default_rng(21)
    h, w = 23, 31
    rgba = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    rgb, grey, la = np.ascontiguousarray(rgba[..., :3]), np.ascontiguousarray(rgba[..., 0]), np.ascontiguousarray(rgba[..., [0, 3]])
    pal = Image.fromarray(rgb).quantize(16)
    files = []

    def save(name, im, **kw):
        im.save(str(tmp_path / name), **kw)
        files.append(name)
    save("rgb.png", Image.fromarray(rgb)); save("rgba.png", Image.fromarray(rgba)); save("grey.png", Image.fromarray(grey)); save("la.png", Image.fromarray(la, "LA"))
    save("pal.png", pal); save("pal_trns.png", pal, transparency=3); save("bw.png", Image.fromarray(grey).convert("1"))
    save("rgb_trns.png", Image.fromarray(rgb), transparency=tuple(int(v) for v in rgb[0, 0]))
    save("grey_trns.png", Image.fromarray(grey), transparency=int(grey[0, 0]))
    for name, shape in (("rgb16.png", (h, w, 3)), ("grey16.png", (h, w)), ("rgba16.png", (h, w, 4))):
        assert cv2.imwrite(str(tmp_path / name), rng.integers(0, 65536, shape, dtype=np.uint16))
        files.append(name)
    save("rgb.bmp", Image.fromarray(rgb)); save("rgba.bmp", Image.fromarray(rgba)); save("pal.bmp", pal); save("bw.bmp", Image.fromarray(grey).convert("1")); save("grey.bmp", Image.fromarray(grey))
    save("rgb.tga", Image.fromarray(rgb)); save("rgba_rle.tga", Image.fromarray(rgba), compression="tga_rle"); save("grey_rle.tga", Image.fromarray(grey), compression="tga_rle")
    save("pal.tga", pal); save("la.tga", Image.fromarray(la, "LA")); save("top.tga", Image.fromarray(rgb), orientation=1)
    save("rgb.ppm", Image.fromarray(rgb)); save("grey.pgm", Image.fromarray(grey))
    # 16-bit PNM (stb_image keeps the low-order byte of each big-endian sample), comments and odd white space in the header
    (tmp_path / "rgb16.ppm").write_bytes(b"P6\n# c\n%d %d\n65535\n" % (w, h) + rng.integers(0, 65536, (h, w, 3), dtype=np.uint16).astype(">u2").tobytes())
    (tmp_path / "grey16.pgm").write_bytes(b"P5 %d\t%d #x\n 65535\n" % (w, h) + rng.integers(0, 65536, (h, w), dtype=np.uint16).astype(">u2").tobytes())
    (tmp_path / "max99.ppm").write_bytes(b"P6 #x\n%d\t%d #y\n 99\n" % (w, h) + rng.integers(0, 100, (h, w, 3), dtype=np.uint8).tobytes())
    files += ["rgb16.ppm", "grey16.pgm", "max99.ppm"]
    for name in files:
        want = _ref_stbi(ref, tmp_path / name)
        assert not isinstance(want, str), (name, want)
        got = _decode_image(lib, tmp_path / name)
        assert got.shape == want.shape and np.array_equal(got, want), name


def test_exr_reader_against_the_references_tinyexr(ngp, tmp_path):
    """tinyexr's LoadEXR as vendored by the reference (what load_exr reads the image mode's file with) against csrc/host_exr.cu on
    every scanline compression at HALF and FLOAT, on the committed fixtures, on a one-channel file and on the staged albert.exr."""
    ref = _ref_images()
    os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
    cv2 = pytest.importorskip("cv2")
    lib = ngp._lib.lib
    rng = np.random.default_rng(22)
    paths = sorted(str(p) for p in __import__("pathlib").Path(os.path.dirname(os.path.abspath(__file__)), "golden", "exr").glob("*.exr"))
    staged = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "data_stage", "albert.exr")
    if os.path.exists(staged):
        paths.append(staged)
    img = rng.standard_normal((37, 53, 4)).astype(np.float32)
    try:
        for comp in ("NO", "RLE", "ZIPS", "ZIP", "PIZ"):
            for tname, t in (("half", cv2.IMWRITE_EXR_TYPE_HALF), ("float", cv2.IMWRITE_EXR_TYPE_FLOAT)):
                for channels in (3, 4):
                    p = str(tmp_path / ("%s_%s_%d.exr" % (comp, tname, channels)))
                    if cv2.imwrite(p, img[..., :channels], [cv2.IMWRITE_EXR_COMPRESSION, getattr(cv2, "IMWRITE_EXR_COMPRESSION_" + comp), cv2.IMWRITE_EXR_TYPE, t]):
                        paths.append(p)
        p = str(tmp_path / "one_channel.exr")
        if cv2.imwrite(p, img[..., 0], [cv2.IMWRITE_EXR_TYPE, cv2.IMWRITE_EXR_TYPE_HALF]):
            paths.append(p)
    except cv2.error:
        pass
    assert len(paths) >= 5
    for p in paths:
        want = _ref_tinyexr(ref, p)
        assert not isinstance(want, str), (p, want)
        got = _decode_exr(lib, p)
        assert got.shape == want.shape and np.array_equal(got.view(np.uint32), want.view(np.uint32)), p


def test_bmp_and_tga_variants_against_the_references_stb_image(ngp, tmp_path):
    """Hand-built BMP and TGA files over the header space (Pillow writes only a corner of it) decoded by the reference's stb_image
    and by csrc/host_image.cu: BMP with 40 / 108 / 124-byte headers, 1 / 4 / 8-bit palettes, 16 / 24 / 32 bits, top-down rows,
    5-6-5 / 1-5-5-5 / 4-4-4-4 / 8-8-8-8 bit fields; TGA true colour 15 / 16 / 24 / 32 bits (5-bit channels scale as v * 255 / 31),
    grey 8 / 16, colour maps of 8 / 15 / 16 / 24 / 32-bit entries under 8- and 16-bit indices, run-length packets, both row
    orders, image-id bytes, and stb_image's reading of the colour map's first-entry field as a byte skip.
    Left out, because the reference itself does not decode them to defined texels: OS/2 12-byte-header palettes and indices
    beyond the palette (stb_image reads uninitialised entries), 56-byte headers with bit fields and channel masks wider than
    8 bits (stb_image refuses both; this library reads them)."""
    ref = _ref_images()
    lib = ngp._lib.lib
    rng = np.random.default_rng(33)

    def bmp(w, h, bpp, hsz=40, compression=0, masks=(0, 0, 0, 0), top_down=False):
        stride = (w * bpp + 31) // 32 * 4
        data = bytes(rng.integers(0, 256, stride * h, dtype=np.uint8))
        pal = bytes(rng.integers(0, 256, (1 << bpp) * 4, dtype=np.uint8)) if bpp <= 8 else b""
        info = struct.pack("<IiiHHIIiiII", 40, w, -h if top_down else h, 1, bpp, compression, len(data), 2835, 2835, 0, 0)
        if hsz == 40 and compression == 3:
            info += struct.pack("<III", *masks[:3])
        elif hsz > 40:
            info = struct.pack("<I", hsz) + info[4:] + struct.pack("<IIII", *masks)
            info += b"\0" * (hsz - len(info))
        off = 14 + len(info) + len(pal)
        return b"BM" + struct.pack("<IHHI", off + len(data), 0, 0, off) + info + pal + data

    def tga(w, h, typ, bits, cmap_bits=0, cmap_len=0, cmap_first=0, desc=0, id_len=0, rle=False):
        hdr = struct.pack("<BBBHHBHHHHBB", id_len, 1 if cmap_len else 0, typ | (8 if rle else 0), cmap_first, cmap_len, cmap_bits, 0, 0, w, h, bits, desc)
        ident = bytes(rng.integers(0, 256, id_len, dtype=np.uint8))
        pal = bytes(rng.integers(0, 256, cmap_first + cmap_len * ((cmap_bits + 7) // 8), dtype=np.uint8))
        tex, n = (bits + 7) // 8, w * h
        if typ == 1:
            px = [int(i).to_bytes(tex, "little") for i in rng.integers(0, cmap_len + 3, n)]   # a few indices past the map: entry 0
        else:
            px = [bytes(rng.integers(0, 256, tex, dtype=np.uint8)) for _ in range(n)]
        if not rle:
            return hdr + ident + pal + b"".join(px)
        data, i = b"", 0
        while i < n:
            c = min(int(rng.integers(1, 129)), n - i)
            if rng.random() < 0.5:
                data += bytes([0x80 | (c - 1)]) + px[i]
                px[i:i + c] = [px[i]] * c
            else:
                data += bytes([c - 1]) + b"".join(px[i:i + c])
            i += c
        return hdr + ident + pal + data

    cases = []
    for w, h in ((7, 5), (1, 1), (33, 3)):
        for bpp in (1, 4, 8, 16, 24, 32):
            for hsz in (40, 108, 124):
                for td in (False, True):
                    cases.append(("bmp %dx%d bpp%d hsz%d td%d" % (w, h, bpp, hsz, td), bmp(w, h, bpp, hsz, top_down=td)))
        for bpp, masks in ((16, (0xF800, 0x07E0, 0x001F, 0)), (16, (0x7C00, 0x03E0, 0x001F, 0x8000)), (16, (0x0F00, 0x00F0, 0x000F, 0xF000)),
                           (32, (0xFF0000, 0xFF00, 0xFF, 0xFF000000)), (32, (0xFF, 0xFF00, 0xFF0000, 0))):
            for hsz in (40, 108, 124):
                cases.append(("bmp bit fields bpp%d %s hsz%d" % (bpp, [hex(m) for m in masks], hsz), bmp(w, h, bpp, hsz, 3, masks)))
        for rle in (False, True):
            for desc in (0, 0x20, 0x10, 0x28):
                for bits in (15, 16, 24, 32):
                    cases.append(("tga rgb%d rle%d desc%#x" % (bits, rle, desc), tga(w, h, 2, bits, desc=desc, rle=rle, id_len=int(rng.integers(0, 5)))))
                for bits in (8, 16):
                    cases.append(("tga grey%d rle%d desc%#x" % (bits, rle, desc), tga(w, h, 3, bits, desc=desc, rle=rle)))
                for cb in (8, 15, 16, 24, 32):
                    for ib in (8, 16):
                        cases.append(("tga map%d index%d rle%d desc%#x" % (cb, ib, rle, desc), tga(w, h, 1, ib, cmap_bits=cb, cmap_len=19, desc=desc, rle=rle)))
                cases.append(("tga map first 5 rle%d" % rle, tga(w, h, 1, 8, cmap_bits=24, cmap_len=19, cmap_first=5, desc=desc, rle=rle)))
    assert len(cases) > 500
    path = tmp_path / "case.bin"
    for tag, blob in cases:
        path.write_bytes(blob)
        want = _ref_stbi(ref, path)
        assert not isinstance(want, str), (tag, want)
        got = _decode_image(lib, path)
        assert got.shape == want.shape and np.array_equal(got, want), tag


def test_interlaced_png_against_the_references_stb_image(ngp, tmp_path):
    """Adam7 files (neither Pillow nor OpenCV writes them; the encoder below does, pass by pass with per-line filters) in every
    colour type and bit depth, at sizes where passes are empty, against stb_image."""
    ref = _ref_images()
    lib = ngp._lib.lib
    rng = np.random.default_rng(44)
    x0, y0, dx, dy = (0, 4, 0, 2, 0, 1, 0), (0, 0, 4, 0, 2, 0, 1), (8, 8, 4, 4, 2, 2, 1), (8, 8, 8, 4, 4, 2, 2)

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    def pack_row(samples, depth):           # samples: [n] ints of `depth` bits -> packed bytes
        if depth == 8:
            return bytes(int(v) for v in samples)
        if depth == 16:
            return b"".join(struct.pack(">H", int(v)) for v in samples)
        bits = "".join(format(int(v), "0%db" % depth) for v in samples)
        bits += "0" * (-len(bits) % 8)
        return bytes(int(bits[i:i + 8], 2) for i in range(0, len(bits), 8))

    def encode(w, h, color, depth, extra=b""):
        channels = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[color]
        img = rng.integers(0, 1 << depth, (h, w, channels))
        bpp = max(1, channels * depth // 8)
        raw = bytearray()
        for p in range(7):
            sub = img[y0[p]::dy[p], x0[p]::dx[p]]
            if sub.shape[0] == 0 or sub.shape[1] == 0:
                continue
            prev = None
            for row in sub:
                line = pack_row(row.reshape(-1), depth)
                ft = int(rng.integers(0, 5))
                up = prev if prev is not None else bytes(len(line))
                out = bytearray(len(line))
                for i in range(len(line)):
                    a = line[i - bpp] if i >= bpp else 0
                    b = up[i]
                    c = up[i - bpp] if i >= bpp else 0
                    pa, pb, pc = abs(b - c), abs(a - c), abs(a + b - 2 * c)
                    paeth = a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)
                    out[i] = (line[i] - (0, a, b, (a + b) // 2, paeth)[ft]) & 0xFF
                raw += bytes([ft]) + out
                prev = line
        ihdr = struct.pack(">IIBBBBB", w, h, depth, color, 0, 0, 1)
        return b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", ihdr) + extra + chunk(b"IDAT", zlib.compress(bytes(raw))) + chunk(b"IEND", b"")

    palette = chunk(b"PLTE", bytes(rng.integers(0, 256, 256 * 3, dtype=np.uint8))) + chunk(b"tRNS", bytes(rng.integers(0, 256, 40, dtype=np.uint8)))
    n = 0
    for w, h in ((1, 1), (2, 3), (5, 1), (9, 9), (19, 13), (8, 8)):
        for color, depths in ((0, (1, 2, 4, 8, 16)), (2, (8, 16)), (3, (1, 2, 4, 8)), (4, (8, 16)), (6, (8, 16))):
            for depth in depths:
                path = tmp_path / "a7.png"
                path.write_bytes(encode(w, h, color, depth, palette if color == 3 else b""))
                want = _ref_stbi(ref, path)
                assert not isinstance(want, str), (w, h, color, depth, want)
                got = _decode_image(lib, path)
                assert got.shape == want.shape and np.array_equal(got, want), (w, h, color, depth)
                n += 1
    assert n == 6 * 15


def test_gif_first_frame_against_the_references_stb_image(ngp, tmp_path):
    """GIF is on the reference loader's list of dataset formats (nerf_loader.cu:315-317) and stbi_load returns its first frame.
    Pillow-written files (2 / 16 / 256 colours, interlaced, transparent index, background index, animated) and hand-built ones
    (frame smaller than the screen with and without a background index, local colour tables, extensions, a Graphic Control
    Extension of odd length, rectangles out of bounds, too many / too few pixels, GIF87a, truncations) give stb_image's texels -
    its quirks included - or fail where it fails."""
    ref = _ref_images()
    Image = pytest.importorskip("PIL.Image")
    lib = ngp._lib.lib
    rng = np.random.default_rng(55)
    cw, ch = ctypes.c_int(), ctypes.c_int()

    def check(tag, path):
        want = _ref_stbi(ref, path)
        if isinstance(want, str):
            assert lib.ngpb_host_decode_image(str(path).encode(), None, 0, ctypes.byref(cw), ctypes.byref(ch)) != 0, (tag, want)
            return 0
        got = _decode_image(lib, path)
        assert got.shape == want.shape and np.array_equal(got, want), tag
        return 1
    n_ok = 0
    p = tmp_path / "g.gif"
    for (h, w) in ((13, 17), (1, 1), (64, 50), (200, 3)):
        noise = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        smooth = np.stack([np.add.outer(np.arange(h) * 3, np.arange(w) * k) % 256 for k in (1, 2, 5)], -1).astype(np.uint8)
        for arr in (noise, smooth):
            for colors in (2, 16, 256):
                im = Image.fromarray(arr).quantize(colors)
                for kw in ({}, {"interlace": True}, {"interlace": False}, {"transparency": 0}, {"transparency": 1, "interlace": True}, {"background": 1}, {"optimize": True}):
                    im.save(str(p), **kw)
                    n_ok += check((h, w, colors, kw), p)
        frames = [Image.fromarray(rng.integers(0, 256, (h, w, 3), dtype=np.uint8)).quantize(8) for _ in range(3)]
        frames[0].save(str(p), save_all=True, append_images=frames[1:], duration=40, loop=0)
        n_ok += check((h, w, "animated"), p)
    assert n_ok == 4 * (2 * 3 * 7 + 1)

    def lzw_plain(indices, cs):
        # uncompressed-GIF trick: a clear code often enough that the code size never grows
        clear, end = 1 << cs, (1 << cs) + 1
        size = cs + 1
        bits, nbits, out = 0, 0, bytearray()
        def put(c):
            nonlocal bits, nbits
            bits |= c << nbits; nbits += size
            while nbits >= 8: out.append(bits & 255); bits >>= 8; nbits -= 8
        n = 0
        put(clear)
        for v in indices:
            put(int(v)); n += 1
            if n == (1 << cs) - 2: put(clear); n = 0
        put(end)
        if nbits: out.append(bits & 255)
        data = bytes([cs])
        for i in range(0, len(out), 255): data += bytes([len(out[i:i+255])]) + bytes(out[i:i+255])
        return data + b'\0'
    def gif(gw, gh, bg, gpal, frames, exts=b''):
        flags = (0x80 | (int(np.log2(len(gpal)//3)) - 1)) if gpal else 0
        b = b'GIF89a' + struct.pack('<HHBBB', gw, gh, flags, bg, 0) + (gpal or b'') + exts
        for (x,y,w,h,lpal,interlace,idx,cs,gce) in frames:
            b += gce
            lf = (0x80 | (int(np.log2(len(lpal)//3)) - 1)) if lpal else 0
            if interlace: lf |= 0x40
            b += b'\x2C' + struct.pack('<HHHHB', x,y,w,h,lf) + (lpal or b'') + lzw_plain(idx, cs)
        return b + b'\x3B'
    pal16 = bytes(rng.integers(0,256,48,dtype=np.uint8)); pal4 = bytes(rng.integers(0,256,12,dtype=np.uint8)); pal256=bytes(rng.integers(0,256,768,dtype=np.uint8))
    gce = lambda fl, tr: b'\x21\xF9\x04' + bytes([fl]) + b'\x05\x00' + bytes([tr]) + b'\0'
    cases = {
     'sub-rect bg3': gif(11,9,3,pal16,[(2,1,5,4,None,False,rng.integers(0,16,20),4,b'')]),
     'sub-rect bg0': gif(11,9,0,pal16,[(2,1,5,4,None,False,rng.integers(0,16,20),4,b'')]),
     'sub-rect bg3 transparent': gif(11,9,3,pal16,[(2,1,5,4,None,False,rng.integers(0,16,20),4,gce(1,3))]),
     'local palette': gif(8,8,1,pal16,[(0,0,8,8,pal4,False,rng.integers(0,4,64),2,b'')]),
     'local palette transparent interlaced': gif(8,9,1,pal16,[(0,0,8,9,pal4,True,rng.integers(0,4,72),2,gce(1,2))]),
     'no global table': gif(8,8,0,None,[(0,0,8,8,pal4,False,rng.integers(0,4,64),2,b'')]),
     'no table at all': gif(8,8,0,None,[(0,0,8,8,None,False,rng.integers(0,4,64),2,b'')]),
     'comment + app ext': gif(8,8,0,pal16,[(0,0,8,8,None,False,rng.integers(0,16,64),4,b'')], exts=b'\x21\xFE\x05hello\0' + b'\x21\xFF\x0BNETSCAPE2.0\x03\x01\x00\x00\0'),
     'gce odd length': gif(8,8,0,pal16,[(0,0,8,8,None,False,rng.integers(0,16,64),4,b'\x21\xF9\x02\x01\x02\0')]),
     'rect out of bounds': gif(8,8,0,pal16,[(4,4,8,8,None,False,rng.integers(0,16,64),4,b'')]),
     'zero width': gif(8,8,2,pal16,[(0,0,0,8,None,False,[],4,b'')]),
     'too many pixels': gif(4,4,0,pal16,[(0,0,4,4,None,False,rng.integers(0,16,40),4,b'')]),
     'too few pixels': gif(6,6,5,pal16,[(0,0,6,6,None,False,rng.integers(0,16,20),4,b'')]),
     '256 colours': gif(20,10,200,pal256,[(3,2,10,5,None,True,rng.integers(0,256,50),8,gce(1,7))]),
     'only trailer': b'GIF89a' + struct.pack('<HHBBB', 4,4,0x80|3,0,0) + pal16 + b'\x3B',
     'unknown tag': b'GIF89a' + struct.pack('<HHBBB', 4,4,0x80|3,0,0) + pal16 + b'\x55',
     'gif87': b'GIF87a' + gif(5,5,0,pal4,[(0,0,5,5,None,False,rng.integers(0,4,25),2,b'')])[6:],
    }
    n_ok = 0
    for tag, blob in cases.items():
        p.write_bytes(blob)
        n_ok += check(tag, p)
    base = cases["256 colours"]
    for cut in (20, 100, 790, 800, 810, len(base) - 3):
        p.write_bytes(base[:cut])
        n_ok += check("cut %d" % cut, p)
    assert n_ok >= 14


def test_psd_composite_against_the_references_stb_image(ngp, tmp_path):
    """PSD is on the reference loader's list too; stb_image reads the flattened RGB composite. Hand-built files - 1 to 16 channels,
    8 / 16 bits, raw and PackBits (with no-op bytes), header blocks to skip, every (colour, alpha) byte pair through the
    white-matte removal, refused variants, truncations - give its texels or fail where it fails."""
    ref = _ref_images()
    lib = ngp._lib.lib
    rng = np.random.default_rng(66)
    cw, ch = ctypes.c_int(), ctypes.c_int()
    path = tmp_path / "p.psd"

    def check(tag, blob):
        path.write_bytes(blob)
        want = _ref_stbi(ref, path)
        if isinstance(want, str):
            assert lib.ngpb_host_decode_image(str(path).encode(), None, 0, ctypes.byref(cw), ctypes.byref(ch)) != 0, (tag, want)
            return 0
        got = _decode_image(lib, path)
        assert got.shape == want.shape and np.array_equal(got, want), tag
        return 1

    def packbits(row):
        out=bytearray(); i=0; n=len(row)
        while i<n:
            j=i
            while j+1<n and row[j+1]==row[i] and j-i<127: j+=1
            if j>i: out+=bytes([257-(j-i+1)]) + bytes([row[i]]); i=j+1
            else:
                j=i
                while j+1<n and (j+2>=n or row[j+1]!=row[j+2]) and j-i<127: j+=1
                out+=bytes([j-i]) + bytes(row[i:j+1]); i=j+1
            if rng.random()<0.1: out+=b'\x80'
        return bytes(out)
    def psd(w,h,channels,depth,compression,mode=3,version=1,extra=(0,0,0),planes=None):
        b = b'8BPS' + struct.pack('>H6xHIIHH', version, channels, h, w, depth, mode)
        for n in extra: b += struct.pack('>I', n) + bytes(rng.integers(0,256,n,dtype=np.uint8))
        b += struct.pack('>H', compression)
        if planes is None:
            planes = [rng.integers(0,1<<depth,(h,w)) for _ in range(channels)]
            for p in planes[:3]: p[:h//2] = p[0,0]    # runs
        if compression == 0:
            for p in planes: b += p.astype('>u1' if depth==8 else '>u2').tobytes()
        else:
            rows = [[packbits(list(p[y].astype(np.uint8))) for y in range(h)] for p in planes]
            for r in rows:
                for x in r: b += struct.pack('>H', len(x))
            for r in rows:
                for x in r: b += x
        return b
    n_ok = 0
    for w, h in ((7, 5), (1, 1), (33, 4)):
        for channels in (1, 2, 3, 4, 5, 16):
            for depth in (8, 16):
                for comp in (0, 1):
                    n_ok += check((w, h, channels, depth, comp), psd(w, h, channels, depth, comp, extra=tuple(int(v) for v in rng.integers(0, 9, 3))))
    col, al = np.meshgrid(np.arange(256), np.arange(256))
    n_ok += check("matte sweep", psd(256, 256, 4, 8, 0, planes=[col, (col * 7) % 256, 255 - col, al]))
    assert n_ok == 3 * 6 * 2 * 2 + 1
    plain = psd(4, 4, 3, 8, 0)
    for tag, blob in (("mode 1", psd(4, 4, 1, 8, 0, mode=1)), ("version 2", psd(4, 4, 3, 8, 0, version=2)), ("17 channels", psd(4, 4, 17, 8, 0)),
                      ("depth 32", plain[:22] + b"\0\x20" + plain[24:]), ("compression 2", plain[:38] + b"\0\x02" + plain[40:])):
        assert check(tag, blob) == 0, tag
    base = psd(9, 7, 4, 8, 1)
    for cut in (10, 30, 60, len(base) - 20, len(base) - 1):
        check("cut %d" % cut, base[:cut])
    base = psd(9, 7, 3, 16, 0)
    for cut in (45, 100, len(base) - 1):
        check("raw cut %d" % cut, base[:cut])


def test_softimage_pic_against_the_references_stb_image(ngp, tmp_path):
    """The last dataset format on the reference loader's list (nerf_loader.cu:315-317): Softimage PIC with raw, pure-RLE and
    mixed-RLE channel packets (16-bit run counts included), channels split over several packets, alpha present or not; bad
    packets, overruns and truncations fail where stb_image fails."""
    ref = _ref_images()
    lib = ngp._lib.lib
    rng = np.random.default_rng(77)
    cw, ch = ctypes.c_int(), ctypes.c_int()
    path = tmp_path / "p.pic"

    def check(tag, blob):
        path.write_bytes(blob)
        want = _ref_stbi(ref, path)
        if isinstance(want, str):
            assert lib.ngpb_host_decode_image(str(path).encode(), None, 0, ctypes.byref(cw), ctypes.byref(ch)) != 0, (tag, want)
            return 0
        got = _decode_image(lib, path)
        assert got.shape == want.shape and np.array_equal(got, want), tag
        return 1

    def pic(w, h, packets, tail=b''):
        # packets: list of (type, channel mask); channel bit 0x80 R, 0x40 G, 0x20 B, 0x10 A
        b = b'\x53\x80\xF6\x34' + struct.pack('>f', 0.0) + bytes(80) + b'PICT' + struct.pack('>HHfHH', w, h, 1.0, 3, 0)
        for k,(typ,chan) in enumerate(packets): b += bytes([1 if k+1 < len(packets) else 0, 8, typ, chan])
        for y in range(h):
            for typ,chan in packets:
                n = bin(chan & 0xF0).count('1')
                if typ == 0: b += bytes(rng.integers(0,256,w*n,dtype=np.uint8))
                elif typ == 1:
                    left = w
                    while left > 0:
                        c = int(rng.integers(1, min(left,255)+1)); b += bytes([c]) + bytes(rng.integers(0,256,n,dtype=np.uint8)); left -= c
                else:
                    left = w
                    while left > 0:
                        r = rng.random()
                        if r < 0.4: c = int(rng.integers(1, min(left,128)+1)); b += bytes([c-1]) + bytes(rng.integers(0,256,c*n,dtype=np.uint8))
                        elif r < 0.8 and left >= 2: c = int(rng.integers(2, min(left,128)+1)); b += bytes([c+127]) + bytes(rng.integers(0,256,n,dtype=np.uint8))
                        else: c = int(rng.integers(1, left+1)); b += bytes([128]) + struct.pack('>H', c) + bytes(rng.integers(0,256,n,dtype=np.uint8))
                        left -= c
        return b + tail
    n_ok = 0
    for w, h in ((7, 5), (1, 1), (300, 3)):
        for typ in (0, 1, 2):
            for packets in ([(typ, 0xE0)], [(typ, 0xF0)], [(typ, 0xE0), (2 - typ, 0x10)], [(typ, 0x80), (typ, 0x40), (0, 0x20), (1, 0x10)], [(typ, 0x80)]):
                n_ok += check((w, h, packets), pic(w, h, packets))
    assert n_ok == 45
    plain = pic(4, 4, [(0, 0xE0)])
    for tag, blob in (("bad type", pic(4, 4, [(3, 0xE0)])), ("11 packets", pic(4, 4, [(0, 0x80)] * 11)), ("16 bpp", plain[:105] + b"\x10" + plain[106:]),
                      ("overrun", pic(4, 1, [(2, 0xE0)])[:108] + bytes([127 + 6, 1, 2, 3]))):
        assert check(tag, blob) == 0, tag
    base = pic(9, 7, [(2, 0xE0), (1, 0x10)])
    for cut in (95, 96, 100, 107, 112, 150, len(base) - 1):
        assert check("cut %d" % cut, base[:cut]) == 0


def test_jpeg_decoder_gives_the_references_texels(ngp, tmp_path):
    """csrc/host_jpeg.cu against the reference's stb_image, byte for byte: all 50 JPEGs of the reference's fox scene when staged,
    and written here - baseline / optimised / progressive, 4:4:4 / 4:2:2 / 4:2:0 / 4:1:1 / 4:4:0, restart intervals, grey, CMYK,
    qualities 1 to 100, custom quantisation tables, sizes below one MCU. (JPEG fixes the entropy coding only; IDCT, upsampling
    and colour conversion are the decoder's, and these are stb_image's.)"""
    ref = _ref_images()
    Image = pytest.importorskip("PIL.Image")
    cv2 = pytest.importorskip("cv2")
    lib = ngp._lib.lib
    rng = np.random.default_rng(88)

    def check(tag, path):
        want = _ref_stbi(ref, path)
        assert not isinstance(want, str), (tag, want)
        got = _decode_image(lib, path)
        assert got.shape == want.shape and np.array_equal(got, want), tag
    fox = sorted(__import__("pathlib").Path(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "data_stage", "fox", "images").glob("*.jpg"))
    for p in fox:
        check(p.name, p)
    yy, xx = np.mgrid[0:67, 0:93]
    smooth = np.stack([(np.sin(xx * 0.2) + np.cos(yy * 0.13)) * 90 + 128, xx * 2.5, (yy * 3.7 + xx) % 256], -1).clip(0, 255).astype(np.uint8)
    noisy = (smooth.astype(int) + rng.integers(-40, 40, smooth.shape)).clip(0, 255).astype(np.uint8)
    p = tmp_path / "j.jpg"
    n = 0
    for arr in (smooth, noisy):
        for sz in ((67, 93), (8, 8), (1, 1), (17, 33), (5, 100)):
            a = np.ascontiguousarray(arr[:sz[0], :sz[1]])
            for sub in (0, 1, 2):
                for q in (20, 90, 100):
                    for kw in ({}, {"optimize": True}, {"progressive": True}):
                        Image.fromarray(a).save(str(p), quality=q, subsampling=sub, **kw)
                        check((sz, sub, q, kw), p)
                        n += 1
            Image.fromarray(a[..., 0]).save(str(p), quality=80)
            check((sz, "grey"), p)
            Image.fromarray(a[..., 0]).save(str(p), quality=80, progressive=True)
            check((sz, "grey progressive"), p)
            Image.fromarray(a).convert("CMYK").save(str(p), quality=80)
            check((sz, "cmyk"), p)
    bgr = np.ascontiguousarray(noisy[..., ::-1])
    for factor in ("411", "420", "422", "440", "444"):
        for rst in (0, 1, 7):
            for prog in (0, 1):
                for sz in ((67, 93), (9, 31)):
                    assert cv2.imwrite(str(p), bgr[:sz[0], :sz[1]], [cv2.IMWRITE_JPEG_SAMPLING_FACTOR, getattr(cv2, "IMWRITE_JPEG_SAMPLING_FACTOR_" + factor),
                                                                      cv2.IMWRITE_JPEG_RST_INTERVAL, rst, cv2.IMWRITE_JPEG_PROGRESSIVE, prog, cv2.IMWRITE_JPEG_QUALITY, 60])
                    check((factor, rst, prog, sz), p)
                    n += 1
    for q in (1, 3):
        Image.fromarray(noisy).save(str(p), quality=q)
        check(("quality", q), p)
    Image.fromarray(noisy).save(str(p), qtables=[[255] * 64, [200] * 64])
    check("qtables", p)
    assert n == 2 * 5 * 27 + 60


def test_jpeg_decoder_refuses_broken_streams(ngp, tmp_path):
    """Cut or bit-flipped files end in an error or a picture of the right size, never in a crash."""
    Image = pytest.importorskip("PIL.Image")
    lib = ngp._lib.lib
    rng = np.random.default_rng(89)
    img = rng.integers(0, 256, (40, 56, 3), dtype=np.uint8)
    cw, ch = ctypes.c_int(), ctypes.c_int()
    out = np.zeros((40, 56, 4), np.uint8)
    for kw in ({}, {"progressive": True}):
        Image.fromarray(img).save(str(tmp_path / "b.jpg"), quality=85, **kw)
        blob = (tmp_path / "b.jpg").read_bytes()
        for trial in range(60):
            bad = bytearray(blob)
            if trial % 2:
                bad = bad[:int(rng.integers(2, len(bad)))]
            else:
                for k in rng.integers(2, len(bad), size=3):
                    bad[int(k)] ^= 1 << int(rng.integers(0, 8))
            (tmp_path / "c.jpg").write_bytes(bytes(bad))
            rc = lib.ngpb_host_decode_image(str(tmp_path / "c.jpg").encode(), None, 0, ctypes.byref(cw), ctypes.byref(ch))
            if rc == 0 and cw.value * ch.value * 4 == out.size:
                assert lib.ngpb_host_decode_image(str(tmp_path / "c.jpg").encode(), out.ctypes.data_as(ctypes.c_void_p), out.size, ctypes.byref(cw), ctypes.byref(ch)) == 0


def test_tiled_exr_against_the_references_tinyexr(ngp, tmp_path):
    """Single-part tiled EXR files (full-resolution level), as tinyexr reads them: hand-written here - HALF / FLOAT, 1 / 3 / 4
    channels, NONE / ZIP / ZIPS, tiles larger than the image and clipped at its edges, a data window away from the origin, chunks
    stored in reverse order, the mip flag set - texel for texel against the reference's vendored tinyexr."""
    ref = _ref_images()
    lib = ngp._lib.lib
    rng = np.random.default_rng(99)

    def tiled_exr(w, h, tw, th, names, dtype, comp, mode=0, xmin=0, ymin=0, order=None):
        data = {n: rng.standard_normal((h, w)).astype(dtype) for n in names}
        def attr(name, typ, val): return name.encode() + b"\0" + typ.encode() + b"\0" + struct.pack("<I", len(val)) + val
        chlist = b"".join(n.encode() + b"\0" + struct.pack("<iBBBBii", 1 if dtype == np.float16 else 2, 0, 0, 0, 0, 1, 1) for n in names) + b"\0"
        box = struct.pack("<4i", xmin, ymin, xmin + w - 1, ymin + h - 1)
        header = struct.pack("<II", 20000630, 2 | 0x200) + attr("channels", "chlist", chlist) + attr("compression", "compression", bytes([comp])) + \
            attr("dataWindow", "box2i", box) + attr("displayWindow", "box2i", box) + attr("lineOrder", "lineOrder", b"\0") + attr("pixelAspectRatio", "float", struct.pack("<f", 1.0)) + \
            attr("screenWindowCenter", "v2f", struct.pack("<2f", 0, 0)) + attr("screenWindowWidth", "float", struct.pack("<f", 1.0)) + attr("tiles", "tiledesc", struct.pack("<IIB", tw, th, mode)) + b"\0"
        nx, ny = (w + tw - 1) // tw, (h + th - 1) // th
        chunks = []
        for ty in range(ny):
            for tx in range(nx):
                x0, y0 = tx * tw, ty * th
                raw = b"".join(data[n][y, x0:x0 + tw].tobytes() for y in range(y0, min(y0 + th, h)) for n in names)
                if comp == 0: packed = raw
                else:
                    b = np.frombuffer(raw, dtype=np.uint8); inter = np.concatenate([b[0::2], b[1::2]]).astype(np.int32); pred = inter.copy(); pred[1:] = (inter[1:] - inter[:-1] + 128 + 256) & 0xFF
                    packed = zlib.compress(pred.astype(np.uint8).tobytes())
                    if len(packed) >= len(raw): packed = raw
                chunks.append(struct.pack("<iiiiI", tx, ty, 0, 0, len(packed)) + packed)
        idx = list(range(len(chunks)))
        if order == 'reversed': file_order = idx[::-1]
        else: file_order = idx
        offs = [0] * len(chunks); pos = len(header) + 8 * len(chunks); body = b""
        for k in file_order: offs[k] = pos + len(body) - 0; body += chunks[k]
        offs = [o for o in offs]
        return header + struct.pack("<%dQ" % len(offs), *offs) + body, data
    path = tmp_path / "t.exr"

    def check(tag, blob):
        path.write_bytes(blob)
        want = _ref_tinyexr(ref, path)
        assert not isinstance(want, str), (tag, want)
        got = _decode_exr(lib, path)
        assert got.shape == want.shape and np.array_equal(got.view(np.uint32), want.view(np.uint32)), tag
    for (w, h, tw, th) in ((37, 29, 16, 16), (64, 64, 32, 32), (5, 3, 8, 8), (33, 17, 7, 5), (100, 9, 64, 4)):
        for names, dtype in ((["B", "G", "R"], np.float16), (["A", "B", "G", "R"], np.float32), (["Y"], np.float16)):
            for comp in (0, 3, 2):
                check((w, h, tw, th, names, comp), tiled_exr(w, h, tw, th, names, dtype, comp)[0])
    check("offset window", tiled_exr(37, 29, 16, 16, ["B", "G", "R"], np.float16, 3, xmin=-5, ymin=7)[0])
    check("reversed file order", tiled_exr(37, 29, 16, 16, ["B", "G", "R"], np.float16, 3, order="reversed")[0])
    check("mip mode flag", tiled_exr(32, 32, 16, 16, ["B", "G", "R"], np.float16, 0, mode=1)[0])
    # a tile of a lower level is refused, as tinyexr refuses it
    blob = bytearray(tiled_exr(32, 32, 16, 16, ["B", "G", "R"], np.float16, 0)[0])
    first = struct.unpack_from("<Q", blob, blob.index(b"tiledesc") + 8 + 1 + 4 + 9 + 1)[0]
    struct.pack_into("<i", blob, first + 8, 1)
    path.write_bytes(bytes(blob))
    cw, ch = ctypes.c_int(), ctypes.c_int()
    assert lib.ngpb_host_decode_exr(str(path).encode(), None, 0, ctypes.byref(cw), ctypes.byref(ch)) != 0
    assert isinstance(_ref_tinyexr(ref, path), str)


def test_radiance_hdr_against_the_references_stb_image(ngp, tmp_path):
    """Radiance RGBE files both ways stb_image hands them out: stbi_loadf floats (the image mode keeps them, load_stbi_gpu,
    src/common_host.cu:268-299) and stbi_load bytes (gamma 1 / 2.2). OpenCV-written files, run-length coded and flat, and
    hand-built ones: the #?RGBE magic, flat data under a width of 8 or more, a later line that is not run-length coded (stb_image
    restarts from the first pixel there), refused headers, bad runs, truncation."""
    ref = _ref_images()
    cv2 = pytest.importorskip("cv2")
    lib = ngp._lib.lib
    rng = np.random.default_rng(111)
    cw, ch = ctypes.c_int(), ctypes.c_int()
    path = tmp_path / "h.hdr"

    def ref_float(p):
        w, h, why = ctypes.c_int(), ctypes.c_int(), ctypes.create_string_buffer(200)
        if ref.ref_stbi_loadf(str(p).encode(), ctypes.byref(w), ctypes.byref(h), None, ctypes.c_size_t(0), why, ctypes.c_size_t(200)):
            return why.value.decode()
        out = np.zeros((h.value, w.value, 4), np.float32)
        ref.ref_stbi_loadf(str(p).encode(), ctypes.byref(w), ctypes.byref(h), out.ctypes.data_as(ctypes.c_void_p), ctypes.c_size_t(out.size), why, ctypes.c_size_t(200))
        return out

    def check(tag):
        want_f, want_b = ref_float(path), _ref_stbi(ref, path)
        if isinstance(want_f, str):
            assert isinstance(want_b, str)
            assert lib.ngpb_host_decode_hdr(str(path).encode(), None, 0, ctypes.byref(cw), ctypes.byref(ch)) != 0, (tag, want_f)
            assert lib.ngpb_host_decode_image(str(path).encode(), None, 0, ctypes.byref(cw), ctypes.byref(ch)) != 0, (tag, want_b)
            return 0
        assert lib.ngpb_host_decode_hdr(str(path).encode(), None, 0, ctypes.byref(cw), ctypes.byref(ch)) == 0, (tag, lib.ngpb_last_error().decode())
        got_f = np.zeros((ch.value, cw.value, 4), np.float32)
        assert lib.ngpb_host_decode_hdr(str(path).encode(), got_f.ctypes.data_as(ctypes.c_void_p), got_f.size, ctypes.byref(cw), ctypes.byref(ch)) == 0
        assert got_f.shape == want_f.shape and np.array_equal(got_f.view(np.uint32), want_f.view(np.uint32)), tag
        got_b = _decode_image(lib, path)
        assert got_b.shape == want_b.shape and np.array_equal(got_b, want_b), tag
        return 1
    n_ok = 0
    for (h, w) in ((13, 17), (1, 1), (5, 7), (9, 8), (40, 300)):
        img = np.exp(rng.standard_normal((h, w, 3)) * 2).astype(np.float32)
        img[:h // 2] = 0.5
        img[0, 0] = 0
        for comp in (0, 1):
            if not cv2.imwrite(str(path), img, [cv2.IMWRITE_HDR_COMPRESSION, comp]):
                pytest.skip("this OpenCV build cannot write Radiance HDR")
            n_ok += check((h, w, comp))
    assert n_ok == 10

    def hdr(w,h,body,magic=b'#?RADIANCE', fmt=b'FORMAT=32-bit_rle_rgbe', res=None, extra=b'EXPOSURE=1.0\n'):
        return magic+b'\n'+extra+fmt+b'\n\n'+(res or (b'-Y %d +X %d' % (h,w)))+b'\n'+body
    w,h=12,4
    flat = bytes(rng.integers(0,256,w*h*4,dtype=np.uint8))
    def rle_line(w):
        out = bytes([2,2,w>>8,w&255])
        for k in range(4):
            i=0
            while i<w:
                c=min(int(rng.integers(1,60)), w-i)
                if rng.random()<0.5: out += bytes([128+c, int(rng.integers(0,256))])
                else: out += bytes([c]) + bytes(rng.integers(0,256,c,dtype=np.uint8))
                i+=c
        return out
    cases = {'flat wide': hdr(w,h,flat), 'rgbe magic': hdr(w,h,b''.join(rle_line(w) for _ in range(h)), magic=b'#?RGBE'),
             'line 3 flat': hdr(w,h,rle_line(w)+rle_line(w)+flat), 'no format': hdr(w,h,flat,fmt=b'FORMAT=32-bit_rle_xyze'), '+Y layout': hdr(w,h,flat,res=b'+Y 4 +X 12'),
             'bad magic': hdr(w,h,flat,magic=b'#?RADIANC'), 'wrong len': hdr(w,h,bytes([2,2,0,11])+flat), 'zero run': hdr(w,h,bytes([2,2,0,12,128,0])+flat), 'truncated': hdr(w,h,rle_line(w))}
    n_ok = 0
    for tag, blob in cases.items():
        path.write_bytes(blob)
        n_ok += check(tag)
    assert n_ok == 4
