"""Snapshot container and key handling (instant-ngp-rendering_b200/snapshot.py) against the rules of
Testbed::save_snapshot / load_snapshot / load_network_config (src/testbed.cu:246-272, 4775-4972):
msgpack with `bin` members, a zlib stream for *.ingp only, version / mode / grid-size checks with the
reference's messages. Host logic: runs without a GPU."""
import importlib
import zlib

import numpy as np
import pytest

msgpack = pytest.importorskip("msgpack")
S = importlib.import_module("instant-ngp-rendering_b200.snapshot")


def _config():
    rng = np.random.default_rng(0)
    return {"encoding": {"otype": "HashGrid", "n_levels": 8}, "network": {"otype": "FullyFusedMLP"},
            "snapshot": {"version": 1, "mode": "nerf", "n_params": 16, "params_type": "__half",
                         "params_binary": rng.standard_normal(16).astype(np.float16).tobytes(), "density_grid_size": 128,
                         "density_grid_binary": b"", "training_step": 7, "loss": 0.125,
                         "camera": {"matrix": [[1.0, 0.0, 0.0, 0.5], [0.0, 1.0, 0.0, 0.5], [0.0, 0.0, 1.0, 0.5]]}}}


def test_container_round_trip_and_compression(tmp_path):
    cfg = _config()
    for name, compress in (("a.ingp", True), ("b.ingp", False), ("c.msgpack", True), ("D.INGP", True)):
        p = str(tmp_path / name)
        S.write_container(p, cfg, compress=compress)
        assert S.read_container(p) == cfg
    raw = (tmp_path / "c.msgpack").read_bytes()
    assert msgpack.unpackb(raw, raw=False) == cfg                                  # plain msgpack, no zlib around it
    assert zlib.decompress((tmp_path / "a.ingp").read_bytes()) == raw               # *.ingp: the same bytes inside a zlib stream
    assert zlib.decompress((tmp_path / "b.ingp").read_bytes()) == raw               # Z_NO_COMPRESSION still writes the zlib framing
    assert len((tmp_path / "b.ingp").read_bytes()) > len(raw)
    # the binary members are msgpack `bin` values (0xc4 = bin8 with the length byte), as nlohmann writes json::binary_t
    assert b"params_binary\xc4\x20" in raw
    # zstr hands a stream without a zlib header through unchanged
    (tmp_path / "plain.ingp").write_bytes(raw)
    assert S.read_container(str(tmp_path / "plain.ingp")) == cfg


def test_missing_file_unknown_extension_and_json(tmp_path):
    with pytest.raises(FileNotFoundError, match="Network snapshot '.*nope.ingp' does not exist."):
        S.read_container(str(tmp_path / "nope.ingp"))
    (tmp_path / "x.bin").write_bytes(b"\x00")
    assert S.read_container(str(tmp_path / "x.bin")) == {}
    with pytest.raises(RuntimeError, match="does not contain a snapshot"):
        S.load(None, str(tmp_path / "x.bin"))
    (tmp_path / "base.json").write_text('{"a": {"b": 1, "c": 2}} // parent\n')
    (tmp_path / "child.json").write_text('{"parent": "base.json", "a": {"c": 3}}')
    assert S.read_container(str(tmp_path / "child.json")) == {"a": {"b": 1, "c": 3}}


def test_checks_come_before_any_device_work():
    cfg = _config()
    cfg["snapshot"]["version"] = 0
    with pytest.raises(RuntimeError, match="Snapshot uses an old format and can not be loaded."):
        S.restore(None, cfg)
    del cfg["snapshot"]["version"]
    with pytest.raises(RuntimeError, match="old format"):
        S.restore(None, cfg)
    cfg = _config()
    cfg["snapshot"]["mode"] = "none"
    with pytest.raises(RuntimeError, match="Unknown snapshot mode"):
        S.restore(None, cfg)
    cfg["snapshot"]["mode"] = "sdf"
    with pytest.raises(RuntimeError, match="not part of the path"):
        S.restore(None, cfg)
    cfg = _config()
    cfg["snapshot"]["density_grid_size"] = 64
    with pytest.raises(RuntimeError, match="Incompatible grid size."):
        S.restore(None, cfg)
    with pytest.raises(RuntimeError, match="does not contain a snapshot"):
        S.restore(None, {"encoding": {}})


def test_binary_and_lens_json_forms():
    blob = bytes(range(7))
    assert S._binary(blob) == blob and S._binary(bytearray(blob)) == blob
    assert S._binary({"bytes": list(blob), "subtype": None}) == blob               # nlohmann's text form of a binary
    with pytest.raises(ValueError, match="must be either binary or object"):
        S._binary([1, 2, 3])
    assert S.lens_to_json(0, [0.0] * 4) is None                                    # Perspective writes nothing
    j = S.lens_to_json(1, [0.1, -0.2, 0.003, -0.004])
    assert j["is_fisheye"] is False and S.lens_from_json(j) == (1, [pytest.approx(0.1), pytest.approx(-0.2), pytest.approx(0.003), pytest.approx(-0.004)])
    assert S.lens_from_json(None) == (0, [0.0] * 4)
    for bad in ({"is_fisheye": True, "k1": 0.1, "k2": 0, "k3": 0, "k4": 0}, {"ftheta_p0": 1.0}, {"latlong": True}):
        with pytest.raises(ValueError):
            S.lens_from_json(bad)
    # a tmat goes out as a list of rows (vec_json.h:39-48)
    m = np.arange(12, dtype=np.float32).reshape(3, 4)
    assert S._mat(m) == [[0.0, 1.0, 2.0, 3.0], [4.0, 5.0, 6.0, 7.0], [8.0, 9.0, 10.0, 11.0]]
