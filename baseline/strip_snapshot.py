"""Reduces a snapshot (.ingp / .msgpack) to a small JSON record that can be committed as a fixture: every key and
scalar as written, binary members replaced by {"n_bytes", "sha256", "head": first 32 bytes hex}.

    python baseline/strip_snapshot.py IN.ingp OUT.json
"""
import hashlib
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
# the container reader only (msgpack + zlib); it needs no device
snap = importlib.import_module("instant-ngp-rendering_b200.snapshot")


def strip(x):
    if isinstance(x, (bytes, bytearray, memoryview)):
        b = bytes(x)
        return {"__binary__": True, "n_bytes": len(b), "sha256": hashlib.sha256(b).hexdigest(), "head": b[:32].hex()}
    if isinstance(x, dict):
        return {str(k): strip(v) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return [strip(v) for v in x]
    return x


if __name__ == "__main__":
    cfg = snap.read_container(sys.argv[1])
    with open(sys.argv[2], "w") as f:
        json.dump(strip(cfg), f, indent=1)
    print("stripped", sys.argv[1], os.path.getsize(sys.argv[1]), "->", sys.argv[2], os.path.getsize(sys.argv[2]))
