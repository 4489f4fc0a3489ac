"""Compares what integration/tcnn_caller wrote on its two back ends (the reference's tcnn build and libngp_b200.so behind
integration/tcnn_cpp_api_b200.cpp): python scripts/compare_tcnn_caller.py <dir with run_b200.bin, run_ref.bin, describe_*.txt>"""
import json
import struct
import sys

import numpy as np

CASES = ["encoding_L16F2", "network_with_input_encoding_L8F4_mlp", "image_network_2d", "network_mlp_32_to_16", "network_mlp_identity_5_to_3"]
ARRAYS = ["params", "inference_out", "forward_out", "dL_dparams", "dL_dinput"]


def read(path):
    out, blob, at = {}, open(path, "rb").read(), 0
    for case in CASES:
        for name in ARRAYS:
            n, stride, total, sumsq, nonzero = struct.unpack_from("<QQddQ", blob, at)
            at += 40
            k = (n + stride - 1) // stride if n else 0
            out[case, name] = {"n": n, "stride": stride, "sum": total, "sumsq": sumsq, "nonzero": nonzero, "sample": np.frombuffer(blob, np.float32, k, at)}
            at += 4 * k
    assert at == len(blob)
    return out


def compare(folder):
    ours, ref = read(folder + "/run_b200.bin"), read(folder + "/run_ref.bin")
    report = {}
    for key in ours:
        a, b = ours[key], ref[key]
        row = {"n": a["n"], "n_ref": b["n"], "sampled": int(a["sample"].size)}
        if a["n"] == b["n"] and a["n"]:
            d = np.abs(a["sample"] - b["sample"])
            scale = float(np.abs(b["sample"]).max())
            row.update(identical=float((a["sample"].view(np.uint32) == b["sample"].view(np.uint32)).mean()), max_abs_diff=float(d.max()), ref_max_abs=scale,
                       rel_l2=float(np.linalg.norm(a["sample"] - b["sample"]) / max(np.linalg.norm(b["sample"]), 1e-30)),
                       rms=[float(np.sqrt(a["sumsq"] / a["n"])), float(np.sqrt(b["sumsq"] / b["n"]))], nonzero=[a["nonzero"], b["nonzero"]])
        report["%s.%s" % key] = row
    d_ours = [json.loads(l) for l in open(folder + "/describe_b200.txt") if l.startswith("{")]
    d_ref = [json.loads(l) for l in open(folder + "/describe_ref.txt") if l.startswith("{")]
    facts = []
    for a, b in zip(d_ours, d_ref):
        same = {k: a.get(k) == b.get(k) for k in b if k != "hyperparams"}
        facts.append({"case": a.get("case", "library"), "equal": same, "hyperparams_ours": a.get("hyperparams"), "hyperparams_ref": b.get("hyperparams")})
    return {"arrays": report, "describe": facts}


if __name__ == "__main__":
    print(json.dumps(compare(sys.argv[1]), indent=1))
