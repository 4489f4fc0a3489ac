"""The drop-in boundary exercised from the reference's side: integration/tcnn_caller.cpp is written against the reference's
raw-pointer facade only (tcnn/include/tiny-cuda-nn/cpp_api.h:58-119) and linked twice by integration/build.sh - against
the reference's own tcnn build and against libngp_b200.so through integration/tcnn_cpp_api_b200.cpp (the binding of
INTEGRATION.md section 1). tests/golden/tcnn_caller_ref.{json,_outputs.npz} hold what the reference back end printed and
computed on a B200 (the `describe` lines, whole-array statistics, the first 512 rows of every output).

Bars (GPU): parameters from initialize_params(seed) identical bit for bit; grid-encoding outputs identical bit for bit;
network outputs rel-L2 <= 2e-3 (the reference accumulates in fp16, this library in fp32); gradients rel-L2 <= 3e-2 (the
reference's split-K sums and atomics are fp16 in arbitrary order).
"""
import json
import os
import struct
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "integration", "_build")
OURS = os.path.join(BUILD, "tcnn_caller_b200")
REF = os.path.join(BUILD, "tcnn_caller_ref")
GOLDEN = os.path.join(ROOT, "tests", "golden")
CASES = ["encoding_L16F2", "network_with_input_encoding_L8F4_mlp", "image_network_2d", "network_mlp_32_to_16", "network_mlp_identity_5_to_3"]
ARRAYS = ["params", "inference_out", "forward_out", "dL_dparams", "dL_dinput"]


def _golden():
    with open(os.path.join(GOLDEN, "tcnn_caller_ref.json")) as f:
        return json.load(f)


def _read(path):
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


def _rel_l2(a, b):
    return float(np.linalg.norm(a.astype(np.float64) - b.astype(np.float64)) / max(np.linalg.norm(b.astype(np.float64)), 1e-30))


def _need(path):
    if not os.path.exists(path):
        pytest.skip(os.path.relpath(path, ROOT) + " is not built (integration/build.sh needs the reference's headers)")


def test_describe_matches_the_reference_back_end():
    """Module facts through the facade, no device needed on this side: batch granularity, loss scales, names, dims, padded
    output widths, parameter counts, precisions and the hyperparams JSON equal what the reference's build answered."""
    _need(OURS)
    res = subprocess.run([OURS, "describe"], capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, res.stderr
    ours = [json.loads(line) for line in res.stdout.splitlines() if line.startswith("{")]
    want = _golden()["describe"]
    assert len(ours) == len(want) == 1 + len(CASES)
    for a, b in zip(ours, want):
        assert a == b, (a, b)
    # the text of a float that came from the config survives as nlohmann prints it (a widened float keeps all its digits)
    assert '"per_level_scale":1.3818999528884888' in res.stdout


def test_shim_source_follows_the_facade():
    """The binding implements every pure virtual and every free function of cpp_api.h (a missing one fails the link in
    integration/build.sh; this keeps the list visible where the reference is absent)."""
    text = open(os.path.join(ROOT, "integration", "tcnn_cpp_api_b200.cpp")).read()
    for name in ("inference(", "forward(", "backward(", "backward_backward_input(", "n_input_dims()", "n_output_dims()", "n_params()", "initialize_params(",
                 "hyperparams()", "name()", "batch_size_granularity()", "cuda_device()", "set_cuda_device(", "free_temporary_memory()", "has_networks()",
                 "default_loss_scale(", "preferred_precision()", "set_log_callback(", "create_network_with_input_encoding(", "create_network(", "create_encoding("):
        assert name in text, name
    assert "#include <tiny-cuda-nn/cpp_api.h>" in text and "#include <ngp_b200.h>" in text
    caller = open(os.path.join(ROOT, "integration", "tcnn_caller.cpp")).read()
    includes = [line for line in caller.splitlines() if line.startswith("#include")]
    assert "#include <tiny-cuda-nn/cpp_api.h>" in includes and not any("ngp_b200" in line for line in includes), "the caller must see the reference's header only"


def _check_against(ours, ref_arrays, ref_stats):
    """ours: _read() of this library's run; ref_arrays(case, name) -> (reference values, slice into ours' sample)."""
    for case in CASES:
        for name in ARRAYS:
            assert ours[case, name]["n"] == ref_stats["%s.%s" % (case, name)]["n"], (case, name)
        # initialize_params: the same pcg32 stream as the reference
        st = ref_stats[case + ".params"]
        assert ours[case, "params"]["sum"] == st["sum"] and ours[case, "params"]["sumsq"] == st["sumsq"], case
        want, sl = ref_arrays(case, "params")
        assert np.array_equal(ours[case, "params"]["sample"][sl].view(np.uint32), want.view(np.uint32)), case
        # outputs
        want, sl = ref_arrays(case, "inference_out")
        got = ours[case, "inference_out"]["sample"][sl]
        assert np.array_equal(ours[case, "inference_out"]["sample"], ours[case, "forward_out"]["sample"]), case
        if case.startswith("encoding"):
            assert np.array_equal(got.astype(np.float16).view(np.uint16), want.astype(np.float16).view(np.uint16)), case
        else:
            assert _rel_l2(got, want.astype(np.float32)) <= 2e-3, (case, _rel_l2(got, want.astype(np.float32)))
        # gradients
        st = ref_stats[case + ".dL_dparams"]
        a = ours[case, "dL_dparams"]
        assert abs(np.sqrt(a["sumsq"] / a["n"]) - np.sqrt(st["sumsq"] / st["n"])) <= 1e-2 * np.sqrt(st["sumsq"] / st["n"]), case
        assert abs(int(a["nonzero"]) - int(st["nonzero"])) <= 1e-3 * st["nonzero"] + 2, case
        want, sl = ref_arrays(case, "dL_dinput")
        got = ours[case, "dL_dinput"]["sample"][sl]
        if case.startswith("network_mlp_"):
            assert np.abs(want).max() > 0.1
            assert _rel_l2(got, want) <= 3e-2, (case, _rel_l2(got, want))
        else:
            assert not got.any() and not want.any()   # not requested: the buffers stay as allocated (zero)


@pytest.mark.gpu
def test_same_caller_same_numbers_against_the_committed_reference_run(tmp_path):
    """The b200 back end on this GPU against what the reference back end wrote on a B200 (golden, first 512 rows)."""
    _need(OURS)
    res = subprocess.run([OURS, "run", str(tmp_path / "ours.bin")], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr + res.stdout
    ours = _read(str(tmp_path / "ours.bin"))
    golden = _golden()
    npz = np.load(os.path.join(GOLDEN, "tcnn_caller_ref_outputs.npz"))

    def ref_arrays(case, name):
        if name == "params":
            return npz[case + ".params_sample"], slice(0, 8 * 4096, 8)
        want = npz["%s.%s" % (case, name)]
        return want, slice(0, want.size)
    _check_against(ours, ref_arrays, golden["stats"])
    # gradient samples (every 8th sampled entry): same entries touched, values within the fp16-accumulation bar
    for case in CASES:
        want = npz[case + ".dL_dparams_sample"].astype(np.float32)
        got = ours[case, "dL_dparams"]["sample"][::8][:4096]
        assert _rel_l2(got, want) <= 3e-2, (case, _rel_l2(got, want))


@pytest.mark.gpu
def test_same_caller_on_both_back_ends_live(tmp_path):
    """Both binaries on this GPU, whole arrays."""
    _need(OURS)
    _need(REF)
    for exe, name in ((OURS, "ours.bin"), (REF, "ref.bin")):
        res = subprocess.run([exe, "run", str(tmp_path / name)], capture_output=True, text=True, timeout=300)
        if exe == REF and res.returncode != 0 and "libcuda" in res.stderr:
            pytest.skip("the reference back end cannot load the CUDA driver here")
        assert res.returncode == 0, res.stderr + res.stdout
    ours, ref = _read(str(tmp_path / "ours.bin")), _read(str(tmp_path / "ref.bin"))
    stats = {"%s.%s" % k: {kk: v[kk] for kk in ("n", "stride", "sum", "sumsq", "nonzero")} for k, v in ref.items()}
    _check_against(ours, lambda case, name: (ref[case, name]["sample"], slice(None)), stats)
    for case in CASES:
        assert _rel_l2(ours[case, "dL_dparams"]["sample"], ref[case, "dL_dparams"]["sample"]) <= 3e-2, case


def test_reference_torch_binding_loads_without_a_device():
    """tcnn/bindings/torch/tinycudann/bindings.cpp, compiled unmodified over the shim (integration/build_torch_binding.py): the
    extension imports, builds modules from Python dicts and answers the facade's questions. (Running it needs a GPU:
    tests/test_tinycudann_gpu.py.)"""
    import glob
    import sys
    built = glob.glob(os.path.join(BUILD, "torch", "tinycudann_bindings", "_90_C*.so"))
    if not built:
        pytest.skip("the torch binding is not built (integration/build_torch_binding.py needs the reference's sources)")
    code = (
        "import sys, json, torch\n"
        "sys.path.insert(0, %r)\n"
        "import importlib\n"
        "C = importlib.import_module('tinycudann_bindings._90_C')\n"
        "grid = {'otype': 'HashGrid', 'n_levels': 16, 'n_features_per_level': 2, 'log2_hashmap_size': 19, 'base_resolution': 16, 'per_level_scale': 1.3819}\n"
        "mlp = {'otype': 'FullyFusedMLP', 'activation': 'ReLU', 'output_activation': 'None', 'n_neurons': 64, 'n_hidden_layers': 2}\n"
        "e = C.create_encoding(3, grid, C.Precision.Fp16)\n"
        "n = C.create_network_with_input_encoding(3, 4, grid, mlp)\n"
        "m = C.create_network(32, 16, mlp)\n"
        "print(json.dumps({'enc': [e.n_params(), e.n_input_dims(), e.n_output_dims(), e.name(), e.hyperparams()],\n"
        "                  'net': [n.n_params(), n.n_output_dims(), n.name()], 'mlp': [m.n_params(), m.n_output_dims()],\n"
        "                  'lib': [C.batch_size_granularity(), C.default_loss_scale(C.Precision.Fp16), C.has_networks(), C.preferred_precision() == C.Precision.Fp16,\n"
        "                          e.param_precision() == C.Precision.Fp16, e.output_precision() == C.Precision.Fp16]}))\n"
    ) % os.path.join(BUILD, "torch")
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr[-2000:]
    got = json.loads(res.stdout.strip().splitlines()[-1])
    want = {d.get("case"): d for d in _golden()["describe"]}
    enc = want["encoding_L16F2"]
    assert got["enc"] == [enc["n_params"], 3, enc["n_output_dims"], enc["name"], enc["hyperparams"]]
    assert got["net"] == [12196240 + 32 * 64 + 64 * 64 + 64 * 16, 16, "NetworkWithInputEncoding"]
    assert got["mlp"] == [want["network_mlp_32_to_16"]["n_params"], 16]
    assert got["lib"] == [256, 128.0, True, True, True, True]
