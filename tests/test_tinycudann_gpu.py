"""The reference's own PyTorch package on this library: tinycudann's Python modules and its bindings.cpp, both unmodified
(integration/build_torch_binding.py compiles the binding against integration/tcnn_cpp_api_b200.cpp and installs the Python
files under baseline/_ref/), driven by scripts/tinycudann_on_b200.py in a process of its own. Bars: tcnn.Encoding equals this
package's wrapper bit for bit on a ragged batch and starts from the same parameters; tcnn.Network agrees with a plain PyTorch
fp32 MLP on the same (half-rounded) weights to rel-L2 2e-3 on outputs and 3e-2 on input gradients (fp16 activations and
gradients between layers); 80 Adam steps of tcnn.NetworkWithInputEncoding on an analytic field cut the loss at least 20-fold."""
import glob
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_reference_tinycudann_package_runs_on_this_library(tmp_path):
    if not glob.glob(os.path.join(ROOT, "integration", "_build", "torch", "tinycudann_bindings", "_90_C*.so")):
        pytest.skip("the torch binding is not built (integration/build_torch_binding.py needs the reference's sources)")
    if not os.path.exists(os.path.join(ROOT, "baseline", "_ref", "tinycudann_pkg", "tinycudann", "modules.py")):
        pytest.skip("the reference's tinycudann Python package is not installed under baseline/_ref")
    out = tmp_path / "report.json"
    res = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "tinycudann_on_b200.py"), str(out)], capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    report = json.loads(out.read_text())
    assert "error" not in report, report.get("error")
    assert report["binding"] == "tinycudann_bindings._90_C"
    enc, net, train = report["encoding"], report["network"], report["training"]
    for section in (enc, net, train):
        assert "error" not in section, section.get("error")
    assert enc["identical_to_own_wrapper"] and enc["same_initial_params"] and enc["shape"] == [4113, 32]
    assert net["out_rel_l2_vs_fp32"] <= 2e-3 and net["dinput_rel_l2_vs_fp32"] <= 3e-2 and net["dparams_nonzero"] > 6000
    assert train["loss_last"] * 20 < train["loss_first"]
