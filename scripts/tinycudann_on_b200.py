"""The reference's own PyTorch package (tinycudann: its Python modules and its bindings.cpp, both unmodified, built by
integration/build_torch_binding.py) running on libngp_b200.so. Writes one JSON report.

    python scripts/tinycudann_on_b200.py [out.json]

Sections (each reported on its own, a failure in one does not hide the others):
  encoding   tcnn.Encoding forward on a ragged batch == this package's own wrapper on the same parameters, bit for bit
  network    tcnn.Network (Identity encoding) output and d(output)/d(input) against a plain PyTorch fp32 MLP on the same weights
             (the loss is a SUM over 2048 samples times tinycudann's loss scale of 128, so the fp16 weight gradients overflow -
             `dparams_finite` reports that; it comes with fp16 gradient storage, not with the kernels. The training section uses a mean)
  training   tcnn.NetworkWithInputEncoding fitted to an analytic field with torch.optim.Adam: the loss falls
"""
import importlib
import json
import os
import sys
import time
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "integration", "_build", "torch"), os.path.join(ROOT, "baseline", "_ref", "tinycudann_pkg"), ROOT]

GRID = {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19, "base_resolution": 16, "per_level_scale": 1.3819}
SMALL_GRID = {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 4, "log2_hashmap_size": 15, "base_resolution": 8, "per_level_scale": 1.6}
MLP = {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2}


def section_encoding(torch, tcnn, pkg):
    enc = tcnn.Encoding(3, GRID)
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.rand(4096 + 17, 3, device="cuda", generator=g)
    with torch.no_grad():
        y = enc(x)
    mine = pkg.modules.Encoding(3, GRID)
    params = enc.params.detach().to(torch.float16).contiguous()
    x_pad = torch.nn.functional.pad(x, [0, 0, 0, (-x.shape[0]) % 256]).contiguous()
    y2 = mine.inference(x_pad, params)[: x.shape[0]]
    # parameters: the binding's initial_params(seed) against this package's initialize_params(seed)
    same_init = bool(torch.equal(mine.initialize_params(1337), params))
    return {"shape": list(y.shape), "dtype": str(y.dtype), "identical_to_own_wrapper": bool(torch.equal(y, y2)), "same_initial_params": same_init,
            "n_params": int(enc.params.numel()), "abs_max": float(y.float().abs().max())}


def section_network(torch, tcnn):
    net = tcnn.Network(32, 16, MLP)
    g = torch.Generator(device="cuda").manual_seed(6)
    x = (torch.rand(2048, 32, device="cuda", generator=g) - 0.5).requires_grad_(True)
    y = net(x)
    w = torch.linspace(-1.0, 1.0, 16, device="cuda")
    (y.float() * w).sum().backward()
    p = net.params.detach().float()
    w0, w1, w2 = p[: 64 * 32].view(64, 32), p[64 * 32: 64 * 32 + 64 * 64].view(64, 64), p[64 * 32 + 64 * 64:].view(16, 64)   # row-major [out][in], fully_fused_mlp.cu:656-678
    x32 = x.detach().clone().requires_grad_(True)
    y32 = torch.relu(torch.relu(x32 @ w0.half().float().T) @ w1.half().float().T) @ w2.half().float().T
    (y32 * w).sum().backward()

    def rel(a, b):
        return float((a - b).norm() / b.norm())
    return {"out_rel_l2_vs_fp32": rel(y.detach().float(), y32.detach()), "dinput_rel_l2_vs_fp32": rel(x.grad, x32.grad),
            "dparams_nonzero": int((net.params.grad != 0).sum()), "dparams_finite": bool(torch.isfinite(net.params.grad).all())}


def section_training(torch, tcnn):
    net = tcnn.NetworkWithInputEncoding(3, 3, SMALL_GRID, MLP)
    opt = torch.optim.Adam(net.parameters(), lr=1e-2, eps=1e-15)
    g = torch.Generator(device="cuda").manual_seed(7)

    def field(x):
        return torch.stack([torch.sin(6.0 * x[:, 0]), x[:, 1] * x[:, 2], torch.cos(4.0 * x[:, 2])], -1)
    losses = []
    for _ in range(80):
        x = torch.rand(1 << 14, 3, device="cuda", generator=g)
        loss = ((net(x).float() - field(x)) ** 2).mean()
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(float(loss))
    return {"loss_first": losses[0], "loss_last": sum(losses[-8:]) / 8, "steps": len(losses), "n_params": int(net.params.numel())}


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "tinycudann_on_b200.json")
    report = {"what": "reference tinycudann Python package + bindings.cpp (unmodified) over integration/tcnn_cpp_api_b200.cpp + libngp_b200.so"}
    t0 = time.time()
    try:
        import torch
        report["device"] = torch.cuda.get_device_name(0)
        import tinycudann as tcnn
        report["binding"] = tcnn.modules._C.__name__
        report["tinycudann_file"] = os.path.relpath(tcnn.__file__, ROOT)
        pkg = importlib.import_module("instant-ngp-rendering_b200")
        for name, fn in (("encoding", lambda: section_encoding(torch, tcnn, pkg)), ("network", lambda: section_network(torch, tcnn)),
                         ("training", lambda: section_training(torch, tcnn))):
            try:
                report[name] = fn()
            except Exception:   # noqa: BLE001 - a report, not a test
                report[name] = {"error": traceback.format_exc()[-1500:]}
    except Exception:   # noqa: BLE001
        report["error"] = traceback.format_exc()[-1500:]
    report["seconds"] = round(time.time() - t0, 2)
    os.makedirs(os.path.dirname(out_path), exist_ok=True)
    with open(out_path, "w") as f:
        json.dump(report, f, indent=1)
    print(json.dumps(report, indent=1))


if __name__ == "__main__":
    main()
