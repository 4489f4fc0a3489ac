"""Seeded inputs of the kernel-grain pin between the CUDA path and the REFERENCE'S OWN GPU KERNELS
(tiny-cuda-nn kernel_grid / kernel_grid_backward, kernel_mlp_fused + its backward, ngp::NerfNetwork,
adam_step / ema_step), run on a B200 through baseline/_ref/ref_harness (baseline/harness.cu).

One table, used by baseline/dump_ref_gpu.py (which runs the reference on the GPU box and writes
tests/golden/ref_gpu_vectors.npz) and by tests/test_ref_gpu_pin.py (which feeds the same inputs to
libngp_b200.so). Inputs are regenerated from the seeds on both sides; only reference OUTPUTS are committed.
"""
import numpy as np

GRID_L16F2 = {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19, "base_resolution": 16, "per_level_scale": 1.3819128274917603}
GRID_L8F4 = {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 4, "log2_hashmap_size": 19, "base_resolution": 16, "per_level_scale": 2.4380271434783936}
GRID_2D = {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 15, "base_resolution": 16, "per_level_scale": 1.5}
NERF_DIR = {"otype": "Composite", "nested": [{"n_dims_to_encode": 3, "otype": "SphericalHarmonics", "degree": 4}, {"otype": "Identity"}]}
NERF_DENSITY = {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 1}
NERF_RGB = {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2}
# configs/nerf/base.json:5-22 of the reference
OPTIMIZER = {"otype": "Ema", "decay": 0.95, "nested": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000, "decay_base": 0.33,
             "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}}}

GRID_CASES = {"L16F2": (3, GRID_L16F2, 2048), "L8F4": (3, GRID_L8F4, 2048), "2D": (2, GRID_2D, 2048)}
MLP_CASES = {"rgb_32_2": (32, 3, 2, 1024), "density_32_1": (32, 16, 1, 1024), "wide_64_3": (64, 3, 3, 512)}


def grid_n_params(n_dims, enc):
    """GridOffsetTable sizes (tcnn grid.h:668-730): dense while res^D fits, else 2^log2_hashmap_size; multiples of 8."""
    total, b = 0, np.float32(np.log2(np.float32(enc["per_level_scale"])))
    for l in range(enc["n_levels"]):
        scale = np.float32(np.exp2(np.float32(l) * b)) * np.float32(enc["base_resolution"]) - np.float32(1.0)
        res = int(np.ceil(scale)) + 1
        n = min(res ** n_dims, 1 << 62)
        n = (n + 7) // 8 * 8
        total += min(n, 1 << enc["log2_hashmap_size"])
    return total * enc["n_features_per_level"]


def grid_case(name):
    n_dims, enc, n = GRID_CASES[name]
    g = np.random.default_rng([7, sorted(GRID_CASES).index(name)])
    x = g.random((n, n_dims), dtype=np.float32)
    # a third of the points repeat a neighbour's cell on the coarse levels: the scatter has to merge them
    x[1::3] = np.clip(x[0::3][:len(x[1::3])] + g.normal(0, 1e-3, size=x[1::3].shape).astype(np.float32), 0, np.float32(0.999999))
    params = (g.standard_normal(grid_n_params(n_dims, enc)) * 0.1).astype(np.float16)
    dy = g.standard_normal((n, enc["n_levels"] * enc["n_features_per_level"])).astype(np.float16)
    return n_dims, enc, x, params, dy


def mlp_n_params(n_in, n_hidden):
    return 64 * n_in + (n_hidden - 1) * 64 * 64 + 16 * 64


def mlp_weights(g, n_in, n_hidden):
    parts = []
    for rows, cols in [(64, n_in)] + [(64, 64)] * (n_hidden - 1) + [(16, 64)]:
        s = np.sqrt(6.0 / (rows + cols))
        parts.append(g.uniform(-s, s, size=rows * cols))
    return np.concatenate(parts).astype(np.float16)


def mlp_case(name):
    n_in, n_out, n_hidden, n = MLP_CASES[name]
    g = np.random.default_rng([11, sorted(MLP_CASES).index(name)])
    cfg = {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": n_hidden}
    # inputs exactly representable in fp16: the reference casts the fp32 input to fp16 first (Identity encoding)
    x = g.uniform(-1, 1, size=(n, n_in)).astype(np.float16).astype(np.float32)
    w = mlp_weights(g, n_in, n_hidden)
    dy = np.zeros((n, 16), dtype=np.float16)
    dy[:, :n_out] = (g.standard_normal((n, n_out)) * 0.1).astype(np.float16)
    return n_in, n_out, cfg, x, w, dy


def nerfnet_case():
    g = np.random.default_rng(13)
    n = 1024
    enc = dict(GRID_L8F4, log2_hashmap_size=15)
    x = g.random((n, 7), dtype=np.float32)   # NerfCoordinate rows: pos, dt, dir in [0,1]
    grid = (g.standard_normal(grid_n_params(3, enc)) * 0.1).astype(np.float16)
    params = np.concatenate([mlp_weights(g, 32, 1), mlp_weights(g, 32, 2), grid])   # density MLP | rgb MLP | grid (nerf_network.h:356-371)
    dy = np.zeros((n, 16), dtype=np.float16)
    dy[:, :4] = (g.standard_normal((n, 4)) * 0.1).astype(np.float16)
    return {"encoding": enc, "dir_encoding": NERF_DIR, "network": NERF_DENSITY, "rgb_network": NERF_RGB}, x, params, dy


def adam_case():
    g = np.random.default_rng(17)
    layers = [(64, 32), (16, 64)]
    n_matrix = sum(r * c for r, c in layers)
    n = n_matrix + 4096                     # matrix weights first, then "grid" weights (zero-gradient skip applies there)
    steps = 24
    w = (g.standard_normal(n) * 0.05).astype(np.float32)
    grads = (g.standard_normal((steps, n)) * 128 * 1e-2).astype(np.float16)   # loss-scaled
    mask = g.random((steps, n)) < 0.5
    mask[:, :n_matrix] = False
    grads[mask] = 0                         # untouched hash-table entries of a step
    return {"optimizer": OPTIMIZER, "layers": [list(l) for l in layers], "n": n, "steps": steps, "loss_scale": 128.0}, w, grads
