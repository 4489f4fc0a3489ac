"""NerfNetwork inference time against batch size (CUDA events, no L2 flush): the fixed cost of the fused inference kernels."""
import importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import ref_gpu_cases as rc
ngp = importlib.import_module("instant-ngp-rendering_b200")
tbm = importlib.import_module("instant-ngp-rendering_b200.testbed")
g = np.random.default_rng(1)
m = ngp.NerfNetwork(rc.GRID_L8F4, rc.NERF_DIR, rc.NERF_DENSITY, rc.NERF_RGB)
params = np.concatenate([rc.mlp_weights(g, 32, 1), rc.mlp_weights(g, 32, 2), (g.standard_normal(rc.grid_n_params(3, rc.GRID_L8F4)) * 0.1).astype(np.float16)])
m.set_params(torch.from_numpy(params).cuda())
for log2 in (8, 12, 14, 16, 17, 18, 19, 20, 22):
    n = 1 << log2
    x = torch.rand((n, 7), device="cuda")
    for _ in range(5):
        m.inference(x)
    torch.cuda.synchronize()
    tbm.profile_enable(True)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        m.inference(x)
    b.record(); torch.cuda.synchronize()
    prof = tbm.profile_collect(); tbm.profile_enable(False)
    print(n, "total %.1f us" % (a.elapsed_time(b) / 20 * 1e3), {k: round(ms / cnt * 1e3, 1) for k, (ms, cnt) in prof.items()})
