"""Renders one trained model with the empty-space skip climbing to the scene's last cascade (the reference's rule) and past it
(nerf_device.cuh kSkipMaxMip): the frames and the per-pixel sample counts must agree."""
import importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
ngp = importlib.import_module("instant-ngp-rendering_b200")
syn = importlib.import_module("instant-ngp-rendering_b200.synthetic")
tb = ngp.Testbed()
scene, poses, focal = syn.make_dataset(tb, n_images=50, width=1920, height=1080, aabb_scale=4, amp_scale=8.0)
tb.reload_network_from_json(syn.NERF_BASE_CONFIG)
for i in range(600):
    tb.train(1 << 18)
view = syn.look_at_ngp(np.array([0.5, 0.5, 0.5]) + 1.4 * np.array([0.62, -0.55, 0.56]))
res = {}
for on in (0, 1, 0, 1):
    tb._set("render_skip_past_last_cascade", on)
    tb.render_nerf(1920, 1080, view, focal)
    torch.cuda.synchronize(); t0 = time.time()
    rgba, depth, steps = tb.render_nerf(1920, 1080, view, focal, return_steps=True)
    torch.cuda.synchronize(); dt = time.time() - t0
    res[on] = (rgba.clone(), steps.clone())
    print("skip past last cascade", on, "ms", round(dt * 1e3, 2), "queries/ray", tb.last_render_samples / (1920 * 1080), "composited samples", int(steps.sum()))
a, b = res[0], res[1]
print("pixels with identical sample counts: %.4f %%; max |d rgba| %.3e; mean %.3e" % (100 * float((a[1] == b[1]).float().mean()), float((a[0] - b[0]).abs().max()), float((a[0] - b[0]).abs().mean())))
