"""The reference's shipped network configs, restated as dicts (configs/nerf/base.json and
configs/image/base.json of this fork; the files themselves stay under /root/reference, which does not travel to
the GPU box). Imported by bench.py, the baseline runners and the scripts; imports nothing of the product."""

NERF_BASE_CONFIG = {
    "loss": {"otype": "Huber"},
    "optimizer": {"otype": "Ema", "decay": 0.95, "nested": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000, "decay_base": 0.33,
                  "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}}},
    "encoding": {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 4, "log2_hashmap_size": 19, "base_resolution": 16},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 1},
    "dir_encoding": {"otype": "Composite", "nested": [{"n_dims_to_encode": 3, "otype": "SphericalHarmonics", "degree": 4}, {"otype": "Identity"}]},
    "rgb_network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}

IMAGE_BASE_CONFIG = {
    "loss": {"otype": "L2"},
    "optimizer": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000, "decay_base": 0.33,
                  "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}},
    "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 24, "base_resolution": 16},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}


def look_at_ngp(eye, target=(0.5, 0.5, 0.5), up=(0.0, 0.0, 1.0)):
    """Camera-to-world 3x4 in the NGP convention: columns right, down, forward, origin (numpy only)."""
    import numpy as np
    eye = np.asarray(eye, dtype=np.float64)
    fwd = np.asarray(target, dtype=np.float64) - eye
    fwd /= np.linalg.norm(fwd)
    right = np.cross(fwd, np.asarray(up, dtype=np.float64))
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    return np.ascontiguousarray(np.stack([right, down, fwd, eye], axis=1), dtype=np.float32)


def orbit_poses(n, radius=1.1, center=(0.5, 0.5, 0.5)):
    """n cameras on a sphere around `center`, golden-angle spiral on the upper 3/4 (the poses of bench.py's synthetic set)."""
    import math
    import numpy as np
    poses = []
    for i in range(n):
        z = 0.75 - 1.25 * (i + 0.5) / n
        phi = i * math.pi * (3.0 - math.sqrt(5.0))
        r = math.sqrt(max(1.0 - z * z, 0.0))
        poses.append(look_at_ngp(np.asarray(center) + radius * np.array([r * math.cos(phi), r * math.sin(phi), z]), center))
    return poses
