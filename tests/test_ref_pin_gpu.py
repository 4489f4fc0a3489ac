"""The CUDA path against the reference's OWN outputs, with no oracle in between: hash-grid corner indices
(north_star: "hash indices ... bit-exact") and level tables of the shipped grids from
tests/golden/ref_host_vectors.npz - produced by the reference's pos_fract + grid_index<CoherentPrime>
+ grid_scale + grid_resolution compiled for the host (oracle/ref_host.cu, tests/golden/make_ref_vectors.py)."""
import os
import sys

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ref_pin_cases as rp  # noqa: E402


@pytest.fixture(scope="module")
def golden():
    return dict(np.load(os.path.join(ROOT, "tests", "golden", "ref_host_vectors.npz")))


@pytest.mark.parametrize("config", sorted(rp.GRID_CONFIGS))
def test_device_hash_indices_equal_the_reference(ngp, golden, config):
    n_dims, enc = rp.GRID_CONFIGS[config]
    e = ngp.Encoding(n_dims, enc)
    x = rp.grid_points(n_dims)
    got = e.indices(torch.from_numpy(x).cuda()).cpu().numpy().view(np.uint32).reshape(len(x), -1, 2 ** n_dims)
    want = golden["grid/" + config][:, :, :2 ** n_dims]
    assert got.shape == want.shape
    bad = np.argwhere(got != want)
    assert len(bad) == 0, (config, len(bad), bad[:4].tolist())
    # the fractional positions the interpolation weights are made of travel with the vectors too
    assert want.max() > 1000


def test_device_level_table_equals_the_reference(ngp, golden):
    rows, _ = rp.cases()["grid_level"]
    want = golden["grid_level"]
    level, pls, base = rows[:, 0], rows[:, 1].copy().view(np.float32), rows[:, 2]
    seen = set()
    for i in range(len(rows)):
        key = (float(pls[i]), int(base[i]))
        if key in seen and i % 16:
            continue
        seen.add(key)
        e = ngp.Encoding(3, {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19, "base_resolution": int(base[i]), "per_level_scale": float(pls[i])})
        offs, scales, res = e.level_table()
        l = int(level[i])
        # the resolution (an integer) must be the reference's; the scale is evaluated with the DEVICE's exp2f, as the
        # reference's kernel does on its GPU (grid.h:97), and may sit one ulp from the host libm value of the vectors
        assert int(res[l]) == int(want[i, 1]), (i, l, key)
        got_bits, want_bits = int(np.asarray(scales, dtype=np.float32).view(np.uint32)[l]), int(want[i, 0])
        assert abs(got_bits - want_bits) <= 1, (i, l, key, got_bits, want_bits)
    assert len(seen) >= 10
