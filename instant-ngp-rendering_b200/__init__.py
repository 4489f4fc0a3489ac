"""ngp-b200: B200 (sm_100a) implementation of instant-ngp's hash-grid / fused-MLP / NeRF
ray-march hot path. The compute lives in libngp_b200.so (C ABI: include/ngp_b200.h);
this package is the Python host mirror of the reference's tcnn / pyngp interfaces.

The directory name contains a hyphen, so import it with
`importlib.import_module("instant-ngp-rendering_b200")` (tests/conftest.py does this and
aliases it as `ngp_b200`).
"""
from . import _lib
from ._lib import NgpbError, has_device, lib, LIB_PATH
from .testbed import Testbed, nerf_matrix_to_ngp, load_network_config
from .modules import (Context, Encoding, Module, NerfNetwork, Network, NetworkWithInputEncoding,
                      batch_size_granularity, default_loss_scale)

__all__ = [
    "NgpbError", "has_device", "Testbed", "nerf_matrix_to_ngp", "load_network_config", "lib", "LIB_PATH", "Context", "Encoding", "Module", "NerfNetwork",
    "Network", "NetworkWithInputEncoding", "batch_size_granularity", "default_loss_scale",
]
