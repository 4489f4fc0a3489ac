"""Python mirror of tcnn's module API over the C ABI.

Shapes and names follow tcnn's PyTorch binding (tcnn/bindings/torch/tinycudann/modules.py:
`Encoding`, `Network`, `NetworkWithInputEncoding`; fields n_input_dims / n_output_dims /
params / loss_scale / seed) and the C++ `tcnn::cpp::Module` underneath it
(tcnn/include/tiny-cuda-nn/cpp_api.h:86-115). torch is used only to own device memory and
streams; every computation happens in libngp_b200.so.
"""
import ctypes
import json

import torch

from . import _lib
from ._lib import lib, check

PRECISION_FP32 = 0
PRECISION_FP16 = 1


def _stream_ptr():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return ctypes.c_void_p(0) if t is None else ctypes.c_void_p(t.data_ptr())


def batch_size_granularity():
    return int(lib.ngpb_batch_size_granularity())


def default_loss_scale(precision=PRECISION_FP16):
    return float(lib.ngpb_default_loss_scale(precision))


class Context:
    """Forward context handle (tcnn::cpp::Context)."""

    def __init__(self, handle):
        self._h = handle

    def close(self):
        if self._h:
            lib.ngpb_context_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Module:
    """Owns a `ngpb_module_t*`. Parameters live in `self.params` (fp16, device) with an fp32
    master copy in `self.params_fp32`, as tcnn's Trainer keeps them (trainer.h:322-335)."""

    def __init__(self, handle, seed=1337, device="cuda"):
        self._h = handle
        self.seed = seed
        self.n_input_dims = int(lib.ngpb_module_n_input_dims(handle))
        self.n_output_dims = int(lib.ngpb_module_n_output_dims(handle))
        self.n_output_dims_unpadded = int(lib.ngpb_module_n_output_dims_unpadded(handle))
        self.n_params = int(lib.ngpb_module_n_params(handle))
        self.loss_scale = default_loss_scale(lib.ngpb_module_param_precision(handle))
        self.params = None
        self.params_fp32 = None
        self._device = device

    # -- introspection -------------------------------------------------------------------
    def hyperparams(self):
        n = lib.ngpb_module_hyperparams(self._h, None, 0)
        buf = ctypes.create_string_buffer(n)
        lib.ngpb_module_hyperparams(self._h, buf, n)
        return json.loads(buf.value.decode())

    def name(self):
        n = lib.ngpb_module_name(self._h, None, 0)
        buf = ctypes.create_string_buffer(n)
        lib.ngpb_module_name(self._h, buf, n)
        return buf.value.decode()

    def layer_sizes(self):
        cap = 64
        arr = (ctypes.c_uint32 * (2 * cap))()
        n = lib.ngpb_module_layer_sizes(self._h, arr, cap)
        return [(arr[2 * i], arr[2 * i + 1]) for i in range(min(n, cap))]

    # -- parameters ----------------------------------------------------------------------
    def initialize_params(self, seed=None, scale=1.0):
        seed = self.seed if seed is None else seed
        self.params_fp32 = torch.empty(self.n_params, dtype=torch.float32, device=self._device)
        check(lib.ngpb_module_initialize_params(self._h, _stream_ptr(), seed, _ptr(self.params_fp32), scale))
        self.params = self.params_fp32.to(torch.float16)
        return self.params

    def set_params(self, params_fp16):
        assert params_fp16.dtype == torch.float16 and params_fp16.numel() == self.n_params
        self.params = params_fp16.contiguous()

    # -- execution -----------------------------------------------------------------------
    def _check_input(self, x):
        if x.dtype != torch.float32 or not x.is_contiguous() or x.dim() != 2 or x.shape[1] != self.n_input_dims:
            raise ValueError(f"input must be contiguous fp32 [n, {self.n_input_dims}]")

    def inference(self, x, params=None):
        self._check_input(x)
        params = self.params if params is None else params
        out = torch.empty((x.shape[0], self.n_output_dims), dtype=torch.float16, device=x.device)
        check(lib.ngpb_module_inference(self._h, _stream_ptr(), x.shape[0], _ptr(x), _ptr(out), _ptr(params)))
        return out

    def forward(self, x, params=None, prepare_input_gradients=False):
        self._check_input(x)
        params = self.params if params is None else params
        out = torch.empty((x.shape[0], self.n_output_dims), dtype=torch.float16, device=x.device)
        ctx = ctypes.c_void_p()
        check(lib.ngpb_module_forward(self._h, _stream_ptr(), x.shape[0], _ptr(x), _ptr(out), _ptr(params),
                                      int(prepare_input_gradients), ctypes.byref(ctx)))
        return out, Context(ctx)

    def backward(self, ctx, x, out, dL_dout, params=None, want_input_grad=False, want_param_grad=True):
        params = self.params if params is None else params
        assert dL_dout.dtype == torch.float16 and dL_dout.is_contiguous()
        dL_dx = torch.empty_like(x) if want_input_grad else None
        dL_dp = torch.empty(self.n_params, dtype=torch.float16, device=x.device) if want_param_grad else None
        check(lib.ngpb_module_backward(self._h, _stream_ptr(), ctx._h, x.shape[0], _ptr(dL_dx), _ptr(dL_dout),
                                       _ptr(dL_dp), _ptr(x), _ptr(out), _ptr(params)))
        return dL_dx, dL_dp

    def close(self):
        if self._h:
            lib.ngpb_module_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Encoding(Module):
    """tcnn.Encoding(n_input_dims, encoding_config) (tinycudann/modules.py)."""

    def __init__(self, n_input_dims, encoding_config, seed=1337, device="cuda"):
        h = ctypes.c_void_p()
        check(lib.ngpb_create_encoding(n_input_dims, json.dumps(encoding_config).encode(), PRECISION_FP16, ctypes.byref(h)))
        super().__init__(h, seed, device)
        self.encoding_config = encoding_config

    # grid-only helpers used by the parity tests
    def level_table(self):
        L = 64
        offs = (ctypes.c_uint32 * (L + 1))()
        scales = (ctypes.c_float * L)()
        res = (ctypes.c_uint32 * L)()
        n = lib.ngpb_grid_level_table(self._h, offs, scales, res)
        return list(offs[: n + 1]), list(scales[:n]), list(res[:n])

    def indices(self, x):
        self._check_input(x)
        L = len(self.level_table()[1])
        idx = torch.empty((x.shape[0], L, 2 ** self.n_input_dims), dtype=torch.int32, device=x.device)
        check(lib.ngpb_grid_indices(self._h, _stream_ptr(), x.shape[0], _ptr(x), _ptr(idx)))
        return idx

    def set_max_level(self, v):
        check(lib.ngpb_grid_set_max_level(self._h, float(v)))


class NetworkWithInputEncoding(Module):
    """tcnn.NetworkWithInputEncoding(n_input_dims, n_output_dims, encoding_config, network_config)."""

    def __init__(self, n_input_dims, n_output_dims, encoding_config, network_config, seed=1337, device="cuda"):
        h = ctypes.c_void_p()
        check(lib.ngpb_create_network_with_input_encoding(
            n_input_dims, n_output_dims, json.dumps(encoding_config).encode(), json.dumps(network_config).encode(), ctypes.byref(h)))
        super().__init__(h, seed, device)


class Network(Module):
    """tcnn.Network(n_input_dims, n_output_dims, network_config)."""

    def __init__(self, n_input_dims, n_output_dims, network_config, seed=1337, device="cuda"):
        h = ctypes.c_void_p()
        check(lib.ngpb_create_network(n_input_dims, n_output_dims, json.dumps(network_config).encode(), ctypes.byref(h)))
        super().__init__(h, seed, device)


class NerfNetwork(Module):
    """ngp::NerfNetwork (include/neural-graphics-primitives/nerf_network.h:79-101)."""

    def __init__(self, pos_encoding, dir_encoding, density_network, rgb_network,
                 n_pos_dims=3, n_dir_dims=3, n_extra_dims=0, dir_offset=4, seed=1337, device="cuda"):
        h = ctypes.c_void_p()
        check(lib.ngpb_create_nerf_network(
            n_pos_dims, n_dir_dims, n_extra_dims, dir_offset,
            json.dumps(pos_encoding).encode(), json.dumps(dir_encoding).encode(),
            json.dumps(density_network).encode(), json.dumps(rgb_network).encode(), ctypes.byref(h)))
        super().__init__(h, seed, device)
