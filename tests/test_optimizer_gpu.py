"""Parity of the fused Adam + ExponentialDecay + EMA step (ngpb_optimizer_step) with a numpy
restatement of tcnn adam.h:72-118, exponential_decay.h:60-71 and ema.h:46-78,102-103.
fp32 state must agree to a few ulp (the kernel evaluates beta^step with ex2.approx); fp16 weights
and the EMA weights are compared as the rounded values of the fp32 results."""
import ctypes
import json

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

CONFIG = {"otype": "Ema", "decay": 0.95, "nested": {"otype": "ExponentialDecay", "decay_start": 4, "decay_interval": 2, "decay_base": 0.33,
          "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}}}


class NumpyAdam:
    def __init__(self, n, n_matrix):
        self.m = np.zeros(n, np.float32)
        self.v = np.zeros(n, np.float32)
        self.steps = np.zeros(n, np.uint32)
        self.ema = np.zeros(n, np.float16)
        self.n_matrix = n_matrix
        self.step_count = 0
        self.base_lr = np.float32(1e-2)

    def step(self, loss_scale, w32, w16, g16):
        f = np.float32
        # ExponentialDecay (exponential_decay.h:59-71): on the nested step count before this step
        if self.step_count >= 4 and (self.step_count - 4) % 2 == 0:
            self.base_lr = f(self.base_lr * f(0.33))
        lr = self.base_lr
        g = g16.astype(np.float32) / f(loss_scale)
        is_matrix = np.arange(len(w32)) < self.n_matrix
        skip = np.where(is_matrix, False, g == 0)
        g = np.where(is_matrix, g + f(1e-6) * w32, g)
        upd = ~skip
        self.m[upd] = f(0.9) * self.m[upd] + f(1 - f(0.9)) * g[upd]
        self.v[upd] = f(0.99) * self.v[upd] + f(1 - f(0.99)) * g[upd] * g[upd]
        self.steps[upd] += 1
        st = self.steps[upd].astype(np.float64)
        scale = np.sqrt(1 - 0.99 ** st) / (1 - 0.9 ** st)
        lr_p = (np.float64(lr) * scale).astype(np.float32)
        eff = lr_p / (np.sqrt(self.v[upd]) + f(1e-15))
        w32[upd] = w32[upd] - eff * self.m[upd]
        w16[upd] = w32[upd].astype(np.float16)
        # Ema (ema.h:63-78): debias with the step counts before / after this step
        self.step_count += 1
        old = f(1 - 0.95 ** (self.step_count - 1))
        new = f(1.0 / (1 - 0.95 ** self.step_count))
        filt = (self.ema.astype(np.float32) * f(0.95) * old + w16.astype(np.float32) * f(1 - f(0.95))) * new
        self.ema = filt.astype(np.float16)


def test_adam_ema_decay_matches_numpy(ngp):
    lib, check = ngp.lib, ngp._lib.check
    n_matrix, n = 64 * 32 + 16 * 64, 64 * 32 + 16 * 64 + 40000
    h = ctypes.c_void_p()
    check(lib.ngpb_optimizer_create(json.dumps(CONFIG).encode(), ctypes.byref(h)))
    layers = (ctypes.c_uint32 * 4)(64, 32, 16, 64)
    check(lib.ngpb_optimizer_allocate(h, n, layers, 2))
    rng = np.random.default_rng(0)
    w32 = rng.normal(0, 0.1, n).astype(np.float32)
    w16 = w32.astype(np.float16)
    ref = NumpyAdam(n, n_matrix)
    d32, d16 = torch.from_numpy(w32.copy()).cuda(), torch.from_numpy(w16.copy()).cuda()
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream or 1)
    for step in range(9):
        g = (rng.normal(0, 1.0, n) * 128 * 1e-2).astype(np.float16)
        g[n_matrix:][rng.random(n - n_matrix) < 0.7] = 0  # most hash-grid entries see no gradient in a step
        dg = torch.from_numpy(g).cuda()
        prev32 = d32.cpu().numpy()
        check(lib.ngpb_optimizer_step(h, stream, 128.0, ctypes.c_void_p(d32.data_ptr()), ctypes.c_void_p(d16.data_ptr()), ctypes.c_void_p(dg.data_ptr())))
        torch.cuda.synchronize()
        ref.step(128.0, w32, w16, g)
        assert abs(lib.ngpb_optimizer_learning_rate(h) - float(ref.base_lr)) <= 1e-6 * float(ref.base_lr), step
        got32 = d32.cpu().numpy()
        np.testing.assert_allclose(got32, w32, rtol=3e-6, atol=2e-7, err_msg=f"step {step}")  # an update is ~1e-2: 2e-7 is 2e-5 of it
        got16 = d16.cpu().numpy()
        assert (np.abs(got16.astype(np.float32) - w16.astype(np.float32)) <= np.abs(w16.astype(np.float32)) * 2.0 ** -10 + 1e-7).all()
        # untouched hash entries keep their weights bit for bit
        untouched = np.concatenate([np.zeros(n_matrix, bool), g[n_matrix:] == 0])
        assert np.array_equal(got32[untouched].view(np.uint32), prev32[untouched].view(np.uint32))
    assert lib.ngpb_optimizer_step_count(h) == 9
    ema_ptr = lib.ngpb_optimizer_custom_weights(h)
    assert ema_ptr
    # read the EMA weights through a torch view of the raw pointer
    class _I:
        __cuda_array_interface__ = {"shape": (n,), "typestr": "<f2", "data": (ema_ptr, False), "version": 2}
    got_ema = torch.as_tensor(_I(), device="cuda").cpu().numpy().astype(np.float32)
    want = ref.ema.astype(np.float32)
    assert (np.abs(got_ema - want) <= np.abs(want) * 2.0 ** -9 + 1e-6).all()
    m1, m2, st = ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_void_p()
    check(lib.ngpb_optimizer_state(h, ctypes.byref(m1), ctypes.byref(m2), ctypes.byref(st)))
    class _S:
        __cuda_array_interface__ = {"shape": (n,), "typestr": "<u4", "data": (st.value, False), "version": 2}
    assert np.array_equal(torch.as_tensor(_S(), device="cuda").cpu().numpy().astype(np.uint32), ref.steps)
    lib.ngpb_optimizer_free(h)
