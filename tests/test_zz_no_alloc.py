"""The allocation-free entry point qdc_contract_norm_static (requested by the run's operators; no
counterpart in the reference): called through the C-ABI, its result equals qdc_contract_norm's on
two calls in a row -- the second with other operands and a narrower kdim, so stale contents of
the shared static scratch would show -- and a kdim beyond the static maximum is rejected with a
clear error.  Collected last on purpose."""
import importlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _norms(lib, _lib, torch, dev, rows, cols, kdim, scale, seed):
    rng = np.random.default_rng(seed)
    A = rng.standard_normal((rows, kdim))
    Bt = rng.standard_normal((cols, kdim))
    A_d, B_d = torch.as_tensor(A, device=dev), torch.as_tensor(Bt, device=dev)
    ws = torch.empty(lib.qdc_norm_workspace_doubles(kdim), dtype=torch.float64, device=dev)
    want = torch.full((1,), float("nan"), dtype=torch.float64, device=dev)
    got = torch.full((1,), float("nan"), dtype=torch.float64, device=dev)
    sp = _lib.stream_ptr()
    _lib.check(lib.qdc_contract_norm(_lib.ptr(A_d), rows, _lib.ptr(B_d), cols, kdim, scale, _lib.ptr(ws),
                                     _lib.ptr(want), sp), "qdc_contract_norm")
    _lib.check(lib.qdc_contract_norm_static(_lib.ptr(A_d), rows, _lib.ptr(B_d), cols, kdim, scale,
                                            _lib.ptr(got), sp), "qdc_contract_norm_static")
    torch.cuda.synchronize()
    return float(got), float(want), scale * float(A.sum(0) @ Bt.sum(0))


def test_norm_static_matches_norm_on_two_calls_in_a_row():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    importlib.import_module("dqva-and-circuit-cutting_b200")
    from qdc_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda:0")
    assert lib.qdc_contract_norm_static_max_kdim() == 256
    with torch.cuda.device(dev):
        # first call: the widest kdim the static scratch takes; second: other data, narrower and odd
        for rows, cols, kdim, scale, seed in ((5000, 3000, 256, 0.0625, 1), (777, 1234, 37, 0.5, 2)):
            got, want, direct = _norms(lib, _lib, torch, dev, rows, cols, kdim, scale, seed)
            assert np.isfinite(got)
            assert abs(got - want) <= 1e-10 * max(1.0, abs(want)), (got, want)    # same kernels, same order
            assert abs(got - direct) <= 1e-9 * max(1.0, abs(direct)), (got, direct)
        # beyond the maximum: a clear error, nothing launched
        A_d = torch.zeros(4, 512, dtype=torch.float64, device=dev)
        out = torch.zeros(1, dtype=torch.float64, device=dev)
        rc = lib.qdc_contract_norm_static(_lib.ptr(A_d), 4, _lib.ptr(A_d), 4, 512, 1.0, _lib.ptr(out),
                                          _lib.stream_ptr())
        assert rc != 0
        msg = lib.qdc_last_error().decode()
        assert "512" in msg and "256" in msg and "static" in msg, msg
