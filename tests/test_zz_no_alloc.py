"""GPU test of `ob200_acq_argmax_static`, the entry point for callers that may not allocate.

Such callers (a service that reserved its whole GPU up front and treats any later cudaMalloc as a
fault, or code running after its framework froze allocation) cannot create an `ob200_ctx` or let one
grow a scratch buffer, so this variant keeps its per-block scratch in a fixed array inside the
library and takes every input and output from the caller.  The test calls it through the C ABI
(ctypes) twice in a row -- a large call followed by a smaller one, so the shared scratch still holds
stale records from the first -- and checks both results against the context-based `ob200_acq_argmax` and,
independently of any device code, against the CPU oracle's acquisition functions and argmax.

Collected last on purpose (file name), so a run that stops at the first failure has already run
every other test; it must pass alone and after the others.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from metaopt_b200 import _lib
from oracle import acquisition as oacq

pytestmark = pytest.mark.gpu

RTOL = 1e-12  # same kernels, same per-candidate arithmetic: values agree to rounding, indices exactly


def _call_static(lib, mu, sd, first_index, mask, y_opt, acq=None, stream=0):
    best = torch.zeros(12, dtype=torch.float64, device=mu.device)  # 3 x ob200_best (value, index, mu, sigma)
    err = C.create_string_buffer(512)
    rc = lib.ob200_acq_argmax_static(C.c_void_p(mu.data_ptr()), C.c_void_p(sd.data_ptr()), mu.numel(), first_index, mask,
                                     y_opt, 0.01, 1.96, C.c_void_p(best.data_ptr()),
                                     C.c_void_p(acq.data_ptr()) if acq is not None else None,
                                     C.c_void_p(stream), err, len(err))
    return rc, err.value.decode(), best


def _unpack(best):
    host = best.cpu()
    idx = host.view(torch.int64)
    return [dict(value=float(host[4 * k]), index=int(idx[4 * k + 1]), mu=float(host[4 * k + 2]),
                 sigma=float(host[4 * k + 3])) for k in range(3)]


def test_static_entry_point_matches_context_entry_point_twice(engine):
    lib = _lib.load()
    torch.cuda.set_device(0)
    rs = np.random.RandomState(11)
    for M, first in ((1_500_003, 7), (40_001, 123_456)):  # second call leaves most of the scratch stale
        mu = torch.as_tensor(rs.randn(M)).cuda()
        sd = torch.as_tensor(np.abs(rs.randn(M))).cuda()
        sd[::11] = 0.0
        y_opt = -0.25
        ref = engine.acq_argmax(mu, sd, first_index=first, acq_mask=7, y_opt=y_opt, dense=True)
        acq = torch.zeros((3, M), dtype=torch.float64, device="cuda")
        rc, msg, best = _call_static(lib, mu, sd, first, 7, y_opt, acq=acq,
                                     stream=torch.cuda.current_stream().cuda_stream)
        assert rc == 0, msg
        torch.cuda.synchronize()
        got = _unpack(best)
        for k, name in enumerate(("EI", "PI", "LCB")):
            assert got[k]["index"] == ref[name]["index"]
            np.testing.assert_allclose(got[k]["value"], ref[name]["value"], rtol=RTOL)
            np.testing.assert_allclose(got[k]["mu"], ref[name]["mu"], rtol=RTOL)
            np.testing.assert_allclose(got[k]["sigma"], ref[name]["sigma"], rtol=RTOL)
        np.testing.assert_allclose(acq.cpu().numpy(), ref["acq"].cpu().numpy(), rtol=RTOL, atol=0)
        # ... and against the oracle (restated skopt acquisitions + first-occurrence argmax), no device code involved
        mu_h, sd_h = mu.cpu().numpy(), sd.cpu().numpy()
        for k in range(3):
            util = oacq.utility(k, mu_h, sd_h, y_opt, 0.01, 1.96)
            assert got[k]["index"] == first + oacq.pick(k, mu_h, sd_h, y_opt, 0.01, 1.96)
            np.testing.assert_allclose(acq[k].cpu().numpy(), util, rtol=1e-9, atol=1e-300)
            np.testing.assert_allclose(got[k]["value"], util[got[k]["index"] - first], rtol=1e-9)


def test_static_entry_point_rejects_oversize_and_bad_arguments():
    lib = _lib.load()
    torch.cuda.set_device(0)
    tiny = torch.zeros(8, dtype=torch.float64, device="cuda")
    best = torch.zeros(12, dtype=torch.float64, device="cuda")
    err = C.create_string_buffer(512)
    too_many = (1 << 24) + 1  # validated before any memory is touched
    rc = lib.ob200_acq_argmax_static(C.c_void_p(tiny.data_ptr()), C.c_void_p(tiny.data_ptr()), too_many, 0, 1, 0.0, 0.01,
                                     1.96, C.c_void_p(best.data_ptr()), None, None, err, len(err))
    assert rc == _lib.EINVAL
    assert "16777216" in err.value.decode() and "exceeds" in err.value.decode()
    rc = lib.ob200_acq_argmax_static(C.c_void_p(tiny.data_ptr()), C.c_void_p(tiny.data_ptr()), 8, 0, 0, 0.0, 0.01, 1.96,
                                     C.c_void_p(best.data_ptr()), None, None, err, len(err))
    assert rc == _lib.EINVAL and "acq_mask" in err.value.decode()
    rc = lib.ob200_acq_argmax_static(None, C.c_void_p(tiny.data_ptr()), 8, 0, 1, 0.0, 0.01, 1.96,
                                     C.c_void_p(best.data_ptr()), None, None, None, 0)
    assert rc == _lib.EINVAL
