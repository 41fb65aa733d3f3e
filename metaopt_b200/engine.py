"""Device engine: thin Python wrapper over the C ABI (include/ob200.h) using torch tensors for
device memory and streams.  One engine per CUDA device, cached at module level (the algorithm
object is deep-copied by Orion on every update, `src/orion/core/worker/producer.py:161`, so no
device handle may live inside it)."""
import atexit
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import Best, Kernel, Slot

_ENGINES = {}


def get_engine(device=None):
    """Process-wide engine for a CUDA device index (default: torch's current device)."""
    if not torch.cuda.is_available():
        raise RuntimeError("metaopt_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    if device is None:
        device = torch.cuda.current_device()
    device = int(device)
    if device not in _ENGINES:
        _ENGINES[device] = Engine(device)
    return _ENGINES[device]


@atexit.register
def _shutdown():
    for eng in list(_ENGINES.values()):
        eng.close()
    _ENGINES.clear()


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


class Engine:
    def __init__(self, device):
        self.lib = _lib.load()
        self.device = int(device)
        self.tdev = torch.device("cuda", self.device)
        ctx = C.c_void_p()
        rc = self.lib.ob200_ctx_create(self.device, C.byref(ctx))
        if rc != 0:
            raise _lib.Ob200Error(rc, "ob200_ctx_create(device=%d) failed (needs an sm_100a GPU)" % self.device)
        self.ctx = ctx
        self.d_out = 0
        self.n_slots = 0
        self.n = self.D = 0
        self._fit_host = None   # (kernel key, X) of the last fit from host arrays: lets fit() append one row
        self.n_appends = 0
        # Cache keys of whoever drives this engine (algo.py, compiler.upload).  They live HERE, on the object that
        # owns the device state, and every call that replaces that state clears them, so two holders of the same
        # device (a `device=[0, 1]` and a `device=0` algorithm, or a test calling fit() directly) cannot score
        # against each other's GP or space.  `fit_generation` counts replacements (MultiEngine replicates on it).
        self._fit_token = None
        self._fit_theta = None
        self._space_key = None
        self.fit_generation = 0
        self.use_torch_stream()

    def close(self):
        if self.ctx is not None:
            self.lib.ob200_ctx_destroy(self.ctx)
            self.ctx = None

    def _check(self, rc):
        _lib.check(self.ctx, rc)

    def use_torch_stream(self):
        """Enqueue on torch's current stream of this device, so torch ops and ours are ordered."""
        s = torch.cuda.current_stream(self.tdev).cuda_stream
        self._check(self.lib.ob200_set_stream(self.ctx, C.c_void_p(s)))

    def to_device(self, array):
        """numpy -> float64 tensor on this engine's device (kept resident across repeated fits)."""
        return torch.as_tensor(np.ascontiguousarray(array, dtype=np.float64)).to(self.tdev)

    def sync(self):
        self._check(self.lib.ob200_sync(self.ctx))

    @property
    def launches(self):
        return int(self.lib.ob200_launch_count(self.ctx))

    # ---- space ---------------------------------------------------------------------------
    def set_space(self, slots, cdf):
        """slots: list of dicts with the fields of `ob200_slot`; cdf: flat list of floats."""
        arr = (Slot * len(slots))()
        for i, s in enumerate(slots):
            for k, _ in Slot._fields_:
                setattr(arr[i], k, s[k])
        cdf_arr = (C.c_double * max(len(cdf), 1))(*cdf)
        self._space_key = None
        self._check(self.lib.ob200_space_set(self.ctx, arr, len(slots), cdf_arr, len(cdf)))
        self.n_slots = len(slots)
        self.d_out = max(s["out_col"] + s["width"] for s in slots)

    def sample(self, seed, first_index, count, normalise=True, variates=None, out=None, check=True):
        """[count x D_out] float64 device tensor of sampled+transformed(+normalised) points."""
        self.use_torch_stream()
        if out is None:
            out = torch.empty((count, self.d_out), dtype=torch.float64, device=self.tdev)
        assert out.dtype == torch.float64 and out.is_cuda and out.stride(1) == 1
        var = None
        if variates is not None:
            var = torch.as_tensor(variates, dtype=torch.float64).to(self.tdev).contiguous()
            assert var.shape == (count, self.n_slots)
        self._check(self.lib.ob200_sample_transform(self.ctx, seed, first_index, count, int(bool(normalise)),
                                                    _ptr(out), out.stride(0), _ptr(var)))
        if check:
            self._check(self.lib.ob200_sample_status(self.ctx))
        return out

    def sample_host(self, seed, first_index, count, normalise=True, variates=None):
        out = np.empty((count, self.d_out), dtype=np.float64)
        var = None
        if variates is not None:
            var = np.ascontiguousarray(variates, dtype=np.float64)
            assert var.shape == (count, self.n_slots)
        self._check(self.lib.ob200_sample_transform_host(
            self.ctx, seed, first_index, count, int(bool(normalise)), out.ctypes.data_as(C.c_void_p), self.d_out,
            var.ctypes.data_as(C.c_void_p) if var is not None else None))
        self._check(self.lib.ob200_sample_status(self.ctx))
        return out

    # ---- fit -------------------------------------------------------------------------------
    def fit(self, X, y, kind=_lib.KERNEL_MATERN52, amplitude=1.0, length_scale=1.0, alpha=1e-10, normalize_y=False,
            incremental=True):
        """Fit the GP.  With host arrays and `incremental`, a call whose X is the previous call's X plus ONE row, with
        the same kernel / diagonal term, is served by the O(n^2) append (`ob200_gp_append`) instead of a refit."""
        self.use_torch_stream()
        ls = np.atleast_1d(np.asarray(length_scale, dtype=np.float64)).copy()
        k = Kernel(kind, len(ls), float(amplitude), ls.ctypes.data_as(C.POINTER(C.c_double)))
        key = None
        prev, self._fit_host = self._fit_host, None
        self._fit_token = self._fit_theta = None
        self.fit_generation += 1
        if isinstance(X, torch.Tensor):
            X = X.to(self.tdev, torch.float64).contiguous()
            y = torch.as_tensor(y).to(self.tdev, torch.float64).contiguous()
            n, D = X.shape
            rc = self.lib.ob200_gp_fit(self.ctx, _ptr(X), _ptr(y), n, D, C.byref(k), float(alpha), int(bool(normalize_y)))
        else:
            X = np.ascontiguousarray(X, dtype=np.float64)
            y = np.ascontiguousarray(y, dtype=np.float64)
            n, D = X.shape
            key = (int(kind), float(amplitude), ls.tobytes(), float(alpha), D)
            if (incremental and prev is not None and prev[0] == key and prev[1].shape[0] == n - 1
                    and np.array_equal(prev[1], X[:n - 1])):
                xy = self.to_device(np.concatenate([X[n - 1], y]))
                rc = self.lib.ob200_gp_append(self.ctx, _ptr(xy), _ptr(xy[D:]), n, int(bool(normalize_y)))
                if rc == _lib.ESTATE:  # no free row in the padded factor: refit
                    rc = None
                else:
                    self.n_appends += 1
            else:
                rc = None
            if rc is None:
                rc = self.lib.ob200_gp_fit_host(self.ctx, X.ctypes.data_as(C.c_void_p), y.ctypes.data_as(C.c_void_p), n, D,
                                                C.byref(k), float(alpha), int(bool(normalize_y)))
        self._check(rc)
        self.n, self.D = n, D
        if key is not None:
            self._fit_host = (key, X.copy())

    def fit_search(self, X, y, kind=_lib.KERNEL_MATERN52, amplitude=1.0, length_scale=1.0, alpha=1e-10, normalize_y=False):
        """Fit for ONE log-marginal-likelihood evaluation of a hyper-parameter search (device tensors): `lml()` /
        `lml_grad()` work afterwards, scoring does not until the next `fit`."""
        self.use_torch_stream()
        ls = np.atleast_1d(np.asarray(length_scale, dtype=np.float64)).copy()
        k = Kernel(kind, len(ls), float(amplitude), ls.ctypes.data_as(C.POINTER(C.c_double)))
        self._fit_host = None
        self._fit_token = self._fit_theta = None
        self.fit_generation += 1
        X = X.to(self.tdev, torch.float64).contiguous()
        y = torch.as_tensor(y).to(self.tdev, torch.float64).contiguous()
        n, D = X.shape
        self._check(self.lib.ob200_gp_fit_search(self.ctx, _ptr(X), _ptr(y), n, D, C.byref(k), float(alpha), int(bool(normalize_y))))
        self.n, self.D = n, D

    def ystats(self):
        a, b, c = C.c_double(), C.c_double(), C.c_double()
        self._check(self.lib.ob200_gp_get_ystats(self.ctx, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def lml(self):
        v = C.c_double()
        self._check(self.lib.ob200_gp_lml(self.ctx, C.byref(v)))
        return v.value

    def lml_grad(self):
        """(lml, grad) of the current fit; grad = [dlog(amplitude), dlog(ls_0..D-1), d(diagonal)]."""
        self.use_torch_stream()
        v = C.c_double()
        g = (C.c_double * (self.D + 2))()
        self._check(self.lib.ob200_gp_lml_grad(self.ctx, C.byref(v), g, self.D + 2))
        return v.value, np.asarray(list(g))

    def predict_grad(self, x):
        """(mu [m], sigma [m], dmu [m, D], dsigma [m, D]) at the normalised rows `x` [m, D] (host arrays)."""
        x = np.ascontiguousarray(np.atleast_2d(np.asarray(x, dtype=np.float64)))
        m, d = x.shape
        if d != self.D:
            raise ValueError("x has %d columns, the fit has D=%d" % (d, self.D))
        self.use_torch_stream()
        out = np.empty((m, 2 + 2 * d), dtype=np.float64)
        self._check(self.lib.ob200_gp_predict_grad_host(self.ctx, x.ctypes.data_as(C.c_void_p), d, m,
                                                        out.ctypes.data_as(C.c_void_p)))
        return out[:, 0], out[:, 1], out[:, 2:2 + d], out[:, 2 + d:]

    def get_factor(self):
        L = np.empty((self.n, self.n), dtype=np.float64)
        al = np.empty(self.n, dtype=np.float64)
        self._check(self.lib.ob200_gp_get_factor(self.ctx, L.ctypes.data_as(C.c_void_p), al.ctypes.data_as(C.c_void_p)))
        return L, al

    def export_state(self):
        """Opaque fit state as a uint8 device tensor (to be broadcast to the other ranks)."""
        self.use_torch_stream()
        nbytes = C.c_int64()
        self._check(self.lib.ob200_gp_state_bytes(self.ctx, self.n, self.D, C.byref(nbytes)))
        blob = torch.empty(nbytes.value, dtype=torch.uint8, device=self.tdev)
        self._check(self.lib.ob200_gp_export(self.ctx, _ptr(blob), nbytes.value))
        return blob

    def state_bytes(self, n, D):
        nbytes = C.c_int64()
        self._check(self.lib.ob200_gp_state_bytes(self.ctx, n, D, C.byref(nbytes)))
        return nbytes.value

    def import_state(self, blob, n, D):
        self.use_torch_stream()
        assert blob.dtype == torch.uint8 and blob.is_cuda and blob.is_contiguous()
        self._fit_token = self._fit_theta = None
        self.fit_generation += 1
        self._check(self.lib.ob200_gp_import(self.ctx, _ptr(blob), blob.numel()))
        self.n, self.D = n, D
        self._fit_host = None

    # ---- score -----------------------------------------------------------------------------
    @staticmethod
    def _bests(arr, mask):
        names = {0: "EI", 1: "PI", 2: "LCB"}
        return {names[k]: dict(value=arr[k].value, index=int(arr[k].index), mu=arr[k].mu, sigma=arr[k].sigma)
                for k in range(3) if mask & (1 << k)}

    def score(self, cand=None, M=None, seed=0, first_index=0, acq_mask=1, y_opt=0.0, xi=0.01, kappa=1.96, dense=False):
        """Score candidates on the device.  cand: [M x >=D] float64 CUDA tensor (normalised), or None
        to generate candidate i from (seed, first_index+i) with the registered space."""
        self.use_torch_stream()
        ldc = 0
        if cand is not None:
            assert cand.dtype == torch.float64 and cand.is_cuda and cand.stride(1) == 1
            M, ldc = cand.shape[0], cand.stride(0)
        mu = sg = aq = None
        if dense:
            mu = torch.empty(M, dtype=torch.float64, device=self.tdev)
            sg = torch.empty(M, dtype=torch.float64, device=self.tdev)
            aq = torch.zeros((3, M), dtype=torch.float64, device=self.tdev)
        best = (Best * 3)()
        self._check(self.lib.ob200_score(self.ctx, _ptr(cand), ldc, M, seed, first_index, acq_mask, y_opt, xi, kappa,
                                         best, _ptr(mu), _ptr(sg), _ptr(aq)))
        res = self._bests(best, acq_mask)
        if dense:
            res.update(mu=mu, sigma=sg, acq=aq)
        return res

    def score_pruned(self, cand=None, M=None, seed=0, first_index=0, acq="EI", y_opt=0.0, xi=0.01, kappa=1.96):
        """Exact pruned scoring for one acquisition: same winner as `score`, computed faster by skipping the
        variance of candidates whose mean-only upper bound cannot reach the incumbent.  Returns
        ({acq: record}, n_scored)."""
        self.use_torch_stream()
        k = {"EI": 0, "PI": 1, "LCB": 2}[acq]
        ldc = 0
        if cand is not None:
            assert cand.dtype == torch.float64 and cand.is_cuda and cand.stride(1) == 1
            M, ldc = cand.shape[0], cand.stride(0)
        best = (Best * 1)()
        n_scored = C.c_int64()
        self._check(self.lib.ob200_score_pruned(self.ctx, _ptr(cand), ldc, M, seed, first_index, k, y_opt, xi, kappa,
                                                best, C.byref(n_scored)))
        rec = dict(value=best[0].value, index=int(best[0].index), mu=best[0].mu, sigma=best[0].sigma)
        return {acq: rec}, int(n_scored.value)

    def score_pruned_multi(self, cand=None, M=None, seed=0, first_index=0, acq_mask=7, y_opt=0.0, xi=0.01, kappa=1.96):
        """`score_pruned` for several acquisitions with ONE shared mean-only pass (the gp_hedge portfolio).
        Returns ({acq: record}, {acq: n_scored})."""
        self.use_torch_stream()
        ldc = 0
        if cand is not None:
            assert cand.dtype == torch.float64 and cand.is_cuda and cand.stride(1) == 1
            M, ldc = cand.shape[0], cand.stride(0)
        best = (Best * 3)()
        n_scored = (C.c_int64 * 3)()
        self._check(self.lib.ob200_score_pruned_multi(self.ctx, _ptr(cand), ldc, M, seed, first_index, acq_mask, y_opt, xi,
                                                      kappa, best, n_scored))
        names = {0: "EI", 1: "PI", 2: "LCB"}
        return self._bests(best, acq_mask), {names[k]: int(n_scored[k]) for k in range(3) if acq_mask & (1 << k)}

    def score_bounds(self, cand):
        """(mu_lb [M], sigma_ub [M], used_f32): pass 1 of the pruned path for the CUDA tensor `cand` [M x >=D]."""
        self.use_torch_stream()
        assert cand.dtype == torch.float64 and cand.is_cuda and cand.stride(1) == 1
        M = cand.shape[0]
        mu = torch.empty(M, dtype=torch.float64, device=self.tdev)
        sg = torch.empty(M, dtype=torch.float64, device=self.tdev)
        used = C.c_int()
        self._check(self.lib.ob200_score_bounds(self.ctx, _ptr(cand), cand.stride(0), M, _ptr(mu), _ptr(sg), C.byref(used)))
        return mu, sg, bool(used.value)

    def prune_stats(self):
        """Counters of the last pruned call: dict(f32, batch_a, batch_b, batch_b_refined, batch_b_subset)."""
        out = (C.c_int64 * 5)()
        self._check(self.lib.ob200_prune_stats(self.ctx, out))
        return dict(f32=bool(out[0]), batch_a=int(out[1]), batch_b=int(out[2]), batch_b_refined=int(out[3]),
                    batch_b_subset=int(out[4]))

    def score_host(self, cand, acq_mask=1, y_opt=0.0, xi=0.01, kappa=1.96, dense=False):
        """Same through host buffers (numpy in, numpy out): H2D/D2H copies happen inside the call."""
        self.use_torch_stream()
        cand = np.asarray(cand, dtype=np.float64)
        assert cand.ndim == 2 and cand.strides[1] == 8
        M, ldc = cand.shape[0], cand.strides[0] // 8
        mu = sg = aq = None
        if dense:
            mu, sg, aq = np.empty(M), np.empty(M), np.zeros((3, M))
        best = (Best * 3)()
        vp = lambda x: x.ctypes.data_as(C.c_void_p) if x is not None else None
        self._check(self.lib.ob200_score_host(self.ctx, vp(cand), ldc, M, acq_mask, y_opt, xi, kappa, best,
                                              vp(mu), vp(sg), vp(aq)))
        res = self._bests(best, acq_mask)
        if dense:
            res.update(mu=mu, sigma=sg, acq=aq)
        return res

    def acq_argmax(self, mu, sigma, first_index=0, acq_mask=1, y_opt=0.0, xi=0.01, kappa=1.96, dense=False):
        self.use_torch_stream()
        M = mu.numel()
        aq = torch.zeros((3, M), dtype=torch.float64, device=self.tdev) if dense else None
        best = (Best * 3)()
        self._check(self.lib.ob200_acq_argmax(self.ctx, _ptr(mu), _ptr(sigma), M, first_index, acq_mask, y_opt, xi, kappa,
                                              best, _ptr(aq)))
        res = self._bests(best, acq_mask)
        if dense:
            res["acq"] = aq
        return res
