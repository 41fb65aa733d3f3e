"""Single-process multi-GPU engine: the candidate shards of ONE `suggest()` spread over several devices.

Orion's worker is one Python process (`/root/reference/src/orion/core/worker/__init__.py:42-88`), so the plugin
cannot rely on a torchrun launch.  `MultiEngine` owns one `Engine` (one C-ABI context) per device and drives them
from a small thread pool -- ctypes releases the GIL during the calls, so the per-device launches and
synchronisations overlap.  The fit runs on the first device; its state blob (inv(L) tiles, alpha, X/ls, y statistics)
is copied to the other devices over NVLink peer copies when it is first needed for scoring (`ob200_gp_export` ->
`tensor.to(device)` -> `ob200_gp_import`), lazily, so the many fits of a hyper-parameter search stay on one GPU.
Candidates are partitioned by global index exactly like the multi-process path (`dist.shard_range`), the winners
are merged with the same deterministic rule (`dist.merge_best`), so the result is identical for any device count.
"""
from concurrent.futures import ThreadPoolExecutor

from . import dist as obdist
from .engine import get_engine


class MultiEngine:
    def __init__(self, devices):
        self.devices = [int(d) for d in devices]
        if len(set(self.devices)) != len(self.devices) or not self.devices:
            raise ValueError("devices must be a non-empty list of distinct CUDA device indices: %r" % (devices,))
        self.engines = [get_engine(d) for d in self.devices]
        self.pool = ThreadPoolExecutor(max_workers=len(self.engines))
        self._replicated_gen = None   # primary.fit_generation the other devices hold a copy of

    # ---- pass-throughs to the first device ---------------------------------------------------
    @property
    def primary(self):
        return self.engines[0]

    @property
    def tdev(self):
        return self.primary.tdev

    # The cache keys of the driver (algo.py / compiler.upload) live on the engines that own the device state, so a
    # single-device holder of the primary that refits or re-registers a space invalidates this wrapper's view too.
    @property
    def _fit_token(self):
        return self.primary._fit_token

    @_fit_token.setter
    def _fit_token(self, value):
        self.primary._fit_token = value

    @property
    def _fit_theta(self):
        return self.primary._fit_theta

    @_fit_theta.setter
    def _fit_theta(self, value):
        self.primary._fit_theta = value

    @property
    def _space_key(self):
        keys = {e._space_key for e in self.engines}
        return keys.pop() if len(keys) == 1 else None

    @_space_key.setter
    def _space_key(self, value):
        for e in self.engines:
            e._space_key = value

    @property
    def fit_generation(self):
        return self.primary.fit_generation

    @property
    def launches(self):
        return sum(e.launches for e in self.engines)

    def to_device(self, array):
        return self.primary.to_device(array)

    def sample(self, *args, **kwargs):
        return self.primary.sample(*args, **kwargs)

    def lml(self):
        return self.primary.lml()

    def lml_grad(self):
        return self.primary.lml_grad()

    def predict_grad(self, x):
        return self.primary.predict_grad(x)

    def ystats(self):
        return self.primary.ystats()

    # ---- replicated state ----------------------------------------------------------------------
    def set_space(self, slots, cdf):
        for e in self.engines:
            e.set_space(slots, cdf)
        self.d_out = self.primary.d_out

    def fit(self, X, y, **kwargs):
        self.primary.fit(X, y, **kwargs)

    def fit_search(self, X, y, **kwargs):
        self.primary.fit_search(X, y, **kwargs)

    def _replicate(self):
        src = self.primary
        if self._replicated_gen == src.fit_generation or len(self.engines) == 1:
            return
        blob = src.export_state()
        src.sync()
        for e in self.engines[1:]:
            e.import_state(blob.to(e.tdev), src.n, src.D)
        self._replicated_gen = src.fit_generation

    # ---- sharded scoring -------------------------------------------------------------------------
    def _shards(self, M, first_index):
        """[(engine, first global index, count)] -- contiguous split of the candidate index range."""
        n = len(self.engines)
        jobs = []
        for rank, e in enumerate(self.engines):
            start, count = obdist.shard_range(M, rank, n)
            jobs.append((e, first_index + start, count))
        return jobs

    def score(self, cand=None, M=None, seed=0, first_index=0, acq_mask=1, y_opt=0.0, xi=0.01, kappa=1.96, dense=False):
        if cand is not None or dense:
            return self.primary.score(cand, M=M, seed=seed, first_index=first_index, acq_mask=acq_mask, y_opt=y_opt, xi=xi,
                                      kappa=kappa, dense=dense)
        self._replicate()

        def one(job):
            e, start, count = job
            if count == 0:
                return None
            return e.score(None, M=count, seed=seed, first_index=start, acq_mask=acq_mask, y_opt=y_opt, xi=xi, kappa=kappa)
        return obdist.merge_best(list(self.pool.map(one, self._shards(M, first_index))))

    def score_pruned(self, cand=None, M=None, seed=0, first_index=0, acq="EI", y_opt=0.0, xi=0.01, kappa=1.96):
        if cand is not None:
            return self.primary.score_pruned(cand, M=M, seed=seed, first_index=first_index, acq=acq, y_opt=y_opt, xi=xi,
                                             kappa=kappa)
        self._replicate()

        def one(job):
            e, start, count = job
            if count == 0:
                return None, 0
            return e.score_pruned(None, M=count, seed=seed, first_index=start, acq=acq, y_opt=y_opt, xi=xi, kappa=kappa)
        results = list(self.pool.map(one, self._shards(M, first_index)))
        return obdist.merge_best([r[0] for r in results]), sum(r[1] for r in results)

    def score_pruned_multi(self, cand=None, M=None, seed=0, first_index=0, acq_mask=7, y_opt=0.0, xi=0.01, kappa=1.96):
        if cand is not None:
            return self.primary.score_pruned_multi(cand, M=M, seed=seed, first_index=first_index, acq_mask=acq_mask,
                                                   y_opt=y_opt, xi=xi, kappa=kappa)
        self._replicate()

        def one(job):
            e, start, count = job
            if count == 0:
                return None, {}
            return e.score_pruned_multi(None, M=count, seed=seed, first_index=start, acq_mask=acq_mask, y_opt=y_opt,
                                        xi=xi, kappa=kappa)
        results = list(self.pool.map(one, self._shards(M, first_index)))
        scored = {}
        for _, ns in results:
            for name, v in ns.items():
                scored[name] = scored.get(name, 0) + v
        return obdist.merge_best([r[0] for r in results]), scored


_MULTI = {}


def get_multi_engine(devices):
    """Process-wide MultiEngine for a device list (kept outside algorithm instances, which Orion deep-copies)."""
    key = tuple(int(d) for d in devices)
    if key not in _MULTI:
        _MULTI[key] = MultiEngine(key)
    return _MULTI[key]
