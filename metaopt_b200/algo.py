"""`BayesianOptimizer`: the B200 plugin behind Orion's unchanged algorithm API.

Drop-in for the reference's documented skopt wrapper (`orion.algo.skopt.BayesianOptimizer`,
`/root/reference/docs/src/user/algorithms.rst:136-217`): same class name (so the same
`algorithms: {BayesianOptimizer: {...}}` config key, matched case-insensitively by
`orion.core.utils.Factory`, `src/orion/core/utils/__init__.py:151-153`), same config keys and
defaults (`seed, n_initial_points, acq_func, alpha, n_restarts_optimizer, noise, normalize_y`;
also pinned by `tests/unittests/core/cli/test_info.py:89-98`), `requires = 'real'`.

Contract with Orion:
  * `BaseAlgorithm.__init__(space, **kwargs)` (`src/orion/algo/base.py:87-119`) turns every kwarg
    into an attribute and into `configuration`; a `seed` kwarg calls `seed_rng`.
  * the object is constructed several times at start-up and `copy.deepcopy`'d on every producer
    update (`src/orion/core/worker/producer.py:161`): no CUDA work in `__init__`, no device handle
    in `__dict__` -- the engine lives in `metaopt_b200.engine._ENGINES`.
  * `suggest(num)` returns points of the transformed ('real') space; Orion then checks
    `point in transformed_space` and reverses it (`src/orion/core/worker/primary_algo.py:61-81`).
  * `observe(points, results)` may deliver the whole history at once (`producer.py:107-132`) or one
    point; results are dicts with an `objective` key (`core/utils/format_trials.py:53-72`).
  * `state_dict` / `set_state` carry the Philox stream position (`producer.py:84`).

When Orion is importable the class derives from `orion.algo.base.BaseAlgorithm`; otherwise from a
minimal mirror with the same behaviour, so the GPU box (no Orion install) runs the same code.
"""
import copy
import hashlib

import numpy as np

from . import _lib, compiler

try:  # the real thing, when the plugin runs under Orion
    from orion.algo.base import BaseAlgorithm  # pylint: disable=import-error
    HAVE_ORION = True
except Exception:  # pragma: no cover - exercised on the GPU box
    HAVE_ORION = False

    class BaseAlgorithm:
        """Mirror of `orion.algo.base.BaseAlgorithm` (`base.py:21-270`) for Orion-less hosts."""

        requires = []

        def __init__(self, space, **kwargs):
            self._space = space
            self._param_names = list(kwargs.keys())
            for name, value in kwargs.items():
                if name == "seed":
                    self.seed_rng(value)
                setattr(self, name, value)

        def seed_rng(self, seed):
            pass

        @property
        def state_dict(self):
            return {}

        def set_state(self, state_dict):
            pass

        @property
        def is_done(self):
            return False

        def score(self, point):
            return 0

        def judge(self, point, measurements):
            return None

        @property
        def should_suspend(self):
            return False

        @property
        def configuration(self):
            conf = {k: getattr(self, k) for k in self._param_names if not k.startswith("_")}
            return {self.__class__.__name__.lower(): conf}

        @property
        def space(self):
            return self._space

        @space.setter
        def space(self, space_):
            self._space = space_


_ACQ = {"EI": _lib.ACQ_EI, "PI": _lib.ACQ_PI, "LCB": _lib.ACQ_LCB}
_KERNELS = {"matern52": _lib.KERNEL_MATERN52, "matern": _lib.KERNEL_MATERN52, "rbf": _lib.KERNEL_RBF}


class BayesianOptimizer(BaseAlgorithm):
    """GP-based Bayesian optimiser whose candidate scoring runs on one B200.

    Reference keys (same meaning as the skopt wrapper): `seed`, `n_initial_points`, `acq_func`
    ('EI' | 'PI' | 'LCB' | 'gp_hedge'), `alpha`, `n_restarts_optimizer`, `noise`, `normalize_y`.
    Additive keys (defaults keep reference behaviour): `n_points` (candidates per suggest, skopt's
    default 10000), `xi`, `kappa`, `kernel` ('matern52' | 'rbf'), `length_scale`, `amplitude`,
    `noise_level` (variance folded into the diagonal when `noise == 'gaussian'`),
    `fit_hyperparameters` (maximise the log marginal likelihood over amplitude, per-dimension length
    scales and -- when `noise == 'gaussian'` -- the noise level, like the reference's scikit-learn
    estimator does, with `n_restarts_optimizer` extra starts; `length_scale`/`amplitude`/`noise_level`
    are then the starting point), `acq_optimizer` ('sampling': the proposal is the best of the `n_points`
    candidates; 'lbfgs': skopt's default for this space -- the best candidate of each of `acq_n_restarts`
    (5) slices of the candidate stream starts an L-BFGS-B ascent of the acquisition inside the unit box,
    `metaopt_b200/acq_opt.py`, and the best end point is proposed).

    Host placement is NOT an algorithm hyper-parameter: Orion stores `configuration` as `experiment.algorithms`
    and raises a conflict (branches the experiment) when two workers disagree on it
    (`src/orion/core/evc/conflicts.py:1034-1038`), so two workers pinned to different GPUs must not differ
    there.  `device` (an index, or a list of indices to shard the candidates of every suggest over several GPUs
    of this process) and `prune` (exact mean-only pruning of hopeless candidates; same winner, faster) are
    therefore read from the environment -- `OB200_DEVICE` ("0", "0,1,2,3"), `OB200_PRUNE` ("1"/"0") -- or
    passed as constructor keywords by code that builds the object itself; either way they never enter
    `configuration`.
    """

    requires = "real"

    # pylint: disable=too-many-arguments
    def __init__(self, space, seed=None, n_initial_points=10, acq_func="gp_hedge", alpha=1e-10,
                 n_restarts_optimizer=0, noise="gaussian", normalize_y=False, n_points=10000, xi=0.01,
                 kappa=1.96, kernel="matern52", length_scale=1.0, amplitude=1.0, noise_level=1e-4, device=None,
                 prune=None, fit_hyperparameters=True, acq_optimizer="sampling", acq_n_restarts=5):
        self._seed = None
        self._offset = 0          # next unused global candidate index of the Philox stream
        self._X = []              # observed points, flat + normalised
        self._y = []
        self._compiled = None
        self._hedge_gains = [0.0, 0.0, 0.0]
        self._hedge_pending = None
        self._theta = None        # last fitted (amplitude, length_scale[D], noise): warm start of the next fit
        self._cache = None        # stacked observations + chained digest (see _data)
        self._device = _placement(device)
        self._prune = _env_flag("OB200_PRUNE", True) if prune is None else bool(prune)
        super().__init__(space, seed=seed, n_initial_points=n_initial_points, acq_func=acq_func, alpha=alpha,
                         n_restarts_optimizer=n_restarts_optimizer, noise=noise, normalize_y=normalize_y,
                         n_points=n_points, xi=xi, kappa=kappa, kernel=kernel, length_scale=length_scale,
                         amplitude=amplitude, noise_level=noise_level,
                         fit_hyperparameters=fit_hyperparameters, acq_optimizer=acq_optimizer,
                         acq_n_restarts=acq_n_restarts)
        if str(acq_optimizer) not in ("sampling", "lbfgs"):
            raise ValueError("acq_optimizer must be 'sampling' or 'lbfgs'; got %r" % (acq_optimizer,))
        if str(acq_func) not in ("EI", "PI", "LCB", "gp_hedge"):
            raise ValueError("acq_func must be one of 'EI', 'PI', 'LCB', 'gp_hedge' (EIps/PIps need timings "
                             "that Orion does not pass); got %r" % (acq_func,))
        if str(kernel).lower() not in _KERNELS:
            raise ValueError("kernel must be 'matern52' or 'rbf'; got %r" % (kernel,))
        if self._seed is None:
            self.seed_rng(None)

    @property
    def device(self):
        """CUDA device index (or list of indices) this process scores on; not part of `configuration`."""
        return self._device

    @property
    def prune(self):
        return self._prune

    # ---- BaseAlgorithm API ---------------------------------------------------------------
    def seed_rng(self, seed):
        """Select the Philox stream (None = fresh entropy) and rewind it."""
        self._seed = compiler.fold_seed(seed)
        self._offset = 0

    @property
    def state_dict(self):
        return {"seed": self._seed, "offset": self._offset, "hedge_gains": list(self._hedge_gains),
                "hedge_pending": copy.deepcopy(self._hedge_pending), "theta": copy.deepcopy(self._theta)}

    def set_state(self, state_dict):
        self._seed = state_dict["seed"]
        self._offset = state_dict["offset"]
        self._hedge_gains = list(state_dict.get("hedge_gains", [0.0, 0.0, 0.0]))
        self._hedge_pending = copy.deepcopy(state_dict.get("hedge_pending"))
        if state_dict.get("theta") is not None:  # warm start of the next hyper-parameter fit
            self._theta = copy.deepcopy(state_dict["theta"])

    @property
    def space(self):
        return self._space

    @space.setter
    def space(self, space_):
        """Orion assigns the transformed space right after construction (`primary_algo.py:43`)."""
        self._space = space_
        self._compiled = None

    def __deepcopy__(self, memo):
        new = copy.copy(self)  # shallow: config scalars, the shared space object, the compiled table
        new._X = list(self._X)
        new._y = list(self._y)
        new._hedge_gains = list(self._hedge_gains)
        new._hedge_pending = copy.deepcopy(self._hedge_pending)
        new._theta = copy.deepcopy(self._theta)
        new._cache = self._cache  # never mutated in place
        memo[id(self)] = new
        return new

    def suggest(self, num=1):
        """`num` points of the transformed space (tuples; 1-D arrays for one-hot / shaped dims).

        `num > 1` (the reference wrapper only supports 1) uses skopt's constant-liar strategy "cl_min": after
        each pick the point is added with the best objective seen so far as a stand-in result, the GP is
        refitted with the hyper-parameters kept, and the next point is picked; the lies are dropped afterwards."""
        comp = self._get_compiled()
        eng = self._engine()
        compiler.upload(eng, comp)
        points, rows = [], []
        n_real = len(self._y)
        saved_fit, saved_cache = self.fit_hyperparameters, self._cache
        hedge_state = None
        try:
            for k in range(int(num)):
                if n_real < max(int(self.n_initial_points), 1):
                    row = eng.sample(self._seed, self._offset, 1, normalise=False).cpu().numpy()[0]
                    self._offset += 1
                else:
                    row = self._suggest_gp(eng, comp)
                    if hedge_state is None:  # portfolio state after the pick made on REAL data only
                        hedge_state = (list(self._hedge_gains), copy.deepcopy(self._hedge_pending))
                    if k + 1 < int(num):  # constant liar: pretend this point returned the incumbent
                        self._X.append(comp.normalise(np.asarray(row, dtype=float)))
                        self._y.append(float(min(self._y[:n_real])))
                        if self._theta is not None:
                            self.fit_hyperparameters = False  # lies must not move the hyper-parameters
                rows.append(np.asarray(row, dtype=np.float64))
                points.append(compiler.regroup_row(comp, row, transformed=True))
            # the check Orion makes point by point in Python right after this call (`primary_algo.py:74-81`), done
            # once for the whole batch: a proposal outside the space is a bug here, not something to hand back
            inside = compiler.contains_rows(comp, np.stack(rows))
            if not inside.all():
                bad = int(np.flatnonzero(~inside)[0])
                raise ValueError("Point is not contained in space:\nPoint: {}\nSpace: {}".format(points[bad], self._space))
        finally:
            del self._X[n_real:]
            del self._y[n_real:]
            self.fit_hyperparameters = saved_fit
            if int(num) > 1:
                self._cache = saved_cache
                if hedge_state is not None:  # skopt asks for a batch on a COPY of the optimiser: lies leave no trace
                    self._hedge_gains, self._hedge_pending = hedge_state
        return points

    def observe(self, points, results):
        """Record evaluated points; the GP is refitted lazily at the next `suggest`."""
        comp = self._get_compiled()
        for point, result in zip(points, results):
            objective = result.get("objective") if isinstance(result, dict) else result
            if objective is None:
                continue
            flat = np.asarray(compiler.flatten_point(comp, point), dtype=float)
            self._X.append(comp.normalise(flat))
            self._y.append(float(objective))

    # ---- internals -------------------------------------------------------------------------
    def _engine(self):
        """Process-wide engine: one device (`device: 0`) or several in this process (`device: [0, 1, ...]`)."""
        if isinstance(self.device, (list, tuple)):
            if len(self.device) > 1:
                from .multi import get_multi_engine
                return get_multi_engine(self.device)
            device = self.device[0]
        else:
            device = self.device
        from .engine import get_engine
        return get_engine(int(device))

    def _get_compiled(self):
        if self._compiled is None:
            if self._space is None:
                raise RuntimeError("no space set")
            comp = compiler.compile_space(self._space)
            if not comp.normalisable:
                raise ValueError("BayesianOptimizer needs a finite interval() for every real/integer dimension "
                                 "(unbounded priors need low/high), like the reference skopt wrapper")
            self._compiled = comp
        return self._compiled

    def _diag_noise(self):
        return float(self.alpha) + (float(self.noise_level) if self.noise == "gaussian" else 0.0)

    def _data(self):
        """(X [n, D], y [n], digest) of the observations.  Stacking 2000 rows and hashing them costs ~2 ms, as much as
        a fifth of a pruned suggest, so both are kept and extended row by row; the digest is a chain
        d_i = H(d_{i-1} | x_i | y_i), identical whichever way the rows arrived (one by one, or all at once on a
        restart).  The cache is trusted while its first and last rows are still the same objects."""
        n = len(self._y)
        c = self._cache
        start = 0
        if c is not None and 0 < c["n"] <= n and self._X[0] is c["first"] and self._X[c["n"] - 1] is c["last"]:
            if c["n"] == n:
                return c["X"], c["y"], c["digest"]
            start = c["n"]
        digest = c["digest"] if start else b""
        for i in range(start, n):
            digest = hashlib.blake2b(digest + np.asarray(self._X[i], dtype=np.float64).tobytes()
                                     + np.float64(self._y[i]).tobytes(), digest_size=16).digest()
        Xn = np.asarray(self._X[start:], dtype=np.float64)
        yn = np.asarray(self._y[start:], dtype=np.float64)
        X = np.concatenate([c["X"], Xn]) if start else Xn
        y = np.concatenate([c["y"], yn]) if start else yn
        if n:
            self._cache = dict(n=n, X=X, y=y, digest=digest, first=self._X[0], last=self._X[n - 1])
        return X, y, digest

    def _ensure_fit(self, eng):
        X, y, digest = self._data()
        ls = np.atleast_1d(np.asarray(self.length_scale, dtype=np.float64))
        kind = _KERNELS[str(self.kernel).lower()]
        h = hashlib.blake2b(digest_size=16)
        for part in (digest, ls.tobytes(),
                     repr((self.kernel, self.amplitude, self._diag_noise(), bool(self.normalize_y),
                           bool(self.fit_hyperparameters), int(self.n_restarts_optimizer),
                           None if (self._theta is None or self.fit_hyperparameters) else
                           (self._theta[0], tuple(np.atleast_1d(self._theta[1])), self._theta[2]))).encode()):
            h.update(part)
        token = h.digest()
        if getattr(eng, "_fit_token", None) != token:
            if self.fit_hyperparameters and len(y) >= 2:
                from . import hyperopt
                amp0, ls0, nz0 = self._theta if self._theta is not None else (
                    float(self.amplitude), np.broadcast_to(ls, (X.shape[1],)).copy(), float(self.noise_level))
                gaussian = self.noise == "gaussian"
                rng = np.random.RandomState(self._seed % (2 ** 32))
                Xd, yd = eng.to_device(X), eng.to_device(y)
                res = hyperopt.optimise(eng, Xd, yd, kind, amp0, ls0, nz0 if gaussian else 0.0, float(self.alpha),
                                        bool(self.normalize_y), anisotropic=True, optimise_noise=gaussian,
                                        n_restarts=int(self.n_restarts_optimizer), rng=rng)
                self._theta = (res.amplitude, np.asarray(res.length_scale), res.noise if gaussian else float(self.noise_level))
            elif self._theta is not None and saved_theta_ok(self._theta, X.shape[1]):
                amp_t, ls_t, nz_t = self._theta  # hyper-parameters of the last optimisation (constant-liar refits)
                eng.fit(X, y, kind=kind, amplitude=amp_t, length_scale=ls_t,
                        alpha=float(self.alpha) + (nz_t if self.noise == "gaussian" else 0.0), normalize_y=bool(self.normalize_y))
            else:
                eng.fit(X, y, kind=kind, amplitude=float(self.amplitude), length_scale=ls,
                        alpha=self._diag_noise(), normalize_y=bool(self.normalize_y))
            eng._fit_token = token
            eng._fit_theta = copy.deepcopy(self._theta)
            self._settle_hedge(eng)
        elif self.fit_hyperparameters and getattr(eng, "_fit_theta", None) is not None:
            # another copy of this algorithm (producer.py:161) already fitted these data: adopt its optimum
            self._theta = copy.deepcopy(eng._fit_theta)
        return float(y.min())

    def _score_range(self, eng, first, count, y_opt):
        """Winner(s) {acq name: record} of the candidates [first, first + count) of the stream."""
        hedge = self.acq_func == "gp_hedge"
        mask = 7 if hedge else (1 << _ACQ[self.acq_func])
        if self.prune:  # same winners as one dense pass; the portfolio shares the mean-only pass
            res, _ = eng.score_pruned_multi(None, M=count, seed=self._seed, first_index=first, acq_mask=mask,
                                            y_opt=y_opt, xi=float(self.xi), kappa=float(self.kappa))
            return res
        return eng.score(None, M=count, seed=self._seed, first_index=first, acq_mask=mask, y_opt=y_opt,
                         xi=float(self.xi), kappa=float(self.kappa))

    def _suggest_gp(self, eng, comp):
        from . import dist as obdist
        y_opt = self._ensure_fit(eng)
        M = int(self.n_points)
        hedge = self.acq_func == "gp_hedge"
        refine = self.acq_optimizer == "lbfgs"
        parts = []
        if refine:  # one start per slice of the stream; the global winner is the best of them
            n_parts = max(1, min(int(self.acq_n_restarts), M))
            for r in range(n_parts):
                start, count = obdist.shard_range(M, r, n_parts)
                parts.append(self._score_range(eng, self._offset + start, count, y_opt))
            res = obdist.merge_best(parts)
        else:
            res = self._score_range(eng, self._offset, M, y_opt)
        self._offset += M
        rows = {}
        if refine:
            from . import acq_opt
            for name in res:
                idx = sorted({int(p[name]["index"]) for p in parts if p and name in p},
                             key=lambda i: (i != int(res[name]["index"]), i))  # global winner first
                starts = np.stack([eng.sample(self._seed, i, 1, normalise=True).cpu().numpy()[0] for i in idx])
                best = acq_opt.refine(eng, name, starts, y_opt, xi=float(self.xi), kappa=float(self.kappa))
                # Orion's intervals are half-open (`space.py:356`: low <= x < high): an ascent that stops on the
                # upper face is moved one ulp inside.
                rows[name] = np.clip(comp.denormalise(best.x), comp.nlo, np.nextafter(comp.nhi, comp.nlo))
                res[name] = dict(res[name], value=best.value, mu=best.mu, sigma=best.sigma, index=idx[best.start])
        if hedge:
            # skopt's gp_hedge: draw one of the three proposals with probability softmax(eta * gains);
            # the draw itself uses the Philox stream so that it follows state_dict/set_state.
            names = ("EI", "LCB", "PI")
            gains = np.asarray(self._hedge_gains)
            logits = gains - gains.max()
            probs = np.exp(logits) / np.exp(logits).sum()
            u = float(self._uniform())
            pick = names[int(min(np.searchsorted(np.cumsum(probs), u, side="right"), 2))]
            # [n_obs at the pick, the three proposals as normalised rows]: their gains are settled by the first
            # refit that has seen a newer observation (_settle_hedge)
            self._hedge_pending = [len(self._y)] + [np.asarray(comp.normalise(rows[n]) if refine else
                                              eng.sample(self._seed, res[n]["index"], 1, normalise=True).cpu().numpy()[0],
                                              dtype=np.float64).tolist() for n in names]
        else:
            pick = self.acq_func
        if refine:
            return rows[pick]
        return eng.sample(self._seed, res[pick]["index"], 1, normalise=False).cpu().numpy()[0]

    def _uniform(self):
        """One host-side uniform from the same counter-based stream (stream position = offset)."""
        from .philox_host import uniform01
        u = uniform01(self._seed, self._offset)
        self._offset += 1
        return u

    def _settle_hedge(self, eng):
        """skopt's `tell` does `gains_ -= est.predict(next_xs_)` with the estimator REFITTED on the new
        observation and `next_xs_` the three proposals of the previous round.  Called right after a refit: the
        posterior mean of the pending proposals under the new model is subtracted from the gains.  Two
        departures that follow from Orion's protocol, not from the arithmetic: a second `suggest` without an
        `observe` in between replaces the pending proposals (skopt's `ask` would return the same point again), and
        in a producer whose naive copy observed lies (`producer.py:159-174`, only when other trials are in
        flight) the refit that settles the gains contains those lies."""
        if self._hedge_pending and self.acq_func == "gp_hedge" and len(self._y) > self._hedge_pending[0]:
            mu = eng.predict_grad(np.asarray(self._hedge_pending[1:], dtype=np.float64))[0]
            for k in range(3):
                self._hedge_gains[k] -= float(mu[k])
            self._hedge_pending = None


def _env_flag(name, default):
    import os
    v = os.environ.get(name)
    return default if v is None or v == "" else v.strip().lower() not in ("0", "false", "no", "off")


def _placement(device):
    """Device index or list of indices: the constructor keyword, else `OB200_DEVICE`, else 0."""
    import os
    if device is None:
        env = os.environ.get("OB200_DEVICE", "").strip()
        if not env:
            return 0
        ids = [int(t) for t in env.replace(";", ",").split(",") if t.strip()]
        return ids[0] if len(ids) == 1 else ids
    if isinstance(device, (list, tuple)):
        return [int(d) for d in device]
    return int(device)


def saved_theta_ok(theta, d):
    """A stored (amplitude, length_scale, noise) triple matches the current space width."""
    return len(np.atleast_1d(theta[1])) in (1, d)


def register():
    """Make the plugin discoverable without a pip install (what `tests/conftest.py:114-116` of the
    reference does for its own fake algorithm); an installed distribution uses the entry point."""
    if not HAVE_ORION:
        raise RuntimeError("Orion is not importable")
    from orion.algo.base import OptimizationAlgorithm  # pylint: disable=import-error
    name = BayesianOptimizer.__name__.lower()
    if name not in OptimizationAlgorithm.typenames:
        OptimizationAlgorithm.types.append(BayesianOptimizer)
        OptimizationAlgorithm.typenames.append(name)
