"""CPU stand-in for `metaopt_b200.engine.Engine`, built on the oracle (TEST infrastructure only).

The build container has the unmodified reference but no GPU; the GPU box has a GPU but no reference.
To exercise the plugin's host logic under the real `PrimaryAlgo` / `Producer` / `orion hunt` here, the
tests monkeypatch `metaopt_b200.engine.get_engine` with this class.  It is never importable from the
product package."""
import numpy as np

from oracle import acquisition as oacq
from oracle import gp as ogp
from oracle import sampling as osamp


class _Host:
    """numpy array with the two tensor methods the plugin calls."""

    def __init__(self, a):
        self.a = a

    def cpu(self):
        return self

    def numpy(self):
        return self.a


class OracleEngine:
    def __init__(self):
        self.slots = self.cdf = None
        self.model = None
        self.calls = {"sample": 0, "fit": 0, "score": 0}
        self._space_key = None
        self._fit_token = None

    def set_space(self, slots, cdf):
        self.slots, self.cdf = slots, cdf
        self.d_out = max(s["out_col"] + s["width"] for s in slots)

    def sample(self, seed, first_index, count, normalise=True, variates=None, out=None, check=True):
        self.calls["sample"] += 1
        mat, status = osamp.sample_transform(self.slots, self.cdf, seed, first_index, count, normalise, variates)
        if status:
            raise ValueError("Improbable bounds: a draw was rejected 4 times. Please make interval larger.")
        return _Host(mat)

    def fit(self, X, y, kind=0, amplitude=1.0, length_scale=1.0, alpha=1e-10, normalize_y=False, incremental=True):
        self.calls["fit"] += 1
        X = np.asarray(X, float)
        ls = np.broadcast_to(np.asarray(length_scale, float), (X.shape[1],)).copy()
        self._y_raw = np.asarray(y, float)
        self.model = ogp.NumpyGP(kind, amplitude, ls, alpha, normalize_y).fit(X, self._y_raw)

    def to_device(self, array):
        return np.asarray(array, dtype=float)

    def lml(self):
        return float(self._gpr().log_marginal_likelihood_value_)

    def _gpr(self):
        from sklearn.gaussian_process import GaussianProcessRegressor
        from sklearn.gaussian_process.kernels import RBF, ConstantKernel, Matern
        m = self.model
        base = Matern(length_scale=m.length_scale, nu=2.5) if m.kind == 0 else RBF(length_scale=m.length_scale)
        return GaussianProcessRegressor(kernel=ConstantKernel(m.amplitude) * base, alpha=m.alpha, optimizer=None,
                                        normalize_y=m.normalize_y).fit(m.X, self._y_raw)

    def lml_grad(self):
        """[dlog amp, dlog ls_d..., d(diag)] from sklearn's analytic gradient (+ a numeric diagonal term)."""
        g = self._gpr()
        lml, grad = g.log_marginal_likelihood(g.kernel_.theta, eval_gradient=True)
        m = self.model
        y = (self._y_raw - m.y_mean) / m.y_std
        import numpy.linalg as la
        Kinv = la.inv(m.L @ m.L.T)
        a = Kinv @ y
        d_diag = 0.5 * (a @ a - np.trace(Kinv))
        return float(lml), np.concatenate([grad, [d_diag]])

    def predict_grad(self, x):
        return self.model.predict_grad(x)

    def score(self, cand=None, M=None, seed=0, first_index=0, acq_mask=1, y_opt=0.0, xi=0.01, kappa=1.96, dense=False):
        self.calls["score"] += 1
        assert cand is None
        C, _ = osamp.sample_transform(self.slots, self.cdf, seed, first_index, M, True)
        mu, sd = self.model.predict(C)
        out = {}
        for k, name in ((0, "EI"), (1, "PI"), (2, "LCB")):
            if acq_mask & (1 << k):
                i = oacq.pick(k, mu, sd, y_opt, xi, kappa)
                out[name] = dict(value=float(oacq.utility(k, mu, sd, y_opt, xi, kappa)[i]), index=first_index + i,
                                 mu=float(mu[i]), sigma=float(sd[i]))
        return out

    def score_pruned_multi(self, cand=None, M=None, seed=0, first_index=0, acq_mask=7, y_opt=0.0, xi=0.01, kappa=1.96):
        out = self.score(cand, M=M, seed=seed, first_index=first_index, acq_mask=acq_mask, y_opt=y_opt, xi=xi, kappa=kappa)
        return out, {name: M for name in out}

    def score_pruned(self, cand=None, M=None, seed=0, first_index=0, acq="EI", y_opt=0.0, xi=0.01, kappa=1.96):
        k = {"EI": 0, "PI": 1, "LCB": 2}[acq]
        out = self.score(cand, M=M, seed=seed, first_index=first_index, acq_mask=1 << k, y_opt=y_opt, xi=xi, kappa=kappa)
        return out, M
