"""GP fit / posterior restatement + scikit-learn anchor (test infrastructure).

The reference's GP arithmetic is scikit-learn's `GaussianProcessRegressor` reached through
skopt.  scikit-learn 1.9.0 is on the box, so `SklearnGP` runs the real
third-party code; `NumpyGP` restates the same operations in numpy/scipy in the same order
(sklearn `gaussian_process/_gpr.py`: y normalisation :275-280, K[diag]+=alpha :350,
cholesky :352, cho_solve :363-367; predict :444-500; Matern nu=2.5 `kernels.py:1720-1729`,
RBF `:1569-1570`).  skopt convention: the white-noise level is folded into the diagonal
(alpha + noise) and NOT added back in `predict`.
"""
import numpy as np
from scipy.linalg import cholesky, cho_solve, solve_triangular
from scipy.spatial.distance import cdist

KERNEL_MATERN52, KERNEL_RBF = 0, 1


def cross_kernel(kind, A, B, amplitude, length_scale):
    ls = np.asarray(length_scale, dtype=float)
    if kind == KERNEL_MATERN52:
        d = cdist(A / ls, B / ls, metric='euclidean')
        K = d * np.sqrt(5.0)
        K = (1.0 + K + K ** 2 / 3.0) * np.exp(-K)
    else:
        d = cdist(A / ls, B / ls, metric='sqeuclidean')
        K = np.exp(-0.5 * d)
    return amplitude * K


class NumpyGP:
    def __init__(self, kind, amplitude, length_scale, alpha, normalize_y):
        self.kind, self.amplitude, self.length_scale = kind, float(amplitude), length_scale
        self.alpha, self.normalize_y = float(alpha), bool(normalize_y)

    def fit(self, X, y):
        X = np.asarray(X, float)
        y = np.asarray(y, float)
        if self.normalize_y:
            self.y_mean = y.mean()
            s = y.std()
            self.y_std = 1.0 if s < 10 * np.finfo(float).eps else s  # _handle_zeros_in_scale
            y = (y - self.y_mean) / self.y_std
        else:
            self.y_mean, self.y_std = 0.0, 1.0
        K = cross_kernel(self.kind, X, X, self.amplitude, self.length_scale)
        K[np.diag_indices_from(K)] += self.alpha
        self.L = cholesky(K, lower=True)
        self.alpha_ = cho_solve((self.L, True), y)
        self.X = X
        return self

    def predict(self, Xc):
        Ks = cross_kernel(self.kind, np.asarray(Xc, float), self.X, self.amplitude,
                          self.length_scale)
        mu = self.y_std * (Ks @ self.alpha_) + self.y_mean
        V = solve_triangular(self.L, Ks.T, lower=True)
        var = self.amplitude - np.einsum('ij,ji->i', V.T, V)
        var[var < 0] = 0.0
        var = var * self.y_std ** 2
        return mu, np.sqrt(var)


    def predict_grad(self, Xc):
        """(mu, sigma, dmu/dx, dsigma/dx) per row of Xc: the analytic derivatives of `predict` that skopt's
        GaussianProcessRegressor subclass returns for return_mean_grad / return_std_grad (skopt
        `learning/gaussian_process/gpr.py`, kernels' `gradient_x`; third-party, unpinned).  Checked against
        central differences of `predict` in tests/test_oracle_golden.py."""
        Xc = np.atleast_2d(np.asarray(Xc, float))
        ls = np.broadcast_to(np.asarray(self.length_scale, float), (self.X.shape[1],))
        mus, sds, dmus, dsds = [], [], [], []
        for x in Xc:
            u = (x - self.X) / ls                       # [n, D]
            r2 = np.einsum('ij,ij->i', u, u)
            if self.kind == KERNEL_MATERN52:
                t = np.sqrt(5.0 * r2)
                k = self.amplitude * (1.0 + t + t * t / 3.0) * np.exp(-t)
                g = self.amplitude * (5.0 / 3.0) * (1.0 + t) * np.exp(-t)
            else:
                k = self.amplitude * np.exp(-0.5 * r2)
                g = k
            dk = -(g[:, None] * u) / ls                 # dk_j / dx_d
            v = solve_triangular(self.L, k, lower=True)
            w = solve_triangular(self.L.T, v, lower=False)
            var = max(self.amplitude - v @ v, 0.0)
            sd = np.sqrt(var)
            mus.append(self.y_std * (k @ self.alpha_) + self.y_mean)
            sds.append(self.y_std * sd)
            dmus.append(self.y_std * (dk.T @ self.alpha_))
            dsds.append(self.y_std * (-(dk.T @ w)) / sd if var > 0 else np.zeros(len(ls)))
        return np.array(mus), np.array(sds), np.array(dmus), np.array(dsds)


class SklearnGP:
    """The on-box scikit-learn estimator with fixed hyper-parameters (optimizer=None)."""

    def __init__(self, kind, amplitude, length_scale, alpha, normalize_y):
        from sklearn.gaussian_process import GaussianProcessRegressor
        from sklearn.gaussian_process.kernels import ConstantKernel, Matern, RBF
        ls = np.asarray(length_scale, dtype=float)
        base = Matern(length_scale=ls, nu=2.5) if kind == KERNEL_MATERN52 else RBF(length_scale=ls)
        self.gpr = GaussianProcessRegressor(kernel=ConstantKernel(float(amplitude)) * base,
                                            alpha=float(alpha), optimizer=None,
                                            normalize_y=bool(normalize_y))

    def fit(self, X, y):
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            self.gpr.fit(np.asarray(X, float), np.asarray(y, float))
        self.L = self.gpr.L_
        self.alpha_ = self.gpr.alpha_
        return self

    def predict(self, Xc):
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            return self.gpr.predict(np.asarray(Xc, float), return_std=True)
