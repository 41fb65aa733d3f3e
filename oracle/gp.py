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
