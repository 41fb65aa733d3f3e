"""Entry-point module: exposes the B200 plugin class to Oríon's `OptimizationAlgorithm` factory."""
from metaopt_b200.algo import BayesianOptimizer  # noqa: F401
