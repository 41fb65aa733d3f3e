#!/usr/bin/env python
"""Installation script for `orion.algo.b200`: registers the B200 Bayesian optimiser with Oríon.

Packaging follows the reference's own external-plugin template
(`tests/functional/gradient_descent_algo/setup.py`): a namespace package `orion.algo` with one module and an
`OptimizationAlgorithm` entry point, which `orion.core.utils.Factory` loads at import time."""
from setuptools import setup

setup(
    name="orion.algo.b200",
    version="0.2.0",
    description="B200-native GP-EI Bayesian optimiser for Oríon (drop-in for orion.algo.skopt's BayesianOptimizer)",
    packages=["orion.algo"],
    package_dir={"": "src"},
    entry_points={"OptimizationAlgorithm": ["bayesianoptimizer = orion.algo.b200:BayesianOptimizer"]},
    install_requires=["orion", "metaopt_b200"],
    zip_safe=False,
)
