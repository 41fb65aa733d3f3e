#!/usr/bin/env python
"""Installation script for `metaopt_b200`: the B200-native scoring library (libob200.so, built with nvcc for
sm_100a by `metaopt_b200/build.py`) and its Python host.  The Orion-facing entry point lives in the separate
namespace distribution under `plugin/` (`orion.algo.b200`), modelled on the reference's external-plugin template
(`tests/functional/gradient_descent_algo/setup.py:6-27`); install both:

    pip install --no-build-isolation .  ./plugin
"""
import os
import shutil
import sys

from setuptools import setup
from setuptools.command.build_py import build_py

ROOT = os.path.dirname(os.path.abspath(__file__))


class BuildWithLibrary(build_py):
    """Compile libob200.so in-tree (no-op when it is newer than its sources), then ship it and the C header
    inside the package."""

    def run(self):
        sys.path.insert(0, ROOT)
        try:
            from metaopt_b200 import build as ob_build
            ob_build.build(force=False)
        finally:
            sys.path.pop(0)
        super().run()
        dest = os.path.join(self.build_lib, "metaopt_b200", "include")
        os.makedirs(dest, exist_ok=True)
        shutil.copy2(os.path.join(ROOT, "include", "ob200.h"), os.path.join(dest, "ob200.h"))


setup(
    name="metaopt_b200",
    version="0.2.0",
    description="B200-native GP candidate scoring (sample/transform, GP fit, fused posterior + acquisition + argmax) "
                "behind Orion's BaseAlgorithm API",
    packages=["metaopt_b200"],
    package_data={"metaopt_b200": ["libob200.so", "csrc/*.cu", "csrc/*.cuh"]},
    cmdclass={"build_py": BuildWithLibrary},
    # torch (a CUDA build matching the driver) is a prerequisite of the host, deliberately not listed: the plugin's
    # entry point is loaded through pkg_resources, which re-verifies the whole requirement chain at import time,
    # and a torch wheel pins exact nvidia-* versions that site installs routinely override.
    install_requires=["numpy", "scipy"],
    python_requires=">=3.8",
    zip_safe=False,
)
