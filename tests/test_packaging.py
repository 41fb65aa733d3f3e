"""The two distributions install and Orion finds the plugin through its entry point, with no `register()` hack.

`pip install --no-build-isolation --no-deps --target <tmp>` of the repo root (`metaopt_b200`: package + libob200.so +
header) and of `plugin/` (`orion.algo.b200`, the namespace distribution with the `OptimizationAlgorithm` entry point,
modelled on the reference's template `tests/functional/gradient_descent_algo/setup.py:6-27`); then a FRESH
interpreter whose only knowledge of this repo is that directory lets the reference's factory
(`src/orion/core/utils/__init__.py:111-115`: `pkg_resources.iter_entry_points('OptimizationAlgorithm')`) find
`bayesianoptimizer`."""
import os
import subprocess
import sys
import textwrap

import pytest

from oracle import ref_loader

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

pytestmark = pytest.mark.skipif(not ref_loader.installed_available(),
                                reason="needs the installed reference (baseline/install_ref.sh -> baseline/_ref)")


def _pip_target(src, target):
    cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--no-deps", "--quiet",
           "--target", str(target), src]
    subprocess.run(cmd, check=True, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)


@pytest.mark.timeout(600)
def test_installed_plugin_is_discovered_through_its_entry_point(tmp_path):
    site = tmp_path / "site"
    _pip_target(ROOT, site)
    _pip_target(os.path.join(ROOT, "plugin"), site)
    assert (site / "metaopt_b200" / "libob200.so").exists() and (site / "metaopt_b200" / "include" / "ob200.h").exists()
    assert (site / "orion" / "algo" / "b200.py").exists()
    probe = tmp_path / "probe.py"
    probe.write_text(textwrap.dedent("""\
        import importlib.util, sys, warnings
        warnings.simplefilter("ignore")
        assert not any(p.rstrip("/") == %(root)r for p in sys.path), sys.path
        spec = importlib.util.spec_from_file_location("ref_loader", %(loader)r)   # in-memory shims only (appdirs, pymongo)
        ref_loader = importlib.util.module_from_spec(spec); spec.loader.exec_module(ref_loader)
        ref_loader.load()
        from orion.algo.base import OptimizationAlgorithm
        from orion.algo.space import Real, Space
        assert "bayesianoptimizer" in OptimizationAlgorithm.typenames, OptimizationAlgorithm.typenames
        space = Space(); space.register(Real("x", "uniform", 0, 1))
        algo = OptimizationAlgorithm("BayesianOptimizer", space, seed=1)
        import metaopt_b200, orion.algo.b200
        assert metaopt_b200.__file__.startswith(%(site)r), metaopt_b200.__file__
        assert type(algo).__module__ == "metaopt_b200.algo" and algo.configuration["bayesianoptimizer"]["seed"] == 1
        from metaopt_b200 import _lib
        _lib.load()                                    # the installed libob200.so exports every declared symbol
        assert _lib.LIB_PATH.startswith(%(site)r)
        print("DISCOVERED", type(algo).__name__)
        """ % dict(root=ROOT, loader=os.path.join(ROOT, "oracle", "ref_loader.py"), site=str(site))))
    shims = tmp_path / "shims"
    ref_loader.write_dist_shims(str(shims))            # pymongo / appdirs are not in the offline image
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([str(site), ref_loader.INSTALLED, str(shims)]),
               PYTHONDONTWRITEBYTECODE="1", OB200_REF_SRC=ref_loader.INSTALLED)
    env.pop("OB200_LIB", None)
    out = subprocess.run([sys.executable, str(probe)], cwd=str(tmp_path), env=env, stdout=subprocess.PIPE,
                         stderr=subprocess.STDOUT, text=True)
    assert out.returncode == 0 and "DISCOVERED BayesianOptimizer" in out.stdout, out.stdout
