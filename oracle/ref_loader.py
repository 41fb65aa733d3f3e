"""Import the UNMODIFIED reference (Orion) with harness-side shims: from /root/reference/src in the build
container, else from the pip-installed copy under `baseline/_ref` (made by `baseline/install_ref.sh`; it is
git-ignored but travels to the GPU box, where /root/reference does not exist).

Test infrastructure only (see oracle/__init__.py).  Nothing is copied or edited: the reference's
`src/` is put on `sys.path` and three things it expects from its 2019 environment are supplied
in memory:
  1. `appdirs`  -> the reference's own vendored `src/orion/core/utils/_appdirs.py`, loaded by path;
  2. `pymongo`  -> a stub with the names `core/io/database/mongodb.py` imports (never used with
                   PickledDB / EphemeralDB);
  3. `numpy.object` -> `object` (removed in numpy >= 1.24; used at `algo/space.py:549,572`).
`available()` is False when neither location exists and every user of this module skips.
"""
import importlib.util
import os
import sys
import types

REF_ROOT = "/root/reference"
INSTALLED = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref")
REF_SRC = os.environ.get("OB200_REF_SRC") or next((p for p in (os.path.join(REF_ROOT, "src"), INSTALLED) if os.path.isdir(os.path.join(p, "orion"))),
               os.path.join(REF_ROOT, "src"))


def available():
    return os.path.isdir(os.path.join(REF_SRC, "orion"))


def installed_available():
    """The pip-installed copy (it has dist-info metadata, which entry-point loading checks)."""
    return os.path.isdir(os.path.join(INSTALLED, "orion"))


def write_dist_shims(directory):
    """Empty `pymongo` / `appdirs` distributions for a sys.path directory: the on-disk twin of the in-memory module
    shims above.  `pkg_resources` verifies a distribution's `install_requires` chain when the reference's factory
    loads an entry point (`src/orion/core/utils/__init__.py:111-112`); neither package is in the offline image."""
    for name, version in (("pymongo", "3.99"), ("appdirs", "1.4.4")):
        d = os.path.join(directory, "%s-%s.dist-info" % (name, version))
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, "METADATA"), "w") as f:
            f.write("Metadata-Version: 2.1\nName: %s\nVersion: %s\n" % (name, version))


def source_tree_available():
    """The reference's own tests/docs (only in the build container), as opposed to the installed package."""
    return os.path.isdir(os.path.join(REF_ROOT, "tests"))


def _install_shims():
    import numpy
    if not hasattr(numpy, "object"):
        numpy.object = object
    if "appdirs" not in sys.modules:
        path = os.path.join(REF_SRC, "orion", "core", "utils", "_appdirs.py")
        spec = importlib.util.spec_from_file_location("appdirs", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        sys.modules["appdirs"] = mod
    if "pymongo" not in sys.modules:
        pm = types.ModuleType("pymongo")
        errors = types.ModuleType("pymongo.errors")
        for name in ("DuplicateKeyError", "BulkWriteError", "ConnectionFailure", "OperationFailure", "InvalidURI",
                     "ConfigurationError"):
            setattr(errors, name, type(name, (Exception,), {}))
        uri_parser = types.ModuleType("pymongo.uri_parser")
        uri_parser.parse_uri = lambda uri, *a, **k: {"nodelist": [("localhost", 27017)], "database": None,
                                                     "username": None, "password": None, "options": {}}

        class MongoClient:  # pragma: no cover - never instantiated
            HOST, PORT = "localhost", 27017

            def __init__(self, *a, **k):
                raise errors.ConnectionFailure("pymongo stub: no MongoDB in the oracle harness")

        class ReturnDocument:
            BEFORE, AFTER = False, True

        pm.errors, pm.uri_parser, pm.MongoClient, pm.ReturnDocument = errors, uri_parser, MongoClient, ReturnDocument
        pm.ASCENDING, pm.DESCENDING = 1, -1
        sys.modules["pymongo"] = pm
        sys.modules["pymongo.errors"] = errors
        sys.modules["pymongo.uri_parser"] = uri_parser


def load():
    """Return the reference's `orion` package (or None when /root/reference is absent)."""
    if not available():
        return None
    sys.dont_write_bytecode = True  # the reference tree is read-only
    _install_shims()
    paths = [REF_SRC]
    if os.path.isdir(os.path.join(REF_SRC, "orion-0+unknown.dist-info")) or any(
            name.endswith(".dist-info") and name.startswith("orion") for name in os.listdir(REF_SRC)):
        # The pip-installed copy advertises entry points, and loading one makes pkg_resources verify the `orion`
        # distribution's install_requires (pymongo, AppDirs): give it the on-disk twins of the module shims.
        import tempfile
        shim_dir = tempfile.mkdtemp(prefix="orion_dist_shims_")
        write_dist_shims(shim_dir)
        paths.append(shim_dir)
    for path in paths:
        if path not in sys.path:
            sys.path.insert(0, path)
    if "pkg_resources" in sys.modules:  # imported before the paths existed: tell its working set
        import pkg_resources
        for path in paths:
            pkg_resources.working_set.add_entry(path)
    import orion  # noqa: F401
    import orion.algo.space  # noqa: F401
    import orion.algo.base  # noqa: F401
    import orion.core.worker.transformer  # noqa: F401
    return sys.modules["orion"]
