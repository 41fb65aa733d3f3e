#!/bin/bash
# Install the UNMODIFIED reference (Orion) into baseline/_ref (git-ignored; travels to the GPU box).
#   * /root/reference is read-only and setup.py writes egg-info -> install from a copy under /tmp
#   * --no-deps: pymongo / appdirs are not in the offline wheelhouse (oracle/ref_loader.py shims them in memory)
#   * build-time shims (PYTHONPATH for the pip subprocess only, nothing of them is installed):
#       - baseline/_shim/sitecustomize.py: versioneer.py needs two configparser names Python 3.12 removed
#       - empty `appdirs` / `pytest-runner` dist-info: setup.py lists them in `setup_requires`, which
#         setuptools would otherwise try to download
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
SRC="${1:-/root/reference}"
TMP="$(mktemp -d /tmp/orion_ref.XXXXXX)"
SHIM="$(mktemp -d /tmp/orion_shim.XXXXXX)"
cp -r "$SRC"/. "$TMP"/
cp "$HERE/_shim/sitecustomize.py" "$SHIM"/
for dist in appdirs-1.4.4 pytest_runner-6.0.1; do
    mkdir -p "$SHIM/$dist.dist-info"
    printf 'Metadata-Version: 2.1\nName: %s\nVersion: %s\n' "${dist%-*}" "${dist##*-}" > "$SHIM/$dist.dist-info/METADATA"
done
rm -rf "$HERE/_ref"
PYTHONPATH="$SHIM${PYTHONPATH:+:$PYTHONPATH}" python -m pip install --no-index --no-build-isolation --no-deps \
    --find-links /opt/wheelhouse --target "$HERE/_ref" "$TMP"
rm -rf "$TMP" "$SHIM"
ls "$HERE/_ref"
