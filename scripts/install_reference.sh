#!/bin/bash
# Place the UNMODIFIED reference service (its `app` package) in git-ignored baseline/_ref so that the
# boundary-A tests and bench.py's WebSocket-level run can drive the reference's own server
# (app/core/websocket_server_vllm.py) on the GPU box, where /root/reference does not exist.
# Used only by tests/ and bench.py — never by the product.  baseline/_ref is NOT gpurun-ignored: it travels.
#
#   scripts/install_reference.sh [/root/reference]
#
# 1st choice: pip install --target (what the round contract names).  The reference builds with hatchling,
# which is not on this image (no index access), so that fails here; 2nd choice: copy the pure-Python package
# directory as it is — exactly what its wheel would contain ([tool.hatch.build.targets.wheel] packages = ["app"]).
set -u
REF=${1:-/root/reference}
ROOT=$(cd "$(dirname "$0")/.." && pwd)
DST=$ROOT/baseline/_ref
[ -d "$REF/app" ] || { echo "install_reference: no reference checkout at $REF"; exit 1; }
rm -rf "$DST"; mkdir -p "$DST"
TMP=$(mktemp -d); cp -r "$REF/." "$TMP/"          # the build writes into the source tree; /root/reference is read-only
if python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
      --target "$DST" "$TMP" > "$ROOT/baseline/install.log" 2>&1 && [ -d "$DST/app" ]; then
  echo "install_reference: pip install --target ok"; HOW=pip
else
  cp -r "$REF/app" "$DST/app"
  find "$DST" -name __pycache__ -type d -exec rm -rf {} +
  echo "install_reference: pip failed ($(grep -m1 -o "No module named '[a-z_]*'" "$ROOT/baseline/install.log")); copied the package directory"; HOW=copy
fi
rm -rf "$TMP"
(cd "$REF" && find app -name '*.py' | sort | xargs sha256sum) > "$DST/SOURCE_SHA256"
echo "how=$HOW from=$REF" > "$DST/INSTALLED"
