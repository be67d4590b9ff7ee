#!/usr/bin/env bash
# Installs the UNMODIFIED reference into baseline/_ref (git-ignored, travels with gpurun).
# The reference ships no setup.py / pyproject.toml, so the plain command fails
# ("Directory '/root/reference' is not installable"); the fallback installs from a /tmp copy
# that only ADDS a three-line pyproject.toml naming the script as a py-module.  The installed
# dataParallelTraining_NN_MPI.py is byte-identical to the reference's (checked with cmp).
set -u
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${1:-/root/reference}"
TARGET="$HERE/_ref"
rm -rf "$TARGET"
if python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse \
      --target "$TARGET" "$REF" >/tmp/ref_install.log 2>&1; then
  echo "installed directly"
else
  echo "direct install failed: $(tail -1 /tmp/ref_install.log)"
  TMP="$(mktemp -d /tmp/refcopy.XXXX)"
  cp "$REF"/* "$TMP"/
  cat > "$TMP/pyproject.toml" <<'PY'
[build-system]
requires = ["setuptools"]
build-backend = "setuptools.build_meta"
[project]
name = "nn-parallel-training-with-mpi-reference"
version = "0.0.0"
[tool.setuptools]
py-modules = ["dataParallelTraining_NN_MPI"]
PY
  python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
      --target "$TARGET" "$TMP" >>/tmp/ref_install.log 2>&1 || { echo "fallback install failed"; exit 1; }
  echo "installed from a /tmp copy with an added pyproject.toml"
fi
cmp "$TARGET/dataParallelTraining_NN_MPI.py" "$REF/dataParallelTraining_NN_MPI.py" && echo "byte-identical to the reference"
