#!/usr/bin/env bash
# Offline install of the UNMODIFIED reference into baseline/_ref (git-ignored).
# /root/reference has no setup.py / pyproject.toml, so `pip install /root/reference` fails with
# "Neither 'setup.py' nor 'pyproject.toml' found".  Per the task rules the install is done from a
# copy under /tmp that adds ONLY packaging metadata; multi_proc_single_gpu.py is byte-identical.
set -euo pipefail
REPO="$(cd "$(dirname "$0")/.." && pwd)"
SRC="${1:-/root/reference}"
STAGE="$(mktemp -d /tmp/ref_pkg.XXXXXX)"
cp "$SRC/multi_proc_single_gpu.py" "$SRC/README.md" "$STAGE/"
cat > "$STAGE/pyproject.toml" <<'TOML'
[build-system]
requires = ["setuptools"]
build-backend = "setuptools.build_meta"

[project]
name = "pytorch_distributed_mnist_reference"
version = "0.0.0"
description = "Packaging shim only: installs the unmodified multi_proc_single_gpu.py module"

[tool.setuptools]
py-modules = ["multi_proc_single_gpu"]
TOML
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
    --target "$REPO/baseline/_ref" --upgrade "$STAGE"
cmp "$REPO/baseline/_ref/multi_proc_single_gpu.py" "$SRC/multi_proc_single_gpu.py" && echo "reference installed, byte-identical"
