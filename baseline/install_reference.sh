#!/usr/bin/env bash
# Installs the UNMODIFIED reference into baseline/_ref (git-ignored) for bench.py --impl reference.
# The reference is a script repository without setup.py/pyproject.toml, so `pip install` cannot
# build it (outcome recorded in DESIGN.md); its harness and configs are copied byte for byte instead
# and run through the reference's own CLI (benchmarking/train_harness.py).
set -u
REF="${1:-/root/reference}"
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
DST="$HERE/_ref"
mkdir -p "$DST"
if python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse \
      --target "$DST" "$REF" >"$DST/pip_install.log" 2>&1; then
  echo "pip install succeeded"
else
  echo "pip install failed ($(tail -1 "$DST/pip_install.log")); copying the reference tree verbatim"
  cp -r "$REF/benchmarking" "$REF/configs" "$DST/"
  cp "$REF/LICENSE" "$DST/" 2>/dev/null || true
fi
( cd "$DST" && find benchmarking configs -type f -exec sha256sum {} \; > SHA256SUMS )
echo "reference installed under $DST"
