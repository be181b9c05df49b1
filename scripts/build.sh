#!/usr/bin/env bash
# Builds the native library for sm_100a, then (if docker is available) the container image.
set -euo pipefail
cd "$(dirname "${BASH_SOURCE[0]}")/.."
make -j"$(nproc)" all
if command -v docker > /dev/null; then docker build -f docker/Dockerfile -t "${IMAGE:-b200bench:latest}" .; else echo "docker not found: native library only"; fi
