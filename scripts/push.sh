#!/usr/bin/env bash
# Pushes the image built by scripts/build.sh.  IMAGE must name a registry you are logged in to.
set -euo pipefail
: "${IMAGE:?set IMAGE=registry/path:tag}"
docker tag b200bench:latest "$IMAGE" 2>/dev/null || true
docker push "$IMAGE"
