#!/usr/bin/env bash
# 1-GPU smoke run: the reference's smoke configuration (ddp, ws 1, seq 2048, tier A, 50 steps, warm-up
# 5, batch 1, accum 4; k8s/job-smoke-1gpu.yaml:26-51 there).  --k8s submits k8s/job-smoke-1gpu.yaml.
set -euo pipefail
cd "$(dirname "${BASH_SOURCE[0]}")/.."
if [ "${1:-}" = "--k8s" ]; then
  IMAGE="${IMAGE:-b200bench:latest}"
  kubectl apply -f k8s/namespace.yaml -f k8s/serviceaccount.yaml
  sed "s|{{IMAGE}}|$IMAGE|g" k8s/job-smoke-1gpu.yaml | kubectl apply -f -
  echo "submitted bench-smoke-1gpu; collect with: scripts/collect_results.sh bench bench-smoke-1gpu ./results"
  exit 0
fi
OUT="${RESULTS_DIR:-./results}"; JOB=bench-smoke-1gpu
mkdir -p "$OUT/${JOB}_results"
python benchmarking/train_harness.py --strategy ddp --world-size 1 --rank 0 --tier A --seq-len 2048 --synthetic \
    --steps "${STEPS:-50}" --warmup-steps 5 --per-device-batch 1 --grad-accum 4 \
    --results-dir "$OUT/${JOB}_results" | tee "$OUT/$JOB.log"
scripts/collect_results.sh local "$JOB" "$OUT"
