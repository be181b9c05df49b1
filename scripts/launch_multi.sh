#!/usr/bin/env bash
# Launch ONE multi-GPU configuration.  Same flags and defaults as the reference's launch_multi.sh
# (--strategy --world-size --seq-len --tier --steps --per-device-batch --grad-accum --image; steps 100,
# batch 1, accum 4; world size >= 2), two back-ends:
#   default          local host: one process per GPU through python -m b200bench.launch
#   --k8s            one pod holding all GPUs of the host (k8s/job-multi.template.yaml); the image
#                    entrypoint starts the same local launcher inside the pod
set -euo pipefail
cd "$(dirname "${BASH_SOURCE[0]}")/.."
STRATEGY=ddp; WORLD_SIZE=2; SEQ_LEN=2048; TIER=A; STEPS=100; PER_DEVICE_BATCH=1; GRAD_ACCUM=4
IMAGE="${IMAGE:-b200bench:latest}"; K8S=0; NAMESPACE="${NAMESPACE:-bench}"; RESULTS_DIR="${RESULTS_DIR:-./results}"
while [ $# -gt 0 ]; do
  case "$1" in
    --strategy) STRATEGY=$2; shift 2 ;;
    --world-size) WORLD_SIZE=$2; shift 2 ;;
    --seq-len) SEQ_LEN=$2; shift 2 ;;
    --tier) TIER=$2; shift 2 ;;
    --steps) STEPS=$2; shift 2 ;;
    --per-device-batch) PER_DEVICE_BATCH=$2; shift 2 ;;
    --grad-accum) GRAD_ACCUM=$2; shift 2 ;;
    --image) IMAGE=$2; shift 2 ;;
    --k8s) K8S=1; shift ;;
    *) echo "unknown argument: $1" >&2; exit 1 ;;
  esac
done
if [ "$WORLD_SIZE" -lt 2 ]; then echo "ERROR: --world-size must be >= 2 (use launch_smoke.sh for one GPU)" >&2; exit 1; fi
JOB="bench-${STRATEGY}-ws${WORLD_SIZE}-seq${SEQ_LEN}-tier${TIER}"
if [ "$K8S" = 1 ]; then
  kubectl apply -f k8s/namespace.yaml -f k8s/serviceaccount.yaml
  sed -e "s|{{JOB}}|$JOB|g" -e "s|{{STRATEGY}}|$STRATEGY|g" -e "s|{{WORLD_SIZE}}|$WORLD_SIZE|g" \
      -e "s|{{SEQ_LEN}}|$SEQ_LEN|g" -e "s|{{TIER}}|$TIER|g" -e "s|{{STEPS}}|$STEPS|g" \
      -e "s|{{PER_DEVICE_BATCH}}|$PER_DEVICE_BATCH|g" -e "s|{{GRAD_ACCUM}}|$GRAD_ACCUM|g" \
      -e "s|{{IMAGE}}|$IMAGE|g" k8s/job-multi.template.yaml | kubectl apply -n "$NAMESPACE" -f -
  echo "submitted $JOB; collect with: scripts/collect_results.sh $NAMESPACE $JOB $RESULTS_DIR"
  exit 0
fi
CFG=()
case "$STRATEGY" in
  zero2|zero3) CFG=(--deepspeed-config "configs/deepspeed/${STRATEGY}.json") ;;
  fsdp)        CFG=(--fsdp-config configs/fsdp/fsdp_config.yaml) ;;
esac
mkdir -p "$RESULTS_DIR/${JOB}_results"
python -m b200bench.launch --nproc "$WORLD_SIZE" --timeout "${TIMEOUT:-900}" -- \
    --strategy "$STRATEGY" --tier "$TIER" --seq-len "$SEQ_LEN" --synthetic --steps "$STEPS" \
    --per-device-batch "$PER_DEVICE_BATCH" --grad-accum "$GRAD_ACCUM" \
    --results-dir "$RESULTS_DIR/${JOB}_results" "${CFG[@]}" | tee "$RESULTS_DIR/${JOB}.log"
scripts/collect_results.sh local "$JOB" "$RESULTS_DIR"
