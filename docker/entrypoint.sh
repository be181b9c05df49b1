#!/usr/bin/env bash
# Container entrypoint: environment -> harness flags.
#
# Two modes, chosen by GPUS_PER_POD (default: every GPU visible in the container):
#   * one pod owns the whole host (GPUS_PER_POD >= 2 and WORLD_SIZE == GPUS_PER_POD): start one harness
#     process per GPU with the local launcher — the native single-box deployment;
#   * one rank per pod (GPUS_PER_POD = 1): the reference's env contract (docker/entrypoint.sh:7-36
#     there): RANK from JOB_COMPLETION_INDEX + 1 for indexed worker jobs, rank 0 advertises POD_IP.
set -euo pipefail
echo "=== b200bench entrypoint: $(date -u +%FT%TZ) ==="

: "${STRATEGY:=ddp}" "${WORLD_SIZE:=1}" "${LOCAL_RANK:=0}" "${MASTER_PORT:=29500}" "${SEQ_LEN:=2048}"
: "${TIER:=A}" "${STEPS:=50}" "${WARMUP_STEPS:=5}" "${PER_DEVICE_BATCH:=1}" "${GRAD_ACCUM:=1}"
: "${SYNTHETIC:=true}" "${RESULTS_DIR:=/results}" "${EXTRA_ARGS:=}" "${APP_DIR:=/app}"
VISIBLE_GPUS=$(nvidia-smi -L 2>/dev/null | wc -l || echo 0)
: "${GPUS_PER_POD:=$VISIBLE_GPUS}"

if [ -n "${JOB_COMPLETION_INDEX:-}" ]; then
  RANK=$((JOB_COMPLETION_INDEX + 1))        # indexed worker job: master is rank 0, workers 1..N-1
else
  : "${RANK:=0}"
fi
if [ "$RANK" = "0" ] && [ -n "${POD_IP:-}" ]; then MASTER_ADDR="$POD_IP"; else : "${MASTER_ADDR:=127.0.0.1}"; fi

printf '  %-18s %s\n' STRATEGY "$STRATEGY" WORLD_SIZE "$WORLD_SIZE" RANK "$RANK" LOCAL_RANK "$LOCAL_RANK" \
  MASTER "$MASTER_ADDR:$MASTER_PORT" TIER "$TIER" SEQ_LEN "$SEQ_LEN" STEPS "$STEPS" \
  PER_DEVICE_BATCH "$PER_DEVICE_BATCH" GRAD_ACCUM "$GRAD_ACCUM" GPUS_PER_POD "$GPUS_PER_POD"
nvidia-smi -L || echo "WARNING: nvidia-smi failed"

COMMON=(--strategy "$STRATEGY" --seq-len "$SEQ_LEN" --tier "$TIER" --steps "$STEPS"
        --warmup-steps "$WARMUP_STEPS" --per-device-batch "$PER_DEVICE_BATCH" --grad-accum "$GRAD_ACCUM"
        --results-dir "$RESULTS_DIR")
[ "$SYNTHETIC" = "true" ] && COMMON+=(--synthetic)
case "$STRATEGY" in
  zero2|zero3) COMMON+=(--deepspeed-config "$APP_DIR/configs/deepspeed/${STRATEGY}.json") ;;
  fsdp)        COMMON+=(--fsdp-config "$APP_DIR/configs/fsdp/fsdp_config.yaml") ;;
esac

cd "$APP_DIR"
if [ "$GPUS_PER_POD" -ge 2 ] && [ "$WORLD_SIZE" = "$GPUS_PER_POD" ]; then
  echo "=== single-host launch: $WORLD_SIZE ranks in this pod ==="
  # shellcheck disable=SC2086
  exec python -u -m b200bench.launch --nproc "$WORLD_SIZE" --timeout "${TIMEOUT:-900}" -- "${COMMON[@]}" $EXTRA_ARGS
fi
echo "=== rank-per-pod launch ==="
# shellcheck disable=SC2086
exec python -u "$APP_DIR/benchmarking/train_harness.py" --world-size "$WORLD_SIZE" --rank "$RANK" \
  --local-rank "$LOCAL_RANK" --master-addr "$MASTER_ADDR" --master-port "$MASTER_PORT" "${COMMON[@]}" $EXTRA_ARGS
