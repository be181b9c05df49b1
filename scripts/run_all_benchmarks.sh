#!/usr/bin/env bash
# Benchmark suite for ONE multi-GPU host: runs every row of the matrix through the local launcher
# (python -m b200bench.launch), keeps each run's log + result.json, then builds metrics.csv, the
# plots and BENCHMARK_REPORT.md.  Same matrix row format as the reference's suite
# ("STRATEGY WORLD_SIZE SEQ_LEN TIER STEPS", scripts/run_all_benchmarks.sh:32-52 there), but with the
# 1-GPU rows the reference lacks (so scaling efficiency is relative to one GPU), a per-run timeout
# enforced by the launcher, and a NON-ZERO exit code when any row fails.
#
#   scripts/run_all_benchmarks.sh                 # full matrix on all visible GPUs
#   GPUS="1 2" STEPS=30 scripts/run_all_benchmarks.sh
#   MATRIX_FILE=my_matrix.txt scripts/run_all_benchmarks.sh
set -uo pipefail
cd "$(dirname "${BASH_SOURCE[0]}")/.."

RESULTS_DIR="${RESULTS_DIR:-./results}"
TIMEOUT="${TIMEOUT:-900}"
STEPS="${STEPS:-100}"
WARMUP_STEPS="${WARMUP_STEPS:-5}"
PER_DEVICE_BATCH="${PER_DEVICE_BATCH:-1}"
GRAD_ACCUM="${GRAD_ACCUM:-4}"
SEQ_LEN="${SEQ_LEN:-2048}"
TIER="${TIER:-A}"
EXTRA_ARGS="${EXTRA_ARGS:-}"
NGPU="$(python -c 'import torch; print(torch.cuda.device_count())')"
GPUS="${GPUS:-$(for n in 1 2 4 8; do [ "$n" -le "$NGPU" ] && printf '%s ' "$n"; done)}"

BENCHMARKS=()
if [ -n "${MATRIX_FILE:-}" ]; then
  while read -r line; do [[ -z "$line" || "$line" == \#* ]] || BENCHMARKS+=("$line"); done < "$MATRIX_FILE"
else
  for s in ddp fsdp zero2 zero3; do for n in $GPUS; do BENCHMARKS+=("$s $n $SEQ_LEN $TIER $STEPS"); done; done
  # long-sequence rows (the reference keeps these commented out: they do not fit a 24 GB A10)
  if [ "${LONG_SEQ:-0}" = "1" ]; then
    for n in $GPUS; do BENCHMARKS+=("fsdp $n 4096 B $STEPS" "zero3 $n 8192 B $STEPS"); done
  fi
fi

echo "=================================================================="
echo "  b200bench suite: ${#BENCHMARKS[@]} configurations on $NGPU GPU(s)"
echo "  results: $RESULTS_DIR   timeout: ${TIMEOUT}s per run"
echo "=================================================================="
mkdir -p "$RESULTS_DIR"
DONE=0; FAILED=0; SKIPPED=0; T0=$(date +%s)

run_one() {
  local strategy=$1 ws=$2 seq=$3 tier=$4 steps=$5
  local job="bench-${strategy}-ws${ws}-seq${seq}-tier${tier}"
  local out="$RESULTS_DIR/${job}_results"
  if [ "$ws" -gt "$NGPU" ]; then
    echo "[skip] $job: needs $ws GPUs, $NGPU visible"; SKIPPED=$((SKIPPED + 1)); return 0
  fi
  mkdir -p "$out"
  local cfg=()
  case "$strategy" in
    zero2|zero3) cfg=(--deepspeed-config "configs/deepspeed/${strategy}.json") ;;
    fsdp)        cfg=(--fsdp-config configs/fsdp/fsdp_config.yaml) ;;
  esac
  echo "[run ] $job"
  # shellcheck disable=SC2086
  python -m b200bench.launch --nproc "$ws" --timeout "$TIMEOUT" --log-dir "$out/ranks" -- \
      --strategy "$strategy" --tier "$tier" --seq-len "$seq" --synthetic --steps "$steps" \
      --warmup-steps "$WARMUP_STEPS" --per-device-batch "$PER_DEVICE_BATCH" --grad-accum "$GRAD_ACCUM" \
      --results-dir "$out" "${cfg[@]}" $EXTRA_ARGS > "$RESULTS_DIR/${job}.log" 2>&1
  local rc=$?
  if [ $rc -eq 0 ] && scripts/collect_results.sh local "$job" "$RESULTS_DIR" > /dev/null; then
    local tps
    tps=$(python -c "import json,sys; print(f\"{json.load(open(sys.argv[1]))['tokens_per_sec']:,.0f}\")" "$out/result.json")
    echo "[ ok ] $job: $tps tokens/s"; DONE=$((DONE + 1))
  else
    if [ $rc -eq 124 ]; then echo "[fail] $job: timed out after ${TIMEOUT}s"; else echo "[fail] $job: exit code $rc"; fi
    tail -n 15 "$RESULTS_DIR/${job}.log" | sed 's/^/        /'
    FAILED=$((FAILED + 1))
  fi
}

for row in "${BENCHMARKS[@]}"; do
  # shellcheck disable=SC2086
  run_one $row
done

echo ""
echo "completed: $DONE   failed: $FAILED   skipped: $SKIPPED   wall: $(( $(date +%s) - T0 ))s"
if [ "$DONE" -gt 0 ]; then
  python scripts/parse_metrics.py --results-dir "$RESULTS_DIR" --out "$RESULTS_DIR/summary" &&
  python scripts/plot.py --results "$RESULTS_DIR/summary/metrics.csv" --out "$RESULTS_DIR/summary/plots" &&
  python scripts/make_report.py --csv "$RESULTS_DIR/summary/metrics.csv" --out "$RESULTS_DIR/summary" || FAILED=$((FAILED + 1))
fi
[ "$FAILED" -eq 0 ]
