#!/usr/bin/env bash
# Measurements that round 1 left open, cheapest first.  Each stage prints what it costs in
# box time; run them one at a time on a box with the stated number of GPUs:
#   scripts/measure_open_items.sh comm 2       # K15 / block-gather standalone numbers      (~1 min on 2 GPUs)
#   scripts/measure_open_items.sh sharded 8    # fsdp + zero3 with the block-prefetch gather (~1 min on N GPUs)
#   scripts/measure_open_items.sh tier7b 8     # Mistral-7B-shaped ZeRO-3                    (~1 min on N GPUs)
#   scripts/measure_open_items.sh gather_ab 2  # param_gather stream vs gemm, same box       (~1.5 min on 2 GPUs)
#   scripts/measure_open_items.sh library 1    # our GEMMs / attention vs cuBLAS / SDPA at the same shapes (~0.5 min)
set -uo pipefail
cd "$(dirname "${BASH_SOURCE[0]}")/.."
STAGE="${1:?stage: comm | sharded | tier7b | gather_ab | library}"
N="${2:?number of GPUs}"
OUT="${OUT:-gpurun_out/open_${STAGE}_n$N}"
mkdir -p "$OUT"
PORT=$((29700 + RANDOM % 200))
trun() { local t=$1; shift; PORT=$((PORT + 1)); timeout "$t" python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port "$PORT" "$@"; }

case "$STAGE" in
  comm)
    trun 300 tests/multi/check_comm.py --bench > "$OUT/check_comm_bench.log" 2>&1
    grep -E "^\[(bench|check)\]" "$OUT/check_comm_bench.log" | tail -60 ;;
  sharded)
    STRATEGIES="fsdp zero3" REF=0 CHECK=0 DIAG=0 OUT="$OUT" scripts/validate_multi.sh "$N" ;;
  tier7b)
    timeout 240 python -m b200bench.launch --nproc "$N" --timeout 200 --log-dir "$OUT/ranks" -- --strategy zero3 \
        --deepspeed-config configs/deepspeed/zero3.json --tier 7B --seq-len 2048 --synthetic --steps 14 \
        --warmup-steps 6 --per-device-batch 1 --grad-accum 4 --results-dir "$OUT" > "$OUT/run.log" 2>&1
    echo "exit code $?"
    grep -E '"tokens_per_sec"|"mean_step_time_sec"|"peak_vram_gb"|"mean_loss"' "$OUT/run.log" ;;
  gather_ab)
    for g in stream gemm; do
      B200BENCH_PARAM_GATHER=$g STRATEGIES="fsdp zero3" REF=0 CHECK=0 DIAG=0 OUT="$OUT/$g" scripts/validate_multi.sh "$N"
    done ;;
  library)
    python scripts/compare_library_kernels.py > "$OUT/library_compare.md" 2> "$OUT/library_compare.err"
    echo "exit code $?"; cat "$OUT/library_compare.md" ;;
  *) echo "unknown stage $STAGE" >&2; exit 1 ;;
esac
