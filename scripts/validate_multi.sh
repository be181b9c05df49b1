#!/usr/bin/env bash
# Multi-GPU validation of one build on ONE host: fused-collective checks, then bench.py for every
# strategy, then the unmodified reference arm.  Everything lands in $OUT (default gpurun_out/nN).
#   scripts/validate_multi.sh 8            # all strategies + reference
#   STRATEGIES="ddp" REF=0 scripts/validate_multi.sh 2
set -uo pipefail
cd "$(dirname "${BASH_SOURCE[0]}")/.."
N="${1:?usage: validate_multi.sh N_GPUS}"
OUT="${OUT:-gpurun_out/n$N}"
STEPS="${STEPS:-40}"
STRATEGIES="${STRATEGIES:-ddp zero2 fsdp zero3}"
REF="${REF:-1}"
CHECK="${CHECK:-1}"
DIAG="${DIAG:-1}"
PORT=29540
mkdir -p "$OUT"
# run <timeout-seconds> <script> [args...]: torchrun under a hard time limit, fresh port each time
run() { local t=$1; shift; PORT=$((PORT + 1)); timeout "$t" python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port "$PORT" "$@"; }
FAIL=0

if [ "$CHECK" = "1" ]; then
  run 600 tests/multi/check_comm.py > "$OUT/check_comm.log" 2>&1 || FAIL=$((FAIL + 1))
  grep -E "^\[check\]" "$OUT/check_comm.log" | tail -12
fi
for s in $STRATEGIES; do
  extra=""
  if [ "$DIAG" = "1" ] && [ "$s" = "ddp" ]; then extra="--diag"; fi
  # shellcheck disable=SC2086
  run 300 bench.py --gpus "$N" --steps "$STEPS" --warmup 5 --strategy "$s" $extra \
      > "$OUT/bench_${s}_fused.json" 2> "$OUT/bench_${s}_fused.err" || FAIL=$((FAIL + 1))
  python - "$OUT/bench_${s}_fused.json" "$s" <<'EOF'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(f"[bench] {sys.argv[2]:6s} fused: {d['value']:,.0f} tok/s  {d['ms_per_step']:.2f} ms/step  e2e {d['e2e']['value']:,.0f}")
except Exception as e:
    print(f"[bench] {sys.argv[2]}: no result ({e})")
EOF
  grep -h "^\[diag\]" "$OUT/bench_${s}_fused.err" | head -8
done
if [ "$REF" = "1" ]; then
  run 400 bench.py --impl reference --gpus "$N" --steps 20 --warmup 5 --ref-timeout 240 \
      > "$OUT/bench_reference.json" 2> "$OUT/bench_reference.err" || FAIL=$((FAIL + 1))
  tail -n 1 "$OUT/bench_reference.json" | cut -c1-400
fi
echo "validate_multi: $FAIL failing stage(s)"
[ "$FAIL" -eq 0 ]
