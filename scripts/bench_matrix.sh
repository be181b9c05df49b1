#!/usr/bin/env bash
# Every strategy (and both parameter-gather schedules) through bench.py on N GPUs of this box; one line per run.
#   scripts/bench_matrix.sh N [steps] > profiles/round2/matrix_nN.txt
set -uo pipefail
cd "$(dirname "${BASH_SOURCE[0]}")/.."
N="${1:?number of GPUs}"; STEPS="${2:-20}"
PORT=$((29600 + RANDOM % 300))
run() {  # name, extra env, args...
  local name="$1" envs="$2"; shift 2
  PORT=$((PORT + 1))
  local out
  if [ "$N" -gt 1 ]; then
    out=$(env $envs timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 \
          --master-port "$PORT" bench.py --gpus "$N" --steps "$STEPS" --warmup 5 --no-extras "$@" 2>/dev/null | tail -1)
  else
    out=$(env $envs timeout 120 python bench.py --gpus 1 --steps "$STEPS" --warmup 5 --no-extras "$@" 2>/dev/null | tail -1)
  fi
  python - "$name" "$out" <<'PY'
import json, sys
name, line = sys.argv[1], sys.argv[2]
try:
    d = json.loads(line)
    c = d.get("clocks", {})
    print(f"{name:28s} {d['value']:12.0f} tok/s  {d['ms_per_step']:7.3f} ms/step  e2e {d['e2e']['value']:12.0f}  "
          f"peak {d.get('peak_vram_gb', 0):6.2f} GB  sm {c.get('sm_mhz')} MHz {','.join(c.get('reasons') or []) or '-'}")
except Exception as e:
    print(f"{name:28s} FAILED ({e}): {line[:200]}")
PY
}
run "ddp" "X=1" --strategy ddp
if [ "$N" -gt 1 ]; then
  run "ddp nccl schedule" "X=1" --strategy ddp --comm nccl
  run "ddp p2p (no NVLS)" "B200BENCH_NVLS=0" --strategy ddp
fi
run "zero2" "X=1" --strategy zero2
if [ "$N" -gt 1 ]; then
  run "fsdp gather=gemm (K14)" "B200BENCH_PARAM_GATHER=gemm" --strategy fsdp
  run "fsdp gather=stream" "B200BENCH_PARAM_GATHER=stream" --strategy fsdp
  run "fsdp nccl schedule" "X=1" --strategy fsdp --comm nccl
  run "zero3 gather=gemm (K14)" "B200BENCH_PARAM_GATHER=gemm" --strategy zero3
  run "zero3 gather=stream" "B200BENCH_PARAM_GATHER=stream" --strategy zero3
  run "zero3 nccl schedule" "X=1" --strategy zero3 --comm nccl
fi
