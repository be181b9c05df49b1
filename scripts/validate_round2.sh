#!/usr/bin/env bash
# Round-2 validation on N GPUs of one box: collectives check + bandwidth sweep, then the bench line with all strategies.
N=${1:-8}
cd "$(dirname "${BASH_SOURCE[0]}")/.."
mkdir -p gpurun_out/n$N
P=$((29500 + RANDOM % 300))
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P tests/multi/check_comm.py --bench > gpurun_out/n$N/check_comm.log 2>&1; echo "check_comm rc=$?"
grep -E "^\[check\]" gpurun_out/n$N/check_comm.log | tail -20
grep -E "^\[bench\] allreduce +(25|256) MiB|K15|gather_range|K16" gpurun_out/n$N/check_comm.log
P=$((P + 1))
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/n$N/bench.json 2> gpurun_out/n$N/bench.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/n$N/bench.json"))
    print("ddp", round(d["value"]), round(d["ms_per_step"], 3), d["clocks"], "e2e", round(d["e2e"]["value"]))
    for k, v in d.get("strategies", {}).items():
        print(k, {kk: (round(vv, 3) if isinstance(vv, float) else vv) for kk, vv in v.items() if kk in ("value", "ms_per_step", "peak_vram_gb", "error", "param_gather", "grad_accum")})
except Exception as e:
    print("bench parse failed", e)
PY
tail -3 gpurun_out/n$N/bench.err
[ "${VARIANTS:-1}" = 1 ] || exit 0
source /dev/stdin <<'EOS'
run(){ PORT=$((29700 + RANDOM % 200)); env "$@" timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --steps 20 --warmup 5 --no-extras 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('%-60s %.3f ms' % (sys.argv[1], d['ms_per_step']))" "$*"; }
EOS
run B200BENCH_NVLS=0
run B200BENCH_NVLS=1 B200BENCH_NVLS_CTAS=32
run B200BENCH_BLOCKS_PER_BUCKET=1
