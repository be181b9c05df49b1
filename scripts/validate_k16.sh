#!/usr/bin/env bash
# K16 (all-gather as one collective) on N GPUs: loopback test, check_comm with the K16 cases and bandwidth lines, then
# fsdp / zero3 with the gathers as peer pulls vs as K16.
N=${1:-2}
cd "$(dirname "${BASH_SOURCE[0]}")/.."
mkdir -p gpurun_out/k16_n$N
timeout 200 python -m pytest tests/test_comm_loopback_gpu.py -x -q 2>&1 | tail -3
P=$((29500 + RANDOM % 300))
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P tests/multi/check_comm.py --bench > gpurun_out/k16_n$N/check_comm.log 2>&1; echo "check_comm rc=$?"
grep -E "^\[check\]" gpurun_out/k16_n$N/check_comm.log | tail -24
grep -E "gather" gpurun_out/k16_n$N/check_comm.log | grep bench
run(){ PORT=$((29800 + RANDOM % 200)); env "$@" timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --steps 20 --warmup 5 --no-extras --strategy $S 2>gpurun_out/k16_n$N/last.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('%-8s %-40s %.3f ms  %.0f tok/s' % (sys.argv[1], sys.argv[2], d['ms_per_step'], d['value']))" $S "$*" || tail -5 gpurun_out/k16_n$N/last.err; }
for S in fsdp zero3; do
  run B200BENCH_GATHER_COLL=0
  run B200BENCH_GATHER_COLL=1
  run B200BENCH_GATHER_COLL=1 B200BENCH_GATHER_MC=0
done
