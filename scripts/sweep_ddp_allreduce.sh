#!/usr/bin/env bash
# ddp all-reduce knobs on N GPUs: CTA counts, blocks per bucket, NVLS on / off (one bench.py line each).
#   scripts/sweep_ddp_allreduce.sh N
run(){ PORT=$((29700 + RANDOM % 200)); env "$@" timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --steps 20 --warmup 5 --no-extras 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('%-60s %.3f ms' % (sys.argv[1], d['ms_per_step']))" "$*"; }
N=${1:-2}
for c in 16 32 64; do run B200BENCH_NVLS=0 B200BENCH_COMM_CTAS=$c; done
for b in 2 4 8; do run B200BENCH_NVLS=0 B200BENCH_BLOCKS_PER_BUCKET=$b; done
for c in 8 16 32; do run B200BENCH_NVLS=1 B200BENCH_NVLS_CTAS=$c; done
run B200BENCH_NVLS=1 B200BENCH_NVLS_CTAS=16 B200BENCH_BLOCKS_PER_BUCKET=4
run B200BENCH_OPT_IN_BACKWARD=0 B200BENCH_NVLS=0
