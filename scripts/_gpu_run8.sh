set -x
N=${1:-8}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 400 $T tests/multi/check_comm.py --bench > gpurun_out/comm_n$N.log 2>&1; echo "exit=$?" >> gpurun_out/comm_n$N.log
grep -E "check|bench|Error|exit=" gpurun_out/comm_n$N.log | tail -40
for sc in "ddp fused" "ddp nccl" "zero2 fused" "fsdp fused" "zero3 fused"; do set -- $sc
timeout 300 $T bench.py --gpus $N --steps 10 --warmup 3 --strategy $1 --comm $2 > gpurun_out/bench_n${N}_$1_$2.log 2>&1; tail -1 gpurun_out/bench_n${N}_$1_$2.log | cut -c1-140
done
timeout 400 $T bench.py --impl reference --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_ref_n$N.log 2>&1; tail -1 gpurun_out/bench_ref_n$N.log | cut -c1-200
