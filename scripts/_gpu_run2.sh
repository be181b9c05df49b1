set -x
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $T tests/multi/check_comm.py --bench > gpurun_out/comm2.log 2>&1; echo "exit=$?" >> gpurun_out/comm2.log
grep -E "check|bench|Error|exit=" gpurun_out/comm2.log | tail -40
for s in ddp zero2; do for c in fused nccl; do
timeout 300 $T bench.py --gpus 2 --steps 10 --warmup 3 --strategy $s --comm $c > gpurun_out/bench_n2_${s}_${c}.log 2>&1; tail -1 gpurun_out/bench_n2_${s}_${c}.log | cut -c1-140
done; done
