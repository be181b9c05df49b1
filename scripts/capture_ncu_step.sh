#!/usr/bin/env bash
# ONE `ncu --set full` invocation over a window of a real eager training step at the headline shape (tier A, seq 2048,
# batch 1, dropout 0.1): the forward GEMMs and attention of the last block, the head, and the backward of the last
# block (GEMMs with bias-gradient / rowdot / multiply epilogues, attention backward, fused LayerNorm backward).
# One GPU, after the plain run has exited 0.  Read the report with scripts/ncu_table.py / ncu_summary.py.
set -uo pipefail
cd "$(dirname "${BASH_SOURCE[0]}")/.."
OUT="${OUT:-gpurun_out/ncu}"
mkdir -p "$OUT"
WORK=(python -m b200bench.cli --strategy ddp --world-size 1 --rank 0 --tier A --seq-len 2048 --synthetic --steps 3
      --warmup-steps 1 --per-device-batch 1 --grad-accum 1 --results-dir "$OUT/run" --no-graph --log-interval 0)
"${WORK[@]}" > "$OUT/plain.log" 2>&1 || { echo "plain run failed, not profiling"; tail -5 "$OUT/plain.log"; exit 1; }
RX='regex:gemm_bf16_tcgen05_kernel|attn_fwd_kernel|attn_bwd_kernel|layernorm_bwd_fused_kernel|attn_bwd_convert'
# matching launches per step: 61 forward + ~135 backward; skip the first step and the forward of blocks 0..10
timeout "${NCU_TIMEOUT:-420}" ncu --set full --clock-control none --import-source on -k "$RX" \
    --launch-skip "${SKIP:-251}" --launch-count "${COUNT:-22}" -f -o "$OUT/step" "${WORK[@]}" > "$OUT/step.log" 2>&1
echo "ncu rc=$?"; ls -la "$OUT"/step.ncu-rep 2>/dev/null || tail -5 "$OUT/step.log"
