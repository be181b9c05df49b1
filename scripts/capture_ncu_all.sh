#!/usr/bin/env bash
# One `ncu --set full` capture per single-GPU kernel family, taken from a REAL eager training step at the headline
# shape (tier A, seq 2048, batch 1), summarised into profiles/<name>_ncu.md.  One GPU only (ncu replays every
# kernel ~40 times; never wrap a multi-rank command).  Run the plain command first and only profile if it exits 0.
#
#   scripts/capture_ncu_all.sh                  # every family below (~6-8 minutes of GPU time)
#   scripts/capture_ncu_all.sh adamw layernorm  # a subset
#
# Families -> kernel-name regex, launch to skip to (a steady-state launch, past the two eager warm-up steps).
set -uo pipefail
cd "$(dirname "${BASH_SOURCE[0]}")/.."
OUT="${OUT:-gpurun_out/ncu}"
mkdir -p "$OUT" profiles
declare -A REGEX=(
  [gemm]="gemm_bf16_tcgen05_kernel"   [attn_fwd]="attn_fwd_kernel"        [attn_bwd]="attn_bwd_kernel"
  [layernorm_fwd]="layernorm_fwd_kernel" [layernorm_bwd]="layernorm_bwd_kernel" [colreduce]="colreduce_kernel"
  [cross_entropy]="cross_entropy_kernel" [adamw]="adamw_kernel"           [embedding]="embedding_(fwd|bwd)_kernel"
  [dropout_apply]="dropout_apply_kernel"
)
FAMILIES=("$@")
[ ${#FAMILIES[@]} -eq 0 ] && FAMILIES=(gemm attn_fwd attn_bwd layernorm_fwd layernorm_bwd colreduce cross_entropy adamw embedding dropout_apply)

# the workload: 3 eager steps (no CUDA graph, so every kernel is a separate launch ncu can pick)
WORK=(python -m b200bench.cli --strategy ddp --world-size 1 --rank 0 --tier A --seq-len 2048 --synthetic --steps 3
      --warmup-steps 1 --per-device-batch 1 --grad-accum 1 --results-dir "$OUT/run" --no-graph --log-interval 0)
echo "== plain run first =="
"${WORK[@]}" > "$OUT/plain.log" 2>&1 || { echo "plain run failed, not profiling"; tail -5 "$OUT/plain.log"; exit 1; }

FAIL=0
for fam in "${FAMILIES[@]}"; do
  rx="${REGEX[$fam]:-}"
  [ -z "$rx" ] && { echo "unknown family $fam"; FAIL=$((FAIL + 1)); continue; }
  echo "== $fam ($rx) =="
  # skip the launches of the first (warm-up) step: 40 launches of any one family is past it for every family here
  timeout 600 ncu --set full --clock-control none --import-source on -k "regex:$rx" --launch-skip 40 --launch-count 1 \
      -f -o "$OUT/$fam" "${WORK[@]}" > "$OUT/$fam.log" 2>&1
  if [ -s "$OUT/$fam.ncu-rep" ]; then
    python scripts/ncu_summary.py "$OUT/$fam.ncu-rep" > "profiles/${fam}_ncu.md" && echo "   -> profiles/${fam}_ncu.md"
  else
    # families with fewer than 41 launches in three steps (cross-entropy, embedding, adamw): take the last one instead
    timeout 600 ncu --set full --clock-control none --import-source on -k "regex:$rx" --launch-skip 2 --launch-count 1 \
        -f -o "$OUT/$fam" "${WORK[@]}" > "$OUT/$fam.log" 2>&1
    if [ -s "$OUT/$fam.ncu-rep" ]; then
      python scripts/ncu_summary.py "$OUT/$fam.ncu-rep" > "profiles/${fam}_ncu.md" && echo "   -> profiles/${fam}_ncu.md"
    else
      echo "   no report (see $OUT/$fam.log)"; FAIL=$((FAIL + 1))
    fi
  fi
done
echo "capture_ncu_all: $FAIL famil(ies) without a report"
[ "$FAIL" -eq 0 ]
