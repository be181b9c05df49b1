#!/usr/bin/env bash
# Pre-flight: are the GPUs this benchmark needs actually usable?
#   local (default): driver, GPU count/model, NVLink/NVSwitch topology, peer access, clocks
#   --k8s          : kubectl reachable, NVIDIA device plugin running, GPU capacity per node, GPUs in
#                    use, namespace + service account present (what the reference's script checks)
set -uo pipefail
FAIL=0
ok() { printf '  [ ok ] %s\n' "$*"; }
warn() { printf '  [warn] %s\n' "$*"; }
bad() { printf '  [FAIL] %s\n' "$*"; FAIL=1; }

if [ "${1:-}" = "--k8s" ]; then
  NS="${NAMESPACE:-bench}"
  echo "== Kubernetes GPU pre-flight (namespace $NS) =="
  if kubectl cluster-info > /dev/null 2>&1; then ok "kubectl can reach the cluster"; else bad "kubectl cannot reach the cluster"; exit 1; fi
  PLUGIN=$(kubectl get pods -A 2>/dev/null | grep -ci 'nvidia-device-plugin' || true)
  if [ "$PLUGIN" -gt 0 ]; then ok "$PLUGIN NVIDIA device-plugin pod(s)"; else bad "no NVIDIA device-plugin pods found"; fi
  echo "  GPU capacity per node:"
  kubectl get nodes -o json | jq -r '.items[] | "    \(.metadata.name): \(.status.capacity["nvidia.com/gpu"] // "0") GPU(s)"'
  TOTAL=$(kubectl get nodes -o json | jq '[.items[].status.capacity["nvidia.com/gpu"] // "0" | tonumber] | add')
  USED=$(kubectl get pods -A -o json | jq '[.items[] | select(.status.phase=="Running") | .spec.containers[].resources.limits["nvidia.com/gpu"] // "0" | tonumber] | add')
  echo "  GPUs: total $TOTAL, requested by running pods ${USED:-0}"
  if [ "${TOTAL:-0}" -ge 1 ]; then ok "at least one GPU node"; else bad "no GPU capacity in the cluster"; fi
  if kubectl get ns "$NS" > /dev/null 2>&1; then ok "namespace $NS exists"; else warn "namespace $NS missing: kubectl apply -f k8s/namespace.yaml"; fi
  if kubectl get sa bench-sa -n "$NS" > /dev/null 2>&1; then ok "service account bench-sa exists"; else warn "service account missing: kubectl apply -f k8s/serviceaccount.yaml"; fi
  exit $FAIL
fi

echo "== local GPU pre-flight =="
command -v nvidia-smi > /dev/null || { bad "nvidia-smi not found"; exit 1; }
nvidia-smi --query-gpu=index,name,memory.total,memory.used,clocks.sm,clocks.max.sm,power.limit --format=csv,noheader | sed 's/^/    /'
N=$(nvidia-smi -L | wc -l)
if [ "$N" -ge 1 ]; then ok "$N GPU(s) visible"; else bad "no GPU visible"; fi
if nvidia-smi --query-gpu=name --format=csv,noheader | grep -q B200; then ok "Blackwell B200 present"; else warn "not a B200: the native library is built for sm_100a only"; fi
if [ "$N" -gt 1 ]; then
  echo "  topology:"
  nvidia-smi topo -m 2>/dev/null | head -n $((N + 2)) | sed 's/^/    /'
  python - <<'PY'
import torch
n = torch.cuda.device_count()
bad = [(i, j) for i in range(n) for j in range(n) if i != j and not torch.cuda.can_device_access_peer(i, j)]
print("  [ ok ] all GPU pairs have peer access (needed by the fused collectives)" if not bad
      else f"  [FAIL] no peer access between {bad[:8]}: use --comm nccl")
PY
fi
SLOW=$(nvidia-smi --query-gpu=clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown --format=csv,noheader | grep -c Active || true)
if [ "$SLOW" = 0 ]; then ok "no thermal / hardware slowdown active"; else warn "$SLOW GPU(s) report an active slowdown reason"; fi
exit $FAIL
