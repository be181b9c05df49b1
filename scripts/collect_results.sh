#!/usr/bin/env bash
# collect_results.sh <namespace|local> <job> <outdir>
# Turns one run's log into <outdir>/<job>_results/result.json by cutting the JSON out from between
# the BENCHMARK_RESULT_JSON_START / _END markers the harness prints (the marker protocol and the
# <job>_results/result.json layout are the reference's, scripts/collect_results.sh:46-52 there, so
# parse_metrics.py finds the files either way).
#   namespace = "local": the log is <outdir>/<job>.log (written by run_all_benchmarks.sh)
#   otherwise          : the log is fetched with kubectl from the job's pod in that namespace
set -euo pipefail
NS="${1:?namespace or 'local'}"; JOB="${2:?job name}"; OUT="${3:-./results}"
DEST="$OUT/${JOB}_results"
mkdir -p "$DEST"
LOG="$OUT/${JOB}.log"
if [ "$NS" != "local" ]; then
  POD="$(kubectl get pods -n "$NS" -l "job-name=$JOB" -o jsonpath='{.items[0].metadata.name}' 2>/dev/null || true)"
  if [ -z "$POD" ]; then echo "ERROR: no pod found for job $JOB in namespace $NS" >&2; exit 1; fi
  kubectl logs -n "$NS" "$POD" > "$LOG"
  echo "logs: $LOG (pod $POD)"
fi
[ -s "$LOG" ] || { echo "ERROR: empty or missing log $LOG" >&2; exit 1; }
# extract to a candidate file first; it only becomes result.json if it parses as JSON
awk '/BENCHMARK_RESULT_JSON_END/{on=0} on{print} /BENCHMARK_RESULT_JSON_START/{on=1}' "$LOG" > "$DEST/result.json.candidate"
if python -c "import json,sys; json.load(open(sys.argv[1]))" "$DEST/result.json.candidate" 2>/dev/null; then
  mv "$DEST/result.json.candidate" "$DEST/result.json"
  echo "result: $DEST/result.json"
else
  echo "ERROR: no valid result JSON between the markers in $LOG (candidate kept as result.json.candidate)" >&2
  exit 2
fi
