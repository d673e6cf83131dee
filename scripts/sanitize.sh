#!/bin/bash
# compute-sanitizer over the small-shape sweep (SURVEY.md section 5 "race detection / sanitizers").
#   scripts/sanitize.sh memcheck|racecheck|initcheck|synccheck   -> gpurun_out/sanitize_<tool>.log
# ONE tool per gpurun call (B200_PROFILING.md).  The sweep is run plainly first; the tool only if that exits 0.
set -u
TOOL=${1:-memcheck}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export B2D_DEBUG_KNOBS=1 B2D_HOST_CHUNK_MB=1
python scripts/sanitize_cases.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/sanitize_plain.log; exit 1; }
timeout 780 compute-sanitizer --tool "$TOOL" --error-exitcode 3 --print-limit 40 \
    python scripts/sanitize_cases.py > "gpurun_out/sanitize_${TOOL}.log" 2>&1
echo "compute-sanitizer --tool $TOOL exit code $?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize_cases:" "gpurun_out/sanitize_${TOOL}.log" | tail -5
grep -c "=========     at " "gpurun_out/sanitize_${TOOL}.log" || true
