#!/bin/bash
# compute-sanitizer over every kernel of the library (SURVEY §5 "Race detection / sanitizers").
#   tools/sanitize.sh memcheck|racecheck|synccheck|initcheck [driver sections...]
# ONE tool per gpurun call (B200_PROFILING.md: the four tools in one call left a GPU unusable).  The plain run
# comes first and must exit 0.  Output: gpurun_out/sanitize_<tool>.log (keep the tail under profiles/).
TOOL=${1:-memcheck}; shift
mkdir -p gpurun_out
LOG=gpurun_out/sanitize_${TOOL}.log
timeout 300 python tools/sanitize_driver.py "$@" > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/sanitize_plain.log; exit 1; }
EXTRA=""
[ "$TOOL" = racecheck ] && EXTRA="--racecheck-report all"
[ "$TOOL" = initcheck ] && EXTRA="--track-unused-memory no"
timeout 1500 compute-sanitizer --tool $TOOL $EXTRA --launch-timeout 0 --print-limit 30 --error-exitcode 9 \
    python tools/sanitize_driver.py "$@" > $LOG 2>&1
RC=$?
echo "compute-sanitizer --tool $TOOL rc=$RC" | tee -a $LOG
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|SANITIZE DRIVER OK|ok$" $LOG | tail -12
exit $RC
