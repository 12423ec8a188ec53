#!/bin/bash
# compute-sanitizer on the single-GPU kernels, ONE tool per invocation (B200_PROFILING.md: never several tools in
# one call).  Usage under gpurun:  scripts/sanitize.sh memcheck|racecheck|synccheck
# Runs only after the same small program has exited 0 without the tool.
set -e
TOOL=${1:-memcheck}
PROG="python scripts/sanitize_case.py"
$PROG > gpurun_out/sanitize_plain.log 2>&1
compute-sanitizer --tool "$TOOL" --error-exitcode 9 --log-file gpurun_out/sanitize_${TOOL}.log $PROG > gpurun_out/sanitize_${TOOL}_stdout.log 2>&1 || echo "sanitizer exit $?"
tail -5 gpurun_out/sanitize_${TOOL}.log
