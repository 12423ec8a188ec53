#!/usr/bin/env bash
# One ncu capture of one kernel, the way the profiling recipe wants it (run inside ONE gpurun call, one GPU):
#   1. the case runs to completion WITHOUT ncu first (exit 0 or nothing is captured),
#   2. ncu --set full --clock-control none --import-source on, one launch of the named kernel,
#   3. the report is summarised on the box and deleted (gpurun_out is merged back only while it stays small).
#
#   scripts/ncu_capture.sh <name> <kernel-regex> <skip-launches> <ncu_case.py arguments...>
#   scripts/ncu_capture.sh heat3d_cluster step3d_cluster 1 heat3d --depth 128 --height 512 --width 1024 \
#                          --generations 1 --cluster3d
# Output: gpurun_out/last/<name>.plain.log, <name>.ncu.log, prof_<name>.summary.txt (copy the summary to profiles/).
set -u
name=$1; regex=$2; skip=$3; shift 3
O=gpurun_out/last
mkdir -p "$O"
timeout 60 python scripts/ncu_case.py "$@" > "$O/$name.plain.log" 2>&1 || { echo "case failed without ncu; not profiling"; exit 1; }
timeout 300 ncu --set full --clock-control none --import-source on -k "regex:$regex" -s "$skip" -c 1 \
    -o "$O/prof_$name" -f python scripts/ncu_case.py "$@" > "$O/$name.ncu.log" 2>&1
python scripts/ncu_summary.py "$O/prof_$name.ncu-rep" > "$O/prof_$name.summary.txt" 2>&1
rm -f "$O"/*.ncu-rep
