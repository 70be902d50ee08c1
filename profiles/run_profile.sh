#!/bin/bash
# Profiling recipe of this repo (run under gpurun on one B200); outputs land in gpurun_out/.
#   $1 = tag (e.g. r01a)   $2 = kernel regex for the --set full capture   $3 = launches to skip  $4 = launches to capture
set -u
TAG=${1:-r01}
KRE=${2:-"join_count|join_emit"}
SKIP=${3:-24}
CNT=${4:-8}
OUT=gpurun_out
mkdir -p $OUT
CMD="python bench.py --steps 1 --warmup 3 --e2e-steps 1 --no-cpu --parts= --min-seconds 0"
$CMD > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu_launches.log 2>&1
$CMD > $OUT/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:$KRE" -s $SKIP -c $CNT -f -o $OUT/${TAG}_full $CMD > $OUT/${TAG}_ncu_full.log 2>&1
ls -la $OUT
