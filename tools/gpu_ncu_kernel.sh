#!/bin/bash
# usage: tools/gpu_ncu_kernel.sh <tag> <kernel regex> <skip> <count> -- <command...>   (runs the command plainly first)
TAG=$1; REGEX=$2; SKIP=$3; COUNT=$4; shift 5
mkdir -p gpurun_out/$TAG
"$@" > gpurun_out/$TAG/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:$REGEX" -s $SKIP -c $COUNT -o gpurun_out/$TAG/prof "$@" > gpurun_out/$TAG/ncu.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/$TAG/ncu.log; cat gpurun_out/$TAG/plain.log | head -5
