#!/bin/bash
# usage: run_ncu.sh <name> <kernel-regex> <cmd...>  — plain run first, then ncu --set full on <=3 launches
name=$1; regex=$2; shift 2
mkdir -p gpurun_out
"$@" > gpurun_out/${name}_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$regex -s 3 -c 2 -f -o gpurun_out/$name "$@" > gpurun_out/${name}_ncu.log 2>&1
echo "ncu exit $?" >> gpurun_out/${name}_ncu.log
tail -3 gpurun_out/${name}_ncu.log
