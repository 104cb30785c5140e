#!/bin/bash
# usage: tools/gpu_check.sh <logname> <command...>   — runs a command on the box with output captured to gpurun_out/
mkdir -p gpurun_out
name=$1; shift
"$@" > gpurun_out/$name.log 2>&1
echo "exit=$?" >> gpurun_out/$name.log
tail -30 gpurun_out/$name.log
