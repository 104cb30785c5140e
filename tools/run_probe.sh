#!/bin/bash
# Runs probe experiments, each under its own timeout so one hang cannot eat the box.
mkdir -p gpurun_out
for exp in ${PROBES:-conv time}; do
  echo "=== $exp ==="
  timeout 120 ./build/b2_probe $exp 2>&1 | tail -60
  echo "exit: $?"
done
echo "=== mn v0 ==="; MN_VARIANT=0 timeout 60 ./build/b2_probe mn 2>&1 | tail -5
echo "=== mn v1 ==="; MN_VARIANT=1 timeout 60 ./build/b2_probe mn 2>&1 | tail -5
