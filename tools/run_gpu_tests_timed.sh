#!/bin/bash
# usage: tools/run_gpu_tests_timed.sh <logname> [pytest args...] — `pytest -m gpu -v` with a wall-clock stamp per line, so a
# killed run still shows which test was slow
name=$1; shift
mkdir -p gpurun_out
python -m pytest tests -m gpu -v --durations=40 "$@" 2>&1 | python -u -c '
import sys, time
t0 = time.time()
for line in sys.stdin:
    sys.stdout.write("%7.1f %s" % (time.time() - t0, line))
' > gpurun_out/$name.log
tail -75 gpurun_out/$name.log
