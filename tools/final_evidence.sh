#!/bin/bash
# usage: tools/final_evidence.sh <tag>   — the single-GPU evidence set of a round: default + slow test tiers, smoke, bench (both
# arms), ncu --set full of the dominant kernel and of the trunk conv.  Everything lands in gpurun_out/<tag>_*.
tag=$1
mkdir -p gpurun_out
timeout 600 tools/run_gpu_tests_timed.sh ${tag}_gpu_tests --timeout 300 > /dev/null
grep -E "passed|failed" gpurun_out/${tag}_gpu_tests.log | tail -2
B2_SLOW=1 timeout 900 python -m pytest tests/test_gan_sequences_gpu.py tests/test_training_gpu.py -m gpu -v -s --timeout 600 \
    -k "loop_iterations or teacher_forced_48px or free_running_divergence or config4" > gpurun_out/${tag}_gpu_tests_slow.log 2>&1
grep -E "passed|failed|max per-step|content loss max|worst deviation" gpurun_out/${tag}_gpu_tests_slow.log | tail -20
python __graft_entry__.py smoke > gpurun_out/${tag}_smoke.log 2>&1; tail -1 gpurun_out/${tag}_smoke.log
python bench.py > gpurun_out/${tag}_bench.json.log 2>&1; tail -c 600 gpurun_out/${tag}_bench.json.log; echo
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_reference_arm.json.log 2>&1; tail -c 300 gpurun_out/${tag}_bench_reference_arm.json.log; echo
python tools/bench_quick.py 3 8 > gpurun_out/${tag}_quick.log 2>&1 && {
  ncu --set full --clock-control none --import-source on -k regex:conv_phases2 -s 3 -c 1 -f -o gpurun_out/${tag}_ncu_full_conv_phases2_convT2 python tools/bench_quick.py 3 8 > gpurun_out/${tag}_ncu1.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:conv_stack -s 70 -c 2 -f -o gpurun_out/${tag}_ncu_full_conv_stack python tools/bench_quick.py 3 8 > gpurun_out/${tag}_ncu2.log 2>&1
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_bench_launches.csv python tools/bench_quick.py 3 8 > gpurun_out/${tag}_ncu3.log 2>&1
}
cat gpurun_out/${tag}_quick.log; ls -la gpurun_out | grep ncu-rep
