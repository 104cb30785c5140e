#!/bin/bash
# usage: tools/evidence_n1.sh <tag>  — default + slow test tiers, smoke and both bench arms on one GPU (no ncu captures)
tag=$1
mkdir -p gpurun_out
timeout 600 tools/run_gpu_tests_timed.sh ${tag}_gpu_tests --timeout 300 > /dev/null
grep -E "passed|failed|FAILED" gpurun_out/${tag}_gpu_tests.log | tail -6
B2_SLOW=1 timeout 900 python -m pytest tests/test_gan_sequences_gpu.py tests/test_training_gpu.py tests/test_attention_gpu.py -m gpu -v -s --timeout 600 \
    -k "loop_iterations or teacher_forced_48px or free_running_divergence or config4 or attention_generator" > gpurun_out/${tag}_gpu_tests_slow.log 2>&1
grep -E "passed|failed|FAILED|max per-step|content loss max|sigma|attention k=|attention full" gpurun_out/${tag}_gpu_tests_slow.log | tail -24
python __graft_entry__.py smoke > gpurun_out/${tag}_smoke.log 2>&1; tail -1 gpurun_out/${tag}_smoke.log
python bench.py > gpurun_out/${tag}_bench.json.log 2>&1; tail -c 400 gpurun_out/${tag}_bench.json.log; echo
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_reference_arm.json.log 2>&1; tail -c 200 gpurun_out/${tag}_bench_reference_arm.json.log; echo
