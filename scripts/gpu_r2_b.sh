# round 2, GPU call B (one GPU): the whole GPU test suite, then kernel variants side by side
set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/tests_gpu.log 2>&1; tail -n 25 gpurun_out/tests_gpu.log
timeout 900 python scripts/time_variants.py > gpurun_out/r2b_variants.json 2> gpurun_out/r2b_variants.err; cat gpurun_out/r2b_variants.err | tail -n 12
