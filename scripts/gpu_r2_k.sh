# round 2, GPU call K (one GPU): cache strings inline in the side record vs not, the host call's chunking swept, parity suite
set -x
mkdir -p gpurun_out
timeout 900 python scripts/time_variants.py > gpurun_out/r2k_variants.json 2> gpurun_out/r2k_variants.err; tail -n 6 gpurun_out/r2k_variants.err | cut -c1-250
timeout 900 python scripts/sweep_host_pipeline.py > gpurun_out/r2k_host_pipeline.json 2> gpurun_out/r2k_host_pipeline.err; tail -n 12 gpurun_out/r2k_host_pipeline.err | cut -c1-300
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r2k_tests_gpu.log 2>&1; tail -n 5 gpurun_out/r2k_tests_gpu.log
