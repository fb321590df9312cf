# round 2, GPU call ZZ (one GPU): the native cache loader end to end (four read connections forced) on the last build
set -x
mkdir -p gpurun_out
SMSGATE_DCACHE_THREADS=4 timeout 45 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "native_cache_loader" > gpurun_out/r2zz_tests_loader.log 2>&1; tail -n 3 gpurun_out/r2zz_tests_loader.log
