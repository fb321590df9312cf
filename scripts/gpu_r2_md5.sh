# round 2, last GPU call (one GPU): the MD5 keys kernel after its padding words stopped being read from memory
set -x
mkdir -p gpurun_out
timeout 30 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "md5 or device_key or cross_source" > gpurun_out/r2md5_tests.log 2>&1; tail -n 3 gpurun_out/r2md5_tests.log
