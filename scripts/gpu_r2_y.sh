# round 2, GPU call Y (one GPU): _iter_sms drop-ins against the extended reference run (rows yielded before an exception)
set -x
mkdir -p gpurun_out
timeout 70 python -m pytest tests/test_gpu_legacy.py tests/test_gpu_dropin.py -m gpu -q -k "iter_sms" > gpurun_out/r2y_tests_iter_sms.log 2>&1; tail -n 3 gpurun_out/r2y_tests_iter_sms.log
