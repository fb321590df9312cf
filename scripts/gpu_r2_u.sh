# round 2, GPU call U (one GPU): keys kernel asking L2 for the next tile during the hash phase vs not
set -x
mkdir -p gpurun_out
VARIANT_WORKLOADS=hit100,mixed timeout 600 python scripts/time_variants.py > gpurun_out/r2u_prefetch_variants.json 2> gpurun_out/r2u_prefetch_variants.err || tail -n 30 gpurun_out/r2u_prefetch_variants.err
grep -E "^(hit100|mixed)" gpurun_out/r2u_prefetch_variants.err | cut -c1-200
SMSGATE_B200_LIB=$PWD/smsgate_b200/variants/libb_l2_prefetch.so timeout 600 python -m pytest tests -m gpu -q -k "parity or golden or keys or scan" > gpurun_out/r2u_tests_gpu_prefetch.log 2>&1; tail -n 3 gpurun_out/r2u_tests_gpu_prefetch.log
