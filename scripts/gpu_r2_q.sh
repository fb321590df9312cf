# round 2, GPU call Q (one GPU): upsert slot claimed with one 64-bit CAS {first,last}; with a plain read before it;
# with one lane per key and warp - against the 32-bit CAS of call P.  Then the GPU suite on the default build and
# the dedup-related part of it on the pre-read build.
set -x
mkdir -p gpurun_out
VARIANT_WORKLOADS=hit100,dup10,skew timeout 900 python scripts/time_variants.py > gpurun_out/r2q_dedup_variants.json 2> gpurun_out/r2q_dedup_variants.err || tail -n 30 gpurun_out/r2q_dedup_variants.err
grep -E "^(hit100|dup10|skew)" gpurun_out/r2q_dedup_variants.err | cut -c1-200
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r2q_tests_gpu.log 2>&1; tail -n 4 gpurun_out/r2q_tests_gpu.log
SMSGATE_B200_LIB=$PWD/smsgate_b200/variants/libc_cas64_preread.so timeout 1200 python -m pytest tests -m gpu -q -k "parity or multi or dedup or no_alloc or dlq" > gpurun_out/r2q_tests_gpu_preread.log 2>&1; tail -n 4 gpurun_out/r2q_tests_gpu_preread.log
