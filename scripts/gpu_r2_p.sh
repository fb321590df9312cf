# round 2, GPU call P (one GPU): upsert-slot atomics on hot keys - three atomics per message / skip the ones that
# cannot change the slot / one lane per key and warp; then the GPU suite on the chosen code and on the warp variant
set -x
mkdir -p gpurun_out
VARIANT_WORKLOADS=hit100,dup10,skew timeout 900 python scripts/time_variants.py > gpurun_out/r2p_dedup_variants.json 2> gpurun_out/r2p_dedup_variants.err || tail -n 30 gpurun_out/r2p_dedup_variants.err
grep -E "^(hit100|dup10|skew)" gpurun_out/r2p_dedup_variants.err | cut -c1-260
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r2p_tests_gpu.log 2>&1; tail -n 4 gpurun_out/r2p_tests_gpu.log
SMSGATE_B200_LIB=$PWD/smsgate_b200/variants/libc_warpagg.so timeout 1200 python -m pytest tests -m gpu -q -k "parity or multi or dedup or no_alloc or dlq" > gpurun_out/r2p_tests_gpu_warpagg.log 2>&1; tail -n 4 gpurun_out/r2p_tests_gpu_warpagg.log
