# round 2, GPU call R (one GPU): upsert slot with the compared key fetched before / after the CAS; the GPU suite,
# the bench line and the launch list on the final code
set -x
mkdir -p gpurun_out
VARIANT_WORKLOADS=hit100,dup10,skew timeout 900 python scripts/time_variants.py > gpurun_out/r2r_dedup_variants.json 2> gpurun_out/r2r_dedup_variants.err || tail -n 30 gpurun_out/r2r_dedup_variants.err
grep -E "^(hit100|dup10|skew)" gpurun_out/r2r_dedup_variants.err | cut -c1-200
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2r_smoke.log 2>&1; tail -n 1 gpurun_out/r2r_smoke.log
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r2r_tests_gpu.log 2>&1; tail -n 4 gpurun_out/r2r_tests_gpu.log
timeout 600 python bench.py > gpurun_out/r2r_bench1.json 2> gpurun_out/r2r_bench1.err || tail -n 30 gpurun_out/r2r_bench1.err
tail -c 600 gpurun_out/r2r_bench1.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r2r_launches.csv python bench.py --steps 3 --warmup 3 --no-extra > gpurun_out/ncu_l.log 2>&1
