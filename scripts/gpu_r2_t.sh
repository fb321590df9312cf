# round 2, GPU call T (one GPU): keys kernel walking its share in full tiles + one short tile vs equal tiles; then
# smoke, the GPU suite and the bench line on the full-tile build
set -x
mkdir -p gpurun_out
VARIANT_WORKLOADS=hit100,dup10,mixed timeout 900 python scripts/time_variants.py > gpurun_out/r2t_tile_variants.json 2> gpurun_out/r2t_tile_variants.err || tail -n 30 gpurun_out/r2t_tile_variants.err
grep -E "^(hit100|dup10|mixed)" gpurun_out/r2t_tile_variants.err | cut -c1-200
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2t_smoke.log 2>&1; tail -n 1 gpurun_out/r2t_smoke.log
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r2t_tests_gpu.log 2>&1; tail -n 4 gpurun_out/r2t_tests_gpu.log
timeout 600 python bench.py > gpurun_out/r2t_bench1.json 2> gpurun_out/r2t_bench1.err || tail -n 30 gpurun_out/r2t_bench1.err
tail -c 600 gpurun_out/r2t_bench1.json
