# round 2, GPU call O (one GPU): final code - smoke, whole GPU suite, the N=1 bench line, launch list
set -x
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2o_smoke.log 2>&1; tail -n 1 gpurun_out/r2o_smoke.log
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r2o_tests_gpu.log 2>&1; tail -n 6 gpurun_out/r2o_tests_gpu.log
timeout 600 python bench.py > gpurun_out/r2o_bench1.json 2> gpurun_out/r2o_bench1.err || tail -n 30 gpurun_out/r2o_bench1.err
tail -c 1500 gpurun_out/r2o_bench1.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r2o_launches.csv python bench.py --steps 3 --warmup 3 --no-extra > gpurun_out/ncu_l.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_join_route -s 4 -c 1 -o gpurun_out/r2o_k_join_route_full -f python bench.py --steps 3 --warmup 3 --no-extra > gpurun_out/ncu_j.log 2>&1
