# round 2, GPU call V (one GPU): the code as committed at round end - smoke, whole GPU suite, bench line, launch list
set -x
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2v_smoke.log 2>&1; tail -n 1 gpurun_out/r2v_smoke.log
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/r2v_tests_gpu.log 2>&1; tail -n 4 gpurun_out/r2v_tests_gpu.log
timeout 400 python bench.py > gpurun_out/r2v_bench1.json 2> gpurun_out/r2v_bench1.err || tail -n 30 gpurun_out/r2v_bench1.err
tail -c 600 gpurun_out/r2v_bench1.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r2v_launches.csv python bench.py --steps 3 --warmup 3 --no-extra > gpurun_out/ncu_l.log 2>&1
