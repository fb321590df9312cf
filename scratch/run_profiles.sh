set -x
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -n 1 gpurun_out/smoke.log
python -m pytest tests -m gpu -x -q 2>&1 | tail -n 5 > gpurun_out/tests_gpu.log
python bench.py > gpurun_out/bench1.json 2> gpurun_out/bench1.err || exit 1
python bench.py --steps 3 --warmup 3 > gpurun_out/b_small.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1_launches.csv python bench.py --steps 3 --warmup 3 > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_scan_keys -s 4 -c 1 -o gpurun_out/k_scan_keys_full -f python bench.py --steps 3 --warmup 3 > gpurun_out/ncu_f.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_join_route -s 4 -c 1 -o gpurun_out/k_join_route_full -f python bench.py --steps 3 --warmup 3 > gpurun_out/ncu_j.log 2>&1
cat gpurun_out/tests_gpu.log gpurun_out/bench1.json
