# round 2, GPU call A (one GPU): tests, smoke, bench line, launch list, full captures of the two top kernels
set -x
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -n 1 gpurun_out/smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/tests_gpu.log 2>&1; tail -n 15 gpurun_out/tests_gpu.log
timeout 600 python bench.py > gpurun_out/r2a_bench1.json 2> gpurun_out/r2a_bench1.err || { tail -n 30 gpurun_out/r2a_bench1.err; exit 1; }
cat gpurun_out/r2a_bench1.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r2a_launches.csv python bench.py --steps 3 --warmup 3 --no-extra > gpurun_out/ncu_l.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_scan_keys -s 4 -c 1 -o gpurun_out/r2a_k_scan_keys_full -f python bench.py --steps 3 --warmup 3 --no-extra > gpurun_out/ncu_f.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_join_route -s 4 -c 1 -o gpurun_out/r2a_k_join_route_full -f python bench.py --steps 3 --warmup 3 --no-extra > gpurun_out/ncu_j.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_scan_keys -s 4 -c 1 -o gpurun_out/r2a_k_scan_keys_mixed_full -f python bench.py --steps 3 --warmup 3 --no-extra --workload mixed > gpurun_out/ncu_m.log 2>&1
ls -la gpurun_out/*.ncu-rep
