# round 2, GPU call W (one GPU): ncu --set full of the keys kernel as committed at round end (with the L2 prefetch)
set -x
mkdir -p gpurun_out
timeout 140 ncu --set full --clock-control none --import-source on -k regex:k_scan_keys -s 4 -c 1 -o gpurun_out/r2w_k_scan_keys_full -f python bench.py --steps 3 --warmup 3 --no-extra > gpurun_out/ncu_w.log 2>&1
tail -n 3 gpurun_out/ncu_w.log
