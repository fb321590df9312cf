nvidia-smi topo -m > gpurun_out/topo8.txt 2>&1; lscpu | grep -E "Model name|Socket|NUMA|^CPU\(s\)" >> gpurun_out/topo8.txt
for N in 4 8; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 2>gpurun_out/bench$N.err | tail -n 1 > gpurun_out/bench$N.json
cat gpurun_out/bench$N.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('N=%d value %.3e step_ms %.4f e2e %.3e e2e_ms %.3f launches %d bound %s' % (d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['gpu_launches'], d['config'].get('host_cpus_bound')))"
done
tail -n 3 gpurun_out/bench8.err; cat gpurun_out/topo8.txt | head -n 30
