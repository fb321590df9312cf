python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3 2>gpurun_out/bench8.err | tail -n 1 > gpurun_out/bench8.json
cat gpurun_out/bench8.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('N=%d value %.3e step_ms %.4f e2e %.3e e2e_ms %.3f launches %d' % (d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['gpu_launches']))"
tail -n 2 gpurun_out/bench8.err
