python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -n 3
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 scratch/prof_p2p.py 2>&1 | grep -E "push|inbox|Error|error" | tail -n 6
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 2>gpurun_out/bench2.err | tail -n 1 > gpurun_out/bench2.json
cat gpurun_out/bench2.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('N=2 value %.3e step_ms %.4f e2e %.3e e2e_ms %.3f launches %d' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['gpu_launches']))"
