# round 2, GPU call S (two GPUs): the code as committed - the multi-GPU tests on real peer memory, the N=2 bench
# line as the driver runs it (configs object with dup10 / skew, exchange parity)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "multi or dist or p2p or exchange" > gpurun_out/r2s_tests_multi_2gpu.log 2>&1; tail -n 4 gpurun_out/r2s_tests_multi_2gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29582 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2s_bench2.json 2> gpurun_out/r2s_bench2.err || tail -n 25 gpurun_out/r2s_bench2.err
tail -c 1200 gpurun_out/r2s_bench2.json
