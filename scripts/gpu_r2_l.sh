# round 2, GPU call L (eight GPUs): the N=8 bench line as the driver runs it (configs object, exchange parity)
set -x
mkdir -p gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r2l_bench8.json 2> gpurun_out/r2l_bench8.err || tail -n 25 gpurun_out/r2l_bench8.err
tail -c 1500 gpurun_out/r2l_bench8.json
