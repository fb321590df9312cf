# round 2, GPU call N (four GPUs): the N=4 bench line as the driver runs it
set -x
mkdir -p gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29581 bench.py --gpus 4 --steps 10 --warmup 3 > gpurun_out/r2n_bench4.json 2> gpurun_out/r2n_bench4.err || tail -n 25 gpurun_out/r2n_bench4.err
tail -c 1200 gpurun_out/r2n_bench4.json
