# round 2, GPU call G (eight GPUs): the N=8 bench line with its configs object and the exchange parity check
set -x
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2g_topo.txt 2>&1
run() { timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 8 "${@:3}" > gpurun_out/$2.json 2> gpurun_out/$2.err || tail -n 25 gpurun_out/$2.err; tail -c 800 gpurun_out/$2.json; }
run 29541 r2g_bench8 --steps 12 --warmup 3
run 29542 r2g_bench8_res8 --steps 12 --warmup 3 --no-extra --reserve-sms 8
