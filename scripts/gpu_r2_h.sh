# round 2, GPU call H (eight GPUs): exchange stream priority and SM reserve at N=8, after the per-message exchange
# kernels got four messages per thread
set -x
mkdir -p gpurun_out
run() { timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 8 "${@:3}" > gpurun_out/$2.json 2> gpurun_out/$2.err || tail -n 25 gpurun_out/$2.err; tail -c 500 gpurun_out/$2.json; }
run 29551 r2h_bench8_xfirst --steps 12 --warmup 3 --no-extra
run 29552 r2h_bench8_mainfirst --steps 12 --warmup 3 --no-extra --main-first
run 29553 r2h_bench8_xfirst_res8 --steps 12 --warmup 3 --no-extra --reserve-sms 8
