# round 2, GPU call F (two GPUs): lanes 2 vs 3, with the one-byte verdicts and the one-pass decimal parser in
set -x
mkdir -p gpurun_out
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 "${@:3}" > gpurun_out/$2.json 2> gpurun_out/$2.err || tail -n 25 gpurun_out/$2.err; tail -c 600 gpurun_out/$2.json; }
run 29531 r2f_bench2_l2 --steps 12 --warmup 3 --no-extra
run 29532 r2f_bench2_l3 --steps 12 --warmup 3 --no-extra --lanes 3
run 29533 r2f_bench2_l3_res8 --steps 12 --warmup 3 --no-extra --lanes 3 --reserve-sms 8
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_multi.py -m gpu -q > gpurun_out/r2f_tests.log 2>&1; tail -n 5 gpurun_out/r2f_tests.log
