# round 2, GPU call E (two GPUs): whole GPU suite on this box (the 2-GPU tests run for real), then N=2 bench lines
set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r2e_tests_gpu.log 2>&1; tail -n 25 gpurun_out/r2e_tests_gpu.log
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 "${@:3}" > gpurun_out/$2.json 2> gpurun_out/$2.err || tail -n 25 gpurun_out/$2.err; tail -c 1500 gpurun_out/$2.json; }
run 29531 r2e_bench2 --steps 10 --warmup 3
run 29532 r2e_bench2_res8 --steps 10 --warmup 3 --no-extra --reserve-sms 8
timeout 600 python bench.py --no-extra > gpurun_out/r2e_bench1.json 2> gpurun_out/r2e_bench1.err; tail -c 1200 gpurun_out/r2e_bench1.json
