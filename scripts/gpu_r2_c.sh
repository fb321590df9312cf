# round 2, GPU call C (two GPUs): the real peer-memory / NCCL tests, the N=2 bench line with its configs object,
# and the keys kernel's SM reserve swept
set -x
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2c_topo.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q > gpurun_out/r2c_tests_multi.log 2>&1; tail -n 8 gpurun_out/r2c_tests_multi.log
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 "${@:3}" > gpurun_out/$2.json 2> gpurun_out/$2.err || tail -n 25 gpurun_out/$2.err; tail -c 3000 gpurun_out/$2.json; }
run 29511 r2c_bench2 --steps 10 --warmup 3
run 29512 r2c_bench2_res0 --steps 10 --warmup 3 --no-extra --reserve-sms 0
run 29513 r2c_bench2_res4 --steps 10 --warmup 3 --no-extra --reserve-sms 4
run 29514 r2c_bench2_res16 --steps 10 --warmup 3 --no-extra --reserve-sms 16
