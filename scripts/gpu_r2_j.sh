# round 2, GPU call J (eight GPUs): bits per key of the prefilter bitmaps (they cross NVLink twice per batch), host-side
# copy bandwidth per rank
set -x
mkdir -p gpurun_out
run() { timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 8 "${@:3}" > gpurun_out/$2.json 2> gpurun_out/$2.err || tail -n 25 gpurun_out/$2.err; tail -c 400 gpurun_out/$2.json; }
run 29561 r2j_bench8_bits8 --steps 12 --warmup 3 --no-extra --prefilter-bits 8
run 29562 r2j_bench8_bits4 --steps 12 --warmup 3 --no-extra --prefilter-bits 4
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29563 scripts/pcie_probe.py > gpurun_out/r2j_pcie_probe_8.json 2> gpurun_out/r2j_pcie.err; cat gpurun_out/r2j_pcie_probe_8.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29564 scripts/pcie_probe.py > gpurun_out/r2j_pcie_probe_4.json 2>> gpurun_out/r2j_pcie.err; cat gpurun_out/r2j_pcie_probe_4.json
timeout 300 python scripts/pcie_probe.py > gpurun_out/r2j_pcie_probe_1.json 2>> gpurun_out/r2j_pcie.err; cat gpurun_out/r2j_pcie_probe_1.json
