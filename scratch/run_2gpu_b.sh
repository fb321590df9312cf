python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "device_keys or records" 2>&1 | tail -n 3
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 scratch/prof_p2p.py 2>&1 | grep -E "scatter|Error|error" | tail -n 5
