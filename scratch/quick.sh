python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -n 2
python bench.py --steps 10 --warmup 3 2>&1 | tail -n 1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('step_ms',d['ms_per_step'],'k1_ms', d['roofline']['kernel_ms'],'e2e_ms', d['e2e']['ms_per_step'])"
