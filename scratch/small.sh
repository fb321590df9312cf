for m in 31250 62500 125000 250000; do
python bench.py --steps 10 --warmup 3 --cpu-sample 1000 --msgs $m 2>/dev/null | tail -n 1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print($m, 'step_ms %.4f k1_ms %.4f e2e_ms %.4f' % (d['ms_per_step'], d['roofline']['kernel_ms'], d['e2e']['ms_per_step']))"
done
