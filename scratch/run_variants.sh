#!/bin/bash
# on the GPU box: time every prebuilt variant with the bench (cpu sample tiny)
cp smsgate_b200/libsmsgate_b200.so /tmp/orig.so
for f in scratch/variants/lib_*.so; do
  cp $f smsgate_b200/libsmsgate_b200.so
  echo -n "$f "
  python bench.py --steps 10 --warmup 3 --cpu-sample 2000 2>/dev/null | tail -n 1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('step_ms %.4f k1_ms %.4f upserts %d' % (d['ms_per_step'], d['roofline']['kernel_ms'], d['upserts_per_gpu']))"
done
cp /tmp/orig.so smsgate_b200/libsmsgate_b200.so
