"""Times the XML stage: ElementTree (what the reference does, watcher.py:37-44) against the GPU
tokenizer (smsgate_xml_scan + smsgate_xml_emit) on the same synthetic backup file."""
import os
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from smsgate_b200 import synth, xml_ingest  # noqa: E402
from smsgate_b200.replay import ReplayEngine  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
backup = synth.generate(n, seed=0x5EED + 2)
path = os.path.join(tempfile.mkdtemp(), "sms.xml")
synth.write_xml(path, backup)
size = os.path.getsize(path)
t0 = time.perf_counter()
host = xml_ingest.read_backup(path)
t_host = time.perf_counter() - t0
import torch
eng = ReplayEngine(0)
data = np.fromfile(path, dtype=np.uint8)
best = 1e9
for _ in range(5):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    d = eng.xml_extract(data)
    torch.cuda.synchronize()
    best = min(best, time.perf_counter() - t0)
assert d["n"] == n and int(d["body_offsets"][-1]) == int(host.body_offsets[-1])
print(f"xml file {size / 1e6:.1f} MB, {n} messages")
print(f"ElementTree + findall + get (host, 1 core): {t_host * 1e3:.0f} ms = {n / t_host / 1e3:.0f} k msgs/s, {size / t_host / 1e6:.0f} MB/s")
print(f"GPU tokenizer incl. host->device copy of the file (pageable): {best * 1e3:.1f} ms = {n / best / 1e6:.1f} M msgs/s, {size / best / 1e9:.2f} GB/s")
