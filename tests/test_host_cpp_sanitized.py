"""The host C++ inside the library that parses bytes someone else wrote - the pickle decoder of the native cache
reader - and the payload emitter, built with AddressSanitizer + UBSan and run on mutated / adversarial input.
No GPU, no CUDA: the two headers are plain C++ (tests/hostcheck/asan_*.cpp include them directly)."""
import os
import pickle
import random
import shutil
import struct
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "smsgate_b200", "csrc")
SAN = ["-std=c++17", "-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined", "-fno-omit-frame-pointer"]


def _build(tmp_path, name):
    gxx = shutil.which("g++")
    if gxx is None:
        pytest.skip("no g++")
    exe = str(tmp_path / name)
    res = subprocess.run([gxx] + SAN + ["-I", CSRC, "-o", exe, os.path.join(ROOT, "tests", "hostcheck", name + ".cpp"), "-ldl", "-lpthread"],
                         capture_output=True, text=True)
    if res.returncode and "sanitize" in res.stderr and "cannot find" in res.stderr:
        pytest.skip("g++ has no sanitizer runtime here")
    assert res.returncode == 0, res.stderr[-2000:]
    return exe


def _run(exe, arg):
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=0:abort_on_error=0", UBSAN_OPTIONS="print_stacktrace=1")
    res = subprocess.run([exe, arg], capture_output=True, text=True, env=env, timeout=600)
    assert res.returncode == 0 and "ERROR: AddressSanitizer" not in res.stderr and "runtime error" not in res.stderr, res.stderr[-3000:]
    return res.stdout


def test_pickle_decoder_on_mutated_blobs(tmp_path):
    rng = random.Random(0xA5A)
    words = ["", "a", "52.00", "1,842.74", "TEST LLC", "Ереван", "☕", "x" * 255, "y" * 256, "z" * 70000,
             "null", "06.05.25 14:23", "***0018", "usd"]

    def value():
        k = rng.random()
        if k < 0.5:
            return rng.choice(words)
        if k < 0.6:
            return None
        if k < 0.7:
            return rng.choice([0, 1, -1, 255, 256, 65535, 65536, 2 ** 31 - 1, 2 ** 31, -2 ** 31, 2 ** 63 - 1, 2 ** 63, -2 ** 63, 10 ** 30])
        if k < 0.75:
            return rng.choice([True, False])
        if k < 0.8:
            return rng.choice([0.0, 12.5, -1e300, float("inf")])
        if k < 0.9:
            return rng.choice([b"bytes", ("t", 1), ["l"], {"d": 1}, {1, 2}, frozenset(), bytearray(b"x")])
        return "".join(rng.choice("aЖ☕ \"\\\n0,.-") for _ in range(rng.randint(0, 30)))
    fields = ["txn_type", "date", "amount", "currency", "card", "merchant", "city", "address", "balance"]
    path = tmp_path / "blobs.bin"
    n = 30000
    with open(path, "wb") as out:
        for _ in range(n):
            keep = [f for f in fields if rng.random() < 0.9]
            rng.shuffle(keep)
            shared = value()
            v = {f: (shared if rng.random() < 0.2 else value()) for f in keep}
            if rng.random() < 0.03:
                v = rng.choice([[], "str", 5, None, (1, 2), {}])
            blob = bytearray(pickle.dumps(v, protocol=rng.choice([0, 1, 2, 3, 4, 5])))
            r = rng.random()
            if r < 0.25:
                blob = blob[:rng.randint(0, len(blob))]                       # cut short
            elif r < 0.7:
                for _ in range(rng.randint(1, 4)):
                    if not blob:
                        break
                    k = rng.randrange(len(blob))
                    op = rng.random()
                    if op < 0.3:
                        del blob[k]
                    elif op < 0.7:
                        blob[k] = rng.randrange(256)
                    else:
                        blob.insert(k, rng.randrange(256))
            elif r < 0.75:
                blob = bytearray(rng.randrange(256) for _ in range(rng.randint(0, 64)))   # noise
            out.write(struct.pack("<I", len(blob)))
            out.write(blob)
    stdout = _run(_build(tmp_path, "asan_pickle"), str(path))
    assert f"decoded {n} blobs" in stdout


def _dump_replay_columns(tmp_path):
    """Columns of a replayed batch with adversarial cache texts, one file per array, for the C++ harnesses."""
    from smsgate_b200 import synth
    from smsgate_b200.batch import ragged_from_strings
    from tests import test_emit_json as T
    rng = random.Random(0xE417)

    def text(maxlen=12):
        return ''.join(rng.choice(T._FUZZ_ALPHABET) for _ in range(rng.randint(0, maxlen)))
    backup = synth.generate(2500, seed=2001, mixed_locale=True)
    cache = dict(backup.cache)
    for key in list(cache):
        if rng.random() < 0.8:
            cache[key] = dict(cache[key], merchant=text() or "m", city=rng.choice([None, text(), ""]),
                              address=rng.choice([None, text(40), "null", ""]),
                              amount=rng.choice(["52.00", "0.0000001", "1e5", "999999999999999999", "0", "-0", "1,234.5", "9" * 25,
                                                 "0." + "0" * 17 + "1"]),
                              balance=rng.choice(["1", "-12,5", None, "000123.4500", "0.00"]),
                              currency=rng.choice(["amd", "USD", "", None, "ß", "a b", 5]),
                              txn_type=rng.choice(["debit", "credit", "otp", "unknown"]))
    senders = [text(8) for _ in backup.senders]
    res, table = T._result(list(backup.bodies), senders, cache)
    d = tmp_path / "cols"
    d.mkdir()

    def w(name, a, dt):
        np.ascontiguousarray(a, dtype=dt).tofile(str(d / name))
    w("n", [len(backup.bodies), table.m], np.int64)
    for name, dt in (("msg_id", np.uint8), ("status", np.uint8), ("flags", np.uint8), ("dec_flags", np.uint8), ("slot", np.int32),
                     ("amount_coef", np.int64), ("balance_coef", np.int64), ("amount_exp", np.int32), ("balance_exp", np.int32),
                     ("date", np.int64), ("card", np.uint8), ("card_len", np.uint8), ("fixed_len", np.int32), ("fixed_bytes", np.uint8)):
        w(name, res.cols[name], dt)
    sb, so = ragged_from_strings(senders)
    w("body_offsets", res.batch.body_offsets, np.int64)
    w("body_bytes", res.batch.body_bytes, np.uint8)
    w("sender_offsets", so, np.int64)
    w("sender_bytes", sb, np.uint8)
    w("text_offsets", table.text[0], np.int64)
    w("text_bytes", table.text[1], np.uint8)
    w("text_kind", table.text[2], np.uint8)
    return d


def test_emitter_writes_exactly_what_it_announced(tmp_path):
    stdout = _run(_build(tmp_path, "asan_emit"), str(_dump_replay_columns(tmp_path)))
    rows = int(stdout.split()[1])
    assert rows > 2000 and "MISMATCH" not in stdout


def test_threaded_host_code_under_tsan(tmp_path):
    """The cache reader on seven connections and the emitter on six threads, built with ThreadSanitizer: no data
    race, and the same arrays / bytes as on one."""
    import diskcache

    from smsgate_b200 import synth
    gxx = shutil.which("g++")
    if gxx is None:
        pytest.skip("no g++")
    exe = str(tmp_path / "tsan_host")
    res = subprocess.run([gxx, "-std=c++17", "-O1", "-g", "-fsanitize=thread", "-fno-omit-frame-pointer", "-I", CSRC, "-o", exe,
                          os.path.join(ROOT, "tests", "hostcheck", "tsan_host.cpp"), "-ldl", "-lpthread"], capture_output=True, text=True)
    if res.returncode and "cannot find" in res.stderr:
        pytest.skip("g++ has no ThreadSanitizer runtime here")
    assert res.returncode == 0, res.stderr[-2000:]
    cache_dir = tmp_path / "cache"
    c = diskcache.Cache(str(cache_dir))
    backup = synth.generate(6000, seed=77, mixed_locale=True)
    with c.transact():
        for k, v in backup.cache.items():
            c[k] = v
        c["f" * 64] = {"txn_type": "debit", "merchant": "X" * 40_000}      # a value diskcache keeps in a side file
        c["e" * 64] = "a plain string"
    c.close()
    cols = _dump_replay_columns(tmp_path)
    res = subprocess.run([exe, str(cache_dir), str(cols)], capture_output=True, text=True, timeout=900,
                         env=dict(os.environ, TSAN_OPTIONS="halt_on_error=0"))
    assert res.returncode == 0, (res.stdout, res.stderr[-3000:])
    assert "ThreadSanitizer" not in res.stderr, res.stderr[-3000:]
    assert "one connection == seven: yes" in res.stdout and "one thread == six: yes" in res.stdout


_DEVICE_LOGIC_SCRIPT = r"""
import random, sys
sys.path.insert(0, sys.argv[1])
import numpy as np
from smsgate_b200 import gemini_cache, synth
from tests import _hostcheck as H
from tests._fuzz import bodies
import datetime as dt
bs = list(bodies(4000, 4242))
sc = H.scan(bs)                                            # keys, filters, rewrite, body date
H.md5(bs)                                                  # MD5 straight off the byte buffer (8 bytes of slack)
H.md5(["", "a" * 55, "b" * 56, "c" * 63, "d" * 64, "e" * 119, "f" * 120])
rng = random.Random(7)
alphabet = "0123456789" * 3 + "..,,  -\u00a0\u202f\u2009\t\nE+eAMD$\u0665\uff11"
H.decimal(["".join(rng.choice(alphabet) for _ in range(rng.choice([0, 1, 2, 3, 5, 8, 12, 18, 24, 40]))) for _ in range(20000)])
H.strict_date(["06.05.25 14:23", "", "x" * 40, "31.02.25 10:00", "06.05.25 14:2", "\u0660\u0666.05.25 14:23"] * 50)
b = synth.generate(3000, seed=9, mixed_locale=True, miss_frac=0.2, dup_frac=0.2)
sc = H.scan(b.bodies)
table = gemini_cache.build_table(b.cache, H.strict_date)
H.join_route(sc, table, gemini_cache.pack_datetime(dt.datetime(2026, 1, 1)))
H.legacy(bs[:2000])
H.legacy(list(b.bodies)[:1500], 3)
from tests._xml_cases import documents
for d in documents()[:600]:
    H.xml_extract(d)
print("device logic ok")
"""


def test_device_logic_under_sanitizers(tmp_path):
    """The kernels' __host__ __device__ cores (scan, keys, decimals, dates, join + routing, legacy matcher, XML
    tokenizer) compiled for the CPU with AddressSanitizer + UBSan and run on adversarial input: a read past a
    buffer is silent on the GPU.  (This is how the MD5 kernel's reads of its padding words were found.)"""
    gxx = shutil.which("g++")
    if gxx is None:
        pytest.skip("no g++")
    lib = str(tmp_path / "libhostcheck_san.so")
    res = subprocess.run([gxx] + SAN + ["-fPIC", "-shared", "-I", CSRC, "-o", lib, os.path.join(ROOT, "tests", "hostcheck", "hostcheck.cpp")],
                         capture_output=True, text=True)
    assert res.returncode == 0, res.stderr[-2000:]
    runtimes = [subprocess.run([gxx, "-print-file-name=" + n], capture_output=True, text=True).stdout.strip()
                for n in ("libasan.so", "libubsan.so")]
    if not all(os.path.isabs(r) and os.path.exists(r) for r in runtimes):
        pytest.skip("sanitizer runtimes not found")
    import sys
    env = dict(os.environ, LD_PRELOAD=":".join(runtimes), ASAN_OPTIONS="detect_leaks=0:abort_on_error=0",
               UBSAN_OPTIONS="print_stacktrace=1", SMSGATE_HOSTCHECK_LIB=lib)
    res = subprocess.run([sys.executable, "-c", _DEVICE_LOGIC_SCRIPT, ROOT], capture_output=True, text=True, env=env, timeout=900)
    assert res.returncode == 0 and "device logic ok" in res.stdout, (res.stdout[-500:], res.stderr[-3000:])
    assert "ERROR: AddressSanitizer" not in res.stderr and "runtime error" not in res.stderr, res.stderr[-3000:]
