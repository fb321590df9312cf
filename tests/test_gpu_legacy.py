"""Legacy offline replay and XML ingestion through the C ABI on the GPU."""
import json
import os
import random
import struct
import xml.etree.ElementTree as ET

import numpy as np
import pytest

from oracle import legacy_oracle as LO
from smsgate_b200 import _lib as L
from smsgate_b200 import legacy, synth, xml_ingest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
    from smsgate_b200.replay import ReplayEngine
    eng = ReplayEngine(0)
    yield eng
    eng.close()


def _bits(d):
    if d is None:
        return None
    out = dict(d)
    for k in ("amount", "balance"):
        out[k] = "%016x" % struct.unpack("<Q", struct.pack("<d", d[k]))[0]
    return out


def _check(res, bodies):
    for i, b in enumerate(bodies):
        st, text, digest, parsed = LO.process_body(b)
        fl = int(res.cols["lflags"][i])
        assert not fl & L.LEG_F_BUDGET, b
        assert res.stripped(i) == text and res.body_hash_hex(i) == digest.hex(), b
        assert int(res.cols["status"][i]) == st, b
        if parsed is not None and not fl & L.LEG_F_INEXACT:
            assert _bits(res.parsed(i)) == _bits(parsed), b


def test_legacy_golden_and_fuzz(engine, golden):
    bodies = [c["body"] for c in golden("legacy.json")["cases"]]
    rnd = random.Random(5)
    base = ("APPROVED PURCHASE DB SALE: TEST LLC, YEREVAN, TEST STR. 29, 24 AREA,06.05.25 14:23,card ***0018. "
            "Amount:52.00 USD, Balance:1842.74 USD")
    for _ in range(3000):
        s = list(base)
        for _ in range(rnd.randint(1, 4)):
            k = rnd.randrange(len(s))
            s[k] = rnd.choice(" ,.:*\n9aAſ٥")
        bodies.append("".join(s))
    _check(legacy.parse_bodies(engine, bodies), bodies)


def test_legacy_on_synthetic_backup(engine):
    backup = synth.generate(20000, seed=77, mixed_locale=True)
    _check(legacy.parse_bodies(engine, backup.bodies), backup.bodies)


def test_process_sms_from_cache_matches_reference_run(engine, golden, tmp_path):
    import diskcache
    g = golden("legacy.json")["process"]
    src, deb, cre = (str(tmp_path / n) for n in ("src", "deb", "cre"))
    with diskcache.Cache(src) as c:
        for i, b in enumerate(g["bodies"]):
            c.set(str(1707900000000 + i), {"address": "BANK", "date": str(1707900000000 + i), "body": b})
        c.set("already", {"body": g["bodies"][0], "status": "processed"})
        c.set("nobody", {"address": "X"})
    stats = legacy.process_sms_from_cache(engine, src, deb, cre)
    with diskcache.Cache(src) as c:
        assert {k: c.get(k).get("status") for k in c} == g["statuses"]
    for path, want in ((deb, g["debit"]), (cre, g["credit"])):
        with diskcache.Cache(path) as c:
            got = {k: dict(_bits({kk: vv for kk, vv in c.get(k).items() if kk not in ("msg_id", "original_body")}),
                           msg_id=c.get(k)["msg_id"], original_body=c.get(k)["original_body"]) for k in c}
        assert got == want
    assert stats["skipped"] == 1 + sum(1 for v in g["statuses"].values() if v == "skipped_otp")
    assert stats["failed"] == sum(1 for v in g["statuses"].values() if v == "failed_to_parse")


def test_iter_sms_golden(engine, golden, tmp_path):
    for c in golden("iter_sms.json"):
        p = tmp_path / (c["name"] + ".xml")
        p.write_text(c["xml"], encoding="utf-8", newline="")
        rows, error = [], None                     # a generator: what it yields before it raises counts too
        try:
            for r in xml_ingest.iter_sms(p, engine):
                rows.append(r.model_dump())
        except Exception as e:                     # noqa: BLE001 - the type is compared with the reference's
            error = type(e).__name__
        assert rows == c["rows"] and error == c.get("error"), c["name"]


def test_parse_and_cache_sms_matches_reference_run(engine, golden, tmp_path):
    """tests/golden/read_xml.json: the DiskCache the reference's own read_xml.parse_and_cache_sms left behind
    for each file (duplicate / missing / empty `date`, reversed attribute order, entities, a malformed file, a
    missing file, a DOCTYPE the tokenizer hands to ElementTree).  Same keys in the same order, same dicts."""
    import diskcache
    fallbacks0 = engine.host_fallbacks
    for c in golden("read_xml.json"):
        xml = tmp_path / (c["name"] + ".xml")
        cdir = tmp_path / (c["name"] + "_cache")
        if c["xml_hex"] is not None:
            xml.write_bytes(bytes.fromhex(c["xml_hex"]))
        n = xml_ingest.parse_and_cache_sms(str(xml), str(cdir), engine)
        with diskcache.Cache(str(cdir)) as cache:
            got = [[k, cache[k]] for k in cache]
        assert got == c["entries"], c["name"]
        for (_, a), (_, b) in zip(got, c["entries"]):
            assert list(a) == list(b), c["name"]                     # attribute order inside the stored dict
        if c["name"] in ("malformed", "missing_file"):
            assert n is None
    assert engine.host_fallbacks == fallbacks0 + 1                   # only the DOCTYPE file left the GPU path


def test_xml_attribs_on_gpu_match_elementtree(engine):
    """Every attribute of every top-level <sms> (read_xml.py:34) for the tokenizer's whole document set."""
    import xml.etree.ElementTree as ET

    from smsgate_b200.replay import UnsupportedXML
    from tests._xml_cases import documents
    checked = 0
    for data in documents():
        try:
            want = [dict(e.attrib) for e in ET.fromstring(data).findall("sms")]
        except Exception:
            want = None
        try:
            got = engine.xml_attribs(np.frombuffer(data, dtype=np.uint8))
        except UnsupportedXML:
            continue
        except ET.ParseError:
            got = None
        assert got == want, data[:200]
        if want:
            assert [list(d) for d in got] == [list(d) for d in want]
            checked += 1
    assert checked > 250


def test_xml_tokenizer_on_gpu_matches_elementtree(engine):
    """smsgate_xml_scan / _emit against ElementTree on the shared documents (synthetic backups, the
    SMS Backup & Restore prolog, malformed and out-of-dialect files, 1 500 mutated documents)."""
    import xml.etree.ElementTree as ET

    from smsgate_b200.replay import UnsupportedXML
    from tests._xml_cases import documents
    from tests.test_xml_logic_on_cpu import compare
    n_unsup = 0
    docs = documents()
    for data in docs:
        try:
            d = engine.xml_extract(data)
        except ET.ParseError:
            res = (1, [], [], [], [])
        except UnsupportedXML:
            res = (2, [], [], [], [])
        else:
            n = d["n"]
            off, soff = d["body_offsets"].cpu().numpy(), d["sender_offsets"].cpu().numpy()
            bb, sb = d["body_bytes"].cpu().numpy(), d["sender_bytes"].cpu().numpy()
            res = (0, [bb[off[i]:off[i + 1]].tobytes().decode() for i in range(n)],
                   [sb[soff[i]:soff[i + 1]].tobytes().decode() for i in range(n)],
                   d["date_ms"].cpu().numpy()[:n].tolist(), d["mflags"].cpu().numpy()[:n].tolist())
        n_unsup += compare(res, data) == "unsupported"
    assert n_unsup < len(docs) * 0.2


def test_xml_to_replay_on_device(engine, tmp_path):
    """A 50 000-message backup: XML bytes -> GPU tokenizer -> replay kernels without the bodies ever
    visiting the host, equal to the host-parsed replay."""
    import datetime as dt

    import torch

    from smsgate_b200.batch import Batch
    now = dt.datetime(2026, 1, 1)
    backup = synth.generate(50000, seed=8, mixed_locale=True, dup_frac=0.1)
    p = tmp_path / "big.xml"
    synth.write_xml(str(p), backup)
    engine.load_cache(backup.cache)
    host = engine.replay(xml_ingest.read_backup(p), now)
    d = engine.xml_extract(np.fromfile(str(p), dtype=np.uint8))
    assert d["n"] == 50000
    total = int(d["body_offsets"][-1])
    cols = engine.device_columns(d["n"], total)
    engine.reserve(d["n"], total)
    engine.replay_device(d["body_bytes"], d["body_offsets"], now, cols)
    torch.cuda.synchronize()
    for name in ("msg_id", "cache_key", "status", "survivor", "amount_coef", "date"):
        assert np.array_equal(cols[name].cpu().numpy(), host.cols[name]), name
    assert np.array_equal(d["date_ms"].cpu().numpy(), np.array(backup.date_ms))


def test_pb_builders_match_reference_vectors(engine, golden):
    """save_to_pocketbase.py:34-78 in bulk: `_dt_str` on the GPU, the payload dicts around it."""
    from smsgate_b200 import legacy
    g = golden("pb_builders.json")
    got = legacy.dt_str_batch(engine, [(d, t) for d, t, _ in g["dt_str"]])
    assert got == [w for _, _, w in g["dt_str"]]
    assert legacy.build_sms_data(engine, [r for r, _ in g["sms_data"]]) == [w for _, w in g["sms_data"]]
    assert legacy.build_transactions(engine, [r for r, _ in g["transactions"]]) == [w for _, w in g["transactions"]]
    assert legacy.dt_str_batch(engine, []) == []
