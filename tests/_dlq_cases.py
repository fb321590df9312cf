"""DLQ payloads shared by the CPU and the GPU tests: what the worker and the writer publish to sms.failed
(services/parser_worker/worker.py:105,144,154,166,177; services/pb_writer/writer.py:80-83) plus damaged ones."""
import hashlib
import json


def payloads(bodies, senders):
    out = []
    for i, (b, s) in enumerate(zip(bodies, senders)):
        sha1 = hashlib.sha1(b.encode()).hexdigest()
        md5 = hashlib.md5(b.encode()).hexdigest()
        kind = i % 10
        raw = {"msg_id": sha1, "sender": s, "body": b, "date": "2024-02-14T09:00:00+00:00", "device_id": "xml_backup",
               "source": "xml"}
        if kind == 0:       # device-sourced: MD5 id, numeric date, no device id
            raw = {"msg_id": md5, "sender": s, "body": b, "date": "1707900000000", "source": "device"}
        elif kind == 1:     # an id shared by several messages (hand-made, not a digest)
            raw["msg_id"] = "dup-%d" % (i % 7)
        elif kind == 2:     # upper-case hex is a different id than lower-case hex
            raw["msg_id"] = sha1.upper()
        if kind == 3:
            out.append(json.dumps({"err": "boom", "entry": raw}).encode())             # no `raw`
        elif kind == 4 and i % 20 == 4:
            out.append(b"\xff\xfe not json")
        elif kind == 4:
            bad = dict(raw)
            bad.pop("body")
            out.append(json.dumps({"reason": "unmatched", "raw": bad}).encode())       # RawSMS validation fails
        elif kind == 5 and i % 20 == 5:
            out.append(json.dumps([1, 2, 3]).encode())                                 # not an object
        elif kind == 5:
            out.append(json.dumps({"reason": "unmatched", "raw": dict(raw, sender="")}).encode())
        elif kind == 6 and i % 20 == 6:
            out.append(json.dumps({"reason": "unmatched", "raw": [raw]}).encode())     # raw is not a mapping
        elif kind == 7 and i % 20 == 7:
            out.append(json.dumps({"reason": "unmatched", "raw": dict(raw, msg_id=123)}).encode())
        elif kind == 8 and i % 20 == 8:
            out.append(json.dumps({"reason": "unmatched", "raw": None}).encode())
        else:
            out.append(json.dumps({"reason": "unmatched", "raw": raw}, ensure_ascii=(i % 3 == 0)).encode())
    # the same message again at the end: last occurrence of an id wins
    out.append(out[9])
    out.append(json.dumps({"reason": "unmatched", "raw": {"msg_id": "x", "sender": "B", "body": "hi", "date": "d",
                                                           "source": "carrier-pigeon"}}).encode())
    return out
