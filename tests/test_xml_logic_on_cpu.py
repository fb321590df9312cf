"""The XML tokenizer's per-tag logic (xml_core.cuh, CPU build) against ElementTree."""
from smsgate_b200 import _lib as L
from tests import _hostcheck as H
from tests._xml_cases import documents, reference


def compare(result, data):
    """result = (code, bodies, senders, dates, mflags) from the tokenizer; returns 'unsupported' or 'checked'."""
    code, bodies, senders, dates, mflags = result
    want = reference(data)
    if code == 2:
        return "unsupported"
    assert want[0] != "Other", data         # only the general parser may see those
    if want[0] == "ParseError":
        assert code == 1, data
        return "checked"
    assert code == 0, data
    assert bodies == want[1], data
    assert senders == want[2], data
    for got, fl, w in zip(dates, mflags, want[3]):
        if w == "ValueError":
            assert fl & L.XML_DATE_BAD, data
        elif abs(w) >= 2 ** 63:
            assert fl & L.XML_DATE_BIG, data
        else:
            assert not (fl & (L.XML_DATE_BAD | L.XML_DATE_BIG)) and got == w, data
    return "checked"


def test_tokenizer_matches_elementtree():
    n_unsup = 0
    docs = documents()
    for data in docs:
        n_unsup += compare(H.xml_extract(data), data) == "unsupported"
    assert n_unsup < len(docs) * 0.2


def test_iter_sms_host_logic_against_reference_run(golden, tmp_path):
    """xml_ingest.iter_sms above the tokenizer (CPU build) against tests/golden/iter_sms.json, a run of the
    reference's `_iter_sms`: the rows it yields, the exception it ends with, and the rows yielded BEFORE that
    exception - a date `int()` refuses in the middle of a file comes after the messages in front of it."""
    from smsgate_b200 import xml_ingest
    engine = H.HostTokenizerEngine()
    for c in golden("iter_sms.json"):
        p = tmp_path / (c["name"] + ".xml")
        p.write_text(c["xml"], encoding="utf-8", newline="")
        rows, error = [], None
        try:
            for r in xml_ingest.iter_sms(p, engine):
                rows.append(r.model_dump())
        except Exception as e:                     # noqa: BLE001
            error = type(e).__name__
        assert rows == c["rows"] and error == c.get("error"), c["name"]
    assert engine.host_fallbacks >= 4              # the files whose dates raise, and the CDATA one
