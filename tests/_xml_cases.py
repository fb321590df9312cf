"""XML documents shared by the CPU and GPU tests of the tokenizer, and the ElementTree reference
evaluation they are compared with."""
import random
import xml.etree.ElementTree as ET

from smsgate_b200 import synth


def reference(data: bytes):
    """What the reference's loop sees (watcher.py:37-44): ('ok', bodies, senders, dates) with
    dates as int or 'ValueError', or ('ParseError',)."""
    try:
        root = ET.fromstring(data)
    except ET.ParseError:
        return ("ParseError",)
    except Exception:                      # e.g. LookupError for an encoding Python does not know
        return ("Other",)
    bodies, senders, dates = [], [], []
    for e in root.findall("sms"):
        bodies.append(e.get("body", ""))
        senders.append(e.get("address", ""))
        try:
            dates.append(int(e.get("date", "0")))
        except ValueError:
            dates.append("ValueError")
    return ("ok", bodies, senders, dates)


def documents():
    docs = []
    b = synth.generate(400, seed=5, mixed_locale=True)
    import io, os, tempfile
    p = os.path.join(tempfile.mkdtemp(), "a.xml")
    synth.write_xml(p, b)
    base = open(p, "rb").read()
    docs.append(base)
    head = b"<?xml version='1.0' encoding='UTF-8' standalone='yes' ?>\n<!--File Created By SMS Backup & Restore v10.05 on 13/06/2025-->\n<?xml-stylesheet type=\"text/xsl\" href=\"sms.xsl\"?>\n"
    docs.append(head + b'<smses count="2" backup_set="x">\n  <sms protocol="0" address="+37499" date="1707900464673" type="1" subject="null" '
                b'body="hi &amp; bye &#10;x &#x2022; &lt;&gt;" toa="null" read="1" />\n  <mms date="5"><parts><part seq="0" text="t"/></parts>'
                b'<addrs><addr address="1"/></addrs><sms address="in" date="9" body="nested"/></mms>\n'
                b"  <sms address='A' date=' 1_000 ' body='q\"q'></sms>\n</smses>\n")
    small = [
        b'<smses><sms address="A" date="1" body="x"/></smses>',
        b'\xef\xbb\xbf<smses><sms address="A" date="1" body="x"/></smses>',
        b'<smses><sms address="A" date="1" body="x"/></smses>\n\n',
        b'<smses><sms address="A" date="1" body="x"/></smses>junk',
        b'junk<smses><sms address="A" date="1" body="x"/></smses>',
        b'<smses><sms address="A" date="1" body="x"/></smses><smses/>',
        b'<smses><sms address="A" date="1" body="x"/>',
        b'<smses><sms address="A" date="1" body="x"/></smse>',
        b'<smses><sms address="A" date="1" body="x"></sm></smses>',
        b'<smses><sms address="A" date="1" body="a<b"/></smses>',
        b'<smses><sms address="A" date="1" body="a&b"/></smses>',
        b'<smses><sms address="A" date="1" body="a&nbsp;b"/></smses>',
        b'<smses><sms address="A" date="1" body="a&#0;b"/></smses>',
        b'<smses><sms address="A" date="1" body="a&#xD800;b"/></smses>',
        b'<smses><sms address="A" date="1" body="a&#x10FFFF;b&#1114112;"/></smses>',
        b'<smses><sms address="A" date="1" body="a&#xFFFE;b"/></smses>',
        b'<smses><sms address="A" date="1" body="a&#;b"/></smses>',
        b'<smses><sms address="A" date="1" body="a&#x;b"/></smses>',
        b'<smses><sms address="A" date="1" body="a&#12"/></smses>',
        b'<smses><sms address="A" date="1" body="tab\there\r\nnext\rlone\nnl"/></smses>',
        b'<smses><sms address="A" date="1" body="ctrl\x01"/></smses>',
        b'<smses><sms address="A" date="1" body="bad\xff"/></smses>',
        b'<smses><sms address="A" date="1" body="bad\xc3"/></smses>',
        b'<smses><sms address="A" date="1" body="over\xc0\x80"/></smses>',
        b'<smses><sms address="A" date="1" body="sur\xed\xa0\x80"/></smses>',
        b'<smses><sms address="A" date="1" body="non\xef\xbf\xbe"/></smses>',
        b'<smses><sms address="A"date="1" body="x"/></smses>',
        b'<smses><sms address="A" date="1" body="x" /></smses>',
        b'<smses><sms address = "A" date\n=\n"1" body="x"/></smses>',
        b'<smses><sms address="A" date="1" body=x/></smses>',
        b'<smses><sms address="A" date="1" body="x/></smses>',
        b'<smses><sms address="A" date="1" body/></smses>',
        b'<smses><sms address="A" date="1" body="x" body="y"/></smses>',
        b'<smses><sms address="A" date="1" Body="x"/></smses>',
        b'<smses><sms 1a="A" date="1" body="x"/></smses>',
        b'<smses><Sms address="A" date="1" body="x"/></smses>',
        b'<smses><sms/><sms address="A"/><sms body="only"/></smses>',
        b'<smses><sms address="A" date="+5" body="x"/><sms address="A" date="-5" body="x"/><sms address="A" date="" body="x"/></smses>',
        b'<smses><sms address="A" date="1.5" body="x"/><sms address="A" date="0x10" body="x"/><sms address="A" date="1__0" body="x"/>'
        b'<sms address="A" date="_1" body="x"/><sms address="A" date="1_" body="x"/><sms address="A" date="\xd9\xa1\xd9\xa2" body="x"/></smses>',
        b'<smses><sms address="A" date="99999999999999999999999" body="x"/></smses>',
        b'<smses>text &amp; more<sms address="A" date="1" body="x"/>tail &#65; ]] > ok</smses>',
        b'<smses>bad & text<sms address="A" date="1" body="x"/></smses>',
        b'<smses>]]><sms address="A" date="1" body="x"/></smses>',
        b'<smses><!-- note --><sms address="A" date="1" body="x"/><!-- a -- b --></smses>',
        b'<smses><!-- has <tag> inside --><sms address="A" date="1" body="x"/></smses>',
        b'<!DOCTYPE smses><smses><sms address="A" date="1" body="x"/></smses>',
        b'<smses><![CDATA[ <sms body="no"/> ]]><sms address="A" date="1" body="x"/></smses>',
        b"<?xml version='1.0' encoding='ISO-8859-1'?><smses><sms address='A' date='1' body='x'/></smses>",
        b"<?xml version='1.0' encoding='utf-8'?><smses><sms address='A' date='1' body='x'/></smses>",
        b" <?xml version='1.0'?><smses><sms address='A' date='1' body='x'/></smses>",
        b"<?xml version=''?><smses><sms address='A' date='1' body='x'/></smses>",
        b"<?xml version='1:0'?><smses><sms address='A' date='1' body='x'/></smses>",
        b"<?xml version='Z9_.-'?><smses><sms address='A' date='1' body='x'/></smses>",
        b"<?xml version='1.0' standalone=''?><smses><sms address='A' date='1' body='x'/></smses>",
        b"<?xml version='1.0'encoding='utf-8'?><smses><sms address='A' date='1' body='x'/></smses>",
        b"<smses><?xml version='1.0'?><sms address='A' date='1' body='x'/></smses>",
        b"<smses><?pi data?><sms address='A' date='1' body='x'/></smses>",
        b'<a:smses xmlns:a="u"><sms address="A" date="1" body="x"/></a:smses>',
        b'<smses><sms address="A" date="1" body="x"/ ></smses>',
        b'<smses>< sms address="A" date="1" body="x"/></smses>',
        b'<smses></smses>', b'<smses/>', b'', b'   ', b'<', b'<>', b'</smses>', b'<smses><sms></smses>',
    ]
    docs += small
    rnd = random.Random(17)
    seed_doc = (b'<?xml version="1.0" encoding="UTF-8"?>\n<smses count="3">\n <sms address="BANK" date="1707900464673" body="Pay 5&#10;now &amp; go" read="1"/>\n'
                b' <sms address="\xd0\x91\xd0\xb0\xd0\xbd\xd0\xba" date="17" body="\xd0\xa1\xd1\x83\xd0\xbc\xd0\xbc\xd0\xb0 1&#160;234,56"/>\n <sms date="3" address="C" body="z"></sms>\n</smses>\n')
    for _ in range(1500):
        d = bytearray(seed_doc)
        for _ in range(rnd.randint(1, 3)):
            k = rnd.randrange(len(d))
            op = rnd.random()
            if op < 0.35:
                del d[k]
            elif op < 0.7:
                d[k] = rnd.choice(b'<>/"\'&;= #xa0\n\r\t-!?s')
            else:
                d.insert(k, rnd.choice(b'<>/"\'&;= #xa0\n\r\t-!?s'))
        docs.append(bytes(d))
    return docs
