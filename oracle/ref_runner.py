"""TEST INFRASTRUCTURE — runs the *unmodified* reference in this container.

Only `oracle/gen_golden.py` (fixture generation, run once here) imports this
module.  `/root/reference` does not exist on the GPU box, so nothing in
`tests/`, `bench.py` or `__graft_entry__.py` may import it at run time.

What it does: puts `/root/reference` on `sys.path`, injects `sys.modules`
stubs for the three third-party packages that are absent from this image
(`google.genai`, `nats`, `async_lru`), sets `GEMINI_API_KEY`, changes into a
scratch directory holding a seeded `.gemini_cache` (the reference opens that
path relative to cwd at import time, libs/gemini_parser.py:33) and imports
the reference modules.  No reference source is copied or edited.
"""
from __future__ import annotations

import contextlib
import io
import os
import sys
import types

REFERENCE_ROOT = "/root/reference"


def _install_stubs() -> None:
    if "google.genai" not in sys.modules:
        google = sys.modules.get("google") or types.ModuleType("google")
        google.__path__ = getattr(google, "__path__", [])
        genai = types.ModuleType("google.genai")
        gtypes = types.ModuleType("google.genai.types")

        class _Anything:
            def __init__(self, *a, **k):
                pass

            @classmethod
            def from_text(cls, *a, **k):
                return cls()

        class _Type:
            STRING = "STRING"
            OBJECT = "OBJECT"

        for name in ("Schema", "Content", "Part", "GenerateContentConfig"):
            setattr(gtypes, name, _Anything)
        gtypes.Type = _Type

        class _Models:
            def generate_content_stream(self, *a, **k):
                raise RuntimeError("offline: Gemini is not reachable")

        class Client:
            def __init__(self, *a, **k):
                self.models = _Models()

        genai.Client = Client
        genai.types = gtypes
        google.genai = genai
        sys.modules["google"] = google
        sys.modules["google.genai"] = genai
        sys.modules["google.genai.types"] = gtypes
    if "async_lru" not in sys.modules:
        m = types.ModuleType("async_lru")
        m.alru_cache = lambda *a, **k: (a[0] if a and callable(a[0]) else (lambda f: f))
        sys.modules["async_lru"] = m
    if "nats" not in sys.modules:
        nats = types.ModuleType("nats")
        nats.__path__ = []
        aio = types.ModuleType("nats.aio")
        aio.__path__ = []
        client = types.ModuleType("nats.aio.client")
        msg = types.ModuleType("nats.aio.msg")
        js = types.ModuleType("nats.js")
        js.__path__ = []
        api = types.ModuleType("nats.js.api")
        errors = types.ModuleType("nats.js.errors")
        nerrors = types.ModuleType("nats.errors")

        class _Dummy:
            def __init__(self, *a, **k):
                pass

        client.Client = _Dummy
        msg.Msg = _Dummy
        for name in ("StreamConfig", "RetentionPolicy", "StorageType", "ConsumerConfig",
                     "DeliverPolicy", "AckPolicy", "DiscardPolicy", "PubAck"):
            setattr(api, name, _Dummy)
        # any other name the reference imports from these modules is a dummy too
        for mod in (api, client, msg):
            mod.__getattr__ = lambda name, _d=_Dummy: _d
        for name in ("NotFoundError", "BadRequestError", "APIError", "Error"):
            setattr(errors, name, type(name, (Exception,), {}))
        for name in ("TimeoutError", "NoServersError", "ConnectionClosedError", "Error"):
            setattr(nerrors, name, type(name, (Exception,), {}))

        async def _connect(*a, **k):
            raise RuntimeError("offline: NATS is not reachable")

        nats.connect = _connect
        nats.aio = aio
        nats.js = js
        nats.errors = nerrors
        aio.client = client
        aio.msg = msg
        js.api = api
        js.errors = errors
        sys.modules.update({
            "nats": nats, "nats.aio": aio, "nats.aio.client": client, "nats.aio.msg": msg,
            "nats.js": js, "nats.js.api": api, "nats.js.errors": errors, "nats.errors": nerrors,
        })


class Reference:
    """Handle on the imported reference modules, bound to one scratch cwd."""

    def __init__(self, workdir: str):
        if not os.path.isdir(REFERENCE_ROOT):
            raise RuntimeError("the reference tree is only mounted in the build container")
        self.workdir = workdir
        os.makedirs(workdir, exist_ok=True)
        _install_stubs()
        os.environ.setdefault("GEMINI_API_KEY", "dummy")
        if REFERENCE_ROOT not in sys.path:
            sys.path.insert(0, REFERENCE_ROOT)
        self._cwd = os.getcwd()
        os.chdir(workdir)
        with contextlib.redirect_stdout(io.StringIO()):
            import libs.decimal_utils as decimal_utils
            import libs.gemini_parser as gemini_parser
            import libs.models as models
            import services.xml_watcher.watcher as watcher
            import read_xml
            import process_cached
        self.decimal_utils = decimal_utils
        self.gemini_parser = gemini_parser
        self.models = models
        self.watcher = watcher
        self.read_xml = read_xml
        self.process_cached = process_cached

    # -- the Gemini response cache the reference opened at import ------------
    def seed_cache(self, entries: dict) -> None:
        cache = self.gemini_parser.cache
        cache.clear()
        for k, v in entries.items():
            cache[k] = v

    def quiet(self):
        return contextlib.redirect_stdout(io.StringIO())

    def close(self):
        os.chdir(self._cwd)
