"""Arrow-style ragged batches: int64 offsets[n+1] + uint8 bytes."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Iterable, Optional, Sequence

import numpy as np


def ragged_from_strings(strings: Sequence[str], offset_dtype=np.int64, errors: str = "strict"):
    """UTF-8 encode and concatenate.  Returns (bytes uint8[total], offsets[n+1])."""
    enc = [s.encode("utf-8", errors) for s in strings]
    offsets = np.zeros(len(enc) + 1, dtype=offset_dtype)
    if enc:
        np.cumsum(np.fromiter((len(b) for b in enc), dtype=np.int64, count=len(enc)), out=offsets[1:])
    data = np.frombuffer(b"".join(enc), dtype=np.uint8)
    return data, offsets


@dataclass
class Batch:
    """One backup's worth of messages in replay order.

    bodies are what `_iter_sms` reads from the `body` attribute
    (services/xml_watcher/watcher.py:41); senders the `address` attribute (:44);
    date_ms `int(date)` (:42)."""
    body_bytes: np.ndarray            # uint8[total]
    body_offsets: np.ndarray          # int64[n+1]
    senders: Optional[list] = None
    date_ms: Optional[np.ndarray] = None

    @property
    def n(self) -> int:
        return len(self.body_offsets) - 1

    @classmethod
    def from_strings(cls, bodies: Sequence[str], senders: Optional[Iterable[str]] = None,
                     date_ms: Optional[Iterable[int]] = None) -> "Batch":
        data, offsets = ragged_from_strings(bodies)
        return cls(data, offsets, list(senders) if senders is not None else None,
                   np.asarray(list(date_ms), dtype=np.int64) if date_ms is not None else None)

    def body(self, i: int) -> str:
        a, b = int(self.body_offsets[i]), int(self.body_offsets[i + 1])
        return self.body_bytes[a:b].tobytes().decode("utf-8")

    def slice(self, lo: int, hi: int) -> "Batch":
        a, b = int(self.body_offsets[lo]), int(self.body_offsets[hi])
        return Batch(self.body_bytes[a:b], self.body_offsets[lo:hi + 1] - a,
                     self.senders[lo:hi] if self.senders is not None else None,
                     self.date_ms[lo:hi] if self.date_ms is not None else None)
