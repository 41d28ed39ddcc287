"""Order-preserving int64 codes for the reference's generic key type K (cmp.Ordered: graph.go:305).

The device keeps neighbour lists sorted by an int64 per row because the reference expands neighbours in ascending key
order (graph.go:137-138).  For integer K that int64 is the key.  For any other ordered K (string, float, ...) the host
-- here as it would in Go -- maps keys to int64 CODES that order exactly as the keys do:

  float : the usual monotone image of the IEEE-754 bits (no table needed)
  other : codes spaced GAP apart over the sorted keys; a new key takes the midpoint between its neighbours' codes, and
          when a gap is used up every code is re-spaced and the device is told (b200hnsw_set_keys)

Nothing here touches the device; `KeyCodec.assign` returns the codes of the new keys and, if it had to re-space, the
new code of every key already stored (for the caller to push down).
"""
from __future__ import annotations

import bisect
import math
import struct
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

GAP = 1 << 32
_LO, _HI = -(1 << 62), (1 << 62)


def _float_code(x: float) -> int:
    if math.isnan(x):
        raise ValueError("NaN cannot be a key (it is not equal to itself: a Go map never finds it again)")
    if x == 0.0:
        x = 0.0  # -0.0 == 0.0 as a Go map key
    u = struct.unpack("<q", struct.pack("<d", x))[0]
    return u if u >= 0 else (-(1 << 63)) - u - 1 + 0  # negative floats: reverse their order, below every positive one


class KeyCodec:
    """kind is fixed by the first key seen, as K is fixed by the Go type parameter."""

    def __init__(self):
        self.kind: Optional[str] = None
        self._sorted: List = []          # 'ordered' kind: keys in ascending order ...
        self._codes: List[int] = []      # ... and their codes, parallel
        self._code_of: Dict = {}

    @staticmethod
    def _kind_of(key) -> str:
        if isinstance(key, (bool, np.bool_)):
            raise TypeError("bool is not cmp.Ordered")
        if isinstance(key, (int, np.integer)):
            return "int"
        if isinstance(key, (float, np.floating)):
            return "float"
        if isinstance(key, (str, bytes)):
            return "ordered"
        raise TypeError(f"key type {type(key).__name__} is not supported (int, float, str, bytes)")

    def check(self, key) -> None:
        k = self._kind_of(key)
        if self.kind is None:
            self.kind = k
        elif k != self.kind or (k == "ordered" and self._sorted and type(key) is not type(self._sorted[0])):
            raise TypeError(f"key {key!r} does not have the graph's key type ({self.kind})")

    def code(self, key) -> Optional[int]:
        """Code of a key already stored (None if unknown); for int / float keys the code needs no table."""
        if self.kind in (None, "int"):
            return int(key)
        if self.kind == "float":
            return _float_code(float(key))
        return self._code_of.get(key)

    def assign(self, new_keys: Sequence) -> Tuple[np.ndarray, Optional[Dict]]:
        """Codes for keys not stored yet (distinct, in any order).  Returns (codes[len(new_keys)], respaced) where
        respaced is None or {stored key: new code} for EVERY key stored before this call."""
        for k in new_keys:
            self.check(k)
        if self.kind == "int":
            return np.asarray([int(k) for k in new_keys], dtype=np.int64), None
        if self.kind == "float":
            return np.asarray([_float_code(float(k)) for k in new_keys], dtype=np.int64), None
        old = set(self._code_of)
        order = sorted(range(len(new_keys)), key=lambda i: new_keys[i])
        merged_keys: List = []
        merged_new: List[bool] = []
        # merge the sorted new keys into the sorted stored keys
        i = j = 0
        S = self._sorted
        while i < len(S) or j < len(order):
            if j >= len(order) or (i < len(S) and S[i] < new_keys[order[j]]):
                merged_keys.append(S[i]); merged_new.append(False); i += 1
            else:
                merged_keys.append(new_keys[order[j]]); merged_new.append(True); j += 1
        codes: List[Optional[int]] = [None] * len(merged_keys)
        for p, k in enumerate(merged_keys):
            if not merged_new[p]:
                codes[p] = self._code_of[k]
        respaced = None
        # every run of new keys gets evenly spaced codes strictly between the stored codes around it
        ok = True
        p = 0
        while p < len(merged_keys) and ok:
            if not merged_new[p]:
                p += 1
                continue
            q = p
            while q < len(merged_keys) and merged_new[q]:
                q += 1
            lo = codes[p - 1] if p > 0 else None
            hi = codes[q] if q < len(merged_keys) else None
            r = q - p
            if lo is None and hi is None:
                base, step = 0, GAP
            elif lo is None:
                base, step = hi - GAP * (r + 1), GAP
            elif hi is None:
                base, step = lo, GAP
            else:
                step = (hi - lo) // (r + 1)
                base = lo
            if step < 1 or base + step < _LO or base + step * r > _HI:
                ok = False
                break
            for t in range(r):
                codes[p + t] = base + step * (t + 1)
            p = q
        if not ok:  # a gap ran out (or the range did): re-space everything
            n = len(merged_keys)
            start = -(n // 2) * GAP
            codes = [start + GAP * t for t in range(n)]
            respaced = {k: codes[p] for p, k in enumerate(merged_keys) if k in old}
        self._sorted = merged_keys
        self._codes = codes  # type: ignore[assignment]
        self._code_of = {k: c for k, c in zip(merged_keys, codes)}
        return np.asarray([self._code_of[k] for k in new_keys], dtype=np.int64), respaced

    def forget_new(self, keys: Sequence) -> None:
        """Undo `assign` for keys whose insertion failed."""
        if self.kind != "ordered":
            return
        for k in keys:
            if k in self._code_of:
                p = bisect.bisect_left(self._sorted, k)
                del self._sorted[p]
                del self._codes[p]
                del self._code_of[k]
