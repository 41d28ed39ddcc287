"""hnsw_b200/keys.py: order-preserving int64 codes for non-integer key types (the reference's K is any cmp.Ordered)."""
import math
import random

import numpy as np
import pytest

from hnsw_b200 import keys as K


def _check_order(codec, stored):
    ks = sorted(stored)
    cs = [codec.code(k) for k in ks]
    assert all(a < b for a, b in zip(cs, cs[1:])), "codes must be strictly increasing with the keys"
    assert all(-(1 << 63) <= c < (1 << 63) for c in cs)


def test_float_codes_are_monotone():
    xs = [-math.inf, -1e300, -2.0, -1.0, -1e-300, -5e-324, 0.0, 5e-324, 1e-300, 1.0, 2.0, 1e300, math.inf]
    c = K.KeyCodec()
    codes, resp = c.assign(xs)
    assert resp is None and list(codes) == sorted(codes) and len(set(codes)) == len(xs)
    assert c.code(-0.0) == c.code(0.0)
    with pytest.raises(ValueError):
        K.KeyCodec().assign([float("nan")])


def test_string_codes_incremental_and_respace(monkeypatch):
    rnd = random.Random(5)
    words = list({"".join(rnd.choice("abcdefgh") for _ in range(rnd.randint(1, 6))) for _ in range(900)})
    rnd.shuffle(words)
    for gap in (1 << 32, 4):  # a gap of 4 forces re-spacing
        monkeypatch.setattr(K, "GAP", gap)
        c = K.KeyCodec()
        stored, current, respaces = [], {}, 0
        for a in range(0, len(words), 37):
            batch = words[a:a + 37]
            codes, resp = c.assign(batch)
            if resp is not None:
                respaces += 1
                assert set(resp) == set(stored)
                current.update(resp)
            current.update(dict(zip(batch, (int(x) for x in codes))))
            stored += batch
            assert all(c.code(k) == current[k] for k in stored)
            _check_order(c, stored)
        assert (respaces > 0) == (gap == 4)
    c.forget_new(words[:5])
    assert all(c.code(k) is None for k in words[:5])
    _check_order(c, words[5:])


def test_kind_is_fixed_by_the_first_key():
    c = K.KeyCodec()
    c.assign(["a"])
    with pytest.raises(TypeError):
        c.assign([3])
    with pytest.raises(TypeError):
        K.KeyCodec().assign([True])
    ci = K.KeyCodec()
    codes, _ = ci.assign([5, -7, np.int64(9)])
    assert list(codes) == [5, -7, 9] and ci.kind == "int"
