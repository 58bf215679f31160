"""The JSON value behind the serialized-schedule format (include/tenzing/json.hpp): anything
Python's json module writes is parsed and written back unchanged (compact, keys sorted - the
reference's nlohmann::json::dump() form), and arbitrary text never does worse than an error."""
import json
import math

import pytest
from hypothesis import given, settings, strategies as st

from tenzing_b200 import _C

values = st.recursive(
    st.none() | st.booleans() | st.integers(min_value=-2 ** 53, max_value=2 ** 53)
    | st.floats(allow_nan=False, allow_infinity=False, width=64) | st.text(max_size=12),
    lambda children: st.lists(children, max_size=4) | st.dictionaries(st.text(max_size=6), children, max_size=4),
    max_leaves=12)


def _norm(v):
    """what a round trip may legitimately change: -0.0 / integral floats print like nlohmann does"""
    if isinstance(v, float):
        return float(v)
    if isinstance(v, list):
        return [_norm(x) for x in v]
    if isinstance(v, dict):
        return {k: _norm(x) for k, x in v.items()}
    return v


@settings(max_examples=300, deadline=None)
@given(values)
def test_round_trip_of_arbitrary_values(v):
    text = json.dumps(v, ensure_ascii=False)
    out = _C.json_roundtrip(text)
    back = json.loads(out)
    assert _norm(back) == _norm(v) or (isinstance(v, float) and math.isclose(back, v, rel_tol=1e-15))
    assert _C.json_roundtrip(out) == out                       # a fixed point
    if isinstance(v, dict):
        assert list(back.keys()) == sorted(back.keys())        # keys come out sorted


@settings(max_examples=400, deadline=None)
@given(st.text(alphabet='{}[]",:\\ntrufalse0123456789.-+eE \t', max_size=40))
def test_garbage_is_rejected_cleanly(text):
    try:
        out = _C.json_roundtrip(text)
    except RuntimeError:
        return
    json.loads(out)                                             # whatever was accepted is JSON
