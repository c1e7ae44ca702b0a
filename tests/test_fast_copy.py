"""The JSON-tree copier behind every object copy of the control plane (`api/common.py::deepcopy`, the stand-in for
the DeepCopy() methods the reference generates, e.g. api/pytorch/v1/zz_generated.deepcopy.go): equal to
copy.deepcopy, shares nothing mutable with its input, and still copies what is not a plain JSON tree."""
import copy

from hypothesis import given, settings, strategies as st

from kubedl_b200.api import common as c

scalars = st.one_of(st.none(), st.booleans(), st.integers(), st.floats(allow_nan=False), st.text(max_size=8))
trees = st.recursive(scalars, lambda kids: st.one_of(st.lists(kids, max_size=4),
                                                     st.dictionaries(st.text(max_size=6), kids, max_size=4)), max_leaves=25)


def _mutable_ids(o, out):
    if isinstance(o, dict):
        out.add(id(o))
        for v in o.values():
            _mutable_ids(v, out)
    elif isinstance(o, list):
        out.add(id(o))
        for v in o:
            _mutable_ids(v, out)
    return out


@settings(max_examples=200, deadline=None)
@given(trees)
def test_equals_deepcopy_and_shares_no_container(tree):
    out = c.deepcopy(tree)
    assert out == copy.deepcopy(tree) and type(out) is type(tree)
    assert not (_mutable_ids(tree, set()) & _mutable_ids(out, set()))


def test_non_json_members_are_still_copied():
    class Box:
        def __init__(self):
            self.items = [1, 2]
    src = {"t": (1, [2, 3]), "b": Box(), "s": {1, 2}}
    out = c.deepcopy(src)
    assert out["t"] == src["t"] and out["t"][1] is not src["t"][1]
    assert out["b"] is not src["b"] and out["b"].items == [1, 2] and out["b"].items is not src["b"].items
    assert out["s"] == src["s"] and out["s"] is not src["s"]
