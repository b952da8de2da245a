"""Comparison helpers: oracle state vs product state."""
import numpy as np

VAR_NAMES = {1: "phi", 2: "rhs", 3: "old", 4: "res"}


def face_slices(nc):
    """index tuples [k, j, i] of the six ghost faces (tangential range 1:nc only)."""
    s = slice(1, nc + 1)
    return [(s, s, 0), (s, s, nc + 1), (s, 0, s), (s, nc + 1, s), (0, s, s), (nc + 1, s, s)]


def _maxdiff(a, b):
    """max |a - b|; any NaN / inf on either side counts as an infinite difference"""
    if not (np.isfinite(a).all() and np.isfinite(b).all()):
        return float("inf")
    return float(np.abs(a - b).max())


def assert_trees_equal(a, b):
    assert a.lowest_lvl == b.lowest_lvl and a.highest_lvl == b.highest_lvl
    assert a.first_normal_lvl == b.first_normal_lvl and a.n_boxes == b.n_boxes
    np.testing.assert_array_equal(a.tree_export(), b.tree_export())
    for lvl in range(a.lowest_lvl, a.highest_lvl + 1):
        assert a.box_size_lvl(lvl) == b.box_size_lvl(lvl)
        np.testing.assert_array_equal(a.dr(lvl), b.dr(lvl))
        assert a.ids(lvl) == b.ids(lvl)
        assert a.leaves(lvl) == b.leaves(lvl)
        assert a.parents(lvl) == b.parents(lvl)
        assert a.ref_bnds(lvl) == b.ref_bnds(lvl)


def state_diff(orc, mg, variables=(1, 2, 3, 4), ghosts_of=(1,), levels=None, ids_of=None):
    """max |difference| per variable over the interiors of all boxes (and the ghost faces of ghosts_of).
    Returns {name: (max_abs_diff, max_abs_value)}."""
    out = {}
    lv = range(orc.lowest_lvl, orc.highest_lvl + 1) if levels is None else levels
    for iv in variables:
        d = 0.0
        m = 0.0
        for lvl in lv:
            nc = orc.box_size_lvl(lvl)
            for id_ in (ids_of(lvl) if ids_of else orc.ids(lvl)):
                a = orc.cc(id_, nc)[iv - 1]
                b = mg.cc(id_, nc)[iv - 1]
                d = max(d, _maxdiff(a[1:-1, 1:-1, 1:-1], b[1:-1, 1:-1, 1:-1]))
                m = max(m, float(np.abs(a[1:-1, 1:-1, 1:-1]).max()))
                if iv in ghosts_of:
                    for sl in face_slices(nc):
                        d = max(d, _maxdiff(a[sl], b[sl]))
        out[VAR_NAMES.get(iv, "var%d" % iv)] = (d, m)
    return out


def assert_state_equal(orc, mg, **kw):
    """Bitwise agreement (difference exactly zero; -0.0 == +0.0)."""
    diff = state_diff(orc, mg, **kw)
    bad = {k: v for k, v in diff.items() if v[0] != 0.0}
    assert not bad, "oracle/product mismatch (max abs diff, max abs value): %r" % bad
