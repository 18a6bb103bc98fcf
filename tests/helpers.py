"""Shared helpers for the parity tests: decode golden-vector shorthand, drive oracle and product the same way."""
import numpy as np

import rdf_fusion_b200 as rf
from rdf_fusion_b200 import _ffi as F

PERMS = {"GSPO": F.GSPO, "GPOS": F.GPOS, "GOSP": F.GOSP, "GPSO": F.GPSO, "GSOP": F.GSOP, "GOPS": F.GOPS}
OPS = {"eq": F.OP_EQ, "gt": F.OP_GT, "gteq": F.OP_GTEQ, "lt": F.OP_LT, "lteq": F.OP_LTEQ}


def decode_pred(d):
    if d is None:
        return None
    if "false" in d:
        return rf.FALSE
    if "in" in d:
        return rf.In(*d["in"])
    if "between" in d:
        return rf.Between(*d["between"])
    if "equal_to" in d:
        return rf.EqualTo(d["equal_to"])
    raise ValueError(d)


def decode_instr(e):
    if "s" in e:
        pred = None
        if "in" in e:
            pred = rf.In(*e["in"])
        elif "between" in e:
            pred = rf.Between(*e["between"])
        return rf.scan(e["s"], pred)
    if "t_in" in e:
        return rf.traverse(rf.In(*e["t_in"]))
    if "t_between" in e:
        return rf.traverse(rf.Between(*e["t_between"]))
    if e.get("t") is None:
        return rf.traverse()
    return rf.traverse(int(e["t"]))


def case_quads(case):
    if "quads_gen" in case:
        g = case["quads_gen"]
        if g["kind"] == "last_level":  # mod.rs:283-285
            return [[1, 2, 3, i + 1] for i in range(g["n"])]
        if g["kind"] == "pred_level":  # mod.rs:306-308
            return [[1, 2, i, 3] for i in range(g["n"])]
        raise ValueError(g)
    return case.get("quads", [])


def canonical(cols):
    """rows sorted lexicographically → (rows, cols) uint32 array; NULL = 0 sorts first"""
    cols = [np.asarray(c, dtype=np.uint32) for c in cols]
    if not cols:
        return np.zeros((0, 0), dtype=np.uint32)
    m = np.stack(cols, axis=1)
    if len(m) == 0:
        return m
    order = np.lexsort(tuple(m[:, i] for i in reversed(range(m.shape[1]))))
    return m[order]


def assert_same_table(names_a, cols_a, names_b, cols_b, ordered=False):
    assert list(names_a) == list(names_b), (names_a, names_b)
    a = np.stack([np.asarray(c, np.uint32) for c in cols_a], axis=1) if cols_a else np.zeros((0, 0), np.uint32)
    b = np.stack([np.asarray(c, np.uint32) for c in cols_b], axis=1) if cols_b else np.zeros((0, 0), np.uint32)
    assert a.shape == b.shape, (a.shape, b.shape)
    if not ordered:
        a, b = canonical(list(a.T)) if a.size else a, canonical(list(b.T)) if b.size else b
    assert np.array_equal(a, b)


def random_quads(rng, n, n_graphs=3, n_s=40, n_p=6, n_o=60, default_graph_frac=0.6):
    g = np.where(rng.random(n) < default_graph_frac, 0, rng.integers(1, n_graphs + 1, n)).astype(np.uint32)
    s = rng.integers(1, n_s + 1, n).astype(np.uint32)
    p = rng.integers(1, n_p + 1, n).astype(np.uint32)
    o = rng.integers(1, n_o + 1, n).astype(np.uint32)
    return np.stack([g, s, p, o], axis=1)


def random_pattern(rng, quads, var_names=("a", "b", "c", "d")):
    """a random scan program in GSPO order: constants drawn from existing quads (so matches are likely), variables
    (sometimes repeated), In sets, Between ranges, and the occasional False."""
    row = quads[rng.integers(0, len(quads))] if len(quads) else np.array([0, 1, 1, 1])
    instr = []
    for lvl in range(4):
        r = rng.random()
        hi = int(quads[:, lvl].max()) if len(quads) else 5
        if r < 0.35:
            instr.append(rf.traverse(int(row[lvl])))
        elif r < 0.70:
            name = var_names[lvl] if rng.random() < 0.8 else var_names[rng.integers(0, 4)]
            pr = rng.random()
            pred = None
            if pr < 0.15:
                pred = rf.In(*[int(x) for x in rng.integers(0, hi + 2, rng.integers(1, 12))])
            elif pr < 0.30:
                a, b = sorted(int(x) for x in rng.integers(0, hi + 2, 2))
                pred = rf.Between(a, b)
            instr.append(rf.scan(name, pred))
        elif r < 0.80:
            instr.append(rf.traverse())
        elif r < 0.88:
            instr.append(rf.traverse(rf.In(*[int(x) for x in rng.integers(0, hi + 2, rng.integers(2, 20))])))
        elif r < 0.97:
            a, b = sorted(int(x) for x in rng.integers(0, hi + 2, 2))
            instr.append(rf.traverse(rf.Between(a, b)))
        else:
            instr.append(rf.traverse(rf.FALSE))
    return instr
