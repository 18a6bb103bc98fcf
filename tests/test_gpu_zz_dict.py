"""GPU: object id -> plain term decode through dict.cu (rfgpu_dict_create / rfgpu_dict_decode / rfgpu_plain_terms_export,
SURVEY §8f-2) against the oracle restatement of MemObjectIdMapping::decode_array, pinned in test_dict_decode_cpu.py on the
reference's own roundtrip vector.

Green on a B200 since the round-1 driver run (GPUTEST_r01.json)."""
import numpy as np
import pytest

import rdf_fusion_b200 as rf
from test_dict_decode_cpu import REF_TERMS, random_terms

pytestmark = [pytest.mark.gpu]


def test_reference_roundtrip_vector(ctx, oracle):
    d = rf.Dictionary(ctx, REF_TERMS)
    assert len(d) == 4
    ids = np.array([1, 2, 3, 4, 0], np.uint32)
    got = d.decode(ctx.import_table({"x": ids}, {"x": True}), "x")
    exp = oracle.decode_plain_terms(REF_TERMS, ids)
    got.validate(full=True)
    assert got.type == exp.type and got.to_pylist() == exp.to_pylist()


@pytest.mark.parametrize("n_terms,n_rows", [(1, 0), (1, 1), (50, 31), (50, 32), (50, 33), (400, 5000), (20000, 300000)])
def test_decode_matches_oracle(ctx, oracle, n_terms, n_rows):
    """empty batch, partial and whole validity words, empty and multi-KB values, multi-byte UTF-8, every term kind, NULLs"""
    rng = np.random.default_rng(n_terms + n_rows)
    terms = random_terms(rng, n_terms)
    ids = rng.integers(0, n_terms + 1, n_rows).astype(np.uint32)
    d = rf.Dictionary(ctx, terms)
    T = ctx.import_table({"s": ids, "o": ids[::-1].copy()}, {"s": True, "o": True})
    for col, want in (("s", ids), ("o", ids[::-1])):
        got, exp = d.decode(T, col), oracle.decode_plain_terms(terms, want)
        got.validate(full=True)
        assert got.null_count == exp.null_count
        assert got.equals(exp)
    if n_rows > 100:  # a row range, as a batch_size-limited stream would ask for
        got = d.decode(T, "s", row_off=37, n_rows=64)
        assert got.equals(oracle.decode_plain_terms(terms, ids[37:101]))


def test_unknown_id_is_an_error(ctx):
    d = rf.Dictionary(ctx, REF_TERMS)
    T = ctx.import_table({"x": np.array([1, 9, 2], np.uint32)})
    with pytest.raises(rf.RfgpuError) as e:
        d.decode(T, "x")
    assert "9" in str(e.value)


def test_decode_after_a_join(ctx, oracle):
    """the ProjectionExec ENC_PT(col) that ends every query plan: decode the columns of a join result"""
    b = rf.generate_bsbm_like(200, seed=3)
    dense, inv = np.unique(b.quads, return_inverse=True)  # dense ids keep the dictionary small; 0 (default graph) stays 0
    assert dense[0] == 0
    quads = inv.reshape(b.quads.shape).astype(np.uint32)
    v = {k: int(np.searchsorted(dense, x)) for k, x in b.vocab.items() if k in ("producer", "label")}
    max_id = int(quads.max())
    terms = [("nn", f"http://example.com/r{i}") if i % 3 else ("lit", str(i), "http://www.w3.org/2001/XMLSchema#integer")
             for i in range(1, max_id + 1)]
    d = rf.Dictionary(ctx, terms)
    st = ctx.store().load_quads(quads)
    a = st.plan([rf.traverse(0), rf.scan("product"), rf.traverse(v["producer"]), rf.scan("producer")]).execute()
    c = st.plan([rf.traverse(0), rf.scan("product"), rf.traverse(v["label"]), rf.scan("label")]).execute()
    j = a.join(c)
    cols = j.to_numpy()
    for name, col in zip(j.names, cols):
        assert d.decode(j, name).equals(oracle.decode_plain_terms(terms, col))


def test_value_filter_through_a_rank_lookup(ctx):
    """FILTER(?p1 > c) / FILTER(?p3 < c) of BSBM Explore Q3 without decoding: object ids assigned in arbitrary order, the
    numeric rank of every id looked up on the device (rfgpu_table_lookup), the comparison an id-level BETWEEN on the rank
    column — against numpy on the values themselves; a non-numeric object fails the comparison (SPARQL type error)."""
    from rdf_fusion_b200 import _ffi as F
    rng = np.random.default_rng(17)
    X = "http://www.w3.org/2001/XMLSchema#"
    n_terms = 4000
    values = rng.integers(-500, 500, n_terms)
    kinds = rng.integers(0, 4, n_terms)  # integer, decimal with a fraction digit, double, IRI
    terms = []
    for v, k in zip(values, kinds):
        terms.append([("lit", str(v), X + "integer"), ("lit", f"{v}.0", X + "decimal"), ("lit", f"{v}e0", X + "double"),
                      ("nn", f"http://example.com/{v}")][k])
    rank = rf.numeric_rank_table(terms)
    n = 60000
    obj = rng.integers(0, n_terms + 1, n).astype(np.uint32)
    T = ctx.import_table({"product": np.arange(1, n + 1, dtype=np.uint32), "p1": obj}, {"p1": True})
    withrank = T.lookup("p1", rank, "p1_rank")
    assert withrank.names == ["product", "p1", "p1_rank"] and withrank.nullable[2]
    val = np.concatenate([[0], values]).astype(np.int64)
    numeric = np.concatenate([[False], kinds != 3])
    for c in (-600, -37, 0, 123, 499, 700):
        rank_c = int(np.unique(values[(kinds != 3) & (values <= c)]).size)  # rank of the largest value <= c (0: none)
        gt = withrank.filter([rf.FPred(F.F_BETWEEN, 2, value=rank_c + 1, hi=0xFFFFFFFF)])
        want = numeric[obj] & (val[obj] > c)
        assert gt.num_rows == int(want.sum()), c
        if gt.num_rows:
            assert np.array_equal(gt.to_numpy()[0], np.arange(1, n + 1, dtype=np.uint32)[want])
        if rank_c:
            le = withrank.filter([rf.FPred(F.F_BETWEEN, 2, value=1, hi=rank_c)])
            assert le.num_rows == int((numeric[obj] & (val[obj] <= c)).sum())
