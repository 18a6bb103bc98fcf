"""CPU: object id -> plain term decode.  The oracle (oracle_py.decode_plain_terms) against the vector of the
reference's own test (lib/storage/src/memory/object_id_mapping.rs:411-441: named node, blank node, typed literal,
language-tagged literal, NULL) and the PlainTermEncoding layout rules; and the host half of the product path — pack_terms
(the dictionary arrays rfgpu_dict_create takes) and plain_terms_to_arrow (the buffers rfgpu_plain_terms_export fills) —
against the oracle, with a numpy model of the three device steps (lengths + validity words, exclusive scan, byte copy)
standing in for the kernels.  The kernels themselves are compared with the oracle in test_gpu_zz_dict.py."""
import numpy as np
import pyarrow as pa
import pytest

from rdf_fusion_b200.api import pack_terms, plain_terms_to_arrow

XSD_STRING = "http://www.w3.org/2001/XMLSchema#string"
XSD_INT = "http://www.w3.org/2001/XMLSchema#integer"
REF_TERMS = [("nn", "http://example.com/a"), ("bn", "b1"), ("lit", "hello", XSD_STRING), ("lang", "world", "en")]


def model_decode(terms, ids):
    """numpy model of dict.cu: what rfgpu_dict_decode + rfgpu_plain_terms_export hand to plain_terms_to_arrow"""
    tt, vo, vb, dt, lg, ao, ab = pack_terms(terms)
    ids = np.asarray(ids, np.uint32)
    n = len(ids)
    ok = ids != 0
    assert int(ids.max(initial=0)) <= len(terms)
    lens = [np.where(ok, (vo[ids + 1] - vo[ids]).astype(np.int64), 0),
            np.where(ok & (dt[ids] != 0), ao[dt[ids] + 1].astype(np.int64) - ao[dt[ids]], 0),
            np.where(ok & (lg[ids] != 0), ao[lg[ids] + 1].astype(np.int64) - ao[lg[ids]], 0)]
    flags = [ok, ok & (dt[ids] != 0), ok & (lg[ids] != 0)]
    valid = [np.packbits(f, bitorder="little") for f in flags]
    valid = [np.concatenate([v, np.zeros(1, np.uint8)]) for v in valid]
    offs = [np.concatenate([[0], np.cumsum(l)]).astype(np.int32) for l in lens]
    data = []
    for k, (src, so, idx) in enumerate([(vb, vo, ids), (ab, ao, dt[ids]), (ab, ao, lg[ids])]):
        out = np.zeros(max(int(offs[k][-1]), 1), np.uint8)
        for r in range(n):
            if lens[k][r]:
                a = int(so[idx[r]])
                out[offs[k][r]:offs[k][r + 1]] = src[a:a + int(lens[k][r])]
        data.append(out)
    term_type = np.where(ok, tt[ids], 0).astype(np.uint8)
    return plain_terms_to_arrow(n, np.concatenate([term_type, np.zeros(1, np.uint8)]), valid, offs, data,
                                [int(o[-1]) for o in offs], [int(n - f.sum()) for f in flags])


def test_oracle_on_the_reference_roundtrip_vector(oracle):
    """object_id_mapping.rs:411-441: decode(encode(x)) == x for [named node, blank node, "hello"^^xsd:string, "world"@en,
    NULL]; the expected struct is what PlainTermArrayElementBuilder builds for x (element_builder.rs:29-134)"""
    got = oracle.decode_plain_terms(REF_TERMS, [1, 2, 3, 4, 0])
    assert got.type == pa.struct([pa.field("term_type", pa.uint8(), False), pa.field("value", pa.utf8(), False),
                                  pa.field("data_type", pa.utf8(), True), pa.field("language_tag", pa.utf8(), True)])
    assert got.to_pylist() == [
        {"term_type": 0, "value": "http://example.com/a", "data_type": None, "language_tag": None},
        {"term_type": 1, "value": "b1", "data_type": None, "language_tag": None},
        {"term_type": 2, "value": "hello", "data_type": XSD_STRING, "language_tag": None},
        {"term_type": 2, "value": "world", "data_type": oracle.RDF_LANG_STRING, "language_tag": "en"},
        None]
    assert got.null_count == 1
    # same ids in another order and repeated: the mapping is a function of the id (test_id_uniqueness_and_consistency :443-481)
    again = oracle.decode_plain_terms(REF_TERMS, [4, 1, 1, 0, 3])
    assert again.to_pylist()[1] == again.to_pylist()[2] == got.to_pylist()[0]
    with pytest.raises(oracle.OracleError):
        oracle.decode_plain_terms(REF_TERMS, [5])


def random_terms(rng, n):
    alphabet = ["a", "b", "é", "ß", "日", "本", "🙂", " ", "/", "#"]
    terms = []
    for i in range(n):
        ln = int(rng.integers(0, 90)) if rng.random() < 0.9 else int(rng.integers(200, 3000))
        v = "".join(rng.choice(alphabet, ln))
        k = rng.integers(0, 4)
        terms.append([("nn", "http://x/" + v), ("bn", f"b{i}"), ("lit", v, [XSD_STRING, XSD_INT][i % 2]),
                      ("lang", v, ["en", "de-AT", "ja"][i % 3])][k])
    return terms


@pytest.mark.parametrize("n_terms,n_rows", [(4, 5), (1, 0), (50, 31), (50, 32), (50, 33), (400, 5000)])
def test_host_half_of_the_product_path_matches_the_oracle(oracle, n_terms, n_rows):
    rng = np.random.default_rng(n_terms + n_rows)
    terms = REF_TERMS if n_terms == 4 else random_terms(rng, n_terms)
    ids = rng.integers(0, len(terms) + 1, n_rows).astype(np.uint32) if n_rows != 5 else np.array([1, 2, 3, 4, 0], np.uint32)
    got, exp = model_decode(terms, ids), oracle.decode_plain_terms(terms, ids)
    got.validate(full=True)
    assert got.type == exp.type and got.null_count == exp.null_count
    assert got.to_pylist() == exp.to_pylist() and got.equals(exp)
    for k in range(4):
        assert got.field(k).null_count == exp.field(k).null_count


def test_numeric_rank_table():
    """value order across datatypes and lexical forms; non-numeric and ill-formed terms have no rank"""
    from rdf_fusion_b200 import numeric_rank_table
    X = "http://www.w3.org/2001/XMLSchema#"
    terms = [("lit", "2", X + "integer"), ("nn", "http://e/a"), ("lit", "2.0", X + "decimal"), ("lit", "-1", X + "integer"),
             ("lit", "1e1", X + "double"), ("lit", "x", X + "integer"), ("lit", "3", X + "string"), ("lang", "4", "en"),
             ("lit", "0.1", X + "decimal"), ("lit", "NaN", X + "double"), ("lit", "02", X + "int")]
    t = numeric_rank_table(terms)
    assert t.tolist() == [0, 3, 0, 3, 1, 4, 0, 0, 0, 2, 0, 3]
