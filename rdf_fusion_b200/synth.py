"""Deterministic synthetic graphs in object-id space (G(N, P, s, seed)).

Counter-based RNG: every draw is splitmix64(seed, stream, row) so the graph does not depend on chunking or thread
count.  Layout of the id space (id 0 = default graph = NULL, lib/storage/src/memory/object_id.rs:21-27):
  graphs      1..16          (only used when named_graph_frac > 0)
  predicates  17..16+P       Zipf(s) by rank (rank 1 = id 17)
  subjects    17+P..16+P+S   S = max(N // 10, 1), uniform
  literals    17+P+S..       L = max(N // 4, 1), uniform
Objects: 50 % entity ids from the subject range (Zipf(s_o); s_o = 0 → uniform), 50 % literals.
Duplicates are NOT removed here (the store is a set: the loader de-duplicates).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x: np.ndarray) -> np.ndarray:
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M64
    z = x
    z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
    z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
    return z ^ (z >> np.uint64(31))


def _draw(seed: int, stream: int, rows: np.ndarray) -> np.ndarray:
    base = splitmix64(np.array([seed * 0x100 + stream], dtype=np.uint64))[0]
    with np.errstate(over="ignore"):
        return splitmix64(rows ^ base)


def _zipf_cdf(n: int, s: float) -> np.ndarray:
    w = 1.0 / np.power(np.arange(1, n + 1, dtype=np.float64), s)
    c = np.cumsum(w)
    return c / c[-1]


def _unit(u: np.ndarray) -> np.ndarray:
    return (u >> np.uint64(11)).astype(np.float64) * (1.0 / (1 << 53))


@dataclass
class SynthGraph:
    n: int
    n_predicates: int
    n_subjects: int
    n_literals: int
    g: np.ndarray
    s: np.ndarray
    p: np.ndarray
    o: np.ndarray

    def pred_id(self, rank: int) -> int:
        return 16 + rank

    @property
    def subject_base(self) -> int:
        return 17 + self.n_predicates

    @property
    def literal_base(self) -> int:
        return 17 + self.n_predicates + self.n_subjects

    def quads(self) -> np.ndarray:
        return np.stack([self.g, self.s, self.p, self.o], axis=1)


def generate_rows(n: int, rows: np.ndarray, n_predicates: int = 1024, s: float = 1.0, seed: int = 2, s_o: float = 0.0,
                  named_graph_frac: float = 0.0, _cdfs=None):
    """(g, s, p, o) of the rows `rows` (uint64 row numbers < n) of G(n, P, s, seed): any subset, in any order — what makes
    the graph independent of chunking, and what lets a rank (or a sampler) draw only the rows it needs."""
    S = max(n // 10, 1)
    L = max(n // 4, 1)
    P = n_predicates
    pcdf, ocdf = _cdfs if _cdfs is not None else (_zipf_cdf(P, s), _zipf_cdf(S, s_o) if s_o > 0 else None)
    rows = np.asarray(rows, dtype=np.uint64)
    with np.errstate(over="ignore"):
        sub = (17 + P + (_draw(seed, 1, rows) % np.uint64(S))).astype(np.uint32)
        pr = np.searchsorted(pcdf, _unit(_draw(seed, 2, rows)), side="right")
        pred = (17 + np.minimum(pr, P - 1)).astype(np.uint32)
        kind = _draw(seed, 3, rows) & np.uint64(1)
        u = _draw(seed, 4, rows)
        if ocdf is None:
            ent = (u % np.uint64(S)).astype(np.int64)
        else:
            ent = np.minimum(np.searchsorted(ocdf, _unit(u), side="right"), S - 1)
        lit = (u % np.uint64(L)).astype(np.int64)
        obj = np.where(kind == 0, 17 + P + ent, 17 + P + S + lit).astype(np.uint32)
        g = np.zeros(len(rows), dtype=np.uint32)
        if named_graph_frac > 0:
            gu = _draw(seed, 5, rows)
            named = _unit(gu) < named_graph_frac
            g = np.where(named, 1 + (gu % np.uint64(16)).astype(np.int64), 0).astype(np.uint32)
    return g, sub, pred, obj


def generate_graph(n: int, n_predicates: int = 1024, s: float = 1.0, seed: int = 2, s_o: float = 0.0,
                   named_graph_frac: float = 0.0, chunk: int = 1 << 24) -> SynthGraph:
    S = max(n // 10, 1)
    L = max(n // 4, 1)
    P = n_predicates
    cdfs = (_zipf_cdf(P, s), _zipf_cdf(S, s_o) if s_o > 0 else None)
    g = np.zeros(n, dtype=np.uint32)
    sub = np.empty(n, dtype=np.uint32)
    pred = np.empty(n, dtype=np.uint32)
    obj = np.empty(n, dtype=np.uint32)
    for lo in range(0, n, chunk):
        hi = min(n, lo + chunk)
        g[lo:hi], sub[lo:hi], pred[lo:hi], obj[lo:hi] = generate_rows(
            n, np.arange(lo, hi, dtype=np.uint64), P, s, seed, s_o, named_graph_frac, cdfs)
    return SynthGraph(n, P, S, L, g, sub, pred, obj)


def graph_shape(n: int, n_predicates: int = 1024) -> SynthGraph:
    """The id-space layout of G(n, P, ..) without any rows (pred_id, subject_base, literal_base)."""
    z = np.zeros(0, np.uint32)
    return SynthGraph(n, n_predicates, max(n // 10, 1), max(n // 4, 1), z, z, z, z)


# ---------------------------------------------------------------- BSBM-shaped graph in id space
@dataclass
class BsbmLikeGraph:
    """Quads with the vocabulary and fan-outs of the BSBM e-commerce graph, generated directly as object ids (BASELINE
    configs[0] / [2] need the Java bsbm-tools generator, which this box does not have; SURVEY §8d substitutes an id-space
    graph that reproduces the join trees of the reference's plan snapshots, bench/tests/plans/snapshots/)."""
    quads: np.ndarray
    vocab: dict

    def __getattr__(self, k):
        try:
            return self.vocab[k]
        except KeyError:
            raise AttributeError(k)


def generate_bsbm_like(n_products: int = 300, seed: int = 7) -> BsbmLikeGraph:
    """Per product: 1–2 types, a producer, a label, 3–6 features, two numeric properties, ~10 reviews; producers and
    reviewers live in one of 10 countries.  Numeric literal ids ascend with their value, so `?p1 > c` on ENC_TV is an id
    range here (the reference decodes; §8f-2)."""
    v = dict(rdf_type=1, producer=2, country=3, review_for=4, reviewer=5, label=6, feature=7, num1=8, num3=9,
             product_type_class=20, country_base=30, n_countries=10, type_base=100, feature_base=300,
             producer_base=1000, product_base=10000, reviewer_base=100000, review_base=200000,
             label_base=2000000, numeric_base=3000000)
    n_types = max(4, n_products // 75)
    n_features = max(10, n_products // 30)
    n_producers = max(3, n_products // 20)
    n_reviews = 10 * n_products
    n_reviewers = max(5, n_reviews // 8)
    v.update(n_types=n_types, n_features=n_features, n_producers=n_producers, n_reviews=n_reviews,
             n_reviewers=n_reviewers, n_products=n_products)

    def u(stream, rows, mod):
        return (_draw(seed, stream, np.arange(rows, dtype=np.uint64)) % np.uint64(mod)).astype(np.int64)
    q = []

    def add(s, p, o):
        s, o = np.asarray(s, np.int64), np.asarray(o, np.int64)
        s, o = np.broadcast_arrays(s, o)
        q.append(np.stack([np.zeros(len(s), np.int64), s, np.full(len(s), p, np.int64), o], axis=1))
    products = v["product_base"] + np.arange(n_products)
    add(v["type_base"] + np.arange(n_types), v["rdf_type"], v["product_type_class"])
    add(products, v["rdf_type"], v["type_base"] + u(1, n_products, n_types))
    second = u(2, n_products, 2) == 0
    add(products[second], v["rdf_type"], v["type_base"] + u(3, n_products, n_types)[second])
    add(products, v["producer"], v["producer_base"] + u(4, n_products, n_producers))
    add(v["producer_base"] + np.arange(n_producers), v["country"], v["country_base"] + u(5, n_producers, v["n_countries"]))
    add(products, v["label"], v["label_base"] + np.arange(n_products))
    for k in range(6):
        keep = (u(10 + k, n_products, 6) >= k - 2)  # 3 always, then thinning
        add(products[keep], v["feature"], v["feature_base"] + u(20 + k, n_products, n_features)[keep])
    add(products, v["num1"], v["numeric_base"] + u(30, n_products, 2000))
    add(products, v["num3"], v["numeric_base"] + u(31, n_products, 2000))
    reviews = v["review_base"] + np.arange(n_reviews)
    add(reviews, v["review_for"], v["product_base"] + u(40, n_reviews, n_products))
    add(reviews, v["reviewer"], v["reviewer_base"] + u(41, n_reviews, n_reviewers))
    add(v["reviewer_base"] + np.arange(n_reviewers), v["country"], v["country_base"] + u(42, n_reviewers, v["n_countries"]))
    return BsbmLikeGraph(np.concatenate(q).astype(np.uint32), v)
