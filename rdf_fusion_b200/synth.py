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


def generate_graph(n: int, n_predicates: int = 1024, s: float = 1.0, seed: int = 2, s_o: float = 0.0,
                   named_graph_frac: float = 0.0, chunk: int = 1 << 24) -> SynthGraph:
    S = max(n // 10, 1)
    L = max(n // 4, 1)
    P = n_predicates
    pcdf = _zipf_cdf(P, s)
    ocdf = _zipf_cdf(S, s_o) if s_o > 0 else None
    g = np.zeros(n, dtype=np.uint32)
    sub = np.empty(n, dtype=np.uint32)
    pred = np.empty(n, dtype=np.uint32)
    obj = np.empty(n, dtype=np.uint32)
    with np.errstate(over="ignore"):
        for lo in range(0, n, chunk):
            hi = min(n, lo + chunk)
            rows = np.arange(lo, hi, dtype=np.uint64)
            sub[lo:hi] = (17 + P + (_draw(seed, 1, rows) % np.uint64(S))).astype(np.uint32)
            pr = np.searchsorted(pcdf, _unit(_draw(seed, 2, rows)), side="right")
            pred[lo:hi] = (17 + np.minimum(pr, P - 1)).astype(np.uint32)
            kind = _draw(seed, 3, rows) & np.uint64(1)
            u = _draw(seed, 4, rows)
            if ocdf is None:
                ent = (u % np.uint64(S)).astype(np.int64)
            else:
                ent = np.minimum(np.searchsorted(ocdf, _unit(u), side="right"), S - 1)
            lit = (u % np.uint64(L)).astype(np.int64)
            obj[lo:hi] = np.where(kind == 0, 17 + P + ent, 17 + P + S + lit).astype(np.uint32)
            if named_graph_frac > 0:
                gu = _draw(seed, 5, rows)
                named = _unit(gu) < named_graph_frac
                g[lo:hi] = np.where(named, 1 + (gu % np.uint64(16)).astype(np.int64), 0).astype(np.uint32)
    return SynthGraph(n, P, S, L, g, sub, pred, obj)
