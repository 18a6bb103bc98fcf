#!/usr/bin/env python
"""Generates tests/golden/minus_golden.json from the W3C SPARQL 1.1 negation tests the reference runs
(testsuite/tests/sparql.rs:36-43, manifest-sparql11-query.ttl -> sparql11/negation, sparql11/exists): the MINUS and FILTER [NOT] EXISTS cases that need no
value-level FILTER.  Same conventions as make_join_golden.py (whose Turtle / result parsers and plan DSL are reused): data dictionary-
encoded to object ids, the query HAND-LOWERED to the ID-level plan of the reference (MinusNode -> LeftAnti join,
lib/logical/src/minus/rewrite.rs:58-93; EXISTS -> correlated sub-query -> semi / anti join, expr_builder_context.rs:223-338), expected solutions encoded with the same dictionary (unbound = 0).

Run in the build container only (needs /root/reference):  python tests/golden/make_minus_golden.py
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_join_golden as G  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "minus_golden.json")
NEG = "../sparql11/negation/"
EXI = "../sparql11/exists/"
EXW = {"ex": "http://www.example.org/", "rdf": G.RDF}
EX = {"": "http://example/", "rdf": G.RDF}
EXN = {"ex": "http://www.w3.org/2009/sparql/docs/tests/data-sparql11/negation#", "rdf": G.RDF}


def minus(l, r): return {"op": "minus", "l": l, "r": r}
def exists(l, r, negated=False): return {"op": "exists", "l": l, "r": r, "negated": negated}


def _subtrahend(q):
    d = q.scan("?d", "a", ":Sub")
    return G.leftjoin(G.leftjoin(d, q.scan("?d", ":q1", "?b")), q.scan("?d", ":q2", "?c"))


def full_minuend(q):
    return minus(G.join(q.scan("?a", ":p1", "?b"), q.scan("?a", ":p2", "?c")), _subtrahend(q))


def part_minuend(q):
    a = G.leftjoin(G.leftjoin(q.scan("?a", "a", ":Min"), q.scan("?a", ":p1", "?b")), q.scan("?a", ":p2", "?c"))
    return minus(a, _subtrahend(q))


def subset_by_excl_02(q):
    ids = [q.term("ex:Reptile")["id"], q.term("ex:Insect")["id"]]
    rhs = G.filt(q.scan("?animal", "a", "?type"), [{"op": "in", "var": "type", "ids": ids}])
    return minus(q.scan("?animal", "a", "ex:Animal"), rhs)


def exists_01(q):
    return exists(q.scan("?set", "a", ":Set"), q.scan("?set", ":member", "9"))


def exists_02(q):  # no set has the member 7: the constant is not in the dictionary, the inner scan is empty
    return exists(q.scan("?set", "a", ":Set"), q.scan("?set", ":member", "7"))


def subset_by_excl_01(q):
    return exists(q.scan("?animal", "a", "ex:Animal"), q.scan("?animal", "a", "ex:Insect"), negated=True)


def rename(x, m): return {"op": "rename", "in": x, "map": m}


def _sets(q, first, second): return G.join(q.scan(first, "rdf:type", ":Set"), q.scan(second, "rdf:type", ":Set"))


def _not_subset(q, a, b):
    """pairs (a, b) where a has a member that b has not:  a :member ?x  FILTER NOT EXISTS { b :member ?x }"""
    return exists(G.join(_sets(q, "?s1", "?s2"), q.scan(a, ":member", "?x")), q.scan(b, ":member", "?x"), negated=True)


def subset_01(q):
    ne = [{"op": "ne_col", "var": "s1", "var2": "s2"}]
    lhs = G.filt(_sets(q, "?s2", "?s1"), ne)
    inner = G.filt(G.join(_sets(q, "?s1", "?s2"), q.scan("?s1", ":member", "?x")), ne)
    rhs = exists(inner, q.scan("?s2", ":member", "?x"), negated=True)
    return rename(minus(lhs, rhs), {"s1": "subset", "s2": "superset"})


def subset_03(q):
    equal_sets = minus(minus(_sets(q, "?s2", "?s1"), _not_subset(q, "?s1", "?s2")), _not_subset(q, "?s2", "?s1"))
    return rename(minus(minus(_sets(q, "?s2", "?s1"), _not_subset(q, "?s1", "?s2")), equal_sets),
                  {"s1": "subset", "s2": "superset"})


def exists01(q):
    return exists(q.scan("?s", "?p", "?o"), q.scan("?s", "?p", "ex:o"))


def exists03(q):  # both patterns inside GRAPH <exists02.ttl>
    g = "<exists02.ttl>"
    return exists(q.scan("?s", "?p", "ex:o1", graph=g), q.scan("?s", "?p", "ex:o2", graph=g))


def exists04(q):
    return exists(q.scan("?s", "?p", "ex:o"), exists(q.scan("?s", "?p", "ex:o1"), q.scan("?s", "?p", "ex:o2")))


def exists05(q):
    return exists(q.scan("?s", "?p", "ex:o"), exists(q.scan("?s", "?p", "ex:o1"), q.scan("?s", "?p", "ex:o2"), negated=True))


CASES = [
    ("negation/full-minuend", [NEG + "full-minuend.ttl"], EX, full_minuend, NEG + "full-minuend.srx", NEG + "full-minuend.rq"),
    ("negation/part-minuend", [NEG + "part-minuend.ttl"], EX, part_minuend, NEG + "part-minuend.srx", NEG + "part-minuend.rq"),
    ("negation/subsetByExcl02", [NEG + "subsetByExcl.ttl"], EXN, subset_by_excl_02, NEG + "subsetByExcl02.srx",
     NEG + "subsetByExcl02.rq"),
    ("negation/subsetByExcl01", [NEG + "subsetByExcl.ttl"], EXN, subset_by_excl_01, NEG + "subsetByExcl01.srx",
     NEG + "subsetByExcl01.rq"),
    ("negation/exists-01", [NEG + "set-data.ttl"], EX, exists_01, NEG + "exists-01.srx", NEG + "exists-01.rq"),
    ("negation/exists-02", [NEG + "set-data.ttl"], EX, exists_02, NEG + "exists-02.srx", NEG + "exists-02.rq"),
    ("negation/subset-01", [NEG + "set-data.ttl"], EX, subset_01, NEG + "subset-01.srx", NEG + "subset-01.rq"),
    ("negation/subset-03", [NEG + "set-data.ttl"], EX, subset_03, NEG + "subset-03.srx", NEG + "subset-03.rq"),
    # sparql11/exists (exists02 has a variable-free inner pattern: a zero-column table, not expressible here)
    ("exists/exists01", [EXI + "exists01.ttl"], EXW, exists01, EXI + "exists01.srx", EXI + "exists01.rq"),
    ("exists/exists03", [EXI + "exists01.ttl"], EXW, exists03, EXI + "exists03.srx", EXI + "exists03.rq", [EXI + "exists02.ttl"]),
    ("exists/exists04", [EXI + "exists01.ttl"], EXW, exists04, EXI + "exists04.srx", EXI + "exists04.rq"),
    ("exists/exists05", [EXI + "exists01.ttl"], EXW, exists05, EXI + "exists05.srx", EXI + "exists05.rq"),
]


def main():
    out = []
    for name, data, px, plan_fn, result, rq, *rest in CASES:
        named = rest[0] if rest else []
        d = G.Dict()
        quads = []
        for k, rel in enumerate(data):
            for s, p, o in G.Turtle(G.read(rel), "d%d_" % k).parse():
                quads.append([0, d.id(s), d.id(p), d.id(o)])
        for k, rel in enumerate(named):  # a named graph is called by the relative IRI the queries use
            gid = d.id("<%s>" % os.path.basename(rel))
            for s, p, o in G.Turtle(G.read(rel), "g%d_" % k).parse():
                quads.append([gid, d.id(s), d.id(p), d.id(o)])
        plan = plan_fn(G.Q(d, px))
        vars_, sols = G.srx_solutions(result)
        rows = []
        for sol in sols:
            row = []
            for v in vars_:
                t = sol.get(v)
                i = 0 if t is None else d.get(t)
                assert i is not None, (name, v, t)
                row.append(i)
            rows.append(row)
        out.append({"name": name, "reference_files": {"query": rq, "data": data, "graph_data": named, "result": result},
                    "n_terms": len(d.ids), "bnode_ids": [], "quads": quads, "plan": plan, "select": vars_, "expected": rows})
    doc = {"_comment": "Generated by tests/golden/make_minus_golden.py from the W3C SPARQL 1.1 negation tests under "
                       "testsuite/rdf-tests/sparql/sparql11/negation of the reference (run by testsuite/tests/sparql.rs:36-43). "
                       "Same layout as join_golden.json; plan op 'minus' = MinusNode, 'exists' = FILTER [NOT] EXISTS.",
           "cases": out}
    with open(OUT, "w") as f:
        json.dump(doc, f, indent=None, separators=(",", ":"))
        f.write("\n")
    print("wrote", OUT, len(out), "cases;", sum(len(c["expected"]) for c in out), "expected rows")


if __name__ == "__main__":
    main()
