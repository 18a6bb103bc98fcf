#!/usr/bin/env python
"""Generates tests/golden/agg_golden.json: the W3C evaluation tests that pin DISTINCT and GROUP BY / COUNT at the SPARQL
level — sparql10/distinct (run by testsuite/tests/sparql.rs:16-34) and sparql11/aggregates agg02 / agg05 (:36-43).
Same conventions as make_join_golden.py (whose parsers and plan DSL are reused).  DISTINCT is by TERM, not by value
("1" and "01" are different terms), which is exactly a DISTINCT on object ids (ReplaceDistinctWithAggregate over the id
columns, lib/logical/src/logical_plan_builder.rs:235-264); "abc" and "abc"^^xsd:string are ONE term (RDF 1.1), so both are
canonicalised to the plain form before the dictionary assigns ids.  COUNT columns hold numbers, not ids.

Run in the build container only (needs /root/reference):  python tests/golden/make_agg_golden.py
"""
import json
import os
import re
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_join_golden as G  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "agg_golden.json")
DIS = "distinct/"
AGG = "../sparql11/aggregates/"
EX = {"": "http://example/", "rdf": G.RDF}
EXO = {"": "http://www.example.org/", "rdf": G.RDF}
XSD_STRING = "^^<%sstring>" % G.XSD


def canon(t):
    return t[:-len(XSD_STRING)] if t.startswith('"') and t.endswith(XSD_STRING) else t


def distinct(x, vars_): return {"op": "distinct", "in": x, "vars": vars_}
def project(x, vars_): return {"op": "project", "in": x, "vars": vars_}
def group_count(x, keys, counted, name): return {"op": "group_count", "in": x, "keys": keys, "counted": counted, "name": name}


def all_v(q): return project(q.scan("?x", "?p", "?v"), ["v"])
def distinct_v(q): return distinct(q.scan("?x", "?p", "?v"), ["v"])
def _opt(q): return G.leftjoin(q.scan(":x1", "?p", "?o"), q.scan("?o", ":q", "?v"))
def opt_v(q): return project(_opt(q), ["v"])
def distinct_opt_v(q): return distinct(_opt(q), ["v"])
def distinct_star(q): return distinct(G.union(q.scan("?s", ":p", "?o"), q.scan("?s", ":q", "?o")), ["s", "o"])
def group01(q): return distinct(q.scan("?s", ":p", "?v"), ["s"])  # GROUP BY ?s without an aggregate = DISTINCT ?s
def agg02(q): return group_count(q.scan("?S", "?P", "?O"), ["P"], "O", "C")
def agg05(q): return group_count(q.scan("?S", "?P", "?O"), ["P"], None, "C")


CASES = [
    # name, data, prefixes, plan, result, query, numeric (COUNT) columns
    ("distinct/no-distinct-num", [DIS + "data-num.ttl"], EX, all_v, DIS + "no-distinct-num.srx", DIS + "no-distinct-1.rq", []),
    ("distinct/distinct-num", [DIS + "data-num.ttl"], EX, distinct_v, DIS + "distinct-num.srx", DIS + "distinct-1.rq", []),
    ("distinct/no-distinct-str", [DIS + "data-str.ttl"], EX, all_v, DIS + "no-distinct-str.srx", DIS + "no-distinct-1.rq", []),
    ("distinct/distinct-str", [DIS + "data-str.ttl"], EX, distinct_v, DIS + "distinct-str.srx", DIS + "distinct-1.rq", []),
    ("distinct/no-distinct-node", [DIS + "data-node.ttl"], EX, all_v, DIS + "no-distinct-node.srx", DIS + "no-distinct-1.rq", []),
    ("distinct/distinct-node", [DIS + "data-node.ttl"], EX, distinct_v, DIS + "distinct-node.srx", DIS + "distinct-1.rq", []),
    ("distinct/no-distinct-opt", [DIS + "data-opt.ttl"], EX, opt_v, DIS + "no-distinct-opt.srx", DIS + "no-distinct-2.rq", []),
    ("distinct/distinct-opt", [DIS + "data-opt.ttl"], EX, distinct_opt_v, DIS + "distinct-opt.srx", DIS + "distinct-2.rq", []),
    ("distinct/no-distinct-all", [DIS + "data-all.ttl"], EX, all_v, DIS + "no-distinct-all.srx", DIS + "no-distinct-1.rq", []),
    ("distinct/distinct-all", [DIS + "data-all.ttl"], EX, distinct_v, DIS + "distinct-all.srx", DIS + "distinct-1.rq", []),
    ("distinct/distinct-star-1", [DIS + "data-star.ttl"], EX, distinct_star, DIS + "distinct-star-1.srx", DIS + "distinct-star-1.rq", []),
    ("grouping/group01", ["../sparql11/grouping/group-data-1.ttl"], EX, group01, "../sparql11/grouping/group01.srx",
     "../sparql11/grouping/group01.rq", []),
    ("aggregates/agg02", [AGG + "agg01.ttl"], EXO, agg02, AGG + "agg02.srx", AGG + "agg02.rq", ["C"]),
    ("aggregates/agg05", [AGG + "agg01.ttl"], EXO, agg05, AGG + "agg05.srx", AGG + "agg05.rq", ["C"]),
]


def main():
    out = []
    for name, data, px, plan_fn, result, rq, numeric in CASES:
        d = G.Dict()
        quads = []
        for k, rel in enumerate(data):
            for s, p, o in G.Turtle(G.read(rel), "d%d_" % k).parse():
                quads.append([0, d.id(canon(s)), d.id(canon(p)), d.id(canon(o))])
        plan = plan_fn(G.Q(d, px))
        vars_, sols = G.srx_solutions(result)
        bnodes = sorted(i for t, i in d.ids.items() if t.startswith("_:"))
        rows = []
        for sol in sols:
            row = []
            for v in vars_:
                t = sol.get(v)
                if t is None:
                    row.append(0)
                elif v in numeric:
                    m = re.fullmatch(r'"(\d+)"\^\^<%sinteger>' % re.escape(G.XSD), t)
                    assert m, (name, v, t)
                    row.append(int(m.group(1)))
                elif t.startswith("_:"):
                    row.append(-1)
                else:
                    i = d.get(canon(t))
                    assert i is not None, (name, v, t)
                    row.append(i)
            rows.append(row)
        out.append({"name": name, "reference_files": {"query": rq, "data": data, "graph_data": [], "result": result},
                    "n_terms": len(d.ids), "bnode_ids": bnodes, "quads": quads, "plan": plan, "select": vars_,
                    "numeric": numeric, "expected": rows})
    doc = {"_comment": "Generated by tests/golden/make_agg_golden.py from the W3C tests sparql10/distinct and "
                       "sparql11/aggregates (agg02, agg05) of the reference's testsuite.  Same layout as join_golden.json; "
                       "plan ops 'project', 'distinct', 'group_count'; columns listed in 'numeric' hold counts, not ids.",
           "cases": out}
    with open(OUT, "w") as f:
        json.dump(doc, f, indent=None, separators=(",", ":"))
        f.write("\n")
    print("wrote", OUT, len(out), "cases;", sum(len(c["expected"]) for c in out), "expected rows")


if __name__ == "__main__":
    main()
