#!/usr/bin/env python
"""Generates tests/golden/tv_golden.json: the W3C SPARQL 1.0 evaluation tests on VALUE comparisons that the reference runs
and passes (testsuite/tests/sparql.rs:16-34; the ignored ones — open-eq-07/08/10/11/12, date-2 — are left out for the reasons
given there): sparql10/expr-equals (eq-1..5, eq-bool, eq-float, eq-2-1 = eq-2-2) and sparql10/open-world (open-eq-03..06, 09,
open-cmp-01/02, date-1, date-3).

Every case is `BGP + FILTER(e)` with e a comparison of typed values: ENC_TV(?v) = 1, ?v1 = ?v2, ?v != "a"^^t:type1,
?v1 < ?v2 || ?v1 > ?v2, ?v > "2006-08-22"^^xsd:date.  The plan is the BGP's joins (make_join_golden's DSL) under a `tv_filter`
node: a postfix program of rfgpu_tv_op instructions over variables and constant terms, kept when its effective boolean value
is true.  The case also carries the dictionary (`terms`, the term of every object id) since the filter needs the values.

Run in the build container only (needs /root/reference):  python tests/golden/make_tv_golden.py
"""
import json
import os
import re
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_join_golden as G  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "tv_golden.json")
XSD = G.XSD
EQ, LT, LE, GT, GE, AND, OR, NOT, BOUND = range(9)  # rfgpu_tv_op


def term_tuple(t):
    """the N-Triples-like strings of make_join_golden -> the term tuples of rdf_fusion_b200.api.pack_terms"""
    if t.startswith("<"):
        return ["nn", t[1:-1]]
    if t.startswith("_:"):
        return ["bn", t[2:]]
    m = re.fullmatch(r'"(.*)"(?:\^\^<([^>]*)>|@([A-Za-z0-9-]+))?', t, flags=re.S)
    v, dt, lang = m.groups()
    if lang:
        return ["lang", v, lang]
    return ["lit", v, dt or XSD + "string"]  # RDF 1.1: a simple literal IS an xsd:string


def cmp(op, lhs, rhs):
    def operand(x):
        return {"var": x[1:]} if isinstance(x, str) and x.startswith("?") else {"const": x}
    return [[op, operand(lhs), operand(rhs)]]


def tv_filter(x, prog, consts=()):
    return {"op": "tv_filter", "in": x, "prog": prog, "consts": [term_tuple(c) for c in consts]}


NOT_I, OR_I = [[NOT, None, None]], [[OR, None, None]]
THINGS = {"": "http://example.org/things#", "xsd": XSD}
OW1 = {"": "http://example/ns#", "t": "http://example/t#", "xsd": XSD}
OW = {"": "http://example/", "xsd": XSD}


def one_var(const, negate=False):
    def plan(q):
        return tv_filter(q.scan("?x", ":p", "?v"), cmp(EQ, "?v", 0) + (NOT_I if negate else []), [const])
    return plan


def eq_pl_pr(q):
    return tv_filter(G.join(q.scan("?x", ":pr", "?vr"), q.scan("?x", ":pl", "?vl")), cmp(EQ, "?vr", "?vl"))


def eq2(negate):
    def plan(q):
        return tv_filter(G.join(q.scan("?x1", ":p", "?v1"), q.scan("?x2", ":p", "?v2")), cmp(EQ, "?v1", "?v2") + (NOT_I if negate else []))
    return plan


def open_eq_09(q):
    return tv_filter(G.join(q.scan("?x", ":p", "?v1"), q.scan("?y", ":q", "?v2")), cmp(EQ, "?v1", "?v2"))


def _cmp_bgp(q):
    return G.join(G.join(q.scan("?x", ":p", "?_b"), q.scan("?_b", ":v1", "?v1")), q.scan("?_b", ":v2", "?v2"))


def open_cmp_01(q):
    return tv_filter(_cmp_bgp(q), cmp(LT, "?v1", "?v2") + cmp(GT, "?v1", "?v2") + OR_I)


def open_cmp_02(q):
    return tv_filter(_cmp_bgp(q), cmp(LT, "?v1", "?v2") + cmp(EQ, "?v1", "?v2") + OR_I + cmp(GT, "?v1", "?v2") + OR_I)


def date(op, const):
    def plan(q):
        return tv_filter(q.scan("?x", ":r", "?v"), cmp(op, "?v", 0), [const])
    return plan


def lit(v, t): return '"%s"^^<%s%s>' % (v, XSD, t)


EE, OWD = "expr-equals/", "open-world/"
CASES = [
    # name, data, prefixes, plan, result, query
    ("expr-equals/eq-1", EE + "data-eq.ttl", THINGS, one_var(lit("1", "integer")), EE + "result-eq-1.ttl", EE + "query-eq-1.rq"),
    ("expr-equals/eq-2", EE + "data-eq.ttl", THINGS, one_var(lit("1.0e0", "double")), EE + "result-eq-2.ttl", EE + "query-eq-2.rq"),
    ("expr-equals/eq-3", EE + "data-eq.ttl", THINGS, one_var('"1"'), EE + "result-eq-3.ttl", EE + "query-eq-3.rq"),
    ("expr-equals/eq-4", EE + "data-eq.ttl", THINGS, one_var('"zzz"'), EE + "result-eq-4.ttl", EE + "query-eq-4.rq"),
    ("expr-equals/eq-5", EE + "data-eq.ttl", THINGS, one_var("<http://example.org/things#z>"), EE + "result-eq-5.ttl", EE + "query-eq-5.rq"),
    ("expr-equals/eq-bool", EE + "data-eq-bool.ttl", THINGS, eq_pl_pr, EE + "result-eq-bool.ttl", EE + "query-eq-bool.rq"),
    ("expr-equals/eq-float", EE + "data-eq-float.ttl", THINGS, eq_pl_pr, EE + "result-eq-float.ttl", EE + "query-eq-float.rq"),
    ("expr-equals/eq2-1", EE + "data-eq.ttl", THINGS, eq2(False), EE + "result-eq2-1.ttl", EE + "query-eq2-1.rq"),
    # query-eq2-2.rq (`!=`) is in the directory but in no manifest: :eq-2-1 and :eq-2-2 both run query-eq2-1.rq (manifest.ttl:70-88)
    ("open-world/open-eq-03", OWD + "data-1.ttl", OW1, one_var(lit("1", "integer")), OWD + "open-eq-03-result.srx", OWD + "open-eq-03.rq"),
    ("open-world/open-eq-04", OWD + "data-1.ttl", OW1, one_var(lit("1", "integer"), True), OWD + "open-eq-04-result.srx", OWD + "open-eq-04.rq"),
    ("open-world/open-eq-05", OWD + "data-1.ttl", OW1, one_var('"a"^^<http://example/t#type1>'), OWD + "open-eq-05-result.srx", OWD + "open-eq-05.rq"),
    ("open-world/open-eq-06", OWD + "data-1.ttl", OW1, one_var('"a"^^<http://example/t#type1>', True), OWD + "open-eq-06-result.srx", OWD + "open-eq-06.rq"),
    ("open-world/open-eq-09", OWD + "data-2.ttl", OW, open_eq_09, OWD + "open-eq-09-result.srx", OWD + "open-eq-09.rq"),
    ("open-world/open-cmp-01", OWD + "data-4.ttl", OW, open_cmp_01, OWD + "open-cmp-01-result.srx", OWD + "open-cmp-01.rq"),
    ("open-world/open-cmp-02", OWD + "data-4.ttl", OW, open_cmp_02, OWD + "open-cmp-02-result.srx", OWD + "open-cmp-02.rq"),
    ("open-world/date-1", OWD + "data-3.ttl", OW, date(EQ, lit("2006-08-23", "date")), OWD + "date-1-result.srx", OWD + "date-1.rq"),
    ("open-world/date-3", OWD + "data-3.ttl", OW, date(GT, lit("2006-08-22", "date")), OWD + "date-3-result.srx", OWD + "date-3.rq"),
]


def main():
    out = []
    for name, data, px, plan_fn, result, rq in CASES:
        d = G.Dict()
        quads = [[0, d.id(s), d.id(p), d.id(o)] for s, p, o in G.Turtle(G.read(data), "d0_").parse()]
        plan = plan_fn(G.Q(d, px))
        vars_, sols = (G.srx_solutions if result.endswith(".srx") else G.ttl_solutions)(result)
        bnodes = sorted(i for t, i in d.ids.items() if t.startswith("_:"))
        rows = []
        for sol in sols:
            row = []
            for v in vars_:
                t = sol.get(v)
                if t is None:
                    row.append(0)
                elif t.startswith("_:"):
                    row.append(-1)
                else:
                    i = d.get(t)
                    assert i is not None, (name, v, t)
                    row.append(i)
            rows.append(row)
        terms = [term_tuple(t) for t, _ in sorted(d.ids.items(), key=lambda kv: kv[1])]
        out.append({"name": name, "reference_files": {"query": rq, "data": [data], "graph_data": [], "result": result},
                    "n_terms": len(d.ids), "terms": terms, "bnode_ids": bnodes, "quads": quads, "plan": plan, "select": vars_,
                    "expected": rows})
    doc = {"_comment": "Generated by tests/golden/make_tv_golden.py from the W3C SPARQL 1.0 evaluation tests sparql10/expr-equals "
                       "and sparql10/open-world of the reference (run by testsuite/tests/sparql.rs:16-34).  Layout of "
                       "join_golden.json plus `terms` (the term of object id i + 1) and the plan node `tv_filter` "
                       "(postfix rfgpu_tv_op program; operands {var} / {const: index into consts}).",
           "cases": out}
    with open(OUT, "w") as f:
        json.dump(doc, f, indent=None, separators=(",", ":"))
        f.write("\n")
    print("wrote", OUT, len(out), "cases;", sum(len(c["expected"]) for c in out), "expected rows")


if __name__ == "__main__":
    main()
