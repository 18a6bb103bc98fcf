#!/usr/bin/env python
"""Generates tests/golden/join_golden.json from the W3C SPARQL 1.0 evaluation tests the reference itself runs
(testsuite/tests/sparql.rs:16-43 over testsuite/rdf-tests/sparql/sparql10/{algebra,optional,bound}).

For every case: the data files are parsed and dictionary-encoded to object ids (default graph = 0, ids from 1, exactly the
layout of lib/storage/src/memory/object_id.rs:21-27), the query is HAND-LOWERED below to the ID-level plan the reference
builds for it (GraphPatternRewriter: BGP = left-deep inner joins, OPTIONAL = SparqlJoinNode(Left), UNION, FILTER(bound)),
and the expected solutions (.srx / result-set .ttl) are encoded with the same dictionary (unbound = 0).  The tests then
run the plan through the oracle and through the CUDA path and compare multisets.

Run in the build container only (needs /root/reference):  python tests/golden/make_join_golden.py
"""
import json
import os
import re
import sys
import xml.etree.ElementTree as ET

REF = "/root/reference/testsuite/rdf-tests/sparql/sparql10"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "join_golden.json")
XSD = "http://www.w3.org/2001/XMLSchema#"
RDF = "http://www.w3.org/1999/02/22-rdf-syntax-ns#"
RS = "http://www.w3.org/2001/sw/DataAccess/tests/result-set#"

TOKEN = re.compile(r'''\s*(?:(\#[^\n]*)|(<[^>]*>)|("(?:[^"\\]|\\.)*")|(\^\^)|(@[a-zA-Z][a-zA-Z0-9-]*)|(_:[A-Za-z0-9_]+)|'''
                   r'''([A-Za-z_][\w.-]*)?:([\w.-]*[\w-])?|(\[)|(\])|([;,.])|(-?\d+)|(a)(?=\s))''')


def tokenize(text):
    pos, out = 0, []
    while pos < len(text):
        if text[pos:].strip() == "":
            break
        m = TOKEN.match(text, pos)
        if not m or m.end() == pos:
            raise ValueError("turtle: cannot tokenize at %r" % text[pos:pos + 40])
        pos = m.end()
        (cmt, iri, lit, dt, lang, bn, pfx, local, lb, rb, punct, num, a) = m.groups()
        if cmt is not None:
            continue
        if iri is not None: out.append(("iri", iri[1:-1]))
        elif lit is not None: out.append(("lit", bytes(lit[1:-1], "utf-8").decode("unicode_escape")))
        elif dt is not None: out.append(("dt", None))
        elif lang is not None:
            out.append(("prefix_kw", None) if lang == "@prefix" else ("lang", lang[1:]))
        elif bn is not None: out.append(("bnode", bn[2:]))
        elif lb is not None: out.append(("[", None))
        elif rb is not None: out.append(("]", None))
        elif punct is not None: out.append((punct, None))
        elif num is not None: out.append(("num", num))
        elif a is not None: out.append(("a", None))
        else: out.append(("pname", ((pfx or ""), (local or ""))))
    return out


class Turtle:
    """Just enough Turtle for the W3C fixtures: prefixes, `a`, ; , lists of objects, typed / tagged literals, labelled and
    anonymous ([ ... ]) blank nodes."""

    def __init__(self, text, bnode_scope):
        self.t = tokenize(text)
        self.i = 0
        self.prefix = {}
        self.triples = []
        self.scope = bnode_scope
        self.anon = 0

    def peek(self):
        return self.t[self.i] if self.i < len(self.t) else (None, None)

    def take(self):
        tok = self.t[self.i]
        self.i += 1
        return tok

    def parse(self):
        while self.i < len(self.t):
            k, _ = self.peek()
            if k == "prefix_kw":
                self.take()
                _, (p, _l) = self.take()
                _, iri = self.take()
                self.prefix[p] = iri
                assert self.take()[0] == "."
                continue
            s = self.node()
            if self.peek()[0] != ".":
                self.pred_obj_list(s)
            assert self.take()[0] == ".", self.t[self.i - 3:self.i + 2]
        return self.triples

    def pred_obj_list(self, s):
        while True:
            k, _ = self.peek()
            p = "<%s>" % (RDF + "type") if k == "a" and self.take() else self.node()
            while True:
                self.triples.append((s, p, self.node()))
                if self.peek()[0] == ",":
                    self.take()
                    continue
                break
            if self.peek()[0] == ";":
                self.take()
                if self.peek()[0] in (".", "]"):
                    break
                continue
            break

    def node(self):
        k, v = self.take()
        if k == "iri":
            return "<%s>" % v
        if k == "pname":
            return "<%s%s>" % (self.prefix[v[0]], v[1])
        if k == "bnode":
            return "_:%s%s" % (self.scope, v)
        if k == "num":
            return '"%s"^^<%sinteger>' % (v, XSD)
        if k == "lit":
            nk, nv = self.peek()
            if nk == "dt":
                self.take()
                return '"%s"^^%s' % (v, self.node())
            if nk == "lang":
                self.take()
                return '"%s"@%s' % (v, nv)
            return '"%s"' % v
        if k == "[":
            self.anon += 1
            b = "_:%sanon%d" % (self.scope, self.anon)
            if self.peek()[0] != "]":
                self.pred_obj_list(b)
            assert self.take()[0] == "]"
            return b
        raise ValueError("turtle: unexpected token %r" % ((k, v),))


def read(rel):
    with open(os.path.join(REF, rel), encoding="utf-8") as f:
        return f.read()


class Dict:
    def __init__(self):
        self.ids = {}

    def id(self, term):
        if term not in self.ids:
            self.ids[term] = len(self.ids) + 1
        return self.ids[term]

    def get(self, term):
        return self.ids.get(term)


def srx_solutions(rel):
    ns = {"s": "http://www.w3.org/2005/sparql-results#"}
    root = ET.fromstring(read(rel))
    vars_ = [v.get("name") for v in root.find("s:head", ns).findall("s:variable", ns)]
    sols = []
    for r in root.find("s:results", ns).findall("s:result", ns):
        sol = {}
        for b in r.findall("s:binding", ns):
            c = list(b)[0]
            tag = c.tag.split("}")[1]
            if tag == "uri":
                t = "<%s>" % c.text
            elif tag == "bnode":
                t = "_:RESULT" + c.text
            else:
                t = '"%s"' % (c.text or "")
                if c.get("datatype"):
                    t += "^^<%s>" % c.get("datatype")
                elif c.get("{http://www.w3.org/XML/1998/namespace}lang"):
                    t += "@" + c.get("{http://www.w3.org/XML/1998/namespace}lang")
            sol[b.get("name")] = t
        sols.append(sol)
    return vars_, sols


def ttl_solutions(rel):
    tr = Turtle(read(rel), "RESULT").parse()
    by_s = {}
    for s, p, o in tr:
        by_s.setdefault(s, []).append((p, o))
    rsv = lambda n: "<%s%s>" % (RS, n)
    root = [s for s, po in by_s.items() if ("<%stype>" % RDF, rsv("ResultSet")) in po][0]
    vars_ = [o[1:-1] for p, o in by_s[root] if p == rsv("resultVariable")]
    sols = []
    for p, o in by_s[root]:
        if p != rsv("solution"):
            continue
        sol = {}
        for p2, b in by_s.get(o, []):
            if p2 != rsv("binding"):
                continue
            d = dict(by_s[b])
            sol[d[rsv("variable")][1:-1]] = d[rsv("value")]
        sols.append(sol)
    return vars_, sols


# ---------------------------------------------------------------- plan DSL (serialised as JSON)
class Q:
    """Builds the JSON plan of one query against one dictionary; constants are looked up, never inserted (an unknown
    constant makes the scan provably empty: snapshot.rs:101-110)."""

    def __init__(self, d, prefixes):
        self.d = d
        self.px = prefixes
        self.fresh = 0

    def term(self, t):
        if t.startswith("?"):
            return {"var": t[1:]}
        if t == "[]":
            self.fresh += 1
            return {"var": "_b%d" % self.fresh}
        if t == "a":
            t = "rdf:type"
        if t.startswith("<") or t.startswith('"'):
            full = t
        elif re.fullmatch(r"-?\d+", t):
            full = '"%s"^^<%sinteger>' % (t, XSD)
        else:
            p, l = t.split(":", 1)
            full = "<%s%s>" % (self.px[p], l)
        return {"id": self.d.get(full), "term": full}

    def scan(self, s, p, o, graph="default"):
        g = "default" if graph == "default" else self.term(graph)
        return {"op": "scan", "g": g, "s": self.term(s), "p": self.term(p), "o": self.term(o)}


def join(l, r): return {"op": "join", "l": l, "r": r}
def leftjoin(l, r): return {"op": "leftjoin", "l": l, "r": r}
def union(l, r): return {"op": "union", "l": l, "r": r}
def filt(x, preds): return {"op": "filter", "in": x, "preds": preds}
def any_bound(x, vars_): return {"op": "filter_any_bound", "in": x, "vars": vars_}


FOAF = {"foaf": "http://xmlns.com/foaf/0.1/", "ex": "http://example.org/things#", "rdf": RDF}
EX = {"": "http://example/", "rdf": RDF}
NS = {"": "http://example.org/ns#", "rdf": RDF}


def q_opt_1(q):
    return leftjoin(q.scan("?x", "foaf:mbox", "?mbox"), q.scan("?x", "foaf:name", "?name"))


def q_opt_2(q):
    return leftjoin(q_opt_1(q), q.scan("?x", "foaf:nick", "?nick"))


def q_opt_3(q):
    a = q.scan("?x", "foaf:mbox", "?mbox")
    return union(a, join(q.scan("?x", "foaf:mbox", "?mbox"), q.scan("?x", "foaf:name", "?name")))


def q_opt_complex_1(q):
    p = leftjoin(q.scan("?person", "foaf:nick", "?nick"), q.scan("?person", "foaf:isPrimaryTopicOf", "?page"))
    inner = join(q.scan("?person", "foaf:name", "?name"),
                 union(q.scan("?person", "foaf:depiction", "?img"), q.scan("?person", "foaf:firstName", "?firstN")))
    return any_bound(leftjoin(p, inner), ["page", "img", "firstN"])


def _person_graph(q):
    bgp = join(q.scan("?person", "a", "foaf:Person"), q.scan("?person", "foaf:name", "?name"))
    g = join(q.scan("?_g1", "foaf:name", "?name", graph="?x"), q.scan("?_g1", "foaf:nick", "?nick", graph="?x"))
    return join(bgp, g)


def q_opt_complex_2(q):
    return leftjoin(_person_graph(q), union(q.scan("?person", "ex:empId", "?id"), q.scan("?person", "ex:ssn", "?ssn")))


def q_opt_complex_3(q):
    return leftjoin(_person_graph(q), leftjoin(q.scan("?person", "ex:healthplan", "?plan"),
                                               q.scan("?person", "ex:department", "?dept")))


def q_opt_complex_4(q):
    lhs = join(q.scan("?person", "foaf:name", "?name"),
               union(q.scan("?person", "ex:healthplan", "?plan"), q.scan("?person", "ex:department", "?dept")))
    g = join(q.scan("?_g1", "foaf:name", "?name", graph="?g"), q.scan("?_g1", "foaf:depiction", "?img", graph="?g"))
    return leftjoin(lhs, join(q.scan("?person", "a", "foaf:Person"), g))


def bound1(q):
    j = leftjoin(q.scan("?a", ":b", "?c"), q.scan("?c", ":d", "?e"))
    return filt(j, [{"op": "is_null", "var": "e"}])


def two_nested_opt(q):
    return leftjoin(q.scan(":x1", ":p", "?v"), leftjoin(q.scan(":x3", ":q", "?w"), q.scan(":x2", ":p", "?v")))


def two_nested_opt_alt(q):
    a = leftjoin(q.scan(":x1", ":p", "?v"), q.scan(":x3", ":q", "?w"))
    return leftjoin(a, join(q.scan(":x3", ":q", "?w"), q.scan(":x2", ":p", "?v")))


def var_scope_join_1(q):
    return join(q.scan("?X", ":name", '"paul"'),
                leftjoin(q.scan("?Y", ":name", '"george"'), q.scan("?X", ":email", "?Z")))


def join_combo_1(q):
    a = leftjoin(q.scan("?a", ":p", "?c"), q.scan("?a", ":r", "?d"))
    b = join(a, q.scan("?a", "?p", "1"))
    return join(b, union(q.scan("?p", "a", "?y"), q.scan("?a", "?z", "?p")))


def join_combo_2(q):
    return join(q.scan("?x", "?p", "1", graph="?g"), union(q.scan("?x", ":p", "?y"), q.scan("?p", "a", "?z")))


CASES = [
    # name, default-graph data, named-graph data, prefixes, plan, result file, reference query file
    ("optional/q-opt-1", ["optional/data.ttl"], [], FOAF, q_opt_1, "optional/result-opt-1.ttl", "optional/q-opt-1.rq"),
    ("optional/q-opt-2", ["optional/data.ttl"], [], FOAF, q_opt_2, "optional/result-opt-2.ttl", "optional/q-opt-2.rq"),
    ("optional/q-opt-3", ["optional/data.ttl"], [], FOAF, q_opt_3, "optional/result-opt-3.ttl", "optional/q-opt-3.rq"),
    ("optional/q-opt-complex-1", ["optional/complex-data-1.ttl"], [], FOAF, q_opt_complex_1,
     "optional/result-opt-complex-1.ttl", "optional/q-opt-complex-1.rq"),
    ("optional/q-opt-complex-2", ["optional/complex-data-2.ttl"], ["optional/complex-data-1.ttl"], FOAF, q_opt_complex_2,
     "optional/result-opt-complex-2.ttl", "optional/q-opt-complex-2.rq"),
    ("optional/q-opt-complex-3", ["optional/complex-data-2.ttl"], ["optional/complex-data-1.ttl"], FOAF, q_opt_complex_3,
     "optional/result-opt-complex-3.ttl", "optional/q-opt-complex-3.rq"),
    ("optional/q-opt-complex-4", ["optional/complex-data-2.ttl"], ["optional/complex-data-1.ttl"], FOAF, q_opt_complex_4,
     "optional/result-opt-complex-4.ttl", "optional/q-opt-complex-4.rq"),
    ("bound/bound1", ["bound/data.ttl"], [], NS, bound1, "bound/bound1-result.ttl", "bound/bound1.rq"),
    ("algebra/two-nested-opt", ["algebra/two-nested-opt.ttl"], [], EX, two_nested_opt, "algebra/two-nested-opt.srx",
     "algebra/two-nested-opt.rq"),
    ("algebra/two-nested-opt-alt", ["algebra/two-nested-opt.ttl"], [], EX, two_nested_opt_alt,
     "algebra/two-nested-opt-alt.srx", "algebra/two-nested-opt-alt.rq"),
    ("algebra/var-scope-join-1", ["algebra/var-scope-join-1.ttl"], [], EX, var_scope_join_1,
     "algebra/var-scope-join-1.srx", "algebra/var-scope-join-1.rq"),
    ("algebra/join-combo-1", ["algebra/join-combo-graph-2.ttl"], [], EX, join_combo_1, "algebra/join-combo-1.srx",
     "algebra/join-combo-1.rq"),
    ("algebra/join-combo-2", ["algebra/join-combo-graph-2.ttl"], ["algebra/join-combo-graph-1.ttl"], EX, join_combo_2,
     "algebra/join-combo-2.srx", "algebra/join-combo-2.rq"),
]


def main():
    out = []
    for name, dflt, named, px, plan_fn, result, rq in CASES:
        d = Dict()
        quads = []
        for k, rel in enumerate(dflt):
            for s, p, o in Turtle(read(rel), "d%d_" % k).parse():
                quads.append([0, d.id(s), d.id(p), d.id(o)])
        for k, rel in enumerate(named):
            gid = d.id("<%s>" % rel)  # the graph IRI is the file's IRI; only its identity matters
            for s, p, o in Turtle(read(rel), "g%d_" % k).parse():
                quads.append([gid, d.id(s), d.id(p), d.id(o)])
        plan = plan_fn(Q(d, px))
        vars_, sols = (srx_solutions if result.endswith(".srx") else ttl_solutions)(result)
        bnodes = sorted(i for t, i in d.ids.items() if t.startswith("_:"))
        rows = []
        for sol in sols:
            row = []
            for v in vars_:
                t = sol.get(v)
                if t is None:
                    row.append(0)
                elif t.startswith("_:"):
                    row.append(-1)  # some blank node: compared by kind only (results are defined up to isomorphism)
                else:
                    i = d.get(t)
                    assert i is not None, (name, v, t)
                    row.append(i)
            rows.append(row)
        out.append({"name": name, "reference_files": {"query": rq, "data": dflt, "graph_data": named, "result": result},
                    "n_terms": len(d.ids), "bnode_ids": bnodes, "quads": quads, "plan": plan, "select": vars_,
                    "expected": rows})
    doc = {"_comment": "Generated by tests/golden/make_join_golden.py from the W3C SPARQL 1.0 evaluation tests under "
                       "testsuite/rdf-tests/sparql/sparql10 of the reference (run by testsuite/tests/sparql.rs:16-43). "
                       "Quads are [g,s,p,o] object ids (0 = default graph); plans are the ID-level lowering of each "
                       "query; expected rows use 0 for unbound and -1 for 'some blank node'.",
           "cases": out}
    with open(OUT, "w") as f:
        json.dump(doc, f, indent=None, separators=(",", ":"))
        f.write("\n")
    print("wrote", OUT, len(out), "cases;", sum(len(c["expected"]) for c in out), "expected rows")


if __name__ == "__main__":
    main()
