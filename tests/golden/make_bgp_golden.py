#!/usr/bin/env python
"""Generates tests/golden/bgp_golden.json: two more groups of the W3C SPARQL 1.0 evaluation tests the reference runs
(testsuite/tests/sparql.rs:16-34) —
  * sparql10/dataset (the 12 approved cases): FROM / FROM NAMED build the dataset — several FROM files merge into the
    default graph (blank nodes scoped per file), FROM NAMED files become named graphs — then the same scans, unions and
    joins over the default graph and GRAPH ?g as sparql10/graph;
  * sparql10/basic: multi-pattern BGPs — RDF collections `(1 ?v)` expanded to rdf:first / rdf:rest chains over fresh
    variables (list-1..4: up to five patterns joined on blank-node variables), a variable bound by two patterns (var-1/2:
    a join on two shared variables), object lists (spoo-1) and a BGP with a constant that is not in the data
    (bgp-no-match: an unknown constant makes the scan provably empty, snapshot.rs:101-110).
Same conventions and plan DSL as make_join_golden.py / make_pattern_golden.py; the Turtle reader is extended with
collections here.

Run in the build container only (needs /root/reference):  python tests/golden/make_bgp_golden.py
"""
import json
import os
import re
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_join_golden as G  # noqa: E402
import make_pattern_golden as P  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "bgp_golden.json")
NS = {"": "http://example.org/ns#", "rdf": G.RDF, "xsd": G.XSD}
FOAF = {"": "http://example.org/", "foaf": "http://xmlns.com/foaf/0.1/", "rdf": G.RDF}
PAREN = re.compile(r"[()]")


class Turtle(G.Turtle):
    """make_join_golden.Turtle + collections: `( a b )` is _:l1 rdf:first a ; rdf:rest [ rdf:first b ; rdf:rest rdf:nil ],
    `()` is rdf:nil"""

    def __init__(self, text, bnode_scope):
        # parentheses become stand-alone pseudo-IRIs the base tokenizer passes through (never inside a literal in these files)
        assert not re.search(r'"[^"\n]*[()][^"\n]*"', text), "parenthesis inside a literal"
        super().__init__(PAREN.sub(lambda m: " <%s> " % ("__open__" if m.group() == "(" else "__close__"), text), bnode_scope)

    def node(self):
        if self.peek() == ("iri", "__open__"):
            self.take()
            items = []
            while self.peek() != ("iri", "__close__"):
                items.append(super().node() if self.peek() != ("iri", "__open__") else self.node())
            self.take()
            head = "<%snil>" % G.RDF
            for it in reversed(items):
                self.anon += 1
                b = "_:%slist%d" % (self.scope, self.anon)
                self.triples.append((b, "<%sfirst>" % G.RDF, it))
                self.triples.append((b, "<%srest>" % G.RDF, head))
                head = b
            return head
        return super().node()


def collection(q, subject, pred, items):
    """the patterns of `subject pred ( items… )` with one fresh variable per list cell"""
    if not items:
        return q.scan(subject, pred, "rdf:nil")
    cells = ["?_l%d" % i for i in range(len(items))]
    plan = q.scan(subject, pred, cells[0])
    for i, it in enumerate(items):
        plan = G.join(plan, q.scan(cells[i], "rdf:first", it))
        plan = G.join(plan, q.scan(cells[i], "rdf:rest", cells[i + 1] if i + 1 < len(items) else "rdf:nil"))
    return plan


def list_1(q): return collection(q, ":x", "?p", [])
def list_2(q): return collection(q, ":x", "?p", ["1"])
def list_3(q): return collection(q, ":x", "?p", ["?v"])
def list_4(q): return collection(q, ":x", "?p", ["?v", "?w"])
def var_1(q): return q.scan(":x", "?p", "?v")
def var_2(q): return G.join(q.scan(":x", "?p", "?v"), q.scan(":x", "?p", "?v"))
def spoo_1(q): return G.join(q.scan("?s", ":p1", "1"), q.scan("?s", ":p1", "2"))
def no_match(q): return G.join(q.scan("?x", "foaf:name", '"John Smith"'), q.scan("?x", "a", "foaf:Womble"))


DS, BA = "dataset/", "basic/"
G4 = ["data-g1.ttl", "data-g2.ttl", "data-g3.ttl", "data-g4.ttl"]
CASES = [
    # name, FROM files (merged default graph), FROM NAMED files, prefixes, plan, result, query, directory
    ("dataset/dataset-01", ["data-g1.ttl"], [], P.EX, P.spo, "dataset-01.ttl", "dataset-01.rq", DS),
    ("dataset/dataset-02", [], ["data-g1.ttl"], P.EX, P.spo, "dataset-02.ttl", "dataset-02.rq", DS),
    ("dataset/dataset-03", [], ["data-g1.ttl"], P.EX, P.g_spo, "dataset-03.ttl", "dataset-03.rq", DS),
    ("dataset/dataset-04", ["data-g1.ttl"], [], P.EX, P.g_spo, "dataset-04.ttl", "dataset-04.rq", DS),
    ("dataset/dataset-05", ["data-g1.ttl"], ["data-g2.ttl"], P.EX, P.spo, "dataset-05.ttl", "dataset-05.rq", DS),
    ("dataset/dataset-06", ["data-g1.ttl"], ["data-g2.ttl"], P.EX, P.g_spo, "dataset-06.ttl", "dataset-06.rq", DS),
    ("dataset/dataset-07", ["data-g1.ttl"], ["data-g2.ttl"], P.EX, P.spo_union_g, "dataset-07.ttl", "dataset-07.rq", DS),
    ("dataset/dataset-08", ["data-g1.ttl"], ["data-g2.ttl"], P.EX, P.spo_join_g, "dataset-08.ttl", "dataset-08.rq", DS),
    ("dataset/dataset-09b", ["data-g3-dup.ttl"], ["data-g3.ttl"], P.EX, P.spo_join_g, "dataset-09.ttl", "dataset-09b.rq", DS),
    ("dataset/dataset-10b", ["data-g3-dup.ttl"], ["data-g3.ttl"], P.EX, P.spo_join_g, "dataset-10.ttl", "dataset-10b.rq", DS),
    ("dataset/dataset-11", ["data-g1.ttl"], G4, P.EX, P.spo_union_g, "dataset-11.ttl", "dataset-11.rq", DS),
    ("dataset/dataset-12b", [x.replace(".ttl", "-dup.ttl") for x in G4], G4, P.EX, P.spo_union_g, "dataset-12.ttl",
     "dataset-12b.rq", DS),
    ("basic/list-1", ["data-2.ttl"], [], NS, list_1, "list-1.srx", "list-1.rq", BA),
    ("basic/list-2", ["data-2.ttl"], [], NS, list_2, "list-2.srx", "list-2.rq", BA),
    ("basic/list-3", ["data-2.ttl"], [], NS, list_3, "list-3.srx", "list-3.rq", BA),
    ("basic/list-4", ["data-2.ttl"], [], NS, list_4, "list-4.srx", "list-4.rq", BA),
    ("basic/var-1", ["data-5.ttl"], [], NS, var_1, "var-1.srx", "var-1.rq", BA),
    ("basic/var-2", ["data-5.ttl"], [], NS, var_2, "var-2.srx", "var-2.rq", BA),
    ("basic/spoo-1", ["data-6.ttl"], [], NS, spoo_1, "spoo-1.srx", "spoo-1.rq", BA),
    ("basic/bgp-no-match", ["data-7.ttl"], [], FOAF, no_match, "bgp-no-match.srx", "bgp-no-match.rq", BA),
]


def main():
    out = []
    for name, dflt, named, px, plan_fn, result, rq, base in CASES:
        d = G.Dict()
        quads = []
        for k, rel in enumerate(dflt):
            for s, p, o in Turtle(G.read(base + rel), "d%d_" % k).parse():
                quads.append([0, d.id(s), d.id(p), d.id(o)])
        for k, rel in enumerate(named):
            gid = d.id("<%s>" % rel)
            for s, p, o in Turtle(G.read(base + rel), "g%d_" % k).parse():
                quads.append([gid, d.id(s), d.id(p), d.id(o)])
        plan = plan_fn(G.Q(d, px))
        vars_, sols = G.srx_solutions(base + result) if result.endswith(".srx") else G.ttl_solutions(base + result)
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
        out.append({"name": name, "reference_files": {"query": base + rq, "data": [base + x for x in dflt],
                                                      "graph_data": [base + x for x in named], "result": base + result},
                    "n_terms": len(d.ids), "bnode_ids": bnodes, "quads": quads, "plan": plan, "select": vars_,
                    "expected": rows})
    doc = {"_comment": "Generated by tests/golden/make_bgp_golden.py from the W3C SPARQL 1.0 evaluation tests sparql10/dataset "
                       "and sparql10/basic of the reference's testsuite.  Same layout as join_golden.json.",
           "cases": out}
    with open(OUT, "w") as f:
        json.dump(doc, f, indent=None, separators=(",", ":"))
        f.write("\n")
    print("wrote", OUT, len(out), "cases;", sum(len(c["expected"]) for c in out), "expected rows")


if __name__ == "__main__":
    main()
