"""oracle/ — TEST INFRASTRUCTURE ONLY: CPU restatements of the reference's algorithms for the hot path, each function citing
the reference file:line it follows.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import, link or execute anything in this package; nothing under rdf_fusion_b200/ does, and the product fails loudly
without its CUDA library (there is no CPU fallback).

  rf_oracle_scan.cpp, rf_oracle_join.cpp, rf_oracle_capi.cpp (C++17, built by the Makefile), oracle_py.py (ctypes wrapper +
      row-at-a-time Python definitions): index permutations, scan instructions, choose_index, range pruning, predicate
      push-down, SPARQL joins (equi / compatible-mapping / cross, OPTIONAL), MINUS / EXISTS, ID-level filters, GROUP BY /
      COUNT / DISTINCT, plain-term decode.  Pinned on tests/golden/scan_golden.json (the reference's own unit-test vectors) and
      on 90 W3C evaluation cases (tests/golden/*_golden.json).
  typed_value.py (pure Python): term -> typed value (the xsd lexical mappings), TypedValueRef::partial_cmp including duration
      order, the comparison operators under SPARQL's three-valued logic, the dense union of ENC_TV.  Pinned on the reference's
      unit-test vectors (date_time.rs, decimal.rs, duration.rs, float.rs, double.rs, int.rs, integer.rs) and on 17 W3C
      expr-equals / open-world cases (tests/golden/tv_golden.json).

The reference itself (Rust, DataFusion 52 / arrow 57 not vendored) cannot be built in this image: there is no oracle/_ref.
"""
