"""GPU: rfgpu_filter_noalloc — the ID-level filter for callers that cannot allocate (no counterpart in the reference).
Its look-back scratch is a static device array; inputs, outputs and the row count are
caller-provided device buffers (torch tensors here, created BEFORE the call: the entry point itself allocates nothing).
Checked through the C ABI against rfgpu_filter, bit for bit, on two calls in a row (the second one smaller, with other
predicates: the array keeps the first call's look-back words and must be cleared), then a third, larger one.

Collected last on purpose (file name): a failure here must not hide another test under pytest -x."""
import ctypes as C

import numpy as np
import pytest

import rdf_fusion_b200 as rf
from rdf_fusion_b200 import _ffi as F

pytestmark = [pytest.mark.gpu]


def _run_noalloc(ctx, cols, preds):
    import torch
    dev = torch.device("cuda", ctx.device)
    n = len(cols[0])
    d_in = [torch.from_numpy(np.ascontiguousarray(c, np.uint32).view(np.int32)).to(dev) for c in cols]
    d_out = [torch.full((max(n, 1),), -1, dtype=torch.int32, device=dev) for _ in cols]
    d_cnt = torch.full((1,), -1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize(dev)  # the ctx stream does not wait for torch's stream
    arr, keep = rf.make_filter_preds(preds)
    inp = (C.c_void_p * len(cols))(*[t.data_ptr() for t in d_in])
    outp = (C.c_void_p * len(cols))(*[t.data_ptr() for t in d_out])
    free_before = torch.cuda.mem_get_info(dev)[0]
    ctx.check(ctx.lib.rfgpu_filter_noalloc(ctx.h, len(cols), inp, n, arr, len(preds), outp, C.c_void_p(d_cnt.data_ptr())))
    ctx.synchronize()
    assert torch.cuda.mem_get_info(dev)[0] >= free_before, "the call took device memory"
    kept = int(d_cnt.cpu()[0])
    return kept, [t[:kept].cpu().numpy().view(np.uint32) for t in d_out]


def _run_existing(ctx, cols, preds):
    names = [f"c{i}" for i in range(len(cols))]
    t = ctx.import_table(dict(zip(names, cols)), {k: True for k in names})
    f = t.filter(preds)
    return f.num_rows, (f.to_numpy() if f.num_rows else [np.zeros(0, np.uint32)] * len(cols))


def test_noalloc_filter_equals_rfgpu_filter_on_consecutive_calls(ctx):
    rng = np.random.default_rng(11)
    calls = [
        (300_000, [rf.FPred(F.F_NE_CONST, 0, value=7), rf.FPred(F.F_IS_NOT_NULL, 1)]),
        (9_001, [rf.FPred(F.F_NE_COL, 0, col2=2), rf.FPred(F.F_IN, 1, ids=[3, 5, 5, 9])]),  # fewer tiles than before
        (1_200_000, [rf.FPred(F.F_BETWEEN, 2, value=4, hi=30)]),
        (0, [rf.FPred(F.F_IS_NULL, 0)]),
    ]
    for n, preds in calls:
        cols = []
        for _ in range(3):
            c = rng.integers(1, 40, n).astype(np.uint32)
            c[rng.random(n) < 0.2] = 0
            cols.append(c)
        want_n, want = _run_existing(ctx, cols, preds)
        got_n, got = _run_noalloc(ctx, cols, preds)
        assert got_n == want_n, (n, got_n, want_n)
        for g, w in zip(got, want):
            assert np.array_equal(g, w), n  # bit-exact, same (input) order
    assert want_n == 0


def test_noalloc_filter_rejects_what_needs_an_allocation(ctx):
    import torch
    dev = torch.device("cuda", ctx.device)
    buf = torch.zeros(16, dtype=torch.int32, device=dev)
    cnt = torch.zeros(1, dtype=torch.int64, device=dev)
    one = (C.c_void_p * 1)(buf.data_ptr())
    arr, keep = rf.make_filter_preds([rf.FPred(F.F_IS_NOT_NULL, 0)])
    too_many = (1 << 26) + 1  # RFGPU_FILTER_NOALLOC_MAX_ROWS + 1: rejected before anything is launched
    rc = ctx.lib.rfgpu_filter_noalloc(ctx.h, 1, one, too_many, arr, 1, one, C.c_void_p(cnt.data_ptr()))
    assert rc == F.E_UNSUPPORTED and b"static scratch" in ctx.lib.rfgpu_last_error(ctx.h)
    arr, keep = rf.make_filter_preds([rf.FPred(F.F_IN, 0, ids=list(range(1, 11)))])
    rc = ctx.lib.rfgpu_filter_noalloc(ctx.h, 1, one, 16, arr, 1, one, C.c_void_p(cnt.data_ptr()))
    assert rc == F.E_UNSUPPORTED and b"8 ids" in ctx.lib.rfgpu_last_error(ctx.h)
