"""igl_fp_run_static: the Filter/Projection launch for callers that may not
allocate device memory while running (engines that pre-reserve all of their HBM,
latency-critical streams that cannot absorb an allocation's implicit
synchronisation).  It must produce exactly what igl_fp_run produces — the two
share the kernel and differ only in where the scan scratch lives (a fixed device
array inside the library, shared by every call on that GPU).

Collected last on purpose (file name), so a failure here cannot hide the result
of any other GPU test in a stop-at-first-failure run.  The checks hold both in a
fresh process and after the rest of the suite has used the GPU.
"""
import ctypes as C

import numpy as np
import pytest

from igloo_b200.expr import Field, col, T_BOOL, T_F64, T_I32

pytestmark = pytest.mark.gpu

STATIC_MAX_ROWS = 4 << 20


@pytest.fixture(scope="module")
def ctx():
    from igloo_b200.device import Context
    c = Context(0)
    yield c
    c.close()


def _run_static(fp, cols, nrows, out):
    from igloo_b200._lib import DevCol, check, lib
    a = (DevCol * len(cols))(*[c.c_struct() for c in cols])
    o = (DevCol * len(out))(*[c.c_struct() for c in out])
    n = C.c_int64(-1)
    check(lib.igl_fp_run_static(fp._h, a, nrows, o, C.byref(n), fp.ctx.stream()))
    return n.value


def test_static_scratch_entry_point_matches_pooled_one_on_two_calls_in_a_row(ctx):
    import torch
    from igloo_b200.device import FilterProject
    seed = 0x16100002
    fields = [Field("a", T_I32), Field("b", T_I32), Field("x", T_F64)]
    proj = [(col("b") + col("a")).alias("s"), (col("x") * 1.5 + 0.25).alias("y")]
    # two different shapes back to back: the second call starts from the status
    # words the first one left in the shared array and must not be fooled by them
    for nrows, thr in ((3_000_017, 500), (777_777, 100)):
        a = ctx.gen(1, seed, 0, 0, nrows, 1000)
        b = ctx.gen(1, seed, 1, 0, nrows, 1 << 31)
        x = ctx.gen(3, seed, 2, 0, nrows)
        fp = FilterProject(ctx, fields, col("a") < thr, proj)
        want, n_want = fp.run([a, b, x], nrows)                      # existing entry point (pooled scratch)
        got = fp.alloc_outputs(nrows)
        for c in got:
            c.data.zero_()
        n_got = _run_static(fp, [a, b, x], nrows, got)
        assert n_got == n_want and 0 < n_got < nrows
        # integer column and float64 column: bit-exact, the tolerance of every parity test here
        assert torch.equal(got[0].data[:n_got], want[0].data[:n_want])
        assert torch.equal(got[1].data[:n_got].view(torch.int64), want[1].data[:n_want].view(torch.int64))
    # and against the CPU oracle (a prefix the oracle finishes quickly; row order is preserved, so
    # the prefix of the output is the output of the prefix)
    from igloo_b200.expr import lower
    from oracle import oracle as orc
    m = 200_000
    ha, hb, hx = (orc.gen_i32_uniform(seed, 0, 0, m, 1000), orc.gen_i32_uniform(seed, 1, 0, m, 1 << 31),
                  orc.gen_f64_unit(seed, 2, 0, m))
    t = orc.Table.scan([orc.Column("a", orc.T_I32, ha), orc.Column("b", orc.T_I32, hb), orc.Column("x", orc.T_F64, hx)])
    ref = t.filter(lower(col("a") < thr, fields)[0]).project([lower(e, fields)[0] for e in proj], ["s", "y"]).columns()
    k = len(ref[0].values)
    assert 0 < k <= n_got
    assert np.array_equal(got[0].data[:k].cpu().numpy(), ref[0].values)
    assert np.array_equal(got[1].data[:k].cpu().numpy().view(np.int64), ref[1].values.view(np.int64))
    # the pooled entry point is unaffected by the static one having run in between
    again, n_again = fp.run([a, b, x], nrows)
    assert n_again == n_want and torch.equal(again[0].data[:n_again], want[0].data[:n_want])


def test_static_scratch_entry_point_rejects_what_it_cannot_hold(ctx):
    from igloo_b200._lib import IglooCudaError
    from igloo_b200.device import DeviceColumn, FilterProject
    fields = [Field("a", T_I32)]
    fp = FilterProject(ctx, fields, col("a") < 5, [])
    too_many = STATIC_MAX_ROWS + 1
    a = DeviceColumn.empty(T_I32, too_many)
    out = fp.alloc_outputs(too_many)
    with pytest.raises(IglooCudaError) as e:
        _run_static(fp, [a], too_many, out)
    assert e.value.code == -1 and "static scratch limit" in e.value.message
    # exactly at the limit is fine
    a.data.zero_()
    assert _run_static(fp, [a], STATIC_MAX_ROWS, out) == STATIC_MAX_ROWS
    # boolean output needs pooled byte planes -> refused, not silently allocated
    fb = FilterProject(ctx, fields, None, [(col("a") < 5).alias("m")])
    ob = fb.alloc_outputs(1000)
    with pytest.raises(IglooCudaError) as e:
        _run_static(fb, [a], 1000, ob)
    assert e.value.code == -2
