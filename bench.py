#!/usr/bin/env python
"""bench.py — hash-join input rows/s (headline) and filter+project HBM GB/s.

Contract (one JSON line on stdout, rank 0):
  metric  : hash_join_input_rows_per_s  — BASELINE.json "hash-join input rows/s"
  workload: --workload c3 (default) — inner join, build 2^27 (k,v) x probe 1e9
            (k,w), int64 uniform keys, every probe row matches once.  c4 = lineitem JOIN
            orders at SF100 row counts, c5a = Zipf(1.1) probe keys on 2^28 unique build
            keys, c5b = Zipf(1.1) build keys (heavy duplicate buckets).  At N>1 the same
            total is range-sharded over the ranks, hash-repartitioned and joined locally
            ("strong" scaling; the exchange is inside the timed step).
  value   : (build rows + probe rows) * steps / device time, inputs resident in HBM
  parity  : after the timed region, the multiset digest of the output (all ranks) against the
            digest of the answer recomputed from the data definition, at the timed size
  e2e     : the same join through the operator API (igl_op_*) with HOST Arrow buffers
  filter_project : the C2 selectivity sweep (1e9 rows) with achieved HBM GB/s
  roofline: dominant kernel = join probe; algorithmic bytes / its CUDA-event time
  cpu_baseline / --impl reference : the reference's algorithm (oracle port,
            printed-string keys, 1024-row batches, one thread) on a bounded sample
"""
import argparse
import functools
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SEED_C2, SEED_C3, SEED_C4, SEED_C5 = 0x16100002, 0x16100003, 0x16100004, 0x16100005
NOMINAL_HBM_GBS = 8000.0   # BASELINE.json "~8 TB/s"
M64 = (1 << 64) - 1


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(kernel_substr):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the dominant kernel, read from
    the committed ncu summary of the same command (profiles/ncu_traffic.json, written by
    tools/ncu_summary.py from an `ncu --set full` capture); None when no capture is on file."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(p):
        return None, None
    try:
        d = json.load(open(p))
    except Exception:
        return None, None
    for k, v in d.get("kernels", {}).items():
        if kernel_substr in k:
            return int(v["dram_bytes"]), f"profiles/ncu_traffic.json <- {d.get('source')}"
    return None, None


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "25"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------
# reference arm / cpu baselines
# ---------------------------------------------------------------------------------
def cpu_reference_join(build_bits, nprobe, mode=0, threads=1):
    from oracle import oracle as orc
    secs, rows = orc.bench_join(SEED_C3, build_bits, nprobe, mode, threads)
    assert rows == nprobe, rows
    return ((1 << build_bits) + nprobe) / secs, secs


def run_reference(args):
    """--impl reference: the reference's CPU hash join (hash_join.rs:100-213 as
    written: printed-string keys into a std map, per-row slices, one output row
    per emit, 1024-row batches) on the host cores.  The Rust crate cannot be
    built here (no cargo), so this is the oracle PORT; the reference's pipeline is one
    sequential stream plus one probe task, so its algorithm uses one core however many
    the box has (`cores_available` is printed beside it)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    bits, nprobe = args.ref_build_bits, args.ref_probe_rows
    for _ in range(args.warmup if args.warmup < 1 else 1):
        cpu_reference_join(max(10, bits - 6), max(1000, nprobe // 64))
    t0 = time.time()
    vals = []
    for _ in range(max(1, args.steps)):
        v, s = cpu_reference_join(bits, nprobe)
        vals.append(v)
    dt = (time.time() - t0) / max(1, args.steps)
    value = statistics.median(vals)
    sample = f"C3 at reduced scale: build 2^{bits} x probe {nprobe} rows per step, faithful string-key port, 1024-row batches"
    line = {"impl": "reference", "metric": "hash_join_input_rows_per_s", "value": value, "unit": "rows/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int64", "data": "synthetic",
            "config": {"workload": "C3 inner hash join 2^27 x 1e9 int64 uniform keys (reference arm runs the bounded sample below)",
                       "sample": sample},
            "cpu_baseline": {"value": value, "unit": "rows/s", "cores": 1, "cores_available": host_cores(), "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def cpu_baselines_extra(args):
    """The baselines BASELINE.md §4 names beside the faithful port: the port's typed multi-threaded
    join on every core, Arrow C++ (pyarrow.compute / Acero) on every core for both workloads, and the
    port's Filter+Projection.  Bounded samples; reported, never a target."""
    from oracle import oracle as orc
    cores = host_cores()
    out = {"cores_available": cores}
    try:
        v, s = cpu_reference_join(22, 31_250_000, mode=1, threads=cores)
        out["port_typed_join_all_cores"] = {"value": v, "unit": "rows/s", "cores": cores, "kind": "port",
                                            "sample": f"C3 / 32 (2^22 x 31.25 M), typed open-addressing oracle, {cores} threads, {s:.2f} s"}
    except Exception as e:   # noqa: BLE001
        out["port_typed_join_all_cores"] = {"error": str(e)[:200]}
    try:
        n = 25_000_000
        s, kept = orc.bench_filter_project(SEED_C2, n, 500)
        out["port_filter_project"] = {"value": n / s, "unit": "rows/s", "gb_per_s": n * (16 + 12 * kept / n) / s / 1e9, "cores": 1, "kind": "port",
                                      "sample": f"C2 / 40 ({n} rows, a < 500), FilterExec -> ProjectionExec port, 1024-row batches, {s:.2f} s"}
    except Exception as e:   # noqa: BLE001
        out["port_filter_project"] = {"error": str(e)[:200]}
    try:
        import pyarrow as pa
        import pyarrow.compute as pc
        pa.set_cpu_count(cores)
        n = 125_000_000
        a = pa.array(orc.gen_i32_uniform(SEED_C2, 0, 0, n, 1000)); b = pa.array(orc.gen_i32_uniform(SEED_C2, 1, 0, n, 1 << 31))
        x = pa.array(orc.gen_f64_unit(SEED_C2, 2, 0, n))
        t = pa.table({"a": a, "b": b, "x": x})
        t0 = time.time()
        f = t.filter(pc.less(t["a"], 500))
        s_ = pc.add(f["b"], f["a"]); y = pc.add(pc.multiply(f["x"], 1.5), 0.25)
        dt = time.time() - t0
        out["arrow_filter_project"] = {"value": n / dt, "unit": "rows/s", "gb_per_s": n * (16 + 12 * len(s_) / n) / dt / 1e9, "cores": cores, "kind": "arrow-cpp",
                                       "sample": f"C2 / 8 ({n} rows, a < 500), pyarrow {pa.__version__} compute kernels, {dt:.2f} s"}
        del t, f, s_, y, a, b, x
        bits, npb = 23, 62_500_000
        R = pa.table({"k": orc.gen_perm_pow2(0, 1 << bits, bits), "v": orc.gen_i64_uniform(SEED_C3, 1, 0, 1 << bits)})
        L = pa.table({"k": orc.gen_i64_uniform(SEED_C3, 2, 0, npb, 1 << bits), "w": orc.gen_i64_uniform(SEED_C3, 3, 0, npb)})
        t0 = time.time()
        J = L.join(R, "k", join_type="inner", use_threads=True)
        dt = time.time() - t0
        assert J.num_rows == npb
        out["arrow_hash_join"] = {"value": ((1 << bits) + npb) / dt, "unit": "rows/s", "cores": cores, "kind": "arrow-cpp",
                                  "sample": f"C3 / 16 (2^{bits} x {npb}), Acero hash join, {dt:.2f} s"}
    except Exception as e:   # noqa: BLE001
        out["arrow"] = {"error": str(e)[:200]}
    return out


# ---------------------------------------------------------------------------------
# workloads: row-range shards and the answer's digest, both from the data
# definition alone
# ---------------------------------------------------------------------------------
class Workload:
    """`shard(rank, world)` -> (right cols, n_right, left cols, n_left) on this rank's GPU;
    `expected(left cols, n_left, rank, world)` -> (digest, rows) of the join output that this
    rank's PROBE shard produces, recomputed without a hash table (the sum over ranks is the
    whole answer, however the join distributes its output)."""

    def __init__(self, ctx, scale):
        self.ctx, self.scale = ctx, scale

    def _range(self, total, rank, world):
        return total * rank // world, total * (rank + 1) // world


class C3(Workload):
    name = "c3"
    check = "digest of (k, w, k, v(perm^-1(k))) regenerated from the data definition"

    def __init__(self, ctx, scale):
        super().__init__(ctx, scale)
        from igloo_b200.expr import Field, T_I64
        self.bits = 27 - max(0, scale.bit_length() - 1) if scale > 1 else 27
        self.nb, self.np = 1 << self.bits, 1_000_000_000 // scale
        self.lf = [Field("k", T_I64), Field("w", T_I64)]
        self.rf = [Field("k", T_I64), Field("v", T_I64)]
        self.lkey = self.rkey = 0
        self.desc = f"C3 inner hash join: build 2^{self.bits} (k,v) x probe {self.np} (k,w), int64 uniform keys, 1 match per probe row"
        self.left_row_bytes, self.right_row_bytes = 16, 16

    def shard(self, rank, world):
        g = self.ctx.gen
        b0, b1 = self._range(self.nb, rank, world); p0, p1 = self._range(self.np, rank, world)
        right = [g(4, 0, 0, b0, b1 - b0, bits=self.bits), g(2, SEED_C3, 1, b0, b1 - b0)]
        left = [g(2, SEED_C3, 2, p0, p1 - p0, m=self.nb), g(2, SEED_C3, 3, p0, p1 - p0)]
        return right, b1 - b0, left, p1 - p0

    def expected(self, left, nl, rank, world):
        v = self.ctx.gen(5, SEED_C3, 1, 0, nl, bits=self.bits, keys=left[0])
        return self.ctx.digest([left[0], left[1], left[0], v], nl), nl


class C5a(C3):
    """Skewed shuffle: 2^28 unique build keys, probe keys Zipf(1.1) over the build domain (inverse
    CDF, rank -> key through the build side's bijection, so hot keys are scattered)."""
    name = "c5a"

    def __init__(self, ctx, scale):
        Workload.__init__(self, ctx, scale)
        from igloo_b200.expr import Field, T_I64
        self.bits = 28 - max(0, scale.bit_length() - 1) if scale > 1 else 28
        self.nb, self.np = 1 << self.bits, 2_000_000_000 // scale
        self.lf = [Field("k", T_I64), Field("w", T_I64)]
        self.rf = [Field("k", T_I64), Field("v", T_I64)]
        self.lkey = self.rkey = 0
        self.desc = f"C5a skewed shuffle: build 2^{self.bits} unique (k,v) x probe {self.np} (k,w), probe keys Zipf(1.1) by inverse CDF"
        self.left_row_bytes, self.right_row_bytes = 16, 16

    def shard(self, rank, world):
        g = self.ctx.gen
        b0, b1 = self._range(self.nb, rank, world); p0, p1 = self._range(self.np, rank, world)
        right = [g(4, 0, 0, b0, b1 - b0, bits=self.bits), g(2, SEED_C5, 1, b0, b1 - b0)]
        left = [g(8, SEED_C5, 2, p0, p1 - p0, m=self.nb, bits=self.bits), g(2, SEED_C5, 3, p0, p1 - p0)]
        return right, b1 - b0, left, p1 - p0

    def expected(self, left, nl, rank, world):
        v = self.ctx.gen(5, SEED_C5, 1, 0, nl, bits=self.bits, keys=left[0])
        return self.ctx.digest([left[0], left[1], left[0], v], nl), nl


class C5b(C5a):
    """Heavy buckets: build keys drawn Zipf(1.1) (the top key owns ~11 % of the build rows),
    probe keys uniform over the domain; every (probe row, build row) pair of equal keys is output."""
    name = "c5b"
    check = "digest of the sort-merge join of the regenerated inputs (torch.sort + searchsorted)"

    def __init__(self, ctx, scale):
        super().__init__(ctx, scale)
        self.desc = f"C5b heavy buckets: build {self.nb} rows with Zipf(1.1) keys x probe {self.np} uniform keys over the 2^{self.bits} domain"

    def _build(self, b0, n):
        g = self.ctx.gen
        return [g(8, SEED_C5, 4, b0, n, m=self.nb, bits=self.bits), g(2, SEED_C5, 1, b0, n)]

    def shard(self, rank, world):
        g = self.ctx.gen
        b0, b1 = self._range(self.nb, rank, world); p0, p1 = self._range(self.np, rank, world)
        left = [g(2, SEED_C5, 5, p0, p1 - p0, m=self.nb), g(2, SEED_C5, 3, p0, p1 - p0)]
        return self._build(b0, b1 - b0), b1 - b0, left, p1 - p0

    def expected(self, left, nl, rank, world):
        """Sort-merge in torch (an independent route to the same multiset): the whole build side is
        regenerated and sorted on this GPU, every probe row's matches are a searchsorted range."""
        import torch
        from igloo_b200.device import DeviceColumn
        bk, bv = self._build(0, self.nb)
        k, order = torch.sort(bk.data[:self.nb])
        v = bv.data[:self.nb][order]
        del order, bk, bv
        dg, rows = [0, 0, 0], 0
        step = 1 << 24
        for s in range(0, nl, step):
            e = min(nl, s + step)
            lk, lw = left[0].data[s:e], left[1].data[s:e]
            lo = torch.searchsorted(k, lk, right=False); hi = torch.searchsorted(k, lk, right=True)
            cnt = hi - lo
            tot = int(cnt.sum().item())
            if tot == 0:
                continue
            pi = torch.repeat_interleave(torch.arange(e - s, device=lk.device), cnt)
            starts = torch.cumsum(cnt, 0) - cnt
            bi = lo[pi] + (torch.arange(tot, device=lk.device) - starts[pi])

            def colz(t):
                return DeviceColumn(3, tot, torch.cat([t, torch.zeros(16, dtype=t.dtype, device=t.device)]))
            d = self.ctx.digest([colz(lk[pi]), colz(lw[pi]), colz(k[bi]), colz(v[bi])], tot)
            dg = [(dg[0] + d[0]) & M64, dg[1] ^ d[1], (dg[2] + d[2]) & M64]
            rows += tot
            del pi, bi, lo, hi, cnt, starts
        return tuple(dg), rows


class C4(Workload):
    """lineitem JOIN orders at SF100 row counts: orders (28 B rows) builds, lineitem (36 B rows)
    probes; 1-7 lines per order; 64 B output rows; build payload gathered by row id."""
    name = "c4"
    check = "digest of lineitem rows ++ the order columns regenerated for each line's order index"

    def __init__(self, ctx, scale):
        super().__init__(ctx, scale)
        from igloo_b200.expr import Field, T_F64, T_I32, T_I64
        self.nb, self.np = 150_000_000 // scale, 600_037_902 // scale
        self.lf = [Field("l_orderkey", T_I64), Field("l_partkey", T_I64), Field("l_quantity", T_F64),
                   Field("l_extendedprice", T_F64), Field("l_shipdate", T_I32)]
        self.rf = [Field("o_orderkey", T_I64), Field("o_custkey", T_I64), Field("o_totalprice", T_F64), Field("o_orderdate", T_I32)]
        self.lkey = self.rkey = 0
        self.desc = f"C4 TPC-H-shaped: orders {self.nb} rows x 28 B (build) JOIN lineitem {self.np} rows x 36 B (probe), 1-7 lines per order"
        self.left_row_bytes, self.right_row_bytes = 36, 28

    def _orders(self, b0, n, keys=None):
        g = self.ctx.gen
        return [g(6, 0, 0, b0, n, keys=keys), g(2, SEED_C4, 1, b0, n, m=15_000_000, keys=keys),
                g(3, SEED_C4, 2, b0, n, keys=keys), g(1, SEED_C4, 3, b0, n, m=2500, keys=keys)]

    def shard(self, rank, world):
        g = self.ctx.gen
        b0, b1 = self._range(self.nb, rank, world); p0, p1 = self._range(self.np, rank, world)
        n = p1 - p0
        idx = g(7, SEED_C4, 4, p0, n, m=self.nb)
        left = [g(6, 0, 0, 0, n, keys=idx), g(2, SEED_C4, 5, p0, n, m=20_000_000), g(3, SEED_C4, 6, p0, n),
                g(3, SEED_C4, 7, p0, n), g(1, SEED_C4, 8, p0, n, m=2500)]
        self._p0 = p0
        return self._orders(b0, b1 - b0), b1 - b0, left, n

    def expected(self, left, nl, rank, world):
        idx = self.ctx.gen(7, SEED_C4, 4, self._p0, nl, m=self.nb)
        return self.ctx.digest(list(left) + self._orders(0, nl, keys=idx), nl), nl


WORKLOADS = {"c3": C3, "c4": C4, "c5a": C5a, "c5b": C5b}


# ---------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------
def run_filter_project(ctx, nrows, steps, warmup, peak):
    """C2 sweep on the run-time specialised kernels (and the interpreter at two selectivities), with
    the compile time a plan pays once, and parity at the timed size: fused == Filter -> Projection,
    and the output prefix equals the oracle's output for the input prefix."""
    import numpy as np
    import torch
    from igloo_b200.device import FilterProject
    from igloo_b200.expr import Field, col, lower, T_F64, T_I32
    a = ctx.gen(1, SEED_C2, 0, 0, nrows, 1000)
    b = ctx.gen(1, SEED_C2, 1, 0, nrows, 1 << 31)
    x = ctx.gen(3, SEED_C2, 2, 0, nrows)
    fields = [Field("a", T_I32), Field("b", T_I32), Field("x", T_F64)]
    proj = [(col("b") + col("a")).alias("s"), (col("x") * 1.5 + 0.25).alias("y")]
    sweep = []
    out = None

    def time_plan(fp):
        for _ in range(warmup):
            fp.launch([a, b, x], nrows, out); fp.finish()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        torch.cuda.synchronize()
        ev[0].record()
        for i in range(steps):
            fp.launch([a, b, x], nrows, out)
            ev[i + 1].record()
        torch.cuda.synchronize()
        kept = fp.finish()
        return statistics.median(ev[i].elapsed_time(ev[i + 1]) for i in range(steps)), kept

    compile_s = None
    for thr in (10, 50, 100, 250, 500, 750, 900):
        fp = FilterProject(ctx, fields, col("a") < thr, proj)
        if out is None:
            out = fp.alloc_outputs(nrows)
        if compile_s is None:   # the first launch of a plan that has seen >= 64 Mi rows compiles its kernels (NVRTC)
            t0 = time.perf_counter()
            fp.launch([a, b, x], nrows, out); fp.finish()
            t1 = time.perf_counter()
            fp.launch([a, b, x], nrows, out); fp.finish()
            compile_s = (t1 - t0) - (time.perf_counter() - t1)
        ms, kept = time_plan(fp)
        sel = kept / nrows
        gb = nrows * (16 + 12 * sel) / 1e9
        sweep.append({"threshold": thr, "selectivity": round(sel, 4), "ms": round(ms, 4), "algorithmic_gb": round(gb, 3),
                      "gb_per_s": round(gb / ms * 1e3, 1), "frac_of_measured_peak": round(gb / ms * 1e3 / peak, 4),
                      "frac_of_nominal_8tbs": round(gb / ms * 1e3 / NOMINAL_HBM_GBS, 4),
                      "rows_per_s": nrows / ms * 1e3, "specialised": fp.is_specialised})
        del fp
    # the interpreter (what a plan runs on before it has seen 64 Mi rows, or without libnvrtc)
    os.environ["IGL_FP_JIT"] = "0"
    interp = []
    for thr in (100, 500):
        fp = FilterProject(ctx, fields, col("a") < thr, proj)
        ms, kept = time_plan(fp)
        gb = nrows * (16 + 12 * kept / nrows) / 1e9
        interp.append({"threshold": thr, "ms": round(ms, 4), "gb_per_s": round(gb / ms * 1e3, 1),
                       "frac_of_measured_peak": round(gb / ms * 1e3 / peak, 4), "specialised": fp.is_specialised})
        del fp
    del os.environ["IGL_FP_JIT"]
    # ---- parity at the timed size
    fused = FilterProject(ctx, fields, col("a") < 250, proj)
    o1, n1 = fused.run([a, b, x], nrows, out)
    d1 = ctx.digest(o1, n1)
    from oracle import oracle as orc
    m = 1 << 20
    ha, hb, hx = orc.gen_i32_uniform(SEED_C2, 0, 0, m, 1000), orc.gen_i32_uniform(SEED_C2, 1, 0, m, 1 << 31), orc.gen_f64_unit(SEED_C2, 2, 0, m)
    t = orc.Table.scan([orc.Column("a", orc.T_I32, ha), orc.Column("b", orc.T_I32, hb), orc.Column("x", orc.T_F64, hx)])
    ref = t.filter(lower(col("a") < 250, fields)[0]).project([lower(e, fields)[0] for e in proj], ["s", "y"]).columns()
    k = len(ref[0].values)
    prefix_ok = bool(np.array_equal(o1[0].data[:k].cpu().numpy(), ref[0].values) and
                     np.array_equal(o1[1].data[:k].cpu().numpy().view(np.int64), ref[1].values.view(np.int64)))
    del o1, out
    filt = FilterProject(ctx, fields, col("a") < 250, [])
    mid, nm = filt.run([a, b, x], nrows)
    pr = FilterProject(ctx, fields, None, proj)
    o2, n2 = pr.run(mid, nm)
    d2 = ctx.digest(o2, n2)
    parity = {"rows": n1, "fused_equals_filter_then_projection": bool(n1 == n2 and d1 == d2),
              "oracle_prefix_rows": k, "oracle_prefix_bit_exact": prefix_ok}
    assert parity["fused_equals_filter_then_projection"] and prefix_ok, parity
    del a, b, x, o2, mid
    torch.cuda.empty_cache()
    return {"workload": f"C2: {nrows} rows (a:i32,b:i32,x:f64), SELECT b+a, x*1.5+0.25 WHERE a<T; bytes = N*(16+12*s)",
            "rows": nrows, "sweep": sweep, "interpreter_IGL_FP_JIT_0": interp,
            "nvrtc_compile_s_first_large_launch": None if compile_s is None else round(compile_s, 3), "parity": parity}


def fold_digests(ds):
    return (sum(d[0] for d in ds) & M64, functools.reduce(lambda p, q: p ^ q, (d[1] for d in ds)), sum(d[2] for d in ds) & M64)


def run_ours(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    from igloo_b200.device import Context, HashJoin
    from igloo_b200.expr import col
    peak, peak_src = load_peaks()
    ctx = Context(local)
    wl = WORKLOADS[args.workload](ctx, args.scale)
    right, nr, left, nl = wl.shard(rank, world)
    rows_per_step = wl.nb + wl.np

    if world > 1:
        from igloo_b200.shuffle import PartitionedJoin, init_comm
        init_comm(ctx, rank, world)
        pj = PartitionedJoin(ctx, rank, world, wl.lf, wl.rf, wl.lkey, wl.rkey, mode=args.exchange)
        pj.prepare(right, nr, left, nl)

        def step():
            return pj.run()
    else:
        join = HashJoin(ctx, 0, wl.lf, wl.rf, col(wl.lf[wl.lkey].name), col(wl.rf[wl.rkey].name))
        join._right_cols = right
        state = {"out": None, "cap": 0}
        probe_ms = []

        def step():
            join.build(right, nr)
            cap = join.probe_count(left, nl) if join.has_duplicates else nl   # duplicate build keys: the output is sized by a count pass
            if state["out"] is None or cap > state["cap"]:
                state["out"] = None
                state["cap"] = cap
                state["out"] = join.alloc_outputs(cap, left)
            _, n = join.probe(left, nl, out=state["out"], capacity=state["cap"])
            probe_ms.append(join.phase_ms())   # CUDA events recorded by the library on the launching stream
            return n

    for _ in range(args.warmup):
        n_out = step()
    if world == 1:
        probe_ms.clear()
    sampler = ClockSampler(local)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    l0 = ctx.launches
    start = torch.cuda.Event(enable_timing=True); end = torch.cuda.Event(enable_timing=True)
    if rank == 0:
        sampler.start()
    torch.cuda.synchronize()
    start.record()
    for _ in range(args.steps):
        n_out = step()
    end.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms_total = start.elapsed_time(end)
    clocks = sampler.stop() if rank == 0 else None
    launches = ctx.launches - l0
    n_local = n_out
    if world > 1:
        t = torch.tensor([ms_total], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
        cnt = torch.tensor([n_out], device=f"cuda:{local}", dtype=torch.int64)
        dist.all_reduce(cnt)
        n_out = int(cnt.item())
    # ---- parity at the timed size, outside the timed region: the multiset digest of what the join
    # wrote (all ranks combined) against the digest of the answer recomputed from the data definition
    # alone — no hash table, no exchange involved (SURVEY.md §8d "parity check at scale")
    got = ctx.digest(pj.out if world > 1 else state["out"], n_local)
    if world == 1:
        state["out"] = None   # room for the expected columns
        torch.cuda.empty_cache()
    want, want_rows = wl.expected(left, nl, rank, world)
    if world > 1:
        both = [None] * world
        dist.all_gather_object(both, (got, want, want_rows))
        got, want, want_rows = fold_digests([b[0] for b in both]), fold_digests([b[1] for b in both]), sum(b[2] for b in both)
    parity = {"digest_equal": got == want, "rows": n_out, "rows_expected": want_rows, "check": wl.check,
              "digest": "sum / xor / second-mix sum of a splitmix64 hash of every output row (igl_digest_rows)"}
    assert n_out == want_rows, (n_out, want_rows)
    assert parity["digest_equal"], (got, want)
    ms_step = ms_total / args.steps
    value = rows_per_step / (ms_step / 1e3)
    out_row_bytes = wl.left_row_bytes + wl.right_row_bytes
    alg_bytes = wl.nb * wl.right_row_bytes + wl.np * wl.left_row_bytes + n_out * out_row_bytes

    line = {"metric": "hash_join_input_rows_per_s", "value": value, "unit": "rows/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "int64", "data": "synthetic",
            "config": {"workload": wl.desc,
                       "l2_policy": "inputs (>= 18 GB) and the hash table are far larger than the 126 MB L2",
                       "parallelism": f"hash-partitioned over {world} GPUs, exchange={args.exchange}" if world > 1 else "1 GPU"},
            "gpu_launches": launches, "clocks": clocks, "parity": parity,
            "whole_join": {"algorithmic_gb": alg_bytes / 1e9, "gb_per_s_all_gpus": alg_bytes / ms_step / 1e6,
                           "frac_of_measured_peak_per_gpu": alg_bytes / ms_step / 1e6 / peak / world,
                           "frac_of_nominal_8tbs_per_gpu": alg_bytes / ms_step / 1e6 / NOMINAL_HBM_GBS / world}}
    if world > 1:
        phases = [None] * world
        dist.all_gather_object(phases, (pj.phase_ms(), pj.stats()))
        line["phases_ms_per_rank"] = [p[0] for p in phases]
        line["exchange_stats_per_rank"] = [p[1] for p in phases]
        # NVLink: bytes each GPU read from its peers (build rows pulled + probe rows fetched by the
        # probe kernel) over the time those reads had (build pull + probe phase), against 900 GB/s
        per = []
        for ph, stt in phases:
            t = (ph["build"] + ph["probe"]) / 1e3
            per.append(stt["bytes_read_from_peers"] / t / 1e9 if t > 0 else 0.0)
        line["nvlink"] = {"bytes_in_per_gpu": [p[1]["bytes_read_from_peers"] for p in phases],
                          "gb_per_s_per_gpu_over_build_and_probe_phases": [round(x, 1) for x in per],
                          "frac_of_900": round(max(per) / 900.0, 3),
                          "floor_ms_at_900": round(max(p[1]["bytes_read_from_peers"] for p in phases) / 900e9 * 1e3, 3)}
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    if world == 1:
        ph = [statistics.median(p[i] for p in probe_ms) for i in range(4)]
        pm = ph[3]
        if pm > 0:
            probe_bytes = wl.np * wl.left_row_bytes + n_out * out_row_bytes
            kname = "join_probe_flat_kernel"
            traffic, traffic_src = ncu_traffic(kname) if args.scale == 1 and args.workload == "c3" else (None, None)
            line["roofline"] = {"bound": "hbm", "kernel": f"{kname}<key type, dense output> (inner join, unique build keys)", "achieved": probe_bytes / pm / 1e6,
                                "peak": peak, "unit": "GB/s", "frac": probe_bytes / pm / 1e6 / peak,
                                "frac_of_nominal_8tbs": probe_bytes / pm / 1e6 / NOMINAL_HBM_GBS,
                                "traffic": traffic, "traffic_source": traffic_src,
                                "peak_source": peak_src, "kernel_ms": pm,
                                "algorithmic_bytes_per_launch": probe_bytes,
                                "phases_ms": {"build_partition": ph[0], "build": ph[1], "probe_partition": ph[2], "probe_kernel": ph[3]}}
        else:
            line["roofline"] = {"bound": "hbm", "kernel": "whole join (duplicate build keys take the general probe kernels; no single dominant launch is timed)",
                                "achieved": alg_bytes / ms_step / 1e6, "peak": peak, "unit": "GB/s", "frac": alg_bytes / ms_step / 1e6 / peak,
                                "traffic": None, "peak_source": peak_src,
                                "phases_ms": {"build_partition": ph[0], "build": ph[1]}}
        del join, right, left
        torch.cuda.empty_cache()
        if not args.skip_e2e and args.workload == "c3":
            line["e2e"] = run_e2e(ctx, wl, pinned=True, steps=max(1, min(args.steps, 2)))
            line["e2e"]["pageable_host_buffers"] = run_e2e(ctx, wl, pinned=False, steps=1)
        torch.cuda.empty_cache()
        if not args.skip_fp:
            line["filter_project"] = run_filter_project(ctx, 1_000_000_000 // args.scale, max(3, args.steps), max(3, args.warmup), peak)
        if not args.skip_cpu:
            v, secs = cpu_reference_join(args.ref_build_bits, args.ref_probe_rows)
            line["cpu_baseline"] = {"value": v, "unit": "rows/s", "cores": 1, "cores_available": host_cores(), "kind": "port",
                                    "sample": f"C3 at reduced scale: build 2^{args.ref_build_bits} x probe {args.ref_probe_rows} rows, "
                                              f"faithful string-key port of hash_join.rs, 1024-row batches, {secs:.1f} s"}
            line["cpu_baselines_extra"] = cpu_baselines_extra(args)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_e2e(ctx, wl, pinned, steps):
    """The C3 join through the operator API a host engine binds — igl_op_hash_join_create,
    igl_op_push_build / igl_op_finish_build, igl_op_push, igl_op_pull — with HOST Arrow batches in
    and host Arrow batches out: every host->device and device->host copy is inside the timed
    region.  Input batches are 4 Mi rows (the operator coalesces smaller ones the same way);
    `pinned`: the input buffers are page-locked, as after a scan into igl_host_alloc_pinned
    memory; otherwise plain pageable memory."""
    import pyarrow as pa
    import torch
    from igloo_b200.expr import col
    from igloo_b200.operators import HashJoinExec, MemoryScanExec
    nb, npr = wl.nb, wl.np
    right, _, left, _ = wl.shard(0, 1)
    host = {}
    for name, c, n in (("rk", right[0], nb), ("rv", right[1], nb), ("lk", left[0], npr), ("lw", left[1], npr)):
        t = torch.empty(n, dtype=torch.int64, pin_memory=pinned)
        t.copy_(c.data[:n])
        host[name] = t
    del right, left
    torch.cuda.empty_cache()
    batch = 1 << 22
    R = pa.table({"k": pa.array(host["rk"].numpy()), "v": pa.array(host["rv"].numpy())})
    L = pa.table({"k": pa.array(host["lk"].numpy()), "w": pa.array(host["lw"].numpy())})
    times, rows, d2h = [], 0, 0
    for s in range(steps + 1):
        op = HashJoinExec(MemoryScanExec(L, batch), MemoryScanExec(R, batch), [col("k")], [col("k")], 0, ctx=ctx)
        t0 = time.perf_counter()
        rows = d2h = 0
        last = None
        for b in op.execute():
            rows += b.num_rows
            # bytes that crossed PCIe: two columns of a batch may share one host buffer (an inner equi-join's
            # right key equals its left key bit for bit, so the operator copies it down once)
            seen = set()
            for c in b.columns:
                for buf in c.buffers():
                    if buf is not None and buf.address not in seen:
                        seen.add(buf.address)
                        d2h += min(buf.size, (b.num_rows * c.type.bit_width + 7) // 8 if buf.size > b.num_rows else buf.size)
            last = b
        dt = time.perf_counter() - t0
        if s > 0:
            times.append(dt)
        assert rows == npr, (rows, npr)
        assert last.column(0).slice(0, 1000).equals(last.column(2).slice(0, 1000))   # what reached the host: k == right_k
        del op, last
    dt = statistics.median(times)
    return {"value": (nb + npr) / dt, "unit": "rows/s", "h2d_bytes_per_step": 16 * (nb + npr), "d2h_bytes_per_step": d2h,
            "ms_per_step": dt * 1e3, "host_buffers": "pinned" if pinned else "pageable",
            "path": "HashJoinExec over igl_op_hash_join_create / igl_op_push_build / igl_op_finish_build / igl_op_push / igl_op_pull: "
                    f"host Arrow batches of {batch} rows in, host Arrow batches out"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--scale", type=int, default=1, help="divide the workload by this power of two (development only)")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"],
                    help="multi-GPU row exchange: probe kernel reading the peers' staging buffers over NVLink (p2p) or NCCL send/recv first")
    ap.add_argument("--skip-fp", action="store_true")
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--skip-e2e", action="store_true", help="development only")
    ap.add_argument("--ref-build-bits", type=int, default=21)
    ap.add_argument("--ref-probe-rows", type=int, default=15_625_000)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
