#!/usr/bin/env python
"""bench.py — hash-join input rows/s (headline) and filter+project HBM GB/s.

Contract (one JSON line on stdout, rank 0):
  metric  : hash_join_input_rows_per_s  — BASELINE.json "hash-join input rows/s"
  workload: C3 — inner join, build 2^27 (k,v) x probe 1e9 (k,w),
            int64 uniform keys, every probe row matches once.  At N>1 the same
            total is range-sharded over the ranks, hash-repartitioned and joined
            locally ("strong" scaling, NCCL/NVLink exchange inside the timed step).
  value   : (build rows + probe rows) * steps / device time, inputs resident in HBM
  e2e     : the same join through host buffers: pinned H2D of both sides and D2H
            of every output batch inside the timed region
  filter_project : the C2 selectivity sweep (1e9 rows) with achieved HBM GB/s
  roofline: dominant kernel = join probe; algorithmic bytes / its CUDA-event time
  cpu_baseline / --impl reference : the reference's algorithm (oracle port,
            printed-string keys, 1024-row batches, one thread) on a bounded sample
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SEED_C2, SEED_C3 = 0x16100002, 0x16100003


# join_probe_flat_kernel at C3 full scale: 21.51 GB read + 32.07 GB written (profiles/ncu_summary_r1h_join_probe_full.txt)
NCU_PROBE_TRAFFIC_BYTES = 53_579_546_000


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


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
                                          "--format=csv,noheader,nounits", "-lms", "100"],
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
# reference arm / cpu baseline: the oracle port of the reference algorithm
# ---------------------------------------------------------------------------------
def cpu_reference_join(build_bits, nprobe):
    from oracle import oracle as orc
    secs, rows = orc.bench_join(SEED_C3, build_bits, nprobe, 0, 1)
    assert rows == nprobe, rows
    return ((1 << build_bits) + nprobe) / secs, secs


def run_reference(args):
    """--impl reference: the reference's CPU hash join (hash_join.rs:100-213 as
    written: printed-string keys into a std map, per-row slices, one output row
    per emit, 1024-row batches) on the host cores.  The Rust crate cannot be
    built here (no cargo), so this is the oracle PORT; its pipeline is one
    sequential stream plus one probe task, so it uses one core."""
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
            "cpu_baseline": {"value": value, "unit": "rows/s", "cores": 1, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------
def run_filter_project(ctx, nrows, steps, warmup, peak):
    import torch
    from igloo_b200.device import FilterProject
    from igloo_b200.expr import Field, col, T_F64, T_I32
    a = ctx.gen(1, SEED_C2, 0, 0, nrows, 1000)
    b = ctx.gen(1, SEED_C2, 1, 0, nrows, 1 << 31)
    x = ctx.gen(3, SEED_C2, 2, 0, nrows)
    fields = [Field("a", T_I32), Field("b", T_I32), Field("x", T_F64)]
    proj = [(col("b") + col("a")).alias("s"), (col("x") * 1.5 + 0.25).alias("y")]
    sweep = []
    out = None
    for thr in (10, 50, 100, 250, 500, 750, 900):
        fp = FilterProject(ctx, fields, col("a") < thr, proj)
        if out is None:
            out = fp.alloc_outputs(nrows)
        for _ in range(warmup):
            fp.launch([a, b, x], nrows, out); kept = fp.finish()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        torch.cuda.synchronize()
        ev[0].record()
        for i in range(steps):
            fp.launch([a, b, x], nrows, out)
            ev[i + 1].record()
        torch.cuda.synchronize()
        kept = fp.finish()
        ms = statistics.median(ev[i].elapsed_time(ev[i + 1]) for i in range(steps))
        sel = kept / nrows
        gb = nrows * (16 + 12 * sel) / 1e9
        sweep.append({"threshold": thr, "selectivity": round(sel, 4), "ms": round(ms, 4), "algorithmic_gb": round(gb, 3),
                      "gb_per_s": round(gb / ms * 1e3, 1), "frac_of_measured_peak": round(gb / ms * 1e3 / peak, 4),
                      "rows_per_s": nrows / ms * 1e3})
        del fp
    del a, b, x, out
    torch.cuda.empty_cache()
    return {"workload": f"C2: {nrows} rows (a:i32,b:i32,x:f64), SELECT b+a, x*1.5+0.25 WHERE a<T; bytes = N*(16+12*s)",
            "rows": nrows, "sweep": sweep}


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
    from igloo_b200.expr import Field, col, T_I64
    peak, peak_src = load_peaks()
    ctx = Context(local)
    scale = args.scale
    bits = 27 - max(0, scale.bit_length() - 1) if scale > 1 else 27
    nb_total = 1 << bits
    np_total = 1_000_000_000 // scale
    lf = [Field("k", T_I64), Field("w", T_I64)]
    rf = [Field("k", T_I64), Field("v", T_I64)]

    if world > 1:
        from igloo_b200.shuffle import PartitionedJoin, init_comm
        init_comm(ctx, rank, world)
        pj = PartitionedJoin(ctx, rank, world, lf, rf, 0, 0, mode=args.exchange)
        # row-range shard of both sides, as after a parallel scan
        b0, b1 = nb_total * rank // world, nb_total * (rank + 1) // world
        p0, p1 = np_total * rank // world, np_total * (rank + 1) // world
        rk = ctx.gen(4, 0, 0, b0, b1 - b0, bits=bits); rv = ctx.gen(2, SEED_C3, 1, b0, b1 - b0)
        lk = ctx.gen(2, SEED_C3, 2, p0, p1 - p0, m=nb_total); lw = ctx.gen(2, SEED_C3, 3, p0, p1 - p0)
        pj.prepare([rk, rv], b1 - b0, [lk, lw], p1 - p0)

        def step():
            return pj.run()
        rows_per_step = nb_total + np_total
    else:
        rk = ctx.gen(4, 0, 0, 0, nb_total, bits=bits); rv = ctx.gen(2, SEED_C3, 1, 0, nb_total)
        lk = ctx.gen(2, SEED_C3, 2, 0, np_total, m=nb_total); lw = ctx.gen(2, SEED_C3, 3, 0, np_total)
        join = HashJoin(ctx, 0, lf, rf, col("k"), col("k"))
        join._right_cols = [rk, rv]
        out = join.alloc_outputs(np_total, [lk, lw])
        probe_ms = []

        def step():
            join.build([rk, rv], nb_total)
            _, n = join.probe([lk, lw], np_total, out=out, capacity=np_total)
            probe_ms.append(join.phase_ms())   # CUDA events recorded by the library on the launching stream
            return n
        rows_per_step = nb_total + np_total

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
    if world > 1:
        t = torch.tensor([ms_total], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
        cnt = torch.tensor([n_out], device=f"cuda:{local}", dtype=torch.int64)
        dist.all_reduce(cnt)
        n_out = int(cnt.item())
    assert n_out == np_total, (n_out, np_total)   # every probe row matches exactly once
    # ---- parity at the timed size, outside the timed region: the multiset digest of what the join
    # wrote (all ranks combined) against the digest of (k, w, k, v(k)) recomputed from the generator
    # definition alone — no hash table, no exchange involved (SURVEY.md §8d "parity check at scale")
    if world > 1:
        n_local, got_cols = pj.out_rows, pj.out
    else:
        n_local, got_cols = n_out, out
    got = ctx.digest(got_cols, n_local)
    n_probe_local = lk.length
    expect_v = ctx.gen(5, SEED_C3, 1, 0, n_probe_local, bits=bits, keys=lk)
    want = ctx.digest([lk, lw, lk, expect_v], n_probe_local)
    del expect_v
    if world > 1:
        both = [None] * world
        dist.all_gather_object(both, (got, want))
        M = (1 << 64) - 1
        fold = lambda ds: (sum(d[0] for d in ds) & M, __import__("functools").reduce(lambda a, b: a ^ b, (d[1] for d in ds)), sum(d[2] for d in ds) & M)  # noqa: E731
        got, want = fold([b[0] for b in both]), fold([b[1] for b in both])
    parity = {"digest_equal": got == want, "rows": n_out, "rows_expected": np_total,
              "check": "sum/xor/sum2 of splitmix64 row hashes of the join output (all ranks) == the same digest of (k, w, k, v(perm^-1(k))) regenerated from the data definition"}
    assert parity["digest_equal"], (got, want)
    ms_step = ms_total / args.steps
    value = rows_per_step / (ms_step / 1e3)

    line = {"metric": "hash_join_input_rows_per_s", "value": value, "unit": "rows/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "int64", "data": "synthetic",
            "config": {"workload": f"C3 inner hash join: build 2^{bits} (k,v) x probe {np_total} (k,w), int64 uniform keys, 1 match per probe row",
                       "l2_policy": "inputs (>= 18 GB) and the 4 GiB table are far larger than the 126 MB L2",
                       "parallelism": f"hash-partitioned over {world} GPUs, exchange={args.exchange}" if world > 1 else "1 GPU"},
            "gpu_launches": launches, "clocks": clocks, "parity": parity}
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
        probe_bytes = np_total * 16 + n_out * 32
        # traffic: dram__bytes_read.sum + dram__bytes_write.sum of this kernel at this size (ncu, profiles/README.md)
        traffic = NCU_PROBE_TRAFFIC_BYTES if scale == 1 else None
        line["roofline"] = {"bound": "hbm", "kernel": "join_probe_flat_kernel<3, 1> (Int64 key, dense output)", "achieved": probe_bytes / pm / 1e6,
                            "peak": peak, "unit": "GB/s", "frac": probe_bytes / pm / 1e6 / peak, "traffic": traffic,
                            "peak_source": peak_src, "kernel_ms": pm,
                            "algorithmic_bytes_per_launch": probe_bytes,
                            "phases_ms": {"build_partition": ph[0], "build": ph[1], "probe_partition": ph[2], "probe_kernel": ph[3]},
                            "whole_join": {"algorithmic_gb": (nb_total * 16 + np_total * 16 + n_out * 32) / 1e9,
                                           "gb_per_s": (nb_total * 16 + np_total * 16 + n_out * 32) / ms_step / 1e6,
                                           "frac": (nb_total * 16 + np_total * 16 + n_out * 32) / ms_step / 1e6 / peak}}
        # ---- e2e: host buffers, pinned H2D in, D2H out, per step ----
        if not args.skip_e2e:
            line["e2e"] = run_e2e(ctx, join, rk, rv, lk, lw, out, nb_total, np_total, max(1, min(args.steps, 3)))
            line["e2e"]["all_columns_downloaded"] = run_e2e(ctx, join, rk, rv, lk, lw, out, nb_total, np_total, 1, share_key=False)
        del out
        torch.cuda.empty_cache()
        if not args.skip_fp:
            line["filter_project"] = run_filter_project(ctx, 1_000_000_000 // scale, max(3, args.steps), max(3, args.warmup), peak)
        if not args.skip_cpu:
            v, secs = cpu_reference_join(args.ref_build_bits, args.ref_probe_rows)
            line["cpu_baseline"] = {"value": v, "unit": "rows/s", "cores": 1, "kind": "port",
                                    "sample": f"C3 at reduced scale: build 2^{args.ref_build_bits} x probe {args.ref_probe_rows} rows, "
                                              f"faithful string-key port of hash_join.rs, 1024-row batches, {secs:.1f} s"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_e2e(ctx, join, rk, rv, lk, lw, out, nb, npr, steps, share_key=True):
    """Join with HOST inputs and HOST outputs: the build side and every probe
    chunk come from pinned host memory, every output chunk is copied back.
    Three streams (upload, join, download) over double-buffered chunks, so the
    two PCIe directions and the kernels overlap; each probe call still returns
    its row count to the host before the chunk's download is queued.
    share_key: the right key column of an inner equi-join on non-null int64 keys is bit-identical to
    the left key column, so (as igl_op_pull does for its Arrow batches) it is not downloaded a second
    time — the host-side right_k column is the k buffer."""
    import torch
    from igloo_b200.device import DeviceColumn
    chunk = min(npr, 1 << 25)
    # host copies of the inputs (pinned), made once outside the timed region
    h_rk = torch.empty(nb, dtype=torch.int64, pin_memory=True); h_rk.copy_(rk.data[:nb])
    h_rv = torch.empty(nb, dtype=torch.int64, pin_memory=True); h_rv.copy_(rv.data[:nb])
    h_lk = torch.empty(npr, dtype=torch.int64, pin_memory=True); h_lk.copy_(lk.data[:npr])
    h_lw = torch.empty(npr, dtype=torch.int64, pin_memory=True); h_lw.copy_(lw.data[:npr])
    h_out = [[torch.empty(chunk, dtype=torch.int64, pin_memory=True) for _ in range(4)] for _ in range(2)]
    d_rk = DeviceColumn.empty(3, nb); d_rv = DeviceColumn.empty(3, nb)
    d_in = [[DeviceColumn.empty(3, chunk), DeviceColumn.empty(3, chunk)] for _ in range(2)]
    d_out = [[DeviceColumn(3, chunk, out[i].data[b * (chunk + 16): (b + 1) * (chunk + 16)], None) for i in range(4)] for b in range(2)]
    out_cols = [0, 1, 3] if share_key else [0, 1, 2, 3]
    if share_key:
        for b in range(2):
            h_out[b][2] = h_out[b][0]
    s_up, s_join, s_down = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()
    starts = list(range(0, npr, chunk))
    torch.cuda.synchronize()
    times = []
    h2d = d2h = 0
    for s in range(steps + 1):
        t0 = time.perf_counter()
        h2d = d2h = 0
        in_ready = [torch.cuda.Event(), torch.cuda.Event()]    # upload of the chunk in buffer b has landed
        in_free = [torch.cuda.Event(), torch.cuda.Event()]     # the probe that read buffer b is done
        out_ready = [torch.cuda.Event(), torch.cuda.Event()]   # probe output in buffer b is complete
        out_free = [torch.cuda.Event(), torch.cuda.Event()]    # download of buffer b is done

        def upload(ci):
            b = ci & 1
            c0 = starts[ci]; n = min(chunk, npr - c0)
            with torch.cuda.stream(s_up):
                if ci >= 2:
                    s_up.wait_event(in_free[b])
                d_in[b][0].data[:n].copy_(h_lk[c0:c0 + n], non_blocking=True)
                d_in[b][1].data[:n].copy_(h_lw[c0:c0 + n], non_blocking=True)
                in_ready[b].record(s_up)
            return 16 * n

        with torch.cuda.stream(s_up):
            d_rk.data[:nb].copy_(h_rk, non_blocking=True); d_rv.data[:nb].copy_(h_rv, non_blocking=True)
            built_in = torch.cuda.Event(); built_in.record(s_up)
        h2d += 16 * nb
        h2d += upload(0)
        with torch.cuda.stream(s_join):
            s_join.wait_event(built_in)
            join.build([d_rk, d_rv], nb)
        total = 0
        for ci, c0 in enumerate(starts):
            b = ci & 1
            n = min(chunk, npr - c0)
            if ci + 1 < len(starts):
                h2d += upload(ci + 1)          # the next chunk travels while this one is probed
            with torch.cuda.stream(s_join):
                s_join.wait_event(in_ready[b])
                if ci >= 2:
                    s_join.wait_event(out_free[b])
                _, m = join.probe(d_in[b], n, out=d_out[b], capacity=chunk)
                in_free[b].record(s_join)
                out_ready[b].record(s_join)
            with torch.cuda.stream(s_down):
                s_down.wait_event(out_ready[b])
                for i in out_cols:
                    h_out[b][i][:m].copy_(d_out[b][i].data[:m], non_blocking=True)
                out_free[b].record(s_down)
            d2h += 8 * len(out_cols) * m
            total += m
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if s > 0:
            times.append(dt)
        assert total == npr
    # spot check of what reached the host: the last chunk's left and right keys agree
    lastb = (len(starts) - 1) & 1
    assert torch.equal(h_out[lastb][0][:1000], h_out[lastb][2][:1000])
    join._right_cols = [rk, rv]
    dt = statistics.median(times)
    return {"value": (nb + npr) / dt, "unit": "rows/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
            "ms_per_step": dt * 1e3,
            "path": "pinned host -> cudaMemcpyAsync (upload stream) -> igl_join_build / igl_join_probe per 32 Mi-row chunk "
                    "(join stream) -> cudaMemcpyAsync (download stream) -> pinned host; double-buffered"
                    + ("; right_k is not downloaded: the host right_k column shares the k buffer (identical for an inner "
                       "equi-join on non-null int64 keys), as in igl_op_pull" if share_key else "; all four output columns downloaded")}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scale", type=int, default=1, help="divide the workload by this power of two (development only)")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"],
                    help="multi-GPU row exchange: partition kernel storing into peer HBM over NVLink (p2p) or NCCL all-to-all")
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
