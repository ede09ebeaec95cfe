"""CPU tests of the multi-GPU join's host logic: igl_shuffle_plan (the pure-host planner
igl_join_shuffle_run uses) must give every source a gap-free, 4-row-aligned staging
layout and every destination a run list that reads each of its rows exactly once,
partition by partition; and (gloo, world_size 2) joining what the run lists deliver
must reproduce the global join.  The device partitioning is emulated in numpy with the
same bin function as partition.cu; the runs are "pulled" from the peers' emulated
staging buffers exactly as seg_copy_kernel / the segmented probe do."""
import os
import subprocess
import sys
import textwrap

import numpy as np

from igloo_b200.shuffle import shuffle_plan

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_plan_layout_is_aligned_gap_free_and_partition_major():
    rng = np.random.default_rng(3)
    for world, lp in ((1, 0), (2, 3), (3, 2), (8, 5)):
        parts, bins = 1 << lp, world << lp
        counts = rng.integers(0, 50, (world, bins))
        counts[0, 0] = 0                                          # empty bins are legal
        plans = [shuffle_plan(counts, r, world, lp) for r in range(world)]
        offs = plans[0][0]
        for r in range(world):
            assert np.array_equal(plans[r][0], offs)               # every rank derives the same layout
        for s in range(world):                                     # a source's bins tile its staging buffer in order
            pos = 0
            for b in range(bins):
                assert offs[s][b] == pos and pos % 4 == 0
                pos += (int(counts[s][b]) + 3) // 4 * 4
        for d in range(world):                                     # destination d walks partition by partition, source by source
            _, segs, total = plans[d]
            assert len(segs) == world * parts
            i = 0
            for part in range(parts):
                for src in range(world):
                    b = d * parts + part
                    assert segs[i] == (src, int(offs[src][b]), int(counts[src][b]))
                    i += 1
            assert total == int(counts[:, d * parts:(d + 1) * parts].sum())
        assert sum(p[2] for p in plans) == int(counts.sum())      # every row has exactly one reader


def test_plan_with_a_broadcast_destination_hands_hot_build_rows_to_every_rank():
    """Hot keys: their build rows go to one extra destination that EVERY rank reads (after its own
    bins, partition by partition); everything else is read by exactly one rank."""
    rng = np.random.default_rng(5)
    world, lp = 4, 2
    parts, bins = 1 << lp, (world + 1) << lp
    counts = rng.integers(0, 40, (world, bins))
    plans = [shuffle_plan(counts, r, world, lp, broadcast=1) for r in range(world)]
    offs = plans[0][0]
    bcast = int(counts[:, world * parts:].sum())
    for d in range(world):
        _, segs, total = plans[d]
        assert len(segs) == 2 * world * parts
        i = 0
        for part in range(parts):
            for k, dest in ((0, d), (1, world)):
                for src in range(world):
                    b = dest * parts + part
                    assert segs[i] == (src, int(offs[src][b]), int(counts[src][b]))
                    i += 1
        assert total == int(counts[:, d * parts:(d + 1) * parts].sum()) + bcast
    assert sum(p[2] for p in plans) == int(counts[:, :world * parts].sum()) + world * bcast


def test_plan_properties_on_random_inputs():
    """Whatever the counts: every rank derives the same layout, bins never overlap, every ordinary row is read by
    exactly one rank and every broadcast row by all of them, and a rank's runs come partition by partition."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=60, deadline=None)
    @given(st.integers(1, 8), st.integers(0, 4), st.integers(0, 1), st.integers(0, 2 ** 31), st.sampled_from([0, 1, 7, 300, 70000]))
    def prop(world, lp, broadcast, seed, hi):
        rng = np.random.default_rng(seed)
        parts, ndest = 1 << lp, world + broadcast
        bins = ndest << lp
        counts = rng.integers(0, hi + 1, (world, bins))
        plans = [shuffle_plan(counts, r, world, lp, broadcast=broadcast) for r in range(world)]
        offs = plans[0][0]
        readers = np.zeros((world, bins), np.int64)
        for r, (o, segs, total) in enumerate(plans):
            assert np.array_equal(o, offs)
            assert len(segs) == (1 + broadcast) * world * parts
            assert total == sum(n for _, _, n in segs)
            last_part = -1
            for i, (src, row0, n) in enumerate(segs):
                part = i // ((1 + broadcast) * world)
                assert part >= last_part
                last_part = part
                k = (i // world) % (1 + broadcast)
                b = ((world if k else r) << lp) + part
                assert (src, row0, n) == (i % world, int(offs[i % world][b]), int(counts[i % world][b]))
                readers[src][b] += 1
        for s in range(world):
            ends = offs[s] + (counts[s] + 3) // 4 * 4
            assert (offs[s][1:] == ends[:-1]).all() and offs[s][0] == 0 and (offs[s] % 4 == 0).all()
        assert (readers[:, : world << lp] == 1).all()
        if broadcast:
            assert (readers[:, world << lp:] == world).all()

    prop()


def test_plan_rejects_bad_arguments():
    import pytest
    from igloo_b200._lib import IglooCudaError
    with pytest.raises(IglooCudaError):
        shuffle_plan(np.zeros((2, 4)), 2, 2, 1)                    # rank out of range
    with pytest.raises(IglooCudaError):
        shuffle_plan(-np.ones((2, 4)), 0, 2, 1)                    # negative count
    with pytest.raises(IglooCudaError):
        shuffle_plan(np.zeros((16, 16 << 7)), 0, 16, 7)            # more than 1024 bins


WORKER = textwrap.dedent('''
    import os, sys
    import numpy as np
    import torch, torch.distributed as dist
    sys.path.insert(0, os.environ["IGL_ROOT"])
    from igloo_b200.shuffle import shuffle_plan
    from igloo_b200.expr import Field, col, lower, T_I64
    from oracle import oracle as orc

    def mix64(k):
        k = k.astype(np.uint64); C = np.uint64(0xD6E8FEB86659FD93)
        k ^= k >> np.uint64(32); k *= C; k ^= k >> np.uint64(32); k *= C; k ^= k >> np.uint64(32)
        return k
    LP = int(os.environ["LP"])
    def bins(keys, n):                                # partition.cu radix_bin_of for non-null int64 keys
        h = mix64(keys)
        d = (((h & np.uint64(0xFFFFFFFF)) * np.uint64(n)) >> np.uint64(32)).astype(np.int64)
        part = (h >> np.uint64(64 - LP)).astype(np.int64) if LP else np.zeros(len(keys), np.int64)
        return d * (1 << LP) + part, d, part

    dist.init_process_group("gloo", rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
    rank, world = dist.get_rank(), dist.get_world_size()
    bits, nprobe = 12, 20000
    nb = 1 << bits
    b0, b1 = nb * rank // world, nb * (rank + 1) // world
    p0, p1 = nprobe * rank // world, nprobe * (rank + 1) // world
    rk = orc.gen_perm_pow2(b0, b1 - b0, bits); rv = orc.gen_i64_uniform(7, 1, b0, b1 - b0)
    lk = orc.gen_i64_uniform(7, 2, p0, p1 - p0, nb); lw = orc.gen_i64_uniform(7, 3, p0, p1 - p0)
    Q = world << LP
    sides = []
    for keys, pay in ((rk, rv), (lk, lw)):
        q, d, part = bins(keys, world)
        hist = torch.tensor(np.bincount(q, minlength=Q), dtype=torch.int64)
        allh = [torch.zeros(Q, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(allh, hist)                                # igl_join_shuffle_run: the one ncclAllGather
        counts = np.stack([h.numpy() for h in allh])
        offs, segs, total = shuffle_plan(counts, rank, world, LP)
        # local partition into the staging buffer (radix_scatter_bins_async): bin b at offs[rank][b]
        size = int(offs[rank][-1] + counts[rank][-1]) + 8
        k_stage = np.full(size, -1, np.int64); v_stage = np.zeros(size, np.int64)
        for b in range(Q):
            sel = q == b
            n = int(sel.sum())
            assert n == counts[rank][b]
            o = int(offs[rank][b])
            assert (k_stage[o:o + n] == -1).all()
            k_stage[o:o + n] = keys[sel]; v_stage[o:o + n] = pay[sel]
        staged = [None] * world
        dist.all_gather_object(staged, (k_stage, v_stage))         # stands for the IPC-mapped peer buffers
        # pull this rank's runs in walk order
        ks = [staged[s][0][r0:r0 + n] for s, r0, n in segs]
        vs = [staged[s][1][r0:r0 + n] for s, r0, n in segs]
        k_slab = np.concatenate(ks) if ks else np.zeros(0, np.int64)
        v_slab = np.concatenate(vs) if vs else np.zeros(0, np.int64)
        assert len(k_slab) == total and (k_slab != -1).all()
        _, dd, pp = bins(k_slab, world)
        assert (dd == rank).all()
        assert (np.diff(pp) >= 0).all()                            # ordered by partition: the local join needs no pass of its own
        sides.append((k_slab, v_slab))
    (rks, rvs), (lks, lws) = sides
    L = orc.Table.scan([orc.Column("k", orc.T_I64, lks), orc.Column("w", orc.T_I64, lws)])
    R = orc.Table.scan([orc.Column("k", orc.T_I64, rks), orc.Column("v", orc.T_I64, rvs)])
    key = lower(col("k"), [Field("k", T_I64)])[0]
    J = L.hash_join(R, [key], [key], 0, 1, 1)
    got = [None] * world
    dist.all_gather_object(got, [J.nrows] + list(J.digest()))
    if rank == 0:
        ark = orc.gen_perm_pow2(0, nb, bits); arv = orc.gen_i64_uniform(7, 1, 0, nb)
        alk = orc.gen_i64_uniform(7, 2, 0, nprobe, nb); alw = orc.gen_i64_uniform(7, 3, 0, nprobe)
        GL = orc.Table.scan([orc.Column("k", orc.T_I64, alk), orc.Column("w", orc.T_I64, alw)])
        GR = orc.Table.scan([orc.Column("k", orc.T_I64, ark), orc.Column("v", orc.T_I64, arv)])
        G = GL.hash_join(GR, [key], [key], 0, 1, 1)
        M = (1 << 64) - 1
        assert sum(g[0] for g in got) == G.nrows == nprobe
        assert (sum(g[1] for g in got) & M, got[0][2] ^ got[1][2], sum(g[3] for g in got) & M) == G.digest()
        print("SHUFFLE_HOST_OK")
    dist.destroy_process_group()
''')


def _run_world2(tmp_path, lp, port):
    script = tmp_path / f"worker_lp{lp}.py"
    script.write_text(WORKER)
    env = dict(os.environ, IGL_ROOT=ROOT, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), WORLD_SIZE="2", LP=str(lp))
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r)), stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
    assert "SHUFFLE_HOST_OK" in outs[0]


def test_world_size_2_gloo_destination_only_split_equals_global_join(tmp_path):
    _run_world2(tmp_path, 0, 29613)


def test_world_size_2_gloo_partitioned_pull_lands_partition_ordered(tmp_path):
    _run_world2(tmp_path, 3, 29617)
