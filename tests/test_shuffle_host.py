"""CPU (gloo, world_size 2) test of the multi-GPU join's host logic: the count
exchange and ExchangePlan offsets must route every row to exactly one place, and
joining the repartitioned shards locally must reproduce the global join.  The
device scatter is emulated in numpy with the same dest function as partition.cu."""
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest

from igloo_b200.shuffle import ExchangePlan, RadixExchangePlan

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_exchange_plan_offsets_are_a_partition_of_each_slab():
    counts = [[3, 5, 0], [7, 1, 2], [0, 4, 9]]
    plans = [ExchangePlan.from_counts(counts, r) for r in range(3)]
    for p in range(3):   # destination p: the sources' regions tile [0, total) without gaps or overlap
        regions = sorted((plans[s].dst_row_offset[p], plans[s].dst_row_offset[p] + counts[s][p]) for s in range(3))
        assert regions[0][0] == 0 and regions[-1][1] == plans[p].total_recv
        assert all(a[1] == b[0] for a, b in zip(regions, regions[1:]))
        assert plans[p].recv_counts == [counts[s][p] for s in range(3)]
        assert plans[p].recv_offsets == [plans[s].dst_row_offset[p] for s in range(3)]
    assert [pl.send_counts for pl in plans] == counts


def test_radix_exchange_plan_lays_slabs_out_partition_by_partition():
    rng = np.random.default_rng(3)
    world, parts = 3, 4
    counts = rng.integers(0, 50, (world, world * parts))          # [src][dest * parts + part]
    plans = [RadixExchangePlan.from_counts(counts, r, world, parts) for r in range(world)]
    c = counts.reshape(world, world, parts)
    for d in range(world):
        # walking the destination's slab: partition 0 of source 0, of source 1, ..., then partition 1, ...
        pos = 0
        for part in range(parts):
            for src in range(world):
                assert plans[src].bin_row_offset[d * parts + part] == pos
                pos += int(c[src][d][part])
        assert pos == plans[d].total_recv
    assert [pl.send_counts for pl in plans] == [[int(c[s][d].sum()) for d in range(world)] for s in range(world)]


WORKER = textwrap.dedent('''
    import os, sys, pickle
    import numpy as np
    import torch, torch.distributed as dist
    sys.path.insert(0, os.environ["IGL_ROOT"])
    from igloo_b200.shuffle import ExchangePlan
    from igloo_b200.expr import Field, col, lower, T_I64
    from oracle import oracle as orc

    def mix64(k):
        k = k.astype(np.uint64); C = np.uint64(0xD6E8FEB86659FD93)
        k ^= k >> np.uint64(32); k *= C; k ^= k >> np.uint64(32); k *= C; k ^= k >> np.uint64(32)
        return k
    def dest(keys, n):   # partition.cu part_dest for non-null int64 keys
        return (((mix64(keys) & np.uint64(0xFFFFFFFF)) * np.uint64(n)) >> np.uint64(32)).astype(np.int64)

    dist.init_process_group("gloo", rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
    rank, world = dist.get_rank(), dist.get_world_size()
    bits, nprobe = 12, 20000
    nb = 1 << bits
    b0, b1 = nb * rank // world, nb * (rank + 1) // world
    p0, p1 = nprobe * rank // world, nprobe * (rank + 1) // world
    rk = orc.gen_perm_pow2(b0, b1 - b0, bits); rv = orc.gen_i64_uniform(7, 1, b0, b1 - b0)
    lk = orc.gen_i64_uniform(7, 2, p0, p1 - p0, nb); lw = orc.gen_i64_uniform(7, 3, p0, p1 - p0)
    sides = []
    for keys, pay in ((rk, rv), (lk, lw)):
        d = dest(keys, world)
        hist = torch.tensor(np.bincount(d, minlength=world), dtype=torch.int64)
        allh = [torch.zeros(world, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(allh, hist)
        counts = [h.tolist() for h in allh]
        plan = ExchangePlan.from_counts(counts, rank)
        assert plan.send_counts == hist.tolist()
        # "scatter": every rank publishes what it sends; each receiver places source s at recv_offsets[s]
        outgoing = [(keys[d == p], pay[d == p]) for p in range(world)]
        everything = [None] * world
        dist.all_gather_object(everything, outgoing)
        k_slab = np.zeros(plan.total_recv, np.int64); v_slab = np.zeros(plan.total_recv, np.int64)
        for s in range(world):
            ks, vs = everything[s][rank]
            assert len(ks) == plan.recv_counts[s]
            k_slab[plan.recv_offsets[s]:plan.recv_offsets[s] + len(ks)] = ks
            v_slab[plan.recv_offsets[s]:plan.recv_offsets[s] + len(ks)] = vs
        assert (dest(k_slab, world) == rank).all()
        sides.append((k_slab, v_slab))
    (rks, rvs), (lks, lws) = sides
    L = orc.Table.scan([orc.Column("k", orc.T_I64, lks), orc.Column("w", orc.T_I64, lws)])
    R = orc.Table.scan([orc.Column("k", orc.T_I64, rks), orc.Column("v", orc.T_I64, rvs)])
    key = lower(col("k"), [Field("k", T_I64)])[0]
    J = L.hash_join(R, [key], [key], 0, 1, 1)
    local = [J.nrows] + list(J.digest())
    got = [None] * world
    dist.all_gather_object(got, local)
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


def test_world_size_2_gloo_repartitioned_join_equals_global_join(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, IGL_ROOT=ROOT, MASTER_ADDR="127.0.0.1", MASTER_PORT="29613", WORLD_SIZE="2")
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r)), stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
    assert "SHUFFLE_HOST_OK" in outs[0]


WORKER_RADIX = textwrap.dedent('''
    import os, sys
    import numpy as np
    import torch, torch.distributed as dist
    sys.path.insert(0, os.environ["IGL_ROOT"])
    from igloo_b200.shuffle import RadixExchangePlan
    from igloo_b200.expr import Field, col, lower, T_I64
    from oracle import oracle as orc

    def mix64(k):
        k = k.astype(np.uint64); C = np.uint64(0xD6E8FEB86659FD93)
        k ^= k >> np.uint64(32); k *= C; k ^= k >> np.uint64(32); k *= C; k ^= k >> np.uint64(32)
        return k
    LP = 3                                            # 8 partitions per destination
    def bins(keys, n):                                # partition.cu radix_bin_of for non-null int64 keys
        h = mix64(keys)
        d = (((h & np.uint64(0xFFFFFFFF)) * np.uint64(n)) >> np.uint64(32)).astype(np.int64)
        part = (h >> np.uint64(64 - LP)).astype(np.int64)
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
        dist.all_gather(allh, hist)
        counts = np.stack([h.numpy() for h in allh])
        plan = RadixExchangePlan.from_counts(counts, rank, world, 1 << LP)
        # "scatter": every rank publishes (bin offsets, rows grouped by bin); the receiver places them
        order = np.argsort(q, kind="stable")
        outgoing = (plan.bin_row_offset, q[order], keys[order], pay[order])
        everything = [None] * world
        dist.all_gather_object(everything, outgoing)
        k_slab = np.full(plan.total_recv, -1, np.int64); v_slab = np.zeros(plan.total_recv, np.int64)
        for s in range(world):
            offs, qs, ks, vs = everything[s]
            for b in range(rank << LP, (rank + 1) << LP):       # the bins whose destination is this rank
                sel = qs == b
                n = int(sel.sum())
                assert n == counts[s][b]
                assert (k_slab[offs[b]:offs[b] + n] == -1).all()   # nobody else wrote here
                k_slab[offs[b]:offs[b] + n] = ks[sel]; v_slab[offs[b]:offs[b] + n] = vs[sel]
        assert (k_slab != -1).all() or plan.total_recv == 0
        _, dd, pp = bins(k_slab, world)
        assert (dd == rank).all()
        assert (np.diff(pp) >= 0).all()                             # the slab is ordered by partition: no local pass needed
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
        print("RADIX_SHUFFLE_HOST_OK")
    dist.destroy_process_group()
''')


def test_world_size_2_gloo_fused_exchange_lands_partition_ordered(tmp_path):
    """The exchange fused with the radix partitioning (RadixExchangePlan), scatter emulated in numpy with
    partition.cu's bin function: every row lands exactly once, each slab is ordered by partition, and the
    local joins add up to the global join."""
    script = tmp_path / "worker_radix.py"
    script.write_text(WORKER_RADIX)
    env = dict(os.environ, IGL_ROOT=ROOT, MASTER_ADDR="127.0.0.1", MASTER_PORT="29617", WORLD_SIZE="2")
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r)), stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
    assert "RADIX_SHUFFLE_HOST_OK" in outs[0]
