"""One rank of the multi-GPU join parity check (launched by test_gpu_multi.py under
torch.distributed.run).  Data come from the device generator, row-range sharded (the
workload definitions bench.py times: C3, C4, C5a, C5b shapes at a small scale); the
combined multiset digest of all ranks' outputs must equal the digest of the answer
recomputed from the data definition, for both exchange modes, twice in a row.

usage: multi_gpu_worker.py <workload> <scale> [skew]"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (workload definitions only)
from igloo_b200.device import Context  # noqa: E402
from igloo_b200.shuffle import PartitionedJoin, init_comm  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    ctx = Context(local)
    init_comm(ctx, rank, world)
    wl = bench.WORKLOADS[sys.argv[1]](ctx, int(sys.argv[2]))
    right, nr, left, nl = wl.shard(rank, world)
    if len(sys.argv) > 3 and sys.argv[3] == "skew":
        # 60 % of the probe rows carry one key: its destination GPU reads far more than an even share
        hot = torch.rand(nl, device=f"cuda:{local}") < 0.6
        left[0].data[:nl][hot] = 12345 % wl.nb
    want, want_rows = wl.expected(left, nl, rank, world)
    both = [None] * world
    dist.all_gather_object(both, (want, want_rows))
    want, want_rows = bench.fold_digests([b[0] for b in both]), sum(b[1] for b in both)
    jt = int(os.environ.get("IGL_TEST_JOIN_TYPE", "0"))
    for mode in ("p2p", "nccl"):
        pj = PartitionedJoin(ctx, rank, world, wl.lf, wl.rf, wl.lkey, wl.rkey, join_type=jt, mode=mode)
        pj.prepare(right, nr, left, nl)
        for _ in range(2):          # twice: staging, slabs and scratch are reused
            n = pj.run()
            got = [None] * world
            dist.all_gather_object(got, (ctx.digest(pj.out, n), n))
            tot = sum(g[1] for g in got)
            assert tot == want_rows, (mode, tot, want_rows)
            assert bench.fold_digests([g[0] for g in got]) == want, (mode, got, want)
        st = pj.stats()
        if rank == 0:
            print(f"mode={mode} fused={bool(st['fused_probe'])} log2_parts={st['log2_parts']} hot={'yes' if st['hot_keys'] else 'no'} "
                  f"prefetched={st['prefetched_bytes']} rows={tot} OK", flush=True)
        pj.close()
    # the same join as a FragmentType.Shuffle fragment (exchange.py: the plan node the reference declares
    # in fragment.rs:8-13 and never builds), executed by the worker-side fragment executor
    from igloo_b200.exchange import FragmentType, QueryFragment, ShuffleJoinSpec, execute_fragment
    frag = QueryFragment("shuffle-join", FragmentType.Shuffle, ShuffleJoinSpec(wl.lf, wl.rf, wl.lkey, wl.rkey, jt), f"worker-{rank}")
    pj, n = execute_fragment(frag, ctx, rank, world, (right, nr, left, nl))
    got = [None] * world
    dist.all_gather_object(got, (ctx.digest(pj.out, n), n))
    assert sum(g[1] for g in got) == want_rows and bench.fold_digests([g[0] for g in got]) == want
    pj.close()
    if rank == 0:
        print("shuffle fragment OK", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
