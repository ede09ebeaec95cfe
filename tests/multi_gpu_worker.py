"""One rank of the multi-GPU join parity check (launched by test_gpu_multi.py under
torch.distributed.run).  C3-shaped data from the device generator, row-range sharded;
the combined multiset digest of all ranks' outputs must equal the digest of
(k, w, k, v(k)) recomputed from the generator definition, for both exchange modes."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from igloo_b200.device import Context  # noqa: E402
from igloo_b200.expr import Field, T_I64  # noqa: E402
from igloo_b200.shuffle import PartitionedJoin, init_comm  # noqa: E402

M64 = (1 << 64) - 1


def combine(d, device):
    parts = [None] * dist.get_world_size()
    dist.all_gather_object(parts, d)
    s = x = s2 = 0
    for a, b, c in parts:
        s = (s + a) & M64; x ^= b; s2 = (s2 + c) & M64
    return s, x, s2


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    ctx = Context(local)
    init_comm(ctx, rank, world)
    seed, bits, nprobe = 0x16100003, int(sys.argv[1]), int(sys.argv[2])
    nb = 1 << bits
    b0, b1 = nb * rank // world, nb * (rank + 1) // world
    p0, p1 = nprobe * rank // world, nprobe * (rank + 1) // world
    rk = ctx.gen(4, 0, 0, b0, b1 - b0, bits=bits); rv = ctx.gen(2, seed, 1, b0, b1 - b0)
    lk = ctx.gen(2, seed, 2, p0, p1 - p0, m=nb); lw = ctx.gen(2, seed, 3, p0, p1 - p0)
    if len(sys.argv) > 3 and sys.argv[3] == "skew":
        # 60 % of the probe rows carry one key: its destination GPU reads far more than an even share
        hot = torch.rand(p1 - p0, device=f"cuda:{local}") < 0.6
        lk.data[: p1 - p0][hot] = 12345 % nb
    expect_v = ctx.gen(5, seed, 1, 0, p1 - p0, bits=bits, keys=lk)
    want = combine(ctx.digest([lk, lw, lk, expect_v], p1 - p0), local)
    f = [Field("k", T_I64), Field("w", T_I64)]
    g = [Field("k", T_I64), Field("v", T_I64)]
    for mode in ("p2p", "nccl"):
        pj = PartitionedJoin(ctx, rank, world, f, g, 0, 0, mode=mode)
        pj.prepare([rk, rv], b1 - b0, [lk, lw], p1 - p0)
        for _ in range(2):          # twice: the slabs and scratch are reused
            n = pj.run()
            got = combine(ctx.digest(pj.out, n), local)
            tot = torch.tensor([n], device=f"cuda:{local}")
            dist.all_reduce(tot)
            assert int(tot.item()) == nprobe, (mode, int(tot.item()), nprobe)
            assert got == want, (mode, got, want)
        st = pj.stats()
        if rank == 0:
            print(f"mode={mode} fused={bool(st['fused_probe'])} log2_parts={st['log2_parts']} OK", flush=True)
        pj.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
