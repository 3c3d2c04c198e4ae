"""Worker of tests/test_gpu_multidevice.py::test_peer_exchanges_two_ranks (launched under torchrun, one rank per GPU):
the exchanges of xengine/sharding.py that need NCCL + symmetric memory -- PeerStream (SM stores and copy engines),
PeerGather in one burst (barrier) and in slices closed by a sequence word.  Exits non-zero on any mismatch."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "caltech-bifrost-dsp_b200"))
from xengine import sharding  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.dup2(2, 1)                                   # NCCL banner off stdout
    dist.init_process_group("nccl", device_id=dev)
    g = torch.Generator(device=dev)
    g.manual_seed(100 + rank)
    n = 48 * 4704 * 2
    # what every rank will send, reproducible on root
    def product(r, seq):
        gg = torch.Generator(device=dev)
        gg.manual_seed(1000 * r + seq)
        return torch.randint(-2 ** 31, 2 ** 31 - 1, (n,), dtype=torch.int32, device=dev, generator=gg)

    for via in ("sm", "ce"):
        ps = sharding.PeerStream(n, torch.int32, dev, dist, nslot=4, root=0, via=via)
        for seq in range(10):
            buf = product(rank, seq)
            if via == "sm" and seq % 2:             # the moving form leaves its source zeroed
                ps.push(buf, seq, move=True)
                torch.cuda.synchronize(dev)
                assert int(buf.abs().sum()) == 0
            else:
                ps.push(buf, seq)
        ps.finish()
        torch.cuda.synchronize(dev)
        dist.barrier()
        if rank == 0:
            for seq in range(6, 10):                # the last nslot steps are still in their slots
                payload, words = ps.slot_on_root(seq)
                assert words.tolist() == [seq] * world, (via, seq, words.tolist())
                for r in range(world):
                    assert torch.equal(payload[r], product(r, seq)), (via, seq, r)

    m = 3 * 1000 * 1000 + 16                        # not a multiple of the slice size
    pg = sharding.PeerGather(m, torch.int32, dev, dist, root=0, chunk_bytes=1 << 20)
    def slab(r, k):
        gg = torch.Generator(device=dev)
        gg.manual_seed(77 * r + k)
        return torch.randint(-2 ** 31, 2 ** 31 - 1, (m,), dtype=torch.int32, device=dev, generator=gg)

    pg.push(slab(rank, 0))                          # one burst + barrier
    torch.cuda.synchronize(dev)
    if rank == 0:
        for r, part in enumerate(pg.parts()):
            assert torch.equal(part, slab(r, 0)), ("burst", r)
    dist.barrier()
    src = slab(rank, 1)
    ranges = pg.slices(7)
    assert ranges[0][0] == 0 and ranges[-1][1] == m and all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
    for lo, hi in ranges:                           # slices, then the sequence word: no rank waits for another
        pg.push_range(src, lo, hi)
    pg.commit(5)
    torch.cuda.synchronize(dev)
    dist.barrier()
    if rank == 0:
        assert pg.arrived().tolist() == [5] * world
        for r, part in enumerate(pg.parts()):
            assert torch.equal(part, slab(r, 1)), ("sliced", r)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
