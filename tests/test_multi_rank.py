"""N > 1 host logic on CPU: channel-slab partition, per-rank headers and the gather of dumps to rank 0, exercised
with two gloo ranks (the GPU path uses the same code with the nccl backend, bench.py --gpus N)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from xengine import sharding


def test_slab_partition_and_headers():
    assert [sharding.slab(r, 8, 3072) for r in range(8)] == [(384 * r, 384) for r in range(8)]
    assert sharding.slab(0, 1, 192) == (0, 192)
    with pytest.raises(ValueError):
        sharding.slab(0, 5, 3072)
    with pytest.raises(ValueError):
        sharding.slab(8, 8, 3072)
    base = {'seq0': 0, 'nstand': 352, 'npol': 2, 'fs_hz': 196000000}
    h = sharding.slab_header(base, 3, 8, 3072)
    assert h['chan0'] == 1152 and h['nchan'] == 384 and h['system_nchan'] == 3072 and h['pipeline_id'] == 3
    assert h['sfreq'] == 1152 * 23925.78125 and h['bw_hz'] == 384 * 23925.78125 and h['nstand'] == 352
    w = np.arange(16 * 2 * 3).reshape(16, 2, 3)
    assert np.array_equal(sharding.slab_of(w, 1, 4), w[4:8])
    assert np.array_equal(sharding.slab_of(w.transpose(1, 0, 2), 3, 4, axis=1), w[12:16].transpose(1, 0, 2))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, q):
    os.environ.update({"MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port)})
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        system_nchan, W = 8, 6
        chan0, nchan = sharding.slab(rank, world, system_nchan)
        # a slab "dump": word value encodes (plane, absolute channel, word)
        pl, ch, wd = np.meshgrid(np.arange(2), np.arange(chan0, chan0 + nchan), np.arange(W), indexing="ij")
        mine = torch.from_numpy((pl * 10000 + ch * 100 + wd).astype(np.int32))
        for dump in range(2):                                   # the gather is reused every long dump
            parts = sharding.gather_to_root(mine + dump, dist)
            if rank == 0:
                full = sharding.assemble_visibilities(parts).numpy()
                pl, ch, wd = np.meshgrid(np.arange(2), np.arange(system_nchan), np.arange(W), indexing="ij")
                assert np.array_equal(full, pl * 10000 + ch * 100 + wd + dump)
            else:
                assert parts is None
        # the per-step exchange (fast dump / beam powers pushed to rank 0 every step, two slots)
        gs = sharding.GatherStream(5, torch.int32, "cpu", dist, nslot=2, root=0)
        for step in range(5):
            gs.push(torch.full((5,), 100 * step + rank, dtype=torch.int32), step)
            if rank == 0:
                payload, seqs = gs.slot_on_root(step)
                assert seqs.tolist() == [step] * world
                assert [int(payload[r][0]) for r in range(world)] == [100 * step + r for r in range(world)]
                if step >= 1:                                    # the other slot still holds the previous step
                    prev, pseq = gs.slot_on_root(step - 1)
                    assert pseq.tolist() == [step - 1] * world and int(prev[world - 1][4]) == 100 * (step - 1) + world - 1
        gs.finish()
        # max-over-ranks timing reduction as bench.py does it
        t = torch.tensor([float(rank + 1)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert t.item() == world
        q.put((rank, "ok"))
    except Exception as e:      # pragma: no cover
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_gather_dumps_two_gloo_ranks():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, "ok"), (1, "ok")], res


def test_peer_gather_slices_cover_the_buffer_exactly():
    """PeerGather.slices: the per-step pieces a long dump leaves in -- contiguous, aligned, covering [0, n) once."""
    from xengine import sharding
    for n, k in ((2 * 384 * 249216, 62), (1000, 7), (4096, 1), (5, 3), (3 * 1000 * 1000 + 16, 7)):
        pg = object.__new__(sharding.PeerGather)
        pg.n = n
        r = pg.slices(k)
        assert r[0][0] == 0 and r[-1][1] == n and len(r) <= max(1, k)
        assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
        assert all(lo % 4096 == 0 and hi > lo for lo, hi in r)
