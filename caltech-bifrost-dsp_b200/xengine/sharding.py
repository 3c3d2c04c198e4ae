"""Channel-slab sharding of the X-engine across GPUs (one process per GPU).

Every product of the path (visibility, sub-selected visibility, beam sample, beam power) is per frequency channel,
and CorrSubsel's channel sums stay inside a slab, so ranks never exchange data on the hot path.  The reference
deploys the same way: one independent pipeline process per sub-band (pipeline/scripts/lwa352-start-pipeline.sh:1-31;
header keys as written by capture_block.py:264-280).  The only communication is the gather of finished products to
rank 0, which stands where the reference sends UDP packets to a central receiver.
"""
CHAN_BW_HZ = 23925.78125          # 196 MHz / 8192 (capture_block.py:165)


def slab(rank, world_size, system_nchan):
    """Contiguous channel slab of ``rank``: (chan0, nchan).  ``system_nchan`` must divide evenly."""
    if world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError("bad rank %r / world size %r" % (rank, world_size))
    if system_nchan % world_size:
        raise ValueError("%d channels do not split evenly over %d ranks" % (system_nchan, world_size))
    nchan = system_nchan // world_size
    return rank * nchan, nchan


def slab_header(base_header, rank, world_size, system_nchan, band_chan0=0):
    """Sequence header of rank's slab: ``chan0``, ``nchan``, ``system_nchan``, ``sfreq``, ``bw_hz``, ``pipeline_id``."""
    chan0, nchan = slab(rank, world_size, system_nchan)
    h = dict(base_header)
    h.update({'chan0': band_chan0 + chan0, 'nchan': nchan, 'system_nchan': system_nchan,
              'sfreq': (band_chan0 + chan0) * CHAN_BW_HZ, 'bw_hz': nchan * CHAN_BW_HZ, 'pipeline_id': rank})
    return h


def slab_of(array, rank, world_size, axis=0):
    """This rank's channels of a full-band array (beam weights [chan, beam, input], calibration tables, ...)."""
    chan0, nchan = slab(rank, world_size, array.shape[axis])
    index = [slice(None)] * array.ndim
    index[axis] = slice(chan0, chan0 + nchan)
    return array[tuple(index)]


def _ce_copy(dst, src):
    """dst <- src on the current torch stream through the COPY ENGINES (xeng_memcpy_async).  A tensor.copy_() onto a
    peer buffer that symmetric memory mapped into this process is, to torch, a same-device copy: an SM kernel, which
    has to share the SMs with the persistent correlator (r02t: a 766 MB dump cost 6.4 of its 7.0 ms even 'overlapped')."""
    import torch
    from . import ops
    assert dst.is_contiguous() and src.is_contiguous() and dst.numel() * dst.element_size() == src.numel() * src.element_size()
    ops.memcpy_async(dst, src, dst.numel() * dst.element_size(), torch.cuda.current_stream(dst.device))


def gather_to_root(tensor, dist, root=0, out=None):
    """Gather every rank's product (same shape on all ranks) to ``root``; returns the list ordered by rank (i.e. by
    channel slab) on root, None elsewhere.  ``dist`` is torch.distributed (NCCL over NVLink on GPUs, gloo in the
    CPU tests).  ``out``: optional pre-allocated receive list on root, ``out[root]`` may alias ``tensor``."""
    world = dist.get_world_size()
    if world == 1:
        return [tensor]
    if dist.get_rank() == root:
        if out is None:
            out = [tensor if r == root else tensor.new_empty(tensor.shape) for r in range(world)]
        dist.gather(tensor, out, dst=root)
        return out
    dist.gather(tensor, None, dst=root)
    return None


class PeerGather(object):
    """Gather of equal-sized per-rank device buffers to ``root`` by peer writes over NVLink: ``root`` owns a
    [world][n] buffer in torch symmetric memory, every other rank maps it and stores its slab with one
    device-to-device copy (copy engines, no SMs -- so it runs beside a persistent kernel that occupies every SM,
    where NCCL's send/recv kernels would have to wait for a slot), then all ranks meet at a symmetric-memory
    barrier on the same stream.  Same result as ``gather_to_root``; needs the NCCL backend and one node."""

    def __init__(self, n, dtype, device, dist, root=0, chunk_bytes=16 << 20):
        import torch.distributed._symmetric_memory as symm_mem
        self.dist, self.root, self.n = dist, root, n
        self.chunk_bytes = chunk_bytes
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.buf = symm_mem.empty(self.world * n, dtype=dtype, device=device)
        self.hdl = symm_mem.rendezvous(self.buf, dist.group.WORLD.group_name)
        self.root_buf = self.hdl.get_buffer(root, (self.world * n,), dtype)
        self.slot = self.root_buf[self.rank * n:(self.rank + 1) * n]
        # one sequence word per rank on root: a producer that ends a dump with commit(seq) never waits for its peers
        import torch
        self.seqbuf = symm_mem.empty(self.world, dtype=torch.int64, device=device)
        self.seqbuf.fill_(-1)
        self.seq_hdl = symm_mem.rendezvous(self.seqbuf, dist.group.WORLD.group_name)
        self.root_seq = self.seq_hdl.get_buffer(root, (self.world,), torch.int64)
        self.numbers = torch.arange(1 << 12, dtype=torch.int64, device=device)

    def push(self, src):
        """Enqueue on the current stream: this rank's ``src`` into its slot on root, then the barrier.  After the
        stream reaches the end of this call on root, ``parts()`` holds every rank's data."""
        if src.data_ptr() != self.slot.data_ptr():
            # in pieces: a copy engine serves its queue one copy at a time, and the small per-step pushes of
            # PeerStream share the engine of this link -- behind ONE 766 MB copy (7 ms at 8 ranks, rank 0's ingress
            # being shared) every fast-dump push would wait, and two steps later so would the correlator
            flat = src.reshape(-1)
            step = max(1, self.chunk_bytes // flat.element_size())
            for i in range(0, flat.numel(), step):
                _ce_copy(self.slot[i:i + step], flat[i:i + step])
        self.hdl.barrier(channel=0)

    def slices(self, nslices, align=4096):
        """Element ranges [(lo, hi), ...] that cut one rank's buffer into ``nslices`` pieces (multiples of ``align``)."""
        per = -(-self.n // max(1, nslices))
        per = -(-per // align) * align
        return [(lo, min(self.n, lo + per)) for lo in range(0, self.n, per)]

    def push_range(self, src, lo, hi):
        """Enqueue on the current stream: elements [lo, hi) of this rank's ``src`` into its slot on root (copy engines).
        A long dump need not leave in one burst: sent as one slice per step of the NEXT integration it shares the
        link with the per-step products instead of standing in their way; ``commit()`` after the last slice."""
        flat = src.reshape(-1)
        step = max(1, self.chunk_bytes // flat.element_size())
        for i in range(lo, hi, step):
            j = min(hi, i + step)
            _ce_copy(self.slot[i:j], flat[i:j])

    def commit(self, seq=None):
        """Ends a sliced push.  ``seq=None``: a symmetric-memory barrier -- after it (in stream order, on root)
        ``parts()`` holds every rank's data, but every rank's stream waits for the slowest one, with a spinning
        kernel (r02z, 8 ranks: 5 steps at twice their time on the ranks that got there first).  ``seq=k``: this
        rank's sequence word on root becomes ``k`` after its last slice (copy engine, stream order); nobody waits, the
        consumer on root reads ``arrived()``."""
        if seq is None:
            self.hdl.barrier(channel=0)
        else:
            _ce_copy(self.root_seq[self.rank], self.numbers[seq % self.numbers.numel()])

    def arrived(self):
        """Root only: the sequence word of every rank's last committed dump, int64 [world]."""
        return self.seqbuf if self.rank == self.root else None

    def parts(self):
        """Rank-ordered views of the gathered slabs (root only)."""
        if self.rank != self.root:
            return None
        return [self.buf[r * self.n:(r + 1) * self.n] for r in range(self.world)]


class PeerStream(object):
    """The frequent exchanges of the path -- CorrSubsel's fast dump every integration (corr_subsel_block.py:298-305)
    and BeamformSum's powers every gulp (beamform_sum_beams_block.py:245-247), which the reference sends as UDP
    packets to a central receiver -- as a stream of peer writes to ``root``: ``root`` owns ``nslot`` x [world][n]
    slots in torch symmetric memory plus one sequence word per (slot, rank); ``push(src, seq)`` stores this rank's
    product of step ``seq`` into slot ``seq % nslot`` and then its sequence number, both on the current stream, with
    no cross-rank wait: the producer never blocks on the consumer.  A consumer on root that sees ``seq`` in the word
    knows the payload has landed.  ``finish()`` is the one barrier, at the end of a run.

    ``via="sm"`` (default) sends with xeng_peer_push: 128-thread CTAs that fit on an SM beside the persistent
    correlator and store straight into the peer mapping.  ``via="ce"`` sends with two cudaMemcpyAsync copies; those
    share the copy-engine queue with the long dump (PeerGather) and wait behind it (profiles/r02v_bench_n2.json)."""

    def __init__(self, n, dtype, device, dist, nslot=2, root=0, max_seq=1 << 16, via="sm"):
        import torch
        import torch.distributed._symmetric_memory as symm_mem
        if via not in ("sm", "ce"):
            raise ValueError("via must be 'sm' or 'ce'")
        self.dist, self.root, self.n, self.nslot, self.via = dist, root, n, nslot, via
        self.nbytes = n * torch.empty((), dtype=dtype).element_size()
        if via == "sm" and self.nbytes % 16:
            raise ValueError("PeerStream(via='sm') wants a product that is a multiple of 16 bytes")
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.buf = symm_mem.empty(nslot * self.world * n, dtype=dtype, device=device)
        self.seqbuf = symm_mem.empty(nslot * self.world, dtype=torch.int64, device=device)
        self.seqbuf.fill_(-1)
        self.hdl = symm_mem.rendezvous(self.buf, dist.group.WORLD.group_name)
        self.seq_hdl = symm_mem.rendezvous(self.seqbuf, dist.group.WORLD.group_name)
        self.root_buf = self.hdl.get_buffer(root, (nslot, self.world, n), dtype)
        self.root_seq = self.seq_hdl.get_buffer(root, (nslot, self.world), torch.int64)
        self.numbers = torch.arange(max_seq, dtype=torch.int64, device=device)    # sources of the sequence words

    def push(self, src, seq, move=False):
        """``move=True`` (via="sm" only): ``src`` is zero afterwards (xeng_peer_move) -- a fast-dump buffer the
        correlator adds into needs no memset before its next use."""
        slot = seq % self.nslot
        if self.via == "sm":
            import torch
            from . import ops
            src = src.reshape(-1)
            assert src.is_contiguous() and src.numel() == self.n
            (ops.peer_move if move else ops.peer_push)(self.root_buf[slot, self.rank], src, self.nbytes,
                                                       self.root_seq[slot, self.rank], seq,
                                                       torch.cuda.current_stream(src.device))
            return
        if move:
            raise ValueError("PeerStream.push(move=True) needs via='sm'")
        _ce_copy(self.root_buf[slot, self.rank], src.reshape(-1))
        _ce_copy(self.root_seq[slot, self.rank], self.numbers[seq % self.numbers.numel()])

    def finish(self):
        self.hdl.barrier(channel=0)

    def slot_on_root(self, seq):
        """(payload [world][n], sequence words [world]) of step ``seq`` as root holds them (root only)."""
        if self.rank != self.root:
            return None, None
        slot = seq % self.nslot
        return (self.buf.view(self.nslot, self.world, self.n)[slot], self.seqbuf.view(self.nslot, self.world)[slot])


class GatherStream(object):
    """Portable form of ``PeerStream`` (any torch.distributed backend: NCCL without symmetric memory, gloo in the CPU
    tests): every ``push`` is one ``dist.gather`` into root's slot ``seq % nslot``.  Collective: every rank must call
    ``push`` for every step, in the same order."""

    def __init__(self, n, dtype, device, dist, nslot=2, root=0):
        import torch
        self.dist, self.root, self.n, self.nslot = dist, root, n, nslot
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.buf = self.seqbuf = None
        if self.rank == root:
            self.buf = torch.zeros((nslot, self.world, n), dtype=dtype, device=device)
            self.seqbuf = torch.full((nslot, self.world), -1, dtype=torch.int64, device=device)

    def push(self, src, seq):
        slot = seq % self.nslot
        if self.rank == self.root:
            self.dist.gather(src.reshape(-1), list(self.buf[slot].unbind(0)), dst=self.root)
            self.seqbuf[slot].fill_(seq)
        else:
            self.dist.gather(src.reshape(-1), None, dst=self.root)

    def finish(self):
        self.dist.barrier()

    def slot_on_root(self, seq):
        if self.rank != self.root:
            return None, None
        return self.buf[seq % self.nslot], self.seqbuf[seq % self.nslot]


def assemble_visibilities(parts):
    """Rank-ordered list of slab buffers int32 [2][nchan_slab][W] -> full band [2][nchan_total][W] (a view-level
    concatenation along channel; used by rank 0 after gather_to_root)."""
    import torch
    return torch.cat(list(parts), dim=1)
