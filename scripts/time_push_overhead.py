"""What does shipping the fast dump every step cost the correlator?  One GPU, 384 channels, fused
Corr+CorrAcc+CorrSubsel launches back to back with: nothing else; the per-step event choreography of bench.py but no
copy; the choreography + xeng_peer_push into a LOCAL buffer (SM stores, no NVLink); the same with cudaMemcpyAsync."""
import json
import sys
sys.path.insert(0, "caltech-bifrost-dsp_b200")
import numpy as np
import torch
from xengine import ops

nchan, nstand, ntime, NS = 384, 352, 480, 4
dev = torch.device("cuda")
wpc = ops.words_per_chan(nstand)
gulps = [torch.randint(0, 256, (ntime, nchan, 2 * nstand), dtype=torch.uint8, device=dev) for _ in range(4)]
acc = torch.zeros(2 * nchan * wpc, dtype=torch.int32, device=dev)
a2bl, conj = ops.get_order(np.arange(2 * nstand, dtype=np.int32).reshape(nstand, 2))
rng = np.random.default_rng(1)
st = rng.choice(nstand, 48, replace=False)
sel = np.array([np.ravel_multi_index((s0, s1, p0, p1), a2bl.shape) for i, s0 in enumerate(st) for s1 in st[:i + 1]
                for p0 in range(2) for p1 in range(2)])
xc = ops.XCorr(nstand=nstand, nchan=nchan, ntime_gulp=ntime)
xc.set_subsel(a2bl.reshape(-1)[sel].copy(), conj.reshape(-1)[sel].copy(), 4)
stream, pstream = torch.cuda.Stream(), torch.cuda.Stream()
xc.set_stream(stream)
fast = [torch.zeros((nchan // 4, len(sel), 2), dtype=torch.int32, device=dev) for _ in range(NS)]
dst = torch.zeros((NS,) + tuple(fast[0].shape), dtype=torch.int32, device=dev)
seqw = torch.zeros(NS, dtype=torch.int64, device=dev)
done = [torch.cuda.Event() for _ in range(NS)]
free = [torch.cuda.Event() for _ in range(NS)]
nb = fast[0].numel() * 4


def run(mode, M=100):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        xc.accumulate_subsel(gulps[0], acc, True, fast[0], True)
        e0.record(stream)
        for j in range(M):
            s = j % NS
            if mode != "none":
                stream.wait_event(free[s])
            xc.accumulate_subsel(gulps[j % 4], acc, False, fast[s], True)
            if mode != "none":
                done[s].record(stream)
                with torch.cuda.stream(pstream):
                    pstream.wait_event(done[s])
                    if mode == "sm":
                        ops.peer_push(dst[s], fast[s], nb, seqw[s:], j, pstream)
                    elif mode == "ce":
                        ops.memcpy_async(dst[s], fast[s], nb, pstream)
                    free[s].record(pstream)
        stream.wait_stream(pstream)
        e1.record(stream)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / M


res = {}
for rep in range(2):
    for mode in ("none", "events", "sm", "ce"):
        run(mode, 10)
        res.setdefault(mode, []).append(round(run(mode), 5))
print(json.dumps(res, indent=1))
