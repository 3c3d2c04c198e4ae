"""CUDA-event timing (L2 flushed) of xeng_reorder / xeng_reorder_packets at 352 stands x 192 channels."""
import sys
sys.path.insert(0, "caltech-bifrost-dsp_b200")
import numpy as np
import torch
from xengine import ops

NSTAND, NCHAN = 352, 192
dev = "cuda"
wpc = ops.words_per_chan(NSTAND)
nw = 2 * NCHAN * wpc
vis = torch.randint(-1000, 1000, (nw,), dtype=torch.int32, device=dev)
a2bl, conj = ops.get_order(np.arange(NSTAND * 2, dtype=np.int32).reshape(NSTAND, 2))
da, dc = torch.from_numpy(a2bl).to(dev), torch.from_numpy(conj).to(dev)
ro = torch.zeros(NSTAND * NSTAND * 4 * NCHAN * 2, dtype=torch.int32, device=dev)
pk = torch.zeros(NSTAND * (NSTAND + 1) // 2 * NCHAN * 8, dtype=torch.int32, device=dev)
flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn):
    ts = []
    for i in range(12):
        flush.add_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts[2:]))


print("reorder full ms", timeit(lambda: ops.reorder(vis, ro, da, dc, NSTAND, 2, NCHAN, wpc)))
print("reorder COR payloads ms", timeit(lambda: ops.reorder_packets(vis, pk, da, dc, NSTAND, 2, NCHAN, wpc, ops.PAYLOAD_COR)))
print("reorder PY payloads ms", timeit(lambda: ops.reorder_packets(vis, pk, da, dc, NSTAND, 2, NCHAN, wpc, ops.PAYLOAD_PY)))
