"""Tuning: onesweep sort time by key count for a forced keys-per-thread (GHR_SORT_ITEMS read once per process).
    GHR_SORT_ITEMS=12 python tools/sort_items_sweep.py [N ...]
Also the ncu target for the sort alone (6 kernels per sort: histogram + 5 passes at 45 bits):
    ncu --set full -k regex:'onesweep|sort_histogram' --launch-skip 18 -c 6 ... python tools/sort_items_sweep.py 16777216"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from mygauhuman_b200 import _lib
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
l = _lib.lib()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
bits = 45
out = []
SIZES = [int(a) for a in sys.argv[1:]] or [3 << 17, 1 << 19, 3 << 18, 1 << 20, 3 << 19, 1 << 21, 3 << 20, 1 << 22, 1 << 23]
for N in SIZES:
    g = torch.Generator(device="cpu").manual_seed(N & 0xFFFF)
    tile = torch.randint(0, 8160, (N,), generator=g, dtype=torch.int64)
    depth = (torch.rand(N, generator=g) * 9.8 + 0.2).view(torch.int32).to(torch.int64)
    keys = ((tile << 32) | depth).to(dev)
    vals = torch.arange(N, dtype=torch.int32, device=dev)
    ka, va = keys.clone(), vals.clone(); kb, vb = torch.empty_like(ka), torch.empty_like(va)
    ws = torch.zeros(max(l.ghr_sort_workspace_bytes(N), 256), dtype=torch.uint8, device=dev)
    sel = torch.zeros(1, dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    ts = []
    for i in range(12):
        ka.copy_(keys); va.copy_(vals); flush.fill_(i)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        _lib.check(l.ghr_sort_pairs_u64(N, bits, ka.data_ptr(), kb.data_ptr(), va.data_ptr(), vb.data_ptr(), ws.data_ptr(),
                                        ws.numel(), sel.data_ptr(), st), "sort")
        b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    out.append((N, round(float(np.median(ts[3:])) * 1e3, 1)))
print(os.environ.get("GHR_SORT_ITEMS", "auto"), out)
