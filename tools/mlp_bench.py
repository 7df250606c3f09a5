"""LBS-weight offset network: the tcgen05 kernel against the same layers through torch (cuDNN Conv1d, as the reference
runs them).  One JSON line.  python tools/mlp_bench.py [P]"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mygauhuman_b200.nets import LBSOffsetDecoder  # noqa: E402


def timed(fn, flush, n=30):
    for _ in range(5):
        fn()
    ts = []
    for _ in range(n):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return float(np.median(ts))


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 50_000
    dev = torch.device("cuda")
    torch.manual_seed(0)
    m = LBSOffsetDecoder().to(dev)
    pts = (torch.rand(1, P, 3, device=dev) * 2 - 1)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    out = {"P": P, "flops": 2 * P * (63 * 128 + 2 * 128 * 128 + 191 * 128 + 128 * 24)}
    with torch.no_grad():
        out["ours_forward_us"] = timed(lambda: m(pts), flush)
        def repack():
            m._blob_key = None
            m(pts)
        out["ours_pack_plus_forward_us"] = timed(repack, flush)
        out["torch_default_tf32_us"] = timed(lambda: m.torch_forward(pts), flush)
        torch.backends.cudnn.allow_tf32 = False
        out["torch_fp32_us"] = timed(lambda: m.torch_forward(pts), flush)
        torch.backends.cudnn.allow_tf32 = True
    out["ours_tflops"] = round(out["flops"] / out["ours_forward_us"] / 1e6, 1)
    out["what"] = "median of 30, CUDA events, L2 flushed; torch = embedding + 5 Conv1d + ReLU/cat as the reference module"
    print(json.dumps(out))


if __name__ == "__main__":
    main()
