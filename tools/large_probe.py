"""The large-N launches (BASELINE.json config 5) on their own, for ncu: the 3M-Gaussian 1920x1080 forward and the
2^24-key onesweep sort, `reps` times each (bench.py's measure_large does the timing).
    ncu --set full -k regex:'onesweep|preprocess_kernel|blend_forward_wide|emit_keys|histogram|tile_ranges' \
        --launch-skip N --clock-control none -o gpurun_out/large python tools/large_probe.py 1"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import bench
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
peak = bench.load_peaks()[0]
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 1
for _ in range(reps):
    out = bench.measure_large(dev, flush, peak)
import json
print(json.dumps(out))
