#!/bin/bash
# tools/gpu_evidence.sh TAG [full]: one GPU-box pass that produces everything profiles/ is built from (run under gpurun,
# every step under its own timeout).  The .ncu-rep files stay in /tmp on the box -- only their raw CSV pages come back
# (gpurun copies at most 64 MiB of gpurun_out/):
#   gpurun_out/TAG_bench_ours.json / TAG_bench_reference.json   the two arms, default flags
#   gpurun_out/TAG_launches.csv                                 ncu launch list of the default bench command (graph mode)
#   gpurun_out/TAG_full_raw.csv                                 ncu --set full of one eager zju_512 step (libghr kernels)
#   gpurun_out/TAG_large_fwd_raw.csv / TAG_large_sort_raw.csv   ncu --set full of the 3 M-Gaussian frame / the 2^24-key sort
#   gpurun_out/TAG_aux.json                                     tools/aux_bench.py
# with "full" as the second argument the GPU test-suite runs first (alone-in-a-fresh-process check of the last file too).
set -u
TAG="${1:-r02}"
OUT=gpurun_out
mkdir -p "$OUT"
if [ "${2:-}" = "full" ]; then
  timeout 120 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > "$OUT/${TAG}_pytest_noalloc_alone.log" 2>&1
  echo "no-alloc alone: exit $?"; tail -2 "$OUT/${TAG}_pytest_noalloc_alone.log"
  timeout 900 python -m pytest tests -m gpu -x -q > "$OUT/${TAG}_pytest_gpu.log" 2>&1
  echo "pytest -m gpu: exit $?"; tail -4 "$OUT/${TAG}_pytest_gpu.log"
fi
timeout 400 python bench.py --impl reference --steps 20 --warmup 5 > "$OUT/${TAG}_bench_reference.json" 2> "$OUT/${TAG}_bench_reference.err"
echo "bench reference: exit $?"; tail -c 600 "$OUT/${TAG}_bench_reference.json"
timeout 400 python bench.py > "$OUT/${TAG}_bench_ours.json" 2> "$OUT/${TAG}_bench_ours.err"
rc=$?; echo "bench ours: exit $rc"; tail -c 3000 "$OUT/${TAG}_bench_ours.json"
[ $rc -ne 0 ] && tail -20 "$OUT/${TAG}_bench_ours.err"
if [ $rc -eq 0 ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file "$OUT/${TAG}_launches.csv" \
    python bench.py --steps 4 --warmup 3 --no-extras --no-cpu-baseline > "$OUT/${TAG}_launches.log" 2>&1
  echo "ncu launch list: exit $?"
  timeout 900 ncu --set full --clock-control none -k regex:'blend_|onesweep|emit_keys|tile_|lbs_|sort_histogram|adam_|nearest_vertex|deform_project|project_deform' --launch-skip 60 -c 36 -f \
    -o /tmp/${TAG}_full python bench.py --mode eager --eager-chain fused --batch-views 1 --steps 2 --warmup 3 --no-extras --no-cpu-baseline \
    > "$OUT/${TAG}_full.log" 2>&1
  echo "ncu full: exit $?"
  ncu -i /tmp/${TAG}_full.ncu-rep --page raw --csv > "$OUT/${TAG}_full_raw.csv" 2>/dev/null
  timeout 300 ncu --set full --clock-control none -k regex:'onesweep|preprocess_kernel|blend_forward_wide|emit_keys|sort_histogram|tile_ranges|tile_order' -c 12 -f \
    -o /tmp/${TAG}_large_fwd python tools/large_probe.py 1 > "$OUT/${TAG}_large_fwd.log" 2>&1
  echo "ncu large fwd: exit $?"
  ncu -i /tmp/${TAG}_large_fwd.ncu-rep --page raw --csv > "$OUT/${TAG}_large_fwd_raw.csv" 2>/dev/null
  timeout 300 ncu --set full --clock-control none -k regex:'onesweep|sort_histogram' --launch-skip 18 -c 6 -f \
    -o /tmp/${TAG}_large_sort python tools/sort_items_sweep.py 16777216 > "$OUT/${TAG}_large_sort.log" 2>&1
  echo "ncu large sort: exit $?"
  ncu -i /tmp/${TAG}_large_sort.ncu-rep --page raw --csv > "$OUT/${TAG}_large_sort_raw.csv" 2>/dev/null
  timeout 600 python tools/aux_bench.py > "$OUT/${TAG}_aux.json" 2> "$OUT/${TAG}_aux.err"
  echo "aux: exit $?"
fi
du -sh "$OUT"
