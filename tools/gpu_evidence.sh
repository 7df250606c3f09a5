#!/bin/bash
# tools/gpu_evidence.sh TAG [full]: one GPU-box pass that produces everything profiles/ is built from (run under gpurun):
#   gpurun_out/TAG_bench_ours.json / TAG_bench_reference.json   the two arms, default flags
#   gpurun_out/TAG_launches.csv                                 ncu launch list of the default bench command (graph mode)
#   gpurun_out/TAG_full.ncu-rep + TAG_full_raw.csv              ncu --set full of one eager zju_512 step (libghr kernels)
# with "full" as the second argument the GPU test-suite runs first (alone-in-a-fresh-process check of the last file too).
set -u
TAG="${1:-r02}"
OUT=gpurun_out
mkdir -p "$OUT"
if [ "${2:-}" = "full" ]; then
  python -m pytest tests/test_zz_no_alloc.py -m gpu -q > "$OUT/${TAG}_pytest_noalloc_alone.log" 2>&1
  echo "no-alloc alone: exit $?"; tail -2 "$OUT/${TAG}_pytest_noalloc_alone.log"
  python -m pytest tests -m gpu -x -q > "$OUT/${TAG}_pytest_gpu.log" 2>&1
  echo "pytest -m gpu: exit $?"; tail -4 "$OUT/${TAG}_pytest_gpu.log"
fi
python bench.py --impl reference --steps 20 --warmup 5 > "$OUT/${TAG}_bench_reference.json" 2> "$OUT/${TAG}_bench_reference.err"
echo "bench reference: exit $?"; tail -c 600 "$OUT/${TAG}_bench_reference.json"
python bench.py > "$OUT/${TAG}_bench_ours.json" 2> "$OUT/${TAG}_bench_ours.err"
rc=$?; echo "bench ours: exit $rc"; tail -c 3000 "$OUT/${TAG}_bench_ours.json"
[ $rc -ne 0 ] && tail -20 "$OUT/${TAG}_bench_ours.err"
if [ $rc -eq 0 ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file "$OUT/${TAG}_launches.csv" \
    python bench.py --steps 4 --warmup 3 --no-extras --no-cpu-baseline > "$OUT/${TAG}_launches.log" 2>&1
  echo "ncu launch list: exit $?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'blend_|onesweep|emit_keys|tile_|lbs_|sort_histogram|adam_|nearest_vertex|deform_project|project_deform' --launch-skip 60 -c 36 -f \
    -o "$OUT/${TAG}_full" python bench.py --mode eager --eager-chain fused --batch-views 1 --steps 2 --warmup 3 --no-extras --no-cpu-baseline \
    > "$OUT/${TAG}_full.log" 2>&1
  echo "ncu full: exit $?"
  ncu -i "$OUT/${TAG}_full.ncu-rep" --page raw --csv > "$OUT/${TAG}_full_raw.csv" 2>/dev/null
  wc -l "$OUT/${TAG}_full_raw.csv"
fi
