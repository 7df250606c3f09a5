run() {
  echo "== items=${1:-auto}"
  if [ -n "$1" ]; then export GHR_SORT_ITEMS=$1; else unset GHR_SORT_ITEMS; fi
  timeout 120 python tools/aux_bench.py --quick --only sort_sweep 2>&1 | grep "ours_ms\|bit_exact" | tr -d '\n '; echo
  timeout 120 python tools/large_probe.py 1 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); s=d['stress_1080p_forward']; print('stress', s['fps'], s['stages']['sort']['us'], 'sort2p24', d['sort_2p24']['ms'])"
}
timeout 300 python -m pytest tests/test_sort_knn_gpu.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -2
run ""
run 12
run 16
