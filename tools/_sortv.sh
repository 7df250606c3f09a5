timeout 300 python -m pytest tests/test_sort_knn_gpu.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -2 | tee /tmp/t.log
grep -q failed /tmp/t.log && exit 1
for it in 12 16; do GHR_SORT_ITEMS=$it timeout 120 python -m pytest tests/test_sort_knn_gpu.py -x -q 2>&1 | tail -1; GHR_SORT_ITEMS=$it timeout 120 python tools/sort_items_sweep.py 2>&1 | tail -1; done
for it in "" 16; do
  echo "== items=${it:-auto}"
  if [ -n "$it" ]; then export GHR_SORT_ITEMS=$it; else unset GHR_SORT_ITEMS; fi
  timeout 120 python tools/large_probe.py 1 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); s=d['stress_1080p_forward']; print('stress', s['fps'], s['stages']['sort']['us'], 'sort2p24', d['sort_2p24']['ms'])"
done
