timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2 | tee /tmp/t.log
grep -q failed /tmp/t.log && exit 1
timeout 120 python tools/large_probe.py 1 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); s=d['stress_1080p_forward']; print('stress', s['fps'], {k:v['us'] for k,v in s['stages'].items()}, 'sort2p24', d['sort_2p24']['ms'])"
timeout 200 python bench.py --no-cpu-baseline --no-extras --steps 60 --warmup 5 2>/dev/null | python -c "
import json,sys
b=json.loads([l for l in sys.stdin if l.startswith('{')][-1])
print('zju', b['value'], b['ms_per_step'], {k:v['us'] for k,v in b['stages'].items() if k in ('sort','emit_keys')})"
