python bench.py --profile-mode --steps 10 --warmup 3 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['kernels']
print('step', round(d['ms_per_step'],4), {n: round(v['ms']/v['calls'],4) for n,v in k.items() if n.startswith('stats')}, 'calls', {n: v['calls'] for n,v in k.items() if n.startswith('stats')})"
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
