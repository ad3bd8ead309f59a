for w in c3_tanks_300x1920x1080_ransac_adaptive c4_4k_500x3840x2160_ransac_adaptive; do
python bench.py --workload $w --profile-mode --steps 5 --warmup 3 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['kernels']
print('$w', 'step', round(d['ms_per_step'],4), {n: round(v['ms']/v['calls'],4) for n,v in k.items() if v['ms']/v['calls']>0.05}, 'views', d.get('views'))"
done
