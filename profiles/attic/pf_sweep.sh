for pf in 0 1 2 3 4 8; do
  MDI_STATS_PREFETCH=$pf python bench.py --profile-mode --steps 10 --warmup 3 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['kernels']
print('pf=$pf', 'step', round(d['ms_per_step'],4), 'stats_main', round(k['stats_main_kernel']['ms']/k['stats_main_kernel']['calls'],4))"
done
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "quantile" 2>&1 | tail -2
