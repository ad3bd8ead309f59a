#!/bin/bash
# Sweep of the bulk-L2-prefetch distance of keep_lattice_kernel (MDI_KEEP_PREFETCH); the prefetching variant of the
# kernel was measured (slower at every distance) and is not in the tree.
for p in 0 4 8 16 32; do
  MDI_KEEP_PREFETCH=$p python bench.py --profile-mode --steps 20 --warmup 3 > gpurun_out/pm_$p.json
  python -c "
import json; d=json.load(open('gpurun_out/pm_$p.json')); k=d['kernels']; print('prefetch $p', round(d['ms_per_step'],4), 'keep', round(k['keep_lattice_kernel']['ms']/k['keep_lattice_kernel']['calls'],4), 'emit', round(k['emit_points_kernel']['ms']/k['emit_points_kernel']['calls'],4))"
done
