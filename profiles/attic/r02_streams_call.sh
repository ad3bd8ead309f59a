#!/bin/bash
# one gpurun call: full GPU test suite, streams sweep on C2 / C3, bench line + ncu launch list at the best setting
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python profiles/attic/streams_sweep.py --best-file gpurun_out/best_streams.txt > gpurun_out/sweep_c2.log 2>&1; echo "sweep rc=$?"
cat gpurun_out/sweep_c2.log
timeout 300 python profiles/attic/streams_sweep.py --workload c3_tanks_300x1920x1080_ransac_adaptive --steps 10 --settings 1,2,3,4,6,1,2,3 > gpurun_out/sweep_c3.log 2>&1
cat gpurun_out/sweep_c3.log
B=$(cat gpurun_out/best_streams.txt 2>/dev/null || echo 1)
timeout 600 python bench.py --steps 20 --warmup 3 --streams $B > gpurun_out/bench_streams.json 2> gpurun_out/bench_streams.err; echo "bench rc=$?"
cat gpurun_out/bench_streams.json | cut -c1-1500
