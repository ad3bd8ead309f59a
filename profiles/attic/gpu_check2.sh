# new statistics loop first (fails fast), then the full GPU suite, then the per-kernel bench
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "quantile" > gpurun_out/gpu_quant.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gpu_quant.log
python bench.py --profile-mode --steps 10 --warmup 3 > gpurun_out/bench_pm.log 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gpu_tests.log
