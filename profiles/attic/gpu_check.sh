python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gpu_tests.log
python bench.py --profile-mode --steps 5 --warmup 2 > gpurun_out/bench_pm.log 2>&1
python bench.py --profile-mode --steps 1 --warmup 1 > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"stats_|keep_|emit_|scan_tiles|ransac|project_filter|gather_fit|setup_cam|factor_table" -c 200 --csv --log-file gpurun_out/launches.csv python bench.py --profile-mode --steps 1 --warmup 1 > gpurun_out/ncu_list.log 2>&1
