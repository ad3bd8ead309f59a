# full GPU suite + bench (both arms), the ncu launch list of the same bench command, one --set full capture of the statistics pass
python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gpu_tests.log
python bench.py --steps 40 > gpurun_out/bench_full.log 2>&1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1
python bench.py --profile-mode --steps 2 --warmup 1 > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"stats_|keep_|emit_|scan_view|ransac|project_filter|gather_fit|setup_cam|factor_table" -c 300 --csv --log-file gpurun_out/launches.csv python bench.py --profile-mode --steps 2 --warmup 1 > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"stats_main_kernel|stats_bracket_kernel|stats_finish_kernel|stats_sample_kernel" --launch-skip 8 -c 4 -o gpurun_out/stats_r01b -f python bench.py --profile-mode --steps 2 --warmup 2 > gpurun_out/ncu_full.log 2>&1
ncu -i gpurun_out/stats_r01b.ncu-rep --page raw --csv > gpurun_out/stats_r01b_raw.csv 2>/dev/null
