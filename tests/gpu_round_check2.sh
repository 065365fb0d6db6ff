mkdir -p gpurun_out/v12
timeout 1500 python -m pytest tests -m gpu -q -rs > gpurun_out/v12/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/v12/pytest_gpu.log
tail -4 gpurun_out/v12/pytest_gpu.log
timeout 900 python bench.py --steps 100 --warmup 5 --no-cpu > gpurun_out/v12/bench_50M.json 2> gpurun_out/v12/bench_50M.err
python tests/bench_summary.py gpurun_out/v12/bench_50M.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/v12/launches_50M.csv python bench.py --steps 12 --warmup 3 --no-cpu --no-e2e > gpurun_out/v12/ncu_launch.log 2>&1
