set -x
mkdir -p gpurun_out/v6
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > gpurun_out/v6/smi.txt
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/v6/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/v6/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/v6/smoke.log 2>&1
timeout 900 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/v6/bench_ref.json 2> gpurun_out/v6/bench_ref.err
timeout 900 python bench.py > gpurun_out/v6/bench_50M.json 2> gpurun_out/v6/bench_50M.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/v6/launches_50M.csv python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/v6/ncu_launch.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k2_|k3_|k_node|k_elem' -s 24 -c 12 -o gpurun_out/v6/full_6M python bench.py --size 6M --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/v6/ncu_full.log 2>&1
ls -la gpurun_out/v6
tail -3 gpurun_out/v6/pytest_gpu.log; cat gpurun_out/v6/bench_50M.json | head -c 1500
