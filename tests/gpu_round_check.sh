# Round check on one B200: GPU tests, smoke, both bench arms (the reference first, as the driver does)
set -x
mkdir -p gpurun_out/final
timeout 1500 python -m pytest tests -m gpu -q -rs > gpurun_out/final/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/final/pytest_gpu.log
tail -5 gpurun_out/final/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final/smoke.log 2>&1; cat gpurun_out/final/smoke.log
timeout 900 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/final/bench_ref.json 2> gpurun_out/final/bench_ref.err
timeout 900 python bench.py > gpurun_out/final/bench_50M.json 2> gpurun_out/final/bench_50M.err
python tests/bench_summary.py gpurun_out/final/bench_50M.json
