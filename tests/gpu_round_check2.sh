mkdir -p gpurun_out/v9
timeout 1500 python -m pytest tests -m gpu -q -rs > gpurun_out/v9/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/v9/pytest_gpu.log
tail -6 gpurun_out/v9/pytest_gpu.log
for s in 6M 50M; do timeout 900 python tests/avg_timing.py $s 40 > gpurun_out/v9/avg_$s.json 2> gpurun_out/v9/avg_$s.err; cat gpurun_out/v9/avg_$s.json; tail -2 gpurun_out/v9/avg_$s.err; done
