mkdir -p gpurun_out/v8
timeout 1500 python -m pytest tests -m gpu -q -rs --durations=5 > gpurun_out/v8/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/v8/pytest_gpu.log
tail -14 gpurun_out/v8/pytest_gpu.log
for s in 6M 50M; do timeout 900 python tests/topology_timing.py $s > gpurun_out/v8/topology_$s.json 2> gpurun_out/v8/topology_$s.err; cat gpurun_out/v8/topology_$s.json; tail -2 gpurun_out/v8/topology_$s.err; done
