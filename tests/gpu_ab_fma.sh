mkdir -p gpurun_out/v7
timeout 1500 python -m pytest tests -m gpu -x -q -rs --durations=8 > gpurun_out/v7/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/v7/pytest_gpu.log
tail -25 gpurun_out/v7/pytest_gpu.log
for v in v7 fma v7 fma; do
  cp abtest/$v.so dynearthsol_b200/csrc/libdes_b200.so
  timeout 600 python bench.py --steps 100 --warmup 5 --no-cpu --no-e2e > gpurun_out/v7/ab_$v.$RANDOM.json 2> gpurun_out/v7/ab_$v.err
done
cp abtest/v7.so dynearthsol_b200/csrc/libdes_b200.so
for f in gpurun_out/v7/ab_*.json; do python tests/bench_summary.py $f; done
