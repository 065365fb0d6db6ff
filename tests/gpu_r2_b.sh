# Round-2: remaining GPU tests (no -x), bridge performance (reference vs b200 on the named small
# configs), and the deferral A/B bit by bit (DES_DEFER bits: 1 ddp in E-C, 2 surface scale in
# E-B, 4 compute_dt on E-A)
set -x
mkdir -p gpurun_out/r2
timeout 180 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2/smoke.log 2>&1 || { echo "SMOKE FAILED rc=$?"; tail -20 gpurun_out/r2/smoke.log; exit 1; }
timeout 1500 python -m pytest tests -m gpu -q -rs --timeout 400 ${PYTEST_EXTRA} > gpurun_out/r2/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2/pytest_gpu.log
tail -15 gpurun_out/r2/pytest_gpu.log
timeout 900 python tests/bridge_perf.py gpurun_out/r2/bridge > gpurun_out/r2/bridge_perf.log 2>&1; tail -5 gpurun_out/r2/bridge_perf.log
for v in ${DES_AB:-7 6 5 0}; do
  DES_DEFER=$v timeout 400 python bench.py --no-cpu --no-e2e --steps 30 > gpurun_out/r2/bench_50M_defer$v.json 2> gpurun_out/r2/bench_50M_defer$v.err
  python tests/bench_summary.py gpurun_out/r2/bench_50M_defer$v.json
done
