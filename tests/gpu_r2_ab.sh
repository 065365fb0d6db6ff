# Round-2 A/B on one B200: smoke and one bridge run first (a hang must cost two minutes, not the
# call), GPU tests, then the 50M bench with and without the three deferrals (DES_DEFER bits:
# 1 ddp in E-C, 2 surface scale in E-B, 4 compute_dt on E-A)
set -x
mkdir -p gpurun_out/r2
timeout 180 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2/smoke.log 2>&1 || { echo "SMOKE FAILED rc=$?"; tail -20 gpurun_out/r2/smoke.log; exit 1; }
cat gpurun_out/r2/smoke.log
mkdir -p /tmp/t3d && (cd /tmp/t3d && timeout 180 $GRAFT_REPO_ROOT/bridge/_build/dynearthsol3d.b200 $GRAFT_REPO_ROOT/tests/cfg/t3d.cfg > $GRAFT_REPO_ROOT/gpurun_out/r2/bridge_t3d.log 2>&1) || { echo "BRIDGE FAILED rc=$?"; tail -20 gpurun_out/r2/bridge_t3d.log; exit 1; }
tail -4 gpurun_out/r2/bridge_t3d.log
timeout 1200 python -m pytest tests -m gpu -q -rs -x --timeout 400 ${PYTEST_EXTRA} > gpurun_out/r2/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2/pytest_gpu.log
tail -5 gpurun_out/r2/pytest_gpu.log
for v in ${DES_AB:-7 0}; do
  DES_DEFER=$v timeout 400 python bench.py --no-cpu --no-e2e --steps 50 > gpurun_out/r2/bench_50M_defer$v.json 2> gpurun_out/r2/bench_50M_defer$v.err
  python tests/bench_summary.py gpurun_out/r2/bench_50M_defer$v.json
done
