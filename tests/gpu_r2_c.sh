# Round-2: all GPU tests, then bridge performance (reference vs b200 on the named small configs)
set -x
mkdir -p gpurun_out/r2
timeout 180 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2/smoke.log 2>&1 || { echo "SMOKE FAILED rc=$?"; tail -20 gpurun_out/r2/smoke.log; exit 1; }
timeout 1500 python -m pytest tests -m gpu -q -rs --timeout 400 ${PYTEST_EXTRA} > gpurun_out/r2/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2/pytest_gpu.log
tail -8 gpurun_out/r2/pytest_gpu.log
timeout 900 python tests/bridge_perf.py gpurun_out/r2/bridge > gpurun_out/r2/bridge_perf.log 2>&1; grep -n "B200\]" gpurun_out/r2/bridge_perf.log | head; tail -3 gpurun_out/r2/bridge_perf.log
