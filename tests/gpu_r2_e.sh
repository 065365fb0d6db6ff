# Round-2: Morton storage order + CUDA-Graph step -- full GPU suite, the suite again with
# DES_REORDER=morton, bridge performance with the graph, ordering A/B with the library doing
# the reordering, the 50M box in both orders
set -x
O=gpurun_out/r2
mkdir -p $O
timeout 180 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1 || { echo "SMOKE FAILED rc=$?"; tail -20 $O/smoke.log; exit 1; }
timeout 1500 python -m pytest tests -m gpu -q -rs --timeout 400 > $O/pytest_gpu.log 2>&1; tail -8 $O/pytest_gpu.log
DES_REORDER=morton timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_branches_gpu.py -m gpu -q --timeout 400 > $O/pytest_env_morton.log 2>&1; tail -4 $O/pytest_env_morton.log
timeout 900 python tests/bridge_perf.py $O/bridge > $O/bridge_perf.log 2>&1; tail -3 $O/bridge_perf.log
DES_GRAPH=0 timeout 900 python tests/bridge_perf.py $O/bridge_nograph t3d.cfg orift-full.cfg > $O/bridge_perf_nograph.log 2>&1; tail -3 $O/bridge_perf_nograph.log
[ -e abtest/tetgen_1M.npz ] && timeout 600 python tests/order_ab.py abtest/tetgen_1M.npz $O/order_ab_tetgen_1M.json 40 > $O/order_ab.log 2>&1; tail -5 $O/order_ab.log | cut -c1-260
DES_REORDER=morton timeout 400 python bench.py --no-cpu --no-e2e --steps 30 > $O/bench_50M_morton.json 2> $O/bench_50M_morton.err; python tests/bench_summary.py $O/bench_50M_morton.json
timeout 400 python bench.py --no-cpu --no-e2e --steps 30 > $O/bench_50M_plain.json 2> $O/bench_50M_plain.err; python tests/bench_summary.py $O/bench_50M_plain.json
