# The A/B switches must all give the reference's results: the GPU suite under each of them
# (interior-first element ranges need the caller's element order: those cases are left out
# under a storage-order switch); then what the in-library partition costs at 6.2M tets
set -x
O=gpurun_out/r2
mkdir -p $O
timeout 180 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1 || { echo "SMOKE FAILED rc=$?"; tail -20 $O/smoke.log; exit 1; }
for v in ${SWITCHES:-DES_REORDER=morton DES_GRAPH=0 DES_DEFER=0 DES_DEFER=5 DES_DEFER=6 DES_REORDER=1}; do
  env $v timeout 1200 python -m pytest tests -m gpu -q --timeout 400 --deselect tests/test_bridge_gpu.py::test_bench_driver_mode -k "not nccl and not (parts_on_one_gpu and True)" > $O/pytest_switch_$v.log 2>&1; echo "$v: $(tail -1 $O/pytest_switch_$v.log)"
done
timeout 900 python tests/multi_partition_timing.py $O/multi_partition_6M.json 6M 4 > $O/multi_partition_6M.log 2>&1; tail -30 $O/multi_partition_6M.log
