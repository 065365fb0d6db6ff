# the warp-cooperative breadth-first search of the node transfer: GPU tests, then the timings again
set -x
O=gpurun_out/transfer2
mkdir -p $O
timeout 300 python -m pytest tests/test_transfer_gpu.py -m gpu -x -q -s --timeout 200 > $O/pytest_transfer.log 2>&1; echo "rc=$?" >> $O/pytest_transfer.log
tail -12 $O/pytest_transfer.log
grep -q "rc=0" $O/pytest_transfer.log || exit 1
OMP_NUM_THREADS=16 timeout 300 python tests/transfer_timing.py 100k $O/timing_100k.json > $O/timing_100k.log 2>&1; echo "rc=$?" >> $O/timing_100k.log
grep -E "^\{|rc=|Error|error" $O/timing_100k.log | cut -c1-1600
TT_DEVICE_ONLY=1 OMP_NUM_THREADS=16 timeout 300 python tests/transfer_timing.py 1M $O/timing_1M_device.json > $O/timing_1M.log 2>&1; echo "rc=$?" >> $O/timing_1M.log
grep -E "^\{|rc=|Error|error" $O/timing_1M.log | cut -c1-1200
TT_DEVICE_ONLY=1 timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches_transfer_100k.csv python tests/transfer_timing.py 100k > $O/ncu_launches.log 2>&1; echo "rc=$?"
grep -E "k_tr_node_bfs|k_tr_acm_samples" $O/launches_transfer_100k.csv | cut -c1-60,200-400 | head -6
