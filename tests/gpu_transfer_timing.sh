# remesh field transfer at scale: reference vs device at 100k tets (same inputs, outputs compared),
# device alone at 1M tets, ncu launch list + one full capture of the dominant kernel
set -x
O=gpurun_out/transfer
mkdir -p $O
OMP_NUM_THREADS=16 timeout 600 python tests/transfer_timing.py 100k $O/timing_100k.json > $O/timing_100k.log 2>&1; echo "rc=$?" >> $O/timing_100k.log
grep -E "^\{|rc=|Error|error" $O/timing_100k.log | cut -c1-1500
TT_DEVICE_ONLY=1 OMP_NUM_THREADS=16 timeout 400 python tests/transfer_timing.py 1M $O/timing_1M_device.json > $O/timing_1M.log 2>&1; echo "rc=$?" >> $O/timing_1M.log
grep -E "^\{|rc=|Error|error" $O/timing_1M.log | cut -c1-1500
export TT_DEVICE_ONLY=1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches_transfer_100k.csv python tests/transfer_timing.py 100k > $O/ncu_launches.log 2>&1; echo "rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_tr_acm_samples -c 1 -o $O/prof_acm python tests/transfer_timing.py 100k > $O/ncu_full.log 2>&1; echo "rc=$?"
tail -3 $O/ncu_full.log
ls -la $O
