set -x
O=gpurun_out/transfer3
mkdir -p $O
for i in 1 2; do
TT_DEVICE_ONLY=1 OMP_NUM_THREADS=16 timeout 300 python tests/transfer_timing.py 1M $O/timing_1M_device_$i.json > $O/timing_1M_$i.log 2>&1; echo "rc=$?" >> $O/timing_1M_$i.log
grep -E "^\{|rc=|Error|error" $O/timing_1M_$i.log | cut -c1-1200
done
OMP_NUM_THREADS=16 timeout 300 python tests/transfer_timing.py 100k $O/timing_100k.json > $O/timing_100k.log 2>&1
grep -E "^\{" $O/timing_100k.log | cut -c1-1700
