# phase_changes on device-resident markers: the GPU test, the driver on orift-phase.cfg against the
# reference binary
set -x
O=gpurun_out/phase
mkdir -p $O
timeout 900 python -m pytest tests/test_phase_gpu.py -m gpu -x -q --timeout 600 > $O/pytest_phase.log 2>&1; echo "rc=$?" >> $O/pytest_phase.log
tail -25 $O/pytest_phase.log
timeout 900 python -m pytest tests/test_bridge_gpu.py -m gpu -x -q -s --timeout 800 -k "orift-phase or t3d-remesh" > $O/pytest_bridge_phase.log 2>&1; echo "rc=$?" >> $O/pytest_bridge_phase.log
tail -12 $O/pytest_bridge_phase.log
