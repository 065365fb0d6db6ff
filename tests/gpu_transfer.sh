# remesh field transfer on the device: the GPU tests (compute-sanitizer is closed on this pool), the
# remeshing driver with the transfer on the device and through the reference's own functions
set -x
O=gpurun_out/transfer
mkdir -p $O
timeout 900 python -m pytest tests/test_transfer_gpu.py -m gpu -x -q -s --timeout 600 > $O/pytest_transfer.log 2>&1; echo "rc=$?" >> $O/pytest_transfer.log
tail -25 $O/pytest_transfer.log
for mode in dev ref; do
  W=/tmp/tr_$mode; mkdir -p $W
  ( cd $W; if [ $mode = ref ]; then export DES_BRIDGE_TRANSFER=ref; fi
    DES_BRIDGE_TRANSFER_TIMING=1 OMP_NUM_THREADS=16 timeout 600 $GRAFT_REPO_ROOT/bridge/_build/dynearthsol3d.b200 $GRAFT_REPO_ROOT/tests/cfg/t3d-remesh.cfg > run.log 2>&1; echo "rc=$?" >> run.log )
  grep -E "B200\]|Remesh|knn|rc=" $W/run.log > $O/bridge_remesh_$mode.log
  cat $O/bridge_remesh_$mode.log
done
cmp /tmp/tr_dev/t3dr.save.000004 /tmp/tr_ref/t3dr.save.000004 && echo "last frames identical (bytes)" | tee -a $O/bridge_remesh_dev.log
timeout 900 python -m pytest tests/test_bridge_gpu.py -m gpu -x -q --timeout 800 -k "remesh or several" > $O/pytest_bridge_remesh.log 2>&1; echo "rc=$?" >> $O/pytest_bridge_remesh.log
tail -8 $O/pytest_bridge_remesh.log
