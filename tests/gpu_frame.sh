# des_gpu_download_frame + the bridge's direct frame writer, then the whole GPU suite
set -x
O=gpurun_out/frame
mkdir -p $O
timeout 900 python -m pytest tests/test_frame_gpu.py tests/test_bridge_gpu.py -m gpu -x -q --timeout 600 -k "frame or direct or nomarkers" > $O/pytest_frame.log 2>&1; echo "rc=$?" >> $O/pytest_frame.log
tail -25 $O/pytest_frame.log
timeout 1500 python -m pytest tests -m gpu -q -rs --timeout 600 > $O/pytest_gpu_all.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu_all.log
tail -12 $O/pytest_gpu_all.log
