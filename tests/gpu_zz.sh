# des_gpu_compute_dt_static: the test alone in a fresh process, then after every other GPU test
set -x
O=gpurun_out/zz
mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 400 > $O/pytest_zz_alone.log 2>&1; echo "rc=$?" >> $O/pytest_zz_alone.log
tail -15 $O/pytest_zz_alone.log
timeout 1200 python -m pytest tests -m gpu -x -q -rs --timeout 400 > $O/pytest_gpu_all.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu_all.log
tail -8 $O/pytest_gpu_all.log
