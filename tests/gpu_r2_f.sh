# Round-2: the driver path on the 6.2M- and the 51.4M-tet box (bench.py --driver), new tests
set -x
O=gpurun_out/r2
mkdir -p $O
timeout 180 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1 || { echo "SMOKE FAILED rc=$?"; tail -20 $O/smoke.log; exit 1; }
timeout 600 python -m pytest tests -m gpu -q --timeout 400 -k "driver_mode or non_positive" > $O/pytest_f.log 2>&1; tail -3 $O/pytest_f.log
timeout 600 python bench.py --driver --size 6M > $O/bench_driver_6M.json 2> $O/bench_driver_6M.err; cut -c1-300 $O/bench_driver_6M.json
free -g | head -2
timeout 1200 python bench.py --driver --size 50M > $O/bench_driver_50M.json 2> $O/bench_driver_50M.err; cut -c1-900 $O/bench_driver_50M.json; tail -3 $O/bench_driver_50M.err
