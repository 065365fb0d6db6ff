# Round-2, evidence run on one B200: restart test, the bench lines (driver path, reference arm
# on the named box, our arm), the ordering A/B on a tetgen mesh, ncu at 51.4M tets
set -x
O=gpurun_out/r2
mkdir -p $O
timeout 180 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1 || { echo "SMOKE FAILED rc=$?"; tail -20 $O/smoke.log; exit 1; }
timeout 600 python -m pytest tests -m gpu -q -rs --timeout 400 -k "restart" > $O/pytest_restart.log 2>&1; tail -3 $O/pytest_restart.log
free -g | head -2; nproc; df -h /tmp | tail -1
timeout 900 python bench.py --driver --size 6M --steps 50 > $O/bench_driver_6M.json 2> $O/bench_driver_6M.err; cat $O/bench_driver_6M.json | cut -c1-400
timeout 900 python bench.py --impl reference --steps 10 --warmup 3 > $O/bench_ref.json 2> $O/bench_ref.err; cat $O/bench_ref.json | cut -c1-900
timeout 900 python bench.py > $O/bench_50M.json 2> $O/bench_50M.err; python tests/bench_summary.py $O/bench_50M.json
[ -e abtest/tetgen_1M.npz ] && timeout 600 python tests/order_ab.py abtest/tetgen_1M.npz $O/order_ab_tetgen_1M.json 40 > $O/order_ab.log 2>&1; tail -4 $O/order_ab.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches_50M.csv python tests/ncu_run.py 50M 12 > $O/ncu_launches.log 2>&1; tail -2 $O/ncu_launches.log
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k2_elem_a|k3_node_a|k2_elem_b|k_node_b|k_elem_c|k3_node_c' --launch-skip 25 --launch-count 6 -f -o $O/ncu_50M python tests/ncu_run.py 50M 8 > $O/ncu_full.log 2>&1; tail -3 $O/ncu_full.log
ncu -i $O/ncu_50M.ncu-rep --page raw --csv > $O/ncu_50M_raw.csv 2>/dev/null; ls -la $O/ncu_50M*; [ $(stat -c %s $O/ncu_50M.ncu-rep) -gt 30000000 ] && rm -f $O/ncu_50M.ncu-rep
