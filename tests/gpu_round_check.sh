# Round check on one B200: smoke, GPU tests,
# both bench arms (the reference first, as the driver does)
set -x
O=gpurun_out/final
mkdir -p $O
timeout 180 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1 || { echo "SMOKE FAILED rc=$?"; tail -20 $O/smoke.log; exit 1; }
cat $O/smoke.log
timeout 1500 python -m pytest tests -m gpu -q -rs --timeout 400 > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -6 $O/pytest_gpu.log
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref.json 2> $O/bench_ref.err; cut -c1-300 $O/bench_ref.json
timeout 900 python bench.py --steps 20 --warmup 5 > $O/bench_50M_k20.json 2> $O/bench_50M_k20.err; python tests/bench_summary.py $O/bench_50M_k20.json
timeout 900 python bench.py > $O/bench_50M.json 2> $O/bench_50M.err; python tests/bench_summary.py $O/bench_50M.json
