mkdir -p gpurun_out/v10
timeout 1500 python -m pytest tests -m gpu -q -rs > gpurun_out/v10/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/v10/pytest_gpu.log
tail -8 gpurun_out/v10/pytest_gpu.log
timeout 900 python bench.py --steps 100 --warmup 5 --no-cpu > gpurun_out/v10/bench_50M.json 2> gpurun_out/v10/bench_50M.err
python tests/bench_summary.py gpurun_out/v10/bench_50M.json
python - <<'PY'
import json
d=json.load(open('gpurun_out/v10/bench_50M.json'))
print(d['ms_per_step'], d['clocks'], d['e2e']['value'] if d.get('e2e') else None)
for k,v in d['roofline']['kernels'].items(): print(k, round(v['avg_ms'],3))
PY
