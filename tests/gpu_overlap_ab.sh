# multi-GPU: bit-identity tests, then A/B of the exchange/compute overlap: bash tests/gpu_overlap_ab.sh N
N=$1
mkdir -p gpurun_out/ovl
timeout 600 python -m pytest tests/test_multigpu.py -m gpu -q -rs > gpurun_out/ovl/pytest_multigpu_n$N.log 2>&1; tail -4 gpurun_out/ovl/pytest_multigpu_n$N.log
for ov in 0 1 0 1; do
  DES_OVERLAP=$ov timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$ov bench.py --gpus $N --steps 40 --warmup 5 --no-e2e > gpurun_out/ovl/n${N}_ov$ov.$RANDOM.json 2> gpurun_out/ovl/n${N}_ov$ov.err
done
python - <<'PY'
import glob, json
for f in sorted(glob.glob('gpurun_out/ovl/n*_ov*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d['ms_per_step'], 3), round(d['value'] / 1e9, 2), d['clocks']['sm_mhz'])
    except Exception as e:
        print(f, 'ERR', e)
PY
tail -3 gpurun_out/ovl/n${N}_ov1.err
