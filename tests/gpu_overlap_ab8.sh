N=$1
mkdir -p gpurun_out/ovl
for ov in 0 1; do
  DES_OVERLAP=$ov timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$ov bench.py --gpus $N --steps 40 --warmup 5 --no-e2e > gpurun_out/ovl/n${N}_ov$ov.json 2> gpurun_out/ovl/n${N}_ov$ov.err
done
python - <<'PY'
import glob, json
for f in sorted(glob.glob('gpurun_out/ovl/n8_ov*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d['ms_per_step'], 3), round(d['value'] / 1e9, 2), d['clocks']['sm_mhz'])
    except Exception as e:
        print(f, 'ERR', e)
PY
