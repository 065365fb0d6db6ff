# strong scaling of the 200M-tet box: bash tests/gpu_strong.sh N
N=$1
mkdir -p gpurun_out/s200
if [ "$N" = "8" ]; then timeout 600 python -m pytest tests/test_multigpu.py -m gpu -q -rs > gpurun_out/s200/pytest_multigpu.log 2>&1; tail -3 gpurun_out/s200/pytest_multigpu.log; fi
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --size 200M --strong --steps 50 --warmup 5 --no-e2e --verbose > gpurun_out/s200/n$N.json 2> gpurun_out/s200/n$N.err
tail -c 1500 gpurun_out/s200/n$N.json; grep "^# rank 0" gpurun_out/s200/n$N.err; tail -3 gpurun_out/s200/n$N.err
