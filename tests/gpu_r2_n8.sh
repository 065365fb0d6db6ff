# 8-GPU box: NCCL bit-identity on 4 and 8 ranks (1.05M-tet box), then bench.py --gpus 8 as the
# driver launches it (the 200M-tet box split over 8 GPUs) and --gpus 4
set -x
O=gpurun_out/r2
mkdir -p $O
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -q -rs --timeout 600 -k "nccl" > $O/pytest_nccl_n8.log 2>&1; tail -6 $O/pytest_nccl_n8.log
for N in 8 4; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 50 --warmup 5 --verbose > $O/bench_200M_strong_n$N.json 2> $O/bench_200M_strong_n$N.err; cat $O/bench_200M_strong_n$N.json | cut -c1-400
done
