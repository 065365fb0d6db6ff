# Round-2 multi-GPU check on N GPUs of one box: bash tests/gpu_r2_n.sh N
#  - NCCL one-process-per-GPU runs bit-identical to one GPU (2 / 4 / 8 ranks as the box allows)
#  - the group handle and the driver over real distinct devices (peer copies)
#  - bench.py --gpus N as the driver launches it: the 200M-tet box split over N GPUs
#    (skipped with a second argument "nobench")
set -x
N=${1:-2}
O=gpurun_out/r2
mkdir -p $O
nvidia-smi topo -m 2>/dev/null | head -12
timeout 1500 python -m pytest tests/test_multigpu.py tests/test_multi_handle_gpu.py tests/test_bridge_gpu.py -m gpu -q -rs --timeout 900 -k "nccl or group_handle or several_devices or morton" > $O/pytest_multigpu_n$N.log 2>&1; tail -8 $O/pytest_multigpu_n$N.log
[ "$2" = "nobench" ] && exit 0
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 50 --warmup 5 --verbose > $O/bench_200M_strong_n$N.json 2> $O/bench_200M_strong_n$N.err; cat $O/bench_200M_strong_n$N.json | cut -c1-700; grep "^# rank" $O/bench_200M_strong_n$N.err | head -8
