# multi-GPU check (run under gpurun --gpus N): the peer test, then bench.py exactly as the driver launches it
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_peer.py -m gpu -x -q > gpurun_out/pytest_peer_n$N.log 2>&1; tail -2 gpurun_out/pytest_peer_n$N.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n$N.log 2>&1
echo "rc=$?"; tail -1 gpurun_out/bench_n$N.log | python scripts/bench_digest.py || tail -30 gpurun_out/bench_n$N.log
