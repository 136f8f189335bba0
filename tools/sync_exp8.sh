# 8-GPU session: the whole GPU test suite (nothing skipped), then the encoder bench with the
# gradient-sum variants. Output under gpurun_out/.
N=${1:-8}
python -m pytest tests -m gpu -q -rs > gpurun_out/r02_pytest_${N}gpu.log 2>&1
tail -5 gpurun_out/r02_pytest_${N}gpu.log
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 10 --warmup 3 2>gpurun_out/r02_bench_n${N}_$1.err | tee gpurun_out/r02_bench_n${N}_$1.json | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['ms_per_step'],2), round(d['value'],1), d.get('grad_sync_check'), d['clocks']['sm_mhz'], round(d['e2e']['ms_per_step'],2))"; }
run default
FFK_BENCH_GRAD_SYNC=nccl run nccl
FFK_BENCH_SYNC_BLOCKS=0 FFK_BENCH_SYNC_MIN_MB=200 run peer_full_m200
FFK_BENCH_SYNC_BLOCKS=64 FFK_BENCH_SYNC_MIN_MB=100 run peer_b64_m100
FFK_BENCH_SYNC_BLOCKS=0 FFK_BENCH_SYNC_MIN_MB=4096 run peer_full_once
