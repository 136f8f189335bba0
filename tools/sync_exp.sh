run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 6 --warmup 3 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', d['ms_per_step'], d['value'], d['clocks']['sm_mhz'])"; }
FFK_BENCH_GRAD_SYNC=nccl run nccl
run default_b32_m48
FFK_BENCH_SYNC_BLOCKS=8 run b8_m48
FFK_BENCH_SYNC_BLOCKS=64 FFK_BENCH_SYNC_MIN_MB=24 run b64_m24
FFK_BENCH_SYNC_BLOCKS=16 FFK_BENCH_SYNC_MIN_MB=96 run b16_m96
FFK_BENCH_SYNC_BLOCKS=148 FFK_BENCH_SYNC_MIN_MB=1024 run b148_once
