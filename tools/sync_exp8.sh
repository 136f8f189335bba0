# 8-GPU session: the whole GPU test suite (nothing skipped), then the encoder bench with two
# gradient-sum variants and the BERT graph tensor-parallel (from the v1 graph) and data-parallel.
# Every step runs under its own timeout. Output under gpurun_out/.
N=${1:-8}
timeout 400 python -m pytest tests -m gpu -q -rs > gpurun_out/r02_pytest_${N}gpu.log 2>&1
tail -4 gpurun_out/r02_pytest_${N}gpu.log
run() { name=$1; shift; timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 10 --warmup 3 "$@" 2>gpurun_out/r02_bench_n${N}_$name.err > gpurun_out/r02_bench_n${N}_$name.json; python - "$name" gpurun_out/r02_bench_n${N}_$name.json <<'PY'
import sys, json
try:
    d = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d['ms_per_step'], 2), round(d['value'], 1), d.get('grad_sync_check'), d['clocks']['sm_mhz'], round(d['e2e']['ms_per_step'], 2))
except Exception as e:
    print(sys.argv[1], 'FAILED', e)
PY
}
run enc_peer
FFK_BENCH_GRAD_SYNC=nccl run enc_nccl
FFK_BENCH_SYNC_BLOCKS=0 FFK_BENCH_SYNC_MIN_MB=150 run enc_peer_full_m150
run bert_tp2_dp4 --workload bert --strategy strategies/bert_base_tp2_dp4.v1.json
FFK_BENCH_GRAD_SYNC=nccl run bert_dp8 --workload bert
