timeout 300 python -m pytest tests/test_p2p_gpu.py tests/test_parallel_ops_gpu.py tests/test_tensor_parallel_gpu.py tests/test_dlrm_sharded_gpu.py -m gpu -q -rs > gpurun_out/r02_pytest_8gpu_rerun.log 2>&1
tail -4 gpurun_out/r02_pytest_8gpu_rerun.log
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --steps 10 --warmup 3 --workload bert --strategy strategies/bert_base_tp2_dp4.v1.json 2>gpurun_out/r02_bench_n8_bert_tp2_dp4.err > gpurun_out/r02_bench_n8_bert_tp2_dp4.json
cut -c1-300 gpurun_out/r02_bench_n8_bert_tp2_dp4.json; tail -3 gpurun_out/r02_bench_n8_bert_tp2_dp4.err
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --steps 10 --warmup 3 2>gpurun_out/r02_bench_n8_enc_default.err > gpurun_out/r02_bench_n8_enc_default.json
cut -c1-300 gpurun_out/r02_bench_n8_enc_default.json
