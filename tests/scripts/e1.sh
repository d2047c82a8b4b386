python -m pytest tests/test_gpu_shard.py -m gpu -q -x 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29516 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/s4_bench_n8.json 2> gpurun_out/s4_bench_n8.err; echo rc=$?
head -c 300 gpurun_out/s4_bench_n8.json
