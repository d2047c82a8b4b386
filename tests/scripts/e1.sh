python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29515 tests/shard_worker.py 2>&1 | grep "rank 0/\|Error\|error\|Traceback" | cut -c1-200 | head -30
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29516 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/s3_bench_n8.json 2> gpurun_out/s3_bench_n8.err; echo rc=$?
python - <<'P'
import json
d=json.load(open('gpurun_out/s3_bench_n8.json'))
print({k:d[k] for k in ('value','n_gpus','ms_per_step')}, d.get('e2e',{}).get('value'))
print(json.dumps(d['cfg5_sharded'], indent=1))
P
