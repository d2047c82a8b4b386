python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 tests/shard_worker.py 2>&1 | grep "rank\|Error\|error\|Traceback" | cut -c1-220 | head -40
