"""Probe: torch symmetric memory (peer-mapped buffers over NVLink) under torchrun."""
import os, sys, time
import torch
import torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl")
import torch.distributed._symmetric_memory as symm
t = symm.empty(1 << 20, dtype=torch.uint8, device="cuda")
hdl = symm.rendezvous(t, dist.group.WORLD.group_name)
print(rank, "buffer_ptrs", [hex(p) for p in hdl.buffer_ptrs], "signal_pad_ptrs", [hex(p) for p in hdl.signal_pad_ptrs], "pad size", hdl.signal_pad_size, flush=True)
t.fill_(rank + 1)
hdl.barrier()
peer = hdl.get_buffer((rank + 1) % world, (16,), torch.uint8)
print(rank, "peer first bytes", peer[:4].tolist(), flush=True)
# barrier cost, eager and in a graph
torch.cuda.synchronize()
for _ in range(5):
    hdl.barrier()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(100):
    hdl.barrier()
b.record()
torch.cuda.synchronize()
print(rank, f"eager barrier {a.elapsed_time(b) * 10:.2f} us", flush=True)
try:
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for _ in range(100):
                hdl.barrier()
        g.replay()
        torch.cuda.synchronize()
        dist.barrier()
        a.record(st)
        g.replay()
        b.record(st)
        torch.cuda.synchronize()
    print(rank, f"graph barrier {a.elapsed_time(b) * 10:.2f} us", flush=True)
except Exception as e:
    print(rank, "graph capture of barrier failed:", repr(e)[:300], flush=True)
dist.destroy_process_group()
