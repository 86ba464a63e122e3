"""Probe: torch symmetric memory on this box (run under torchrun with >= 2 ranks)."""
import os, sys, time
import torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
t = symm.empty(1 << 20, dtype=torch.float32, device=dev)
hdl = symm.rendezvous(t, dist.group.WORLD)
print(rank, "buffer_ptrs", [hex(p) for p in hdl.buffer_ptrs], "multicast", hdl.has_multicast_support if hasattr(hdl, "has_multicast_support") else None,
      "mc_ptr", hex(hdl.multicast_ptr) if getattr(hdl, "multicast_ptr", 0) else 0, flush=True)
t.fill_(float(rank))
hdl.barrier(channel=0)
peer = hdl.get_buffer((rank + 1) % world, (1 << 20,), torch.float32)
print(rank, "peer value", peer[:4].tolist(), flush=True)
# peer write
peer[100:104] = 1000.0 + rank
hdl.barrier(channel=0)
print(rank, "after peer write", t[100:104].tolist(), flush=True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(100): hdl.barrier(channel=0)
e1.record(); torch.cuda.synchronize()
print(rank, "barrier", e0.elapsed_time(e1) / 100 * 1e3, "us", flush=True)
big = symm.empty(64 << 20, dtype=torch.uint8, device=dev)
h2 = symm.rendezvous(big, dist.group.WORLD)
pb = h2.get_buffer((rank + 1) % world, (64 << 20,), torch.uint8)
for _ in range(3): pb.copy_(big)
torch.cuda.synchronize(); e0.record()
for _ in range(10): pb.copy_(big)
e1.record(); torch.cuda.synchronize()
print(rank, "peer copy GB/s", 10 * (64 << 20) / (e0.elapsed_time(e1) * 1e-3) / 1e9, flush=True)
dist.barrier(); dist.destroy_process_group()
