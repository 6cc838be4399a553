#!/usr/bin/env python
"""Probe: does torch symmetric memory (peer-mapped buffers over NVLink) come up on this box?
    torchrun --nproc-per-node 2 tools/symm_probe.py"""
import os
import torch
import torch.distributed as dist

rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
import torch.distributed._symmetric_memory as symm_mem
print(rank, "symm_mem attrs:", [a for a in dir(symm_mem) if not a.startswith("_")][:40], flush=True)
t = symm_mem.empty(1 << 20, dtype=torch.uint8, device=dev)
t.zero_()
hdl = symm_mem.rendezvous(t, group=dist.group.WORLD)
print(rank, "rendezvous ok; world", hdl.world_size, "rank", hdl.rank, "ptrs", [hex(p) for p in hdl.buffer_ptrs],
      "signal", [hex(p) for p in hdl.signal_pad_ptrs], "multicast", hex(getattr(hdl, "multicast_ptr", 0) or 0), flush=True)
peer = (rank + 1) % world
buf = hdl.get_buffer(peer, (16,), torch.uint8, 0)
buf.fill_(rank + 1)
torch.cuda.synchronize()
hdl.barrier()
torch.cuda.synchronize()
print(rank, "local after peer write:", t[:4].tolist(), flush=True)
dist.destroy_process_group()
