"""Probe: does torch symmetric memory (peer-mapped buffers over NVLink) work on this box?  torchrun --nproc-per-node N"""
import os
import torch
import torch.distributed as dist

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import torch.distributed._symmetric_memory as symm_mem
t = symm_mem.empty(1024, dtype=torch.float32, device="cuda")
hdl = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
t.fill_(float(rank + 1))
hdl.barrier()
peer = hdl.get_buffer((rank + 1) % world, (1024,), torch.float32)
print(rank, "peer value", peer[0].item(), "ptrs", [hex(p) for p in hdl.buffer_ptrs][:3], "signal pads", len(hdl.signal_pad_ptrs), flush=True)
hdl.barrier()
dist.destroy_process_group()
