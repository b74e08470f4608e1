"""Probe: does torch's symmetric memory give peer and multicast (NVLS) mappings on this box?  (torchrun)"""
import os, torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
import torch.distributed._symmetric_memory as symm_mem
try:
    t = symm_mem.empty(1 << 20, dtype=torch.float32, device=dev)
    h = symm_mem.rendezvous(t, group=dist.group.WORLD.group_name)
    if rank == 0:
        print("symm ok: world", h.world_size, "buffer_ptrs", [hex(p) for p in h.buffer_ptrs][:3], "multicast_ptr", hex(h.multicast_ptr),
              "signal_pad_ptrs", len(h.signal_pad_ptrs), "buffer_size", h.buffer_size)
except Exception as ex:
    if rank == 0: print("symm failed:", repr(ex)[:500])
dist.destroy_process_group()
