import os, torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
try:
    import torch.distributed._symmetric_memory as symm_mem
    t = symm_mem.empty(1 << 20, dtype=torch.float32, device=f"cuda:{local}")
    h = symm_mem.rendezvous(t, group=dist.group.WORLD)
    if rank == 0:
        print("SYMM OK", type(h).__name__, "multicast_ptr", hex(h.multicast_ptr), "buffer_ptrs", [hex(p) for p in h.buffer_ptrs][:4],
              "signal_pad", [hex(p) for p in h.signal_pad_ptrs][:2], "has_multicast", getattr(h, "multicast_ptr", 0) != 0)
        print([a for a in dir(h) if not a.startswith("_")])
except Exception as e:
    if rank == 0:
        print("SYMM FAIL", repr(e)[:500])
dist.barrier()
dist.destroy_process_group()
