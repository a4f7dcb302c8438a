import os, sys, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(4096, dtype=torch.float32, device=torch.device("cuda", rank))
    h = symm.rendezvous(t, dist.group.WORLD.group_name)
    mc = getattr(h, "multicast_ptr", None)
    print(rank, "symm ok; buffer_ptrs", [hex(p) for p in h.buffer_ptrs], "multicast_ptr", hex(mc) if mc else mc,
          "signal_pad_ptrs", len(h.signal_pad_ptrs), flush=True)
except Exception as e:
    print(rank, "symm FAILED:", type(e).__name__, str(e)[:300], flush=True)
dist.destroy_process_group()
