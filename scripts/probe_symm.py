import os, torch, torch.distributed as dist
rank=int(os.environ["RANK"]); local=int(os.environ["LOCAL_RANK"]); world=int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(1024, dtype=torch.float32, device=torch.device("cuda", local))
    h = symm.rendezvous(t, dist.group.WORLD.group_name)
    t.fill_(rank + 1)
    h.barrier()
    print(rank, "buffer_ptrs", [hex(p) for p in h.buffer_ptrs], "signal", [hex(p) for p in h.signal_pad_ptrs][:2], "size", h.buffer_size, h.signal_pad_size, flush=True)
    peer = h.get_buffer((rank + 1) % world, (1024,), torch.float32)
    print(rank, "peer value", peer[0].item(), "multicast", hex(h.multicast_ptr) if h.multicast_ptr else None, flush=True)
except Exception as e:
    import traceback; traceback.print_exc()
    print(rank, "SYMM FAILED", repr(e)[:300], flush=True)
print(rank, "p2p", torch.cuda.can_device_access_peer(local, (local+1)%world))
dist.barrier(); dist.destroy_process_group()
