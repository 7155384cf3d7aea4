import os, sys, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import torch.distributed._symmetric_memory as symm
if rank == 0:
    print("symm attrs:", [a for a in dir(symm) if not a.startswith("_")][:60], flush=True)
try:
    t = symm.empty(1 << 20, dtype=torch.float32, device=torch.device("cuda", local))
    t.fill_(rank + 1)
    hdl = symm.rendezvous(t, dist.group.WORLD.group_name)
    if rank == 0:
        print("handle attrs:", [a for a in dir(hdl) if not a.startswith("_")], flush=True)
        print("buffer_ptrs", [hex(p) for p in hdl.buffer_ptrs], "signal_pad_ptrs", [hex(p) for p in hdl.signal_pad_ptrs],
              "multicast_ptr", hex(hdl.multicast_ptr) if getattr(hdl, "multicast_ptr", 0) else None,
              "signal_pad_size", getattr(hdl, "signal_pad_size", None), "buffer_size", getattr(hdl, "buffer_size", None), flush=True)
    hdl.barrier()
    peer = hdl.get_buffer((rank + 1) % world, (16,), torch.float32)
    print(rank, "peer view", peer[:4].tolist(), flush=True)
    torch.cuda.synchronize()
except Exception as e:
    import traceback; traceback.print_exc()
dist.barrier(); dist.destroy_process_group()
