"""Feasibility probe: torch symmetric memory (peer pointers + stream barrier) on this box."""
import os, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(1 << 20, dtype=torch.float32, device=dev)
    t.fill_(-1.0)
    hdl = symm.rendezvous(t, dist.group.WORLD)
    print(rank, "rendezvous ok", type(hdl).__name__, "world", hdl.world_size, "ptrs", [hex(p) for p in hdl.buffer_ptrs][:4], flush=True)
    hdl.barrier()
    peer = hdl.get_buffer((rank + 1) % world, (1 << 20,), torch.float32)
    peer[:16] = float(rank)
    hdl.barrier()
    torch.cuda.synchronize()
    print(rank, "after peer write my buffer holds", t[:2].tolist(), "(expected %d)" % ((rank - 1) % world), flush=True)
    import time
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(200):
        hdl.barrier()
    torch.cuda.synchronize(); print(rank, "barrier us", (time.perf_counter() - t0) / 200 * 1e6, flush=True)
    big = hdl.get_buffer((rank + 1) % world, (1 << 20,), torch.float32)
    src = torch.ones(1 << 20, device=dev)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(100):
        big.copy_(src)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 100
    print(rank, "peer copy 4 MB: %.1f us = %.0f GB/s" % (dt * 1e6, 4.194304e-3 / dt), flush=True)
except Exception as e:
    import traceback; traceback.print_exc()
    print(rank, "SYMM UNAVAILABLE:", repr(e), flush=True)
dist.destroy_process_group()
