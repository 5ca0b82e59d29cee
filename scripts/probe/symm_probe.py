"""torchrun probe: does torch symmetric memory (CUDA IPC / fabric handles) work on this box, and how fast is a
copy-engine pull of a peer's buffer (cudaMemcpyAsync peer-to-peer, zero SMs)?"""
import os, sys, time
import torch, torch.distributed as dist
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
rank, world = dist.get_rank(), dist.get_world_size()
import torch.distributed._symmetric_memory as symm
n = 2048 * 10000
try:
    t = symm.empty(n, dtype=torch.float32, device=dev)
    t.fill_(float(rank + 1))
    h = symm.rendezvous(t, dist.group.WORLD)
    print(f"[rank {rank}] rendezvous ok: world {h.world_size} rank {h.rank}", flush=True)
    dst = torch.empty(world, n, dtype=torch.float32, device=dev)
    torch.cuda.synchronize(); dist.barrier()
    peers = [(rank + o) % world for o in range(1, world)]
    bufs = {p: h.get_buffer(p, (n,), torch.float32) for p in peers}
    for rep in range(3):
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for p in peers[:1]:
            dst[p].copy_(bufs[p])
        e1.record(); torch.cuda.synchronize()
        one = e0.elapsed_time(e1)
        dist.barrier()
        # all peers, each copy on its own stream (several copy engines)
        streams = [torch.cuda.Stream(device=dev) for _ in peers]
        e0.record()
        for s, p in zip(streams, peers):
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                dst[p].copy_(bufs[p])
        for s in streams:
            torch.cuda.current_stream().wait_stream(s)
        e1.record(); torch.cuda.synchronize()
        allp = e0.elapsed_time(e1)
        ok = all(float(dst[p][0]) == p + 1 and float(dst[p][-1]) == p + 1 for p in peers)
        if rank == 0:
            print(f"rep {rep}: one peer {n*4/1e6:.0f} MB in {one:.3f} ms = {n*4/one/1e6:.0f} GB/s; {len(peers)} peers concurrently "
                  f"{allp:.3f} ms = {len(peers)*n*4/allp/1e6:.0f} GB/s ingress; data ok {ok}", flush=True)
except Exception as ex:
    print(f"[rank {rank}] symmetric memory FAILED: {type(ex).__name__}: {ex}", flush=True)
dist.barrier()
dist.destroy_process_group()
