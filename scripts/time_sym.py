"""torchrun script: the C5 sweep point (m x d, default 16384 x 10000) row-sharded over the ranks; prints the unit time
(max over ranks), rank 0's per-kernel times and the stage timeline of the block-symmetric ring.
    torchrun --nproc-per-node 8 scripts/time_sym.py [m] [d] [reps]          (SMI_SHARD_SYM=0: the all-columns path)"""
import json, os, sys
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from smi_experiments_b200.dist import bench_sharded_force

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
m = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
d = int(sys.argv[2]) if len(sys.argv) > 2 else 10000
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
r = bench_sharded_force(m, d, reps, lambda: buf.zero_())
if dist.get_rank() == 0:
    print(json.dumps(dict(world=dist.get_world_size(), m=m, d=d, ms=r["ms"], path=r["path"], h=r["h"], loss=r["loss"],
                          kernels={k: round(v["ms_per_step"], 3) for k, v in r["kernels"].items()},
                          kernel_sum_ms=round(sum(v["ms_per_step"] for v in r["kernels"].values()), 3),
                          timeline_ms=r["timeline"])), flush=True)
dist.destroy_process_group()
