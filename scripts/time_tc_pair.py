"""A/B timing of the tcgen05 Stein force per sweep point: run once with SMI_TC_PAIR=1 (CTA pairs) and once with
SMI_TC_PAIR=0 (one-CTA kernel).  Prints the force time and the share of the two GEMMs per m."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import smi_experiments_b200 as smi

dev = torch.device("cuda")
pk = bench.peaks()
buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
flush = (lambda: None) if os.environ.get("SMI_NOFLUSH") else (lambda: buf.zero_())
ms_list = [int(a) for a in sys.argv[1:]] or [1024, 2048, 4096, 8192, 16384]
for m in ms_list:
    r = bench.bench_stein_sweep(torch, smi, dev, m, 10000, 10 if m <= 8192 else 4, flush, pk)
    g = {"ms_per_step": r.get("gemm_only", {}).get("ms_per_step", 0.0)}
    ks = {k: round(v["ms_per_step"], 4) for k, v in r["kernels"].items()}
    print(json.dumps(dict(pair=os.environ.get("SMI_TC_PAIR", "1"), m=m, ms=round(r["ms"], 4),
                          gemm_ms=round(g.get("ms_per_step", 0.0), 4), frac=round(r["roofline"]["frac"], 4), kernels=ks)))
