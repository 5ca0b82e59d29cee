import numpy as np, torch, sys
sys.path.insert(0, '.')
from smi_experiments_b200 import ops
dev = torch.device('cuda')
def tf32(x):
    return (x.view(np.uint32) & np.uint32(0xFFFFE000)).view(np.float32)
for m, d in [(128, 32), (128, 64), (128, 96), (128, 128), (128, 256), (128, 1024), (256, 64), (200, 64), (100, 64), (300, 64), (512, 64), (1024, 64), (384, 128)]:
    rng = np.random.default_rng(m + d)
    x = rng.standard_normal((m, d)).astype(np.float32)
    D2 = ops.sqdist(torch.from_numpy(x).to(dev), path=ops.PATH_TC).cpu().numpy()
    x64 = x.astype(np.float64)
    sq = (x64 * x64).sum(1)
    ref = sq[:, None] + sq[None, :] - 2 * x64 @ x64.T
    xh = tf32(x).astype(np.float64)
    ref1 = sq[:, None] + sq[None, :] - 2 * xh @ xh.T
    xl = tf32(x - tf32(x)).astype(np.float64)
    ref3 = sq[:, None] + sq[None, :] - 2 * (xh @ xh.T + xh @ xl.T + xl @ xh.T)
    off = ~np.eye(m, dtype=bool)
    e = np.abs(D2 - ref)[off]; e1 = np.abs(ref1 - ref)[off]; e3 = np.abs(ref3 - ref)[off]
    bad = np.argwhere((np.abs(D2 - ref) > 20 * e3.max()) & off)
    print(f"m={m} d={d}: err max {e.max():.3e} mean {e.mean():.3e} | 1xTF32 emu max {e1.max():.3e} | 3x emu max {e3.max():.3e} | nbad {len(bad)} rows {np.unique(bad[:,0])[:8] if len(bad) else ''} cols {np.unique(bad[:,1])[:8] if len(bad) else ''}")
print("---- force")
for m, d in [(256, 1000), (1024, 10000), (4096, 10000)]:
    rng = np.random.default_rng(m + d)
    x = rng.standard_normal((m, d)).astype(np.float32)
    g = rng.standard_normal((m, d)).astype(np.float32)
    xt, gt = torch.from_numpy(x).to(dev), torch.from_numpy(g).to(dev)
    F, loss, h = ops.stein_force(xt, gt, path=ops.PATH_TC)
    hv = float(h.item())
    x64 = xt.double(); g64 = gt.double()
    sq = (x64 * x64).sum(1)
    d2 = (sq[:, None] + sq[None, :] - 2 * x64 @ x64.T).clamp_min(0)
    d2.fill_diagonal_(0)
    a = torch.sort(d2.sqrt().reshape(-1)).values
    n = m * m
    q = np.float32(0.5) * np.float32(n - 1); lo, hi = int(np.floor(q)), int(np.ceil(q))
    h64 = float((0.5 * (a[lo] + a[hi])) ** 2 / np.log(m) + 1.1920929e-07)
    def force(hh):
        K = torch.exp(-d2 / hh); c = 2.0 / (hh * m)
        return K @ (g64 - c * x64) + c * K.sum(1)[:, None] * x64
    F64 = force(h64); F64h = force(hv)
    D2g = ops.sqdist(xt, path=ops.PATH_TC).double()
    e = (D2g - d2).abs()
    rel = lambda A, B: float((A - B).norm() / B.norm())
    print(f"m={m} d={d}: h rel err {abs(hv-h64)/h64:.2e}  F rel err {rel(F.double(), F64):.2e}  (at GPU h: {rel(F.double(), F64h):.2e})  D2 err max {float(e.max()):.3e} mean {float(e.mean()):.3e} bias {float((D2g-d2).mean()):.3e}")
