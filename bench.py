#!/usr/bin/env python
"""bench.py -- SMI steps/s on the UCI-Boston-shape BNN (BASELINE.json configs[1]) and the
Stein-force GFLOP/s sweep point (configs[4]), on N B200s of one node.

    python bench.py --gpus 1 --steps 50 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference --steps 5 --warmup 1      # CPU arm (oracle port, host cores)

One JSON line on stdout (rank 0).  `value` = SMI steps/s with inputs resident in HBM; `e2e` = the
same through engine.update with HOST buffers (H2D of x, y and D2H of the loss inside the timed
region, every step -- the reference's src/engine_run.py:52-56 loop shape).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "SMI steps/sec & Stein-force GFLOP/s (% roofline), 1/2/4/8 B200 vs host CPU"


# ---- workload: BASELINE.json configs[1], SURVEY.md 8(d) C2 ---------------------------------
def make_uci(seed=1):
    """506 x 13 synthetic Boston-shape data, z-scored (src/preprocess.py:4-11); first 455 rows train."""
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((506, 13)).astype(np.float32)
    w = rng.standard_normal(13).astype(np.float32)
    y = (x @ w + 0.1 * rng.standard_normal(506)).astype(np.float32)
    xt, yt = x[:455], y[:455]
    xt = (xt - xt.mean(0)) / xt.std(0)
    yt = (yt - yt.mean()) / yt.std()
    return np.ascontiguousarray(xt, np.float32), np.ascontiguousarray(yt, np.float32)


C2 = dict(m=20, S=100, B=100, in_dim=13, lr=5e-3, radius=0.1)
UNIT = "SMI steps/s"      # the SAME string in both arms (the driver divides them)


def bench_config(world):
    """`config` of the bench line, identical in the GPU arm and the --impl reference arm."""
    return dict(workload="uci_boston_bnn_smi (BASELINE configs[1]): 455x13 train, BNN 13-50-50-1 ReLU, AutoNormal, "
                         "minibatch 100, 100 ELBO draws, Adam",
                particles_per_gpu=C2["m"], particles_total=C2["m"] * world, elbo_draws=C2["S"], minibatch=C2["B"],
                unit_note="one unit = one SMI step on 20 particles; under weak scaling (--gpus N) a step updates "
                          "20 N particles of one mixture and counts N units",
                # how the GPU arm is timed (the same dict is printed by --impl reference so that both arms carry an
                # identical `config`; the CPU arm times whole steps with perf_counter)
                l2="GPU arm: flushed between steps (256 MiB memset, untimed)",
                timing="GPU arm: cuda events per step, max over ranks; CUDA-graph replay of the whole step (NCCL all-gathers included at N > 1)")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=d["hbm_gbs"], bf16=d["bf16_tflops"], bf16_sus=d["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sus=1400.0, src="fallback")


_UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def traffic_from_profile(relpath, kernel_substr):
    """dram__bytes_read.sum + dram__bytes_write.sum of the launches whose kernel name contains
    `kernel_substr`, summed, read from a committed ncu summary under profiles/ (scripts/ncu_summary.py
    format).  None when the file or the kernel is absent."""
    path = os.path.join(ROOT, relpath)
    if not os.path.exists(path):
        return None
    total, cur, found = 0.0, False, False
    with open(path) as f:
        for ln in f:
            t = ln.split()
            if not t:
                continue
            if ln.startswith("Kernel Name"):
                cur = kernel_substr in ln
                found |= cur
            elif cur and t[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum") and len(t) >= 3:
                total += float(t[1]) * _UNIT.get(t[2], 1.0)
    return int(total) if found else None


def tf32_peaks(pk):
    """Dense TF32 tensor peak DERIVED from the measured bf16 GEMM (1100 / 2250 of it): burst and sustained;
    a 3xTF32 kernel's logical FP32-equivalent peak is one third."""
    return 1100.0 * pk["bf16"] / 2250.0, 1100.0 * pk["bf16_sus"] / 2250.0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for nm, v in zip(names, r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["no samples"])
        return dict(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons),
                    samples=len(sm))


# ---- CPU arm: the oracle port on the host cores --------------------------------------------
def cpu_step_runner(threads=None, world=1):
    """One step of the reference algorithm on the host (oracle port, torch-CPU fp32, all threads).
    world = 1: the full C2 step on 20 particles.  world = N > 1 (the weak-scaled workload, m = 20 N particles of one
    mixture): ONE unit of it = what one rank does per step -- the ELBO gradients of 20 of the 20 N particles against
    all 20 N mixture components, the Stein force (evaluated in full here: 0.1 % of the step at these m) and the
    optimizer update of those 20 rows.  A bounded sample of the workload with no extrapolation: units / second."""
    import torch
    from oracle import elbo as oelbo
    from oracle import models as omodels
    from oracle import optim as ooptim
    from oracle import stein as ostein
    from oracle import step as ostep
    if threads:
        torch.set_num_threads(threads)
    xt, yt = make_uci()
    om = omodels.uci_bnn(C2["in_dim"])
    rows, L, d = omodels.layout(om, "normal")
    rng = np.random.default_rng(1)
    n, S, B = C2["m"], C2["S"], C2["B"]
    m = n * world
    psi = rng.uniform(-C2["radius"], C2["radius"], (m, d)).astype(np.float32)
    for r in rows:
        psi[:, r["rho"]:r["rho"] + r["size"]] = np.log(np.expm1(0.1))
    st = dict(psi=psi, opt=(np.zeros_like(psi), np.zeros_like(psi)), it=0)

    def noise():
        # noise drawn with numpy (the threefry chain in pure numpy is not what a JAX CPU run pays for)
        eps = rng.standard_normal((n, S, L)).astype(np.float32)
        idx = np.stack([rng.permutation(455)[:B] for _ in range(n * S)]).reshape(n, S, B).astype(np.int32)
        return eps, idx

    def step():
        eps, idx = noise()
        out = ostep.smi_step(om, st["psi"], st["opt"], st["it"], eps, idx, xt, yt, 455, "normal", 1.0, 1.0, 1.0,
                             "adam", C2["lr"], dtype=np.float32)
        st["psi"], st["opt"], st["it"] = out["psi"], out["opt_state"], st["it"] + 1
        return float(out["loss"])

    def unit():
        eps, idx = noise()
        Gr, _ = oelbo.elbo_grads(om, st["psi"], eps, idx, xt, yt, 455, 1.0, "normal", torch.float32, rows=(0, n))
        G = np.tile(Gr, (world, 1))                    # the other ranks' scores: same shape, values irrelevant to the cost
        F, loss, h, _ = ostein.stein_force(st["psi"], G, 1.0, 1.0, ostein.MEDIAN_JAX_FP32_INDEX, np.float32)
        a, b = st["opt"]
        x2, a2, b2 = ooptim.adam_update(st["it"], -F[:n], st["psi"][:n], a[:n], b[:n], C2["lr"])
        st["psi"][:n], a[:n], b[:n] = x2, a2, b2
        st["it"] += 1
        return float(loss)

    return (step if world == 1 else unit), torch.get_num_threads()


def run_reference(args):
    """--impl reference: the reference's CPU path.  JAX / NumPyro are not installable here
    (DESIGN.md), so this is the oracle PORT of the same step (torch-CPU ops, all host threads)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # the same workload as the GPU arm at this --gpus N: weak scaling = 20 particles per GPU, m = 20 N in ONE
    # mixture; one unit = one rank's share of a step (20 particles); the GPU arm completes N units per step
    world = max(args.gpus, 1)
    m_total = C2["m"] * world
    step, cores = cpu_step_runner(os.cpu_count(), world)
    for _ in range(max(args.warmup, 1)):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    v = args.steps / dt         # units / s: a step of this arm is ONE unit (see cpu_step_runner)
    line = dict(metric=METRIC, value=v, unit=UNIT, n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                ms_per_step=1e3 * dt / args.steps, higher_is_better=True, scaling="weak", vs_baseline=None,
                dtype="f32", data="synthetic", impl="reference",
                config=bench_config(world),
                cpu_baseline=dict(value=v, unit=UNIT, cores=cores, kind="port",
                                  sample=("%d full SMI steps of the C2 workload (20 particles; oracle port, torch-CPU "
                                          "fp32, all host threads)" % args.steps) if world == 1 else
                                         ("%d units of the weak-scaled workload at m = %d particles: per unit the ELBO "
                                          "gradients of 20 particles against all %d mixture components + the Stein force "
                                          "+ Adam on those rows = one rank's share of a step (oracle port, torch-CPU fp32, "
                                          "all host threads)" % (args.steps, m_total, m_total))),
                e2e=dict(value=v, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line), flush=True)


# ---- GPU arm --------------------------------------------------------------------------------
def profile_report(lib):
    buf = C.create_string_buffer(1 << 16)
    lib.smi_profile_report(buf, len(buf))
    out = {}
    for ln in buf.value.decode().splitlines():
        nm, cnt, ms = ln.split()
        out[nm] = (int(cnt), float(ms))
    return out


def l2_flusher(torch, dev):
    buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def flush():
        buf.zero_()
    return flush


def bench_stein_sweep(torch, smi, dev, m, d, reps, flush, pk):
    """Stein-force stage alone at BASELINE configs[4] (m x d = 16384 x 10000 when it fits)."""
    gen = torch.Generator(device=dev).manual_seed(4)
    X = torch.randn(m, d, device=dev, generator=gen)
    G = torch.randn(m, d, device=dev, generator=gen)
    F = torch.empty_like(X)
    lib = smi._lib.load()
    for _ in range(2):
        smi.ops.stein_force(X, G, out=F, path=smi.ops.PATH_TC)
    torch.cuda.synchronize()
    # the evaluation is replayed from a CUDA graph, as the engine replays its whole step: below m ~ 4k the ~17 launches
    # of one evaluation (5 - 60 us each) are otherwise paced by the eager Python / ctypes issue rate
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        smi.ops.stein_force(X, G, out=F, path=smi.ops.PATH_TC)
    graph.replay()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in evs:
        flush()
        a.record()
        graph.replay()
        b.record()
    torch.cuda.synchronize()
    ms = float(np.mean([a.elapsed_time(b) for a, b in evs]))
    # per-kernel durations: one more pass, eager, events on the launching stream (untimed)
    lib.smi_profile(1)
    for _ in range(reps):
        flush()
        smi.ops.stein_force(X, G, out=F, path=smi.ops.PATH_TC)
    torch.cuda.synchronize()
    lib.smi_profile(0)
    prof = profile_report(lib)
    out = stein_point(m, d, ms, 1, pk)
    out["timing"] = "CUDA-graph replay of one smi_stein_force_f32 call, cuda events, L2 flushed between replays"
    out["kernels"] = {k: dict(launches=v[0] // reps if reps else v[0], ms_per_step=v[1] / reps) for k, v in prof.items()}
    gm = sum(v[1] for k, v in prof.items() if k.startswith("gemm3xtf32")) / max(reps, 1)
    if gm > 0:
        W = out["flop_per_step"]
        burst, sus = tf32_peaks(pk)
        ex = executed_mma_flop(m, d)
        out["gemm_only"] = dict(ms_per_step=gm, tflops=W / gm / 1e9, frac=(W / gm / 1e9) / (burst / 3.0),
                                frac_sustained_peak=(W / gm / 1e9) / (sus / 3.0),
                                executed_tf32_tflops=ex / gm / 1e9, executed_frac=(ex / gm / 1e9) / burst,
                                executed_frac_sustained_peak=(ex / gm / 1e9) / sus)
    # untimed correctness record of the timed point: h bits, loss, sampled force rows against the float64 definition
    _, loss, h = smi.ops.stein_force(X, G, out=F, path=smi.ops.PATH_TC)
    out["check"] = force_rows_check(torch, X, G, F, 0, float(h))
    out["check"]["loss"] = float(loss)
    return out


def sharded_point(torch, m, d, reps, flush, pk, world):
    """The sweep point on N GPUs (strong scaling: m fixed, rows sharded).  Every rank generates the SAME full X, G
    as the N = 1 run (seed 4), so h / loss are comparable across world sizes; the untimed check compares sampled
    force rows of every rank with the float64 definition and all-reduces the worst error."""
    import torch.distributed as dist
    from smi_experiments_b200.dist import bench_sharded_force
    r = bench_sharded_force(m, d, reps, flush)
    out = stein_point(m, d, r["ms"], world, pk)
    out["kernels"] = r["kernels"]
    out["collectives"] = r["collectives"]
    out["path"] = r["path"]
    out["timeline_ms_rank0"] = r["timeline"]
    nv = 2.0 * (world - 1) / world * m * d * 4.0
    burst, _ = tf32_peaks(pk)
    out["roofline"].update(nvlink_bytes_per_rank=nv, t_comm_ms_at_770GBs=nv / 770e9 * 1e3,
                           t_flop_ms=4.0 * m * m * d / world / (burst / 3.0 * 1e12) * 1e3)
    chk = force_rows_check(torch, r["X"], r["G"], r["F_rows"], r["row0"], r["h"])
    worst = torch.tensor([chk["force_rows_rel_err_vs_float64"]], dtype=torch.float64, device=r["X"].device)
    dist.all_reduce(worst, op=dist.ReduceOp.MAX)
    hb = torch.tensor([int(chk["h_bits"], 16)], dtype=torch.int64, device=r["X"].device)
    hmin, hmax = hb.clone(), hb.clone()
    dist.all_reduce(hmin, op=dist.ReduceOp.MIN)
    dist.all_reduce(hmax, op=dist.ReduceOp.MAX)
    chk.update(force_rows_rel_err_vs_float64=float(worst), ok=bool(float(worst) <= 1e-5), loss=r["loss"],
               h_bits_equal_on_all_ranks=bool(int(hmin) == int(hmax)), rows_checked=chk["rows_checked"] * world)
    out["check"] = chk
    return out


def executed_mma_flop(m, d):
    """TF32 MMA FLOPs the two pair GEMMs execute: 3 MMAs per product term (3xTF32), padded tiles; the symmetric Gram
    computes only the 256 x 256 pair tiles touching j >= i."""
    T = (m + 255) // 256
    kd, km = (d + 31) // 32 * 32, (m + 31) // 32 * 32
    gram = T * (T + 1) // 2 * 256 * 256 * 2.0 * kd
    force = T * ((d + 255) // 256) * 256 * 256 * 2.0 * km
    return 3.0 * (gram + force)


def stein_point(m, d, ms, world, pk):
    """Roofline record of one Stein-force evaluation (W = 4 m^2 d) on `world` GPUs."""
    W = 4.0 * m * m * d
    burst, sus = tf32_peaks(pk)
    ach = W / ms / 1e9
    ex = executed_mma_flop(m, d)
    traffic = None
    src = "profiles/r02_ncu_c5_pair_gemms.txt"
    if (m, d, world) == (16384, 10000, 1):
        traffic = traffic_from_profile(src, "gemm3xtf32_pair_kernel")
        if traffic is None:
            src = "profiles/r01i_ncu_gemm3xtf32_pair_m16384_d10000.txt"
            traffic = traffic_from_profile(src, "gemm3xtf32_pair_kernel")
    return dict(m=m, d=d, ms=ms, gflops=W / ms / 1e6, flop_per_step=W,
                roofline=dict(bound="tensor", achieved=ach, peak=world * burst / 3.0, unit="TFLOP/s",
                              frac=ach / (world * burst / 3.0),
                              peak_sustained=world * sus / 3.0, frac_sustained_peak=ach / (world * sus / 3.0),
                              executed_tf32_tflops=ex / ms / 1e9, executed_frac=(ex / ms / 1e9) / (world * burst),
                              executed_frac_sustained_peak=(ex / ms / 1e9) / (world * sus),
                              traffic=traffic, traffic_source=src if traffic is not None else None,
                              peak_source=pk["src"],
                              note="logical FP32-equivalent peak = TF32 peak / 3 (3xTF32); TF32 peak DERIVED = 1100 x measured "
                                   "bf16 / 2250 (burst: frac; sustained: frac_sustained_peak); W = 4 m^2 d counts the full Gram, "
                                   "the symmetric kernel executes about half of it, so frac can pass 1 -- executed_* count the "
                                   "TF32 MMA FLOPs actually issued (3 per term, symmetric Gram) over the WHOLE evaluation; "
                                   "traffic = DRAM read + write bytes of the two GEMM launches from the named ncu summary"))


def force_rows_check(torch, X, G, F_rows, row0, h, nsample=16, alpha=1.0, c_att=1.0):
    """Relative error of `nsample` of the force rows this rank holds against the float64 DEFINITION
    F_i = c_att sum_j K_ij g_j + (2 alpha / (h m)) (x_i sum_j K_ij - sum_j K_ij x_j), K_ij = exp(-|x_i - x_j|^2 / h)
    evaluated with torch float64 on the device at the bandwidth the kernels selected (not an oracle/ import:
    a definition check, as tests/test_gpu_stein.py::test_full_size_sweep_point_properties)."""
    m, d = X.shape
    n = F_rows.shape[0]
    loc = torch.linspace(0, n - 1, min(nsample, n), device=X.device).round().long().unique()
    rows = loc + row0
    xr = X[rows].double()
    d2 = torch.empty(len(rows), m, dtype=torch.float64, device=X.device)
    for j0 in range(0, m, 2048):
        xb = X[j0:j0 + 2048].double()
        d2[:, j0:j0 + 2048] = ((xr * xr).sum(1)[:, None] + (xb * xb).sum(1)[None, :] - 2.0 * (xr @ xb.T)).clamp_min(0)
    d2[torch.arange(len(rows), device=X.device), rows] = 0.0
    K = torch.exp(-d2 / h)
    c = 2.0 * alpha / (h * m)
    Fref = torch.zeros(len(rows), d, dtype=torch.float64, device=X.device)
    for j0 in range(0, m, 2048):
        Fref += K[:, j0:j0 + 2048] @ (c_att * G[j0:j0 + 2048].double() - c * X[j0:j0 + 2048].double())
    Fref += c * K.sum(1, keepdim=True) * xr
    err = float((F_rows[loc].double() - Fref).norm() / Fref.norm())
    import struct
    return dict(h=h, h_bits="0x%08x" % struct.unpack("<I", struct.pack("<f", h))[0], rows_checked=len(rows),
                force_rows_rel_err_vs_float64=err, ok=bool(err <= 1e-5))


def bench_other_configs(torch, smi, dev, flush, pk=None):
    """The other BASELINE.json configs (parity-test cases) timed the same
    way as `value` (graph replay, L2 flushed between steps, CUDA events): context, not bench lines."""
    out = {}

    def timed(name, eng, state, args, kwargs, steps):
        for _ in range(4):
            state, loss = eng.update(state, *args, **kwargs)
        torch.cuda.synchronize()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for a, b in evs:
            flush()
            a.record()
            state, loss = eng.update(state, *args, **kwargs)
            b.record()
        torch.cuda.synchronize()
        ms = float(np.mean([a.elapsed_time(b) for a, b in evs]))
        out[name] = dict(ms_per_step=ms, steps_per_s=1e3 / ms, loss=float(loss))

    rng = np.random.default_rng(0)
    model = smi.reg_1d_bnn(50)
    eng = smi.SteinVI(model, smi.AutoNormal(model), smi.Adam(1e-3), smi.RBFKernel(), num_stein_particles=5,
                      num_elbo_particles=100)
    x = torch.linspace(-1.5, 1.7, 40, device=dev)
    y = torch.sin(x)
    timed("C1 sine BNN 1x50 hidden, SMI 5 particles, 100 draws", eng, eng.init(smi.PRNGKey(0), x, y), (x, y), {}, 20)
    model = smi.GaussianTarget(1000)
    eng = smi.SVGD(model, smi.Adam(0.05), smi.RBFKernel(), num_stein_particles=20,
                   guide_kwargs={"init_loc_fn": smi.init_to_uniform(radius=2.0)})
    timed("C3 Gaussian d=1000, SVGD 20 particles", eng, eng.init(smi.PRNGKey(2)), (), {}, 20)
    eng = smi.SteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=2.0)), smi.Adagrad(0.05),
                      smi.RBFKernel(), num_stein_particles=20, num_elbo_particles=100)
    timed("C3 Gaussian d=1000, SMI 20 particles, 100 draws", eng, eng.init(smi.PRNGKey(2)), (), {}, 20)
    model = smi.mnist_mlp()
    eng = smi.SteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=0.05)), smi.Adagrad(0.8),
                      smi.RBFKernel(), num_stein_particles=10, num_elbo_particles=55)
    xb = torch.from_numpy(rng.integers(0, 256, (128, 784)).astype(np.float32) / 255.0).to(dev)
    yb = torch.from_numpy(rng.integers(0, 10, 128).astype(np.int32)).to(dev)
    st = eng.init(smi.PRNGKey(4), xb, yb, n=60000, batch_size=128)
    c4 = "C4 MNIST-shape MLP 784-100-100-10, SMI 10 particles, 55 draws, batch 128"
    timed(c4, eng, st, (xb, yb), dict(n=60000, batch_size=128), 10)
    if pk is not None:
        # HBM roofline of the C4 step: algorithmic bytes = 4 (m S L + 2 m d) for the ELBO gradients with
        # eps as an input + 28 m d for the optimizer; the design's own traffic (eps written + read twice, Z / DZ each
        # way, W1^T operand planes written + read) is about ten times that
        m_, S_, L_, d_ = 10, 55, eng.layout.latent_dim, eng.layout.particle_dim
        ms = out[c4]["ms_per_step"]
        alg = 4.0 * (m_ * S_ * L_ + 2 * m_ * d_) + 28.0 * m_ * d_
        design = 7.0 * 4.0 * m_ * S_ * L_ + 2.0 * 2.0 * 4.0 * m_ * S_ * 100 * 800 + 2.0 * 4.0 * m_ * S_ * 128 * 100 * 3 + 40.0 * m_ * d_
        out[c4]["roofline"] = dict(bound="hbm", achieved=alg / ms / 1e6, peak=pk["hbm"], unit="GB/s",
                                   frac=(alg / ms / 1e6) / pk["hbm"], algorithmic_bytes=alg,
                                   design_traffic_bytes=design, frac_of_design_traffic=(design / ms / 1e6) / pk["hbm"],
                                   peak_source=pk["src"], traffic=None)
    return out


def run_gpu(args):
    import torch
    import torch.distributed as dist
    import smi_experiments_b200 as smi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    smi.ops.device_check()
    lib = smi._lib.load()
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    pk = peaks()
    flush = l2_flusher(torch, dev)

    xt, yt = make_uci()
    m_local = C2["m"]
    m_total = m_local * world      # weak scaling: 20 particles per GPU
    model = smi.uci_bnn(C2["in_dim"])
    guide = smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=C2["radius"]))
    if world == 1:
        eng = smi.SteinVI(model, guide, smi.Adam(C2["lr"]), smi.RBFKernel(), num_stein_particles=m_total,
                          num_elbo_particles=C2["S"])
    else:
        from smi_experiments_b200.dist import ShardedSteinVI
        eng = ShardedSteinVI(model, guide, smi.Adam(C2["lr"]), smi.RBFKernel(), num_stein_particles=m_total,
                             num_elbo_particles=C2["S"])
    xd, yd = torch.from_numpy(xt).to(dev), torch.from_numpy(yt).to(dev)
    kw = dict(subsample=C2["B"])
    state = eng.init(smi.PRNGKey(1), x=xd, y=yd, **kw)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- value: inputs resident in HBM, CUDA-graph replay, L2 flushed between steps (untimed) ----
    for _ in range(max(args.warmup, 3)):
        state, loss = eng.update(state, x=xd, y=yd, **kw)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    for a, b in evs:
        flush()
        a.record()
        state, loss = eng.update(state, x=xd, y=yd, **kw)
        b.record()
    barrier()
    dev_ms = float(sum(a.elapsed_time(b) for a, b in evs))
    # ---- e2e: host buffers, H2D each step, loss read back each step ----
    for _ in range(3):
        state, loss = eng.update(state, x=xt, y=yt, **kw)
        float(loss)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        state, loss = eng.update(state, x=xt, y=yt, **kw)
        last = float(loss)                     # device_get(loss), src/engine_run.py:56
    e1.record()
    barrier()
    e2e_ms = float(e0.elapsed_time(e1))
    h2d = eng.h2d_bytes
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = float(t[0]), float(t[1])

    # ---- per-kernel durations of the same step, eager, events on the launching stream ----
    eng.use_cuda_graph = False
    eng._graph = None
    lib.smi_profile(1)
    for _ in range(args.steps):
        flush()
        state, loss = eng.update(state, x=xd, y=yd, **kw)
    lib.smi_profile(0)
    prof = profile_report(lib)
    launches = eng.launches_per_update
    eng.use_cuda_graph = True

    if rank == 0:
        steps = args.steps
        units = float(world)      # one unit = one SMI step on 20 particles; each rank owns 20
        value = units * steps / (dev_ms / 1e3)
        e2e_v = units * steps / (e2e_ms / 1e3)
        kern = {k: dict(launches=v[0] // steps, ms_per_step=v[1] / steps) for k, v in prof.items()}
        dom = max(prof.items(), key=lambda kv: kv[1][1])
        dom_name, (dom_cnt, dom_ms) = dom
        dom_avg_ms = dom_ms / dom_cnt
        S, B, I, H = C2["S"], C2["B"], C2["in_dim"], 50
        roof = None
        burst, sus = tf32_peaks(pk)
        if dom_name.startswith("bnn_reg"):
            # layer products, forward + backward: 3 x 2 x (B I H + B H H + B H) per draw
            flop = 3.0 * 2.0 * (B * I * H + B * H * H + B * H) * m_local * S
            tc = dom_name.startswith("bnn_reg_tc")
            src = "profiles/r02c_ncu_bnn_reg_tc_v1_C2.txt" if tc else "profiles/r01g_ncu_bnn_reg_mma3_ct_kernel_C2.txt"
            roof = dict(kernel=dom_name, bound="tensor", achieved=flop / dom_avg_ms / 1e9, peak=burst / 3.0,
                        unit="TFLOP/s", frac=(flop / dom_avg_ms / 1e9) / (burst / 3.0),
                        peak_sustained=sus / 3.0, frac_sustained_peak=(flop / dom_avg_ms / 1e9) / (sus / 3.0),
                        traffic=traffic_from_profile(src, "bnn_reg") if world == 1 else None, traffic_source=src,
                        algorithmic_flop=flop, peak_source=pk["src"],
                        note=("tcgen05 3xTF32 per-draw kernel" if tc else "warp-level mma.sync 3xTF32 kernel") +
                             " against the tcgen05 3xTF32 roofline the north star names: peak = 1100 x measured bf16 / 2250 / 3")
        else:
            L_ = eng.layout.latent_dim
            byt = 4.0 * (m_local * S * L_ * 3)
            roof = dict(kernel=dom_name, bound="hbm", achieved=byt / dom_avg_ms / 1e6, peak=pk["hbm"], unit="GB/s",
                        frac=(byt / dom_avg_ms / 1e6) / pk["hbm"], traffic=None, algorithmic_bytes=byt,
                        peak_source=pk["src"])
        # whole step against the same tensor roofline (3.84 GFLOP of layer products per step)
        step_flop = 3.0 * 2.0 * (B * I * H + B * H * H + B * H) * m_local * S
        roof["step_frac"] = (step_flop / (dev_ms / steps) / 1e9) / (burst / 3.0)
        cfg = bench_config(world)
        line = dict(metric=METRIC, value=value, unit=UNIT,
                    n_gpus=world, steps=steps, warmup=max(args.warmup, 3), ms_per_step=dev_ms / steps,
                    higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
                    config=cfg,
                    e2e=dict(value=e2e_v, unit=UNIT, h2d_bytes_per_step=int(h2d), d2h_bytes_per_step=4,
                             ms_per_step=e2e_ms / steps),
                    gpu_launches=int(launches * steps), launches_per_step=int(launches),
                    clocks=clocks, roofline=roof, kernels=kern, last_loss=last)
    # ---- Stein-force sweep point (BASELINE configs[4]) ----
    if not args.no_sweep:
        # configs[4]'s sweep (m = 1k .. 16k at d = 1e4) in ascending order, the m = 16384 point last: the small
        # points are launch/latency-bound and would otherwise be timed on clocks still power-capped by the big one
        pts = []
        if args.sweep_m == 16384 and not args.no_sweep_points:
            for mm in (1024, 2048, 4096, 8192):
                reps = args.sweep_reps * (4 if mm <= 4096 else 1)
                if world == 1:
                    p = bench_stein_sweep(torch, smi, dev, mm, 10000, reps, flush, pk)
                else:
                    p = sharded_point(torch, mm, 10000, reps, flush, pk, world)
                pts.append(dict(m=mm, ms=p["ms"], tflops=p["roofline"]["achieved"], frac=p["roofline"]["frac"]))
        if world == 1:
            sw = bench_stein_sweep(torch, smi, dev, args.sweep_m, 10000, args.sweep_reps, flush, pk)
        else:
            sw = sharded_point(torch, args.sweep_m, 10000, args.sweep_reps, flush, pk, world)
        if pts:
            pts.append(dict(m=sw["m"], ms=sw["ms"], tflops=sw["roofline"]["achieved"], frac=sw["roofline"]["frac"]))
            sw["sweep"] = pts
        if rank == 0:
            line["stein_force"] = sw
    if rank == 0 and world == 1 and not args.no_configs:
        line["other_configs"] = bench_other_configs(torch, smi, dev, flush, pk)
    # ---- CPU baseline (rank 0, N = 1 only): bounded sample of the same workload ----
    if rank == 0 and world == 1 and not args.no_cpu:
        step, cores = cpu_step_runner(os.cpu_count())
        step()
        n, t0 = 0, time.perf_counter()
        while n < 8 and time.perf_counter() - t0 < 20.0:
            step()
            n += 1
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = dict(value=n / dt, unit=UNIT, cores=cores, kind="port",
                                    sample="%d full SMI steps of the same C2 workload (oracle port, torch-CPU fp32)" % n)
    if rank == 0 and world == 1 and not args.no_cpu and not args.no_sweep:
        # CPU baseline of the Stein stage: the oracle's algorithm (difference-form kernel) in
        # torch fp32 on all host threads at m = 1024, d = 1e4; larger m extrapolated with m^2 (labelled)
        from oracle import stein as ostein
        rng = np.random.default_rng(4)
        mc = 1024
        Xc = rng.standard_normal((mc, 10000)).astype(np.float32)
        Gc = rng.standard_normal((mc, 10000)).astype(np.float32)
        t0 = time.perf_counter()
        ostein.stein_force_torch(Xc, Gc, threads=os.cpu_count())
        dt = time.perf_counter() - t0
        sw = line["stein_force"]
        scale = (sw["m"] / mc) ** 2
        line["stein_force"]["cpu_baseline"] = dict(
            value=4.0 * mc * mc * 10000 / dt / 1e9, unit="GFLOP/s (W = 4 m^2 d)", cores=os.cpu_count(), kind="port",
            sample="one Stein-force evaluation at m = 1024, d = 10000 (oracle algorithm, torch-CPU fp32): %.2f s" % dt,
            ms_at_m1024=dt * 1e3, ms_at_bench_m_extrapolated=dt * 1e3 * scale,
            extrapolation="time ~ m^2 at fixed d (labelled extrapolation, BASELINE.md section 3)")
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        eng._graph = None            # the captured step holds NCCL kernels: drop it before the communicator goes
        torch.cuda.synchronize()
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--sweep-m", type=int, default=16384)
    ap.add_argument("--sweep-reps", type=int, default=3)
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--no-sweep-points", action="store_true", help="only the m = 16384 point of the Stein-force sweep")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-configs", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
