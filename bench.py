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


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=d["hbm_gbs"], bf16=d["bf16_tflops"], bf16_sus=d["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sus=1400.0, src="fallback")


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
def cpu_step_runner(threads=None):
    import torch
    from oracle import models as omodels
    from oracle import step as ostep
    if threads:
        torch.set_num_threads(threads)
    xt, yt = make_uci()
    om = omodels.uci_bnn(C2["in_dim"])
    rows, L, d = omodels.layout(om, "normal")
    rng = np.random.default_rng(1)
    m, S, B = C2["m"], C2["S"], C2["B"]
    psi = rng.uniform(-C2["radius"], C2["radius"], (m, d)).astype(np.float32)
    for r in rows:
        psi[:, r["rho"]:r["rho"] + r["size"]] = np.log(np.expm1(0.1))
    st = dict(psi=psi, opt=(np.zeros_like(psi), np.zeros_like(psi)), it=0)

    def step():
        # noise drawn with numpy (the threefry chain in pure numpy is not what a JAX CPU run pays for)
        eps = rng.standard_normal((m, S, L)).astype(np.float32)
        idx = np.stack([rng.permutation(455)[:B] for _ in range(m * S)]).reshape(m, S, B).astype(np.int32)
        out = ostep.smi_step(om, st["psi"], st["opt"], st["it"], eps, idx, xt, yt, 455, "normal", 1.0, 1.0, 1.0,
                             "adam", C2["lr"], dtype=np.float32)
        st["psi"], st["opt"], st["it"] = out["psi"], out["opt_state"], st["it"] + 1
        return float(out["loss"])

    return step, torch.get_num_threads()


def run_reference(args):
    """--impl reference: the reference's CPU path.  JAX / NumPyro are not installable here
    (DESIGN.md), so this is the oracle PORT of the same step (torch-CPU ops, all host threads)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    step, cores = cpu_step_runner(os.cpu_count())
    for _ in range(max(args.warmup, 1)):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    v = args.steps / dt
    line = dict(metric=METRIC, value=v, unit="SMI steps/s", n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                ms_per_step=1e3 * dt / args.steps, higher_is_better=True, scaling="weak", vs_baseline=None,
                dtype="f32", data="synthetic", impl="reference",
                config=dict(workload="uci_boston_bnn_smi (BASELINE configs[1])", particles=C2["m"],
                            elbo_draws=C2["S"], minibatch=C2["B"], optimizer="adam"),
                cpu_baseline=dict(value=v, unit="SMI steps/s", cores=cores, kind="port",
                                  sample="%d full SMI steps of the C2 workload (oracle, torch-CPU fp32)" % args.steps),
                e2e=dict(value=v, unit="SMI steps/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))
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
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    lib.smi_profile(1)
    for a, b in evs:
        flush()
        a.record()
        smi.ops.stein_force(X, G, out=F, path=smi.ops.PATH_TC)
        b.record()
    torch.cuda.synchronize()
    lib.smi_profile(0)
    prof = profile_report(lib)
    ms = float(np.mean([a.elapsed_time(b) for a, b in evs]))
    W = 4.0 * m * m * d
    tf32_peak = 1100.0 * pk["bf16"] / 2250.0      # TF32 dense peak derived from the measured bf16 GEMM
    gemm_ms = (1, sum(v[1] for k, v in prof.items() if k.startswith("gemm3xtf32")))
    out = dict(m=m, d=d, ms=ms, gflops=W / ms / 1e6, flop_per_step=W,
               roofline=dict(bound="tensor", achieved=W / ms / 1e9, peak=tf32_peak / 3.0, unit="TFLOP/s",
                             frac=(W / ms / 1e9) / (tf32_peak / 3.0),
                             # dram read + write bytes of the two CTA-pair GEMM launches (Gram 35.4 GB + force 66.8 GB)
                             traffic=102233679000 if (m, d) == (16384, 10000) else None,
                             traffic_source="profiles/r01i_ncu_gemm3xtf32_pair_m16384_d10000.txt",
                             note="logical FP32-equivalent peak = TF32 peak / 3 (3xTF32); TF32 peak = 1100 x measured bf16 / 2250; "
                                  "W counts the full Gram, the symmetric kernel executes about half of it, so frac can pass 1; "
                                  "ncu: tensor pipe 95-96 % active in both GEMMs"),
               kernels={k: dict(launches=v[0] // reps if reps else v[0], ms_per_step=v[1] / reps) for k, v in prof.items()})
    gm = gemm_ms[1] / max(reps, 1)
    if gm > 0:
        out["gemm_only"] = dict(ms_per_step=gm, tflops=W / gm / 1e9, frac=(W / gm / 1e9) / (tf32_peak / 3.0))
    return out


def bench_other_configs(torch, smi, dev, flush):
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
    timed("C4 MNIST-shape MLP 784-100-100-10, SMI 10 particles, 55 draws, batch 128", eng, st, (xb, yb),
          dict(n=60000, batch_size=128), 10)
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
        if dom_name.startswith("bnn_reg"):
            # layer products, forward + backward: 3 x 2 x (B I H + B H H + B H) per draw
            flop = 3.0 * 2.0 * (B * I * H + B * H * H + B * H) * m_local * S
            tf32_peak = 1100.0 * pk["bf16"] / 2250.0
            roof = dict(kernel=dom_name, bound="tensor", achieved=flop / dom_avg_ms / 1e9, peak=tf32_peak / 3.0,
                        unit="TFLOP/s", frac=(flop / dom_avg_ms / 1e9) / (tf32_peak / 3.0),
                        traffic=27392768 if world == 1 else None,   # dram read + write bytes per launch, ncu capture
                        traffic_source="profiles/r01g_ncu_bnn_reg_mma3_ct_kernel_C2.txt",
                        algorithmic_flop=flop, peak_source=pk["src"],
                        note="warp-level mma.sync 3xTF32 kernel (one CTA per ELBO draw, per-draw weights) against the "
                             "tcgen05 3xTF32 roofline the north star names: peak = 1100 x measured bf16 / 2250 / 3")
        else:
            d_ = eng.layout.particle_dim
            L_ = eng.layout.latent_dim
            byt = 4.0 * (m_local * S * L_ * 3)
            roof = dict(kernel=dom_name, bound="hbm", achieved=byt / dom_avg_ms / 1e6, peak=pk["hbm"], unit="GB/s",
                        frac=(byt / dom_avg_ms / 1e6) / pk["hbm"], traffic=None, algorithmic_bytes=byt,
                        peak_source=pk["src"])
        line = dict(metric=METRIC, value=value, unit="SMI steps/s (20-particle step units; x N under weak scaling)",
                    n_gpus=world, steps=steps, warmup=max(args.warmup, 3), ms_per_step=dev_ms / steps,
                    higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
                    config=dict(workload="uci_boston_bnn_smi (BASELINE configs[1]): 455x13 train, BNN 13-50-50-1 ReLU, "
                                         "AutoNormal, minibatch 100, 100 ELBO draws, Adam",
                                particles_per_gpu=m_local, particles_total=m_total, elbo_draws=S, minibatch=B,
                                l2="flushed between steps (256 MiB memset, untimed)", timing="cuda events per step, max over ranks",
                                cuda_graph=True),
                    e2e=dict(value=e2e_v, unit="SMI steps/s", h2d_bytes_per_step=int(h2d), d2h_bytes_per_step=4,
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
                    from smi_experiments_b200.dist import bench_sharded_force
                    p = bench_sharded_force(mm, 10000, reps, flush, pk)
                pts.append(dict(m=mm, ms=p["ms"], tflops=p["roofline"]["achieved"], frac=p["roofline"]["frac"]))
        if world == 1:
            sw = bench_stein_sweep(torch, smi, dev, args.sweep_m, 10000, args.sweep_reps, flush, pk)
        else:
            from smi_experiments_b200.dist import bench_sharded_force
            sw = bench_sharded_force(args.sweep_m, 10000, args.sweep_reps, flush, pk)
        if pts:
            pts.append(dict(m=sw["m"], ms=sw["ms"], tflops=sw["roofline"]["achieved"], frac=sw["roofline"]["frac"]))
            sw["sweep"] = pts
        if rank == 0:
            line["stein_force"] = sw
    if rank == 0 and world == 1 and not args.no_configs:
        line["other_configs"] = bench_other_configs(torch, smi, dev, flush)
    # ---- CPU baseline (rank 0, N = 1 only): bounded sample of the same workload ----
    if rank == 0 and world == 1 and not args.no_cpu:
        step, cores = cpu_step_runner(os.cpu_count())
        step()
        n, t0 = 0, time.perf_counter()
        while n < 8 and time.perf_counter() - t0 < 20.0:
            step()
            n += 1
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = dict(value=n / dt, unit="SMI steps/s", cores=cores, kind="port",
                                    sample="%d full SMI steps of the same C2 workload (oracle port, torch-CPU fp32)" % n)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
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
