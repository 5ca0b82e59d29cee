"""Build libsmi.so (sm_100a) in-tree with nvcc.  Used by __graft_entry__.build().

    python -m smi_experiments_b200.csrc.build [--force] [--verbose]

Objects are cached per source file (rebuilt when the source or any header is newer).
"""

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
OUT = os.path.join(PKG, "libsmi.so")
OBJ = os.path.join(HERE, "build")

SOURCES = ["common.cu", "stein_small.cu", "stein_mid.cu", "select.cu", "stein_tc.cu", "stein_api.cu", "optim.cu",
           "elbo_guide.cu", "elbo_mlp.cu", "elbo_bnn_mma.cu", "elbo_bnn_tc.cu", "elbo_mlp_cls.cu", "elbo_cls_tc.cu", "elbo_mix_tc.cu", "elbo_api.cu", "prng.cu", "predictive.cu"]

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC,-fvisibility=hidden", "--expt-relaxed-constexpr"]


def _newest_header():
    t = 0.0
    for root in (HERE, os.path.join(os.path.dirname(PKG), "include")):
        for f in os.listdir(root):
            if f.endswith((".cuh", ".h")):
                t = max(t, os.path.getmtime(os.path.join(root, f)))
    return t


def _compile(src, force, verbose):
    obj = os.path.join(OBJ, src.replace(".cu", ".o"))
    sp = os.path.join(HERE, src)
    if (not force and os.path.exists(obj)
            and os.path.getmtime(obj) > max(os.path.getmtime(sp), _newest_header())):
        return obj, ""
    extra = os.environ.get("SMI_NVCC_EXTRA", "").split()   # e.g. -DSMI_STEIN_TRACE for the phase trace
    cmd = [NVCC] + FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", sp, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
    return obj, r.stderr


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    srcs = [s for s in SOURCES if os.path.exists(os.path.join(HERE, s))]
    missing = [s for s in SOURCES if s not in srcs]
    if missing:
        raise RuntimeError("missing sources: %s" % missing)
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        res = list(ex.map(lambda s: _compile(s, force, verbose), srcs))
    objs = [o for o, _ in res]
    if verbose:
        for _, log in res:
            if log:
                print(log)
    need_link = force or not os.path.exists(OUT) or any(os.path.getmtime(o) > os.path.getmtime(OUT) for o in objs)
    if need_link:
        cmd = [NVCC, "-shared", "-o", OUT] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
