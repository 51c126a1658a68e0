#!/usr/bin/env python
"""Build-time kernel variants of libctrfe.so, A/B-measured through bench.py.

    python benchtools/sweep_variants.py --build            # here (nvcc cross-compiles, no GPU)
    python benchtools/sweep_variants.py --run cfg2 [cfg3]  # on the GPU box: one bench.py per variant

Each variant is the same source with different -D switches, written to benchtools/_variants/
(git-ignored, travels to the GPU box) and selected with CTRFE_LIB.  Tuning tool, not product code.
"""
from __future__ import annotations

import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "benchtools", "_variants")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
BASE = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-shared", "-gencode", "arch=compute_100a,code=sm_100a",
        "-ftz=true", "-fmad=false"]

# name -> (extra nvcc flags, environment for the run)
VARIANTS = {
    "base": ([], {}),
    "u8_m8": (["-DCTRFE_UNROLL=8"], {}),
    "u8_m6": (["-DCTRFE_UNROLL=8", "-DCTRFE_MINB16=6", "-DCTRFE_MINB32=6"], {}),
    "u8_m5": (["-DCTRFE_UNROLL=8", "-DCTRFE_MINB16=5", "-DCTRFE_MINB32=5"], {}),
    "u8_m4": (["-DCTRFE_UNROLL=8", "-DCTRFE_MINB16=4", "-DCTRFE_MINB32=4"], {}),
    "u16_m3": (["-DCTRFE_UNROLL=16", "-DCTRFE_MINB16=3", "-DCTRFE_MINB32=3"], {}),
}


def build(names):
    os.makedirs(OUT, exist_ok=True)
    procs = []
    for n in names:
        if not VARIANTS[n][0] and n != "base":
            continue  # an environment-only variant runs the base build
        cmd = [NVCC, *BASE, *VARIANTS[n][0], "-o", os.path.join(OUT, f"libctrfe_{n}.so"),
               os.path.join(ROOT, "paddle_infer_b200", "csrc", "ctrfe_api.cu")]
        procs.append((n, subprocess.Popen(cmd, cwd=ROOT)))
        if len(procs) >= 6:
            for nn, p in procs:
                assert p.wait() == 0, nn
            procs = []
    for nn, p in procs:
        assert p.wait() == 0, nn


def run(workloads, names, steps):
    rows = []
    for wl in workloads:
        for n in names:
            lib = os.path.join(OUT, f"libctrfe_{n if VARIANTS[n][0] else 'base'}.so")
            if not os.path.exists(lib):
                continue
            env = dict(os.environ, CTRFE_LIB=lib, **VARIANTS[n][1])
            cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--workload", wl, "--steps", str(steps), "--warmup", "20",
                   "--quick", "--e2e-steps", "3", "--regions", "5"]
            p = subprocess.run(cmd, env=env, capture_output=True, text=True, cwd=ROOT)
            line = None
            for l in p.stdout.splitlines():
                if l.startswith("{"):
                    line = json.loads(l)
            if line is None:
                print(f"{wl} {n}: FAILED\n{p.stderr[-2000:]}", flush=True)
                continue
            r = line["roofline"]
            rows.append((wl, n, r["kernel_ms_mean"] * 1e3, r["kernel_ms_isolated_median"] * 1e3, r["kernel_ms_isolated_mean"] * 1e3))
            print(f"{wl:5s} {n:14s} overlapped {rows[-1][2]:8.2f} us/step   isolated median {rows[-1][3]:8.2f}  mean {rows[-1][4]:8.2f}",
                  flush=True)
    return rows


if __name__ == "__main__":
    a = sys.argv[1:]
    names = list(VARIANTS)
    if "--only" in a:
        names = a[a.index("--only") + 1].split(",")
    if "--build" in a:
        build(names)
    if "--run" in a:
        i = a.index("--run")
        wls = [w for w in a[i + 1:] if w.startswith("cfg")]
        steps = 1000
        if "--steps" in a:
            steps = int(a[a.index("--steps") + 1])
        run(wls or ["cfg2"], names, steps)
