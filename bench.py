#!/usr/bin/env python
"""bench.py -- pooled rows/sec of the fused gather + sequence_pool + CVM front end.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2] [--impl reference]

One "step" = one pass of the hot path over one synthetic batch (all slots of the batch, one
plan launch).  Headline workload: BASELINE.json configs[1] (DeepFM: 39 slots, one shared
1e8 x 16 fp32 table, batch 8192, Zipf ids, SUM + CVM).  N > 1 (torchrun): every rank holds a
replica of the table and its own batches -- weak scaling, no data-path collective.

Prints ONE JSON line (contract in the task statement):
  value         whole-job pooled rows/s, inputs resident in HBM: the median over several timed
                regions of K back-to-back launches (ctrfe_plan_launch_many, one call into the library
                per region), max over ranks
  survey_8d     the same step as SURVEY.md 8(d) words it: host LoD offsets in pinned memory -> one
                async upload -> launch, as an isolated per-step median and back to back
  e2e           the metric through the host-buffer call (H2D of ids + LoD, D2H of the pooled block
                inside the timed region)
  roofline      for the dominant kernel, from CUDA events; the kernel name comes from the plan
  parity        a sample of the LAST timed batch checked against the CPU oracle
  configs       (N = 1) the other BASELINE configs at full size: cfg1, cfg3, cfg4 on one GPU, cfg5
  sharded_cfg4  BASELINE config 4 on N GPUs: N = 1 the whole 128 GB table on one GPU; N > 1 the table
                row-sharded over the ranks (strong scaling), with and without hot-row replicas
  cpu_baseline  the reference CPU arithmetic timed on this box's host cores (N = 1)
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "pooled rows/sec (gather+seqpool+CVM)"
UNIT = "pooled_rows/s"
# e2e: which form of the host-buffer call `--e2e-mode auto` takes at N GPUs (mode, ids width, requests in
# flight, slot groups) -- measured, see DESIGN.md section 6 -- and the forms reported beside it at N = 1
# (this box's PCIe link is close to half duplex for GPU-initiated traffic -- uploads and SM reads of host memory
# stall while rows stream down, profiles/r2_e2e_sweep.txt -- so a step costs about the SUM of its bytes at
# 56 GB/s: 32-bit ids (the table has < 2^31 rows; lookup_table_v2 takes them) and a second request in flight
# are what is left.  N = 1: zero_copy int64 0.59 ms, int32 0.53, int32 with 2 in flight 0.465; pipe 0.64 - 0.69)
E2E_AUTO = {1: ("zero_copy", "int32", 2, 3), 2: ("zero_copy", "int32", 2, 3), 4: ("pipe", "int32", 1, 3), 8: ("pipe", "int32", 1, 3)}
E2E_VARIANTS = {
    "zero_copy_int64": dict(mode="zero_copy", ids="int64", inflight=1),
    "zero_copy_int32": dict(mode="zero_copy", ids="int32", inflight=1),
    "zero_copy_int32_3_in_flight": dict(mode="zero_copy", ids="int32", inflight=3),
    "pipe_int64_3_groups": dict(mode="pipe", ids="int64", inflight=1, groups=3),
    "pipe_int32_3_groups": dict(mode="pipe", ids="int32", inflight=1, groups=3),
    "pipe_int32_1_group_2_in_flight": dict(mode="pipe", ids="int32", inflight=2, groups=1),
}
N_BATCHES = 16  # distinct pre-generated batches the timed loop rotates through
NVLINK_PEAK_GBS = 770.0  # B200_PROFILING.md: measured peer copy, per direction per GPU


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def _traffic_from_profiles(workload):
    """dram bytes per launch from the committed ncu summary, or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f).get(workload)
    return None


class ClockSampler:
    """Samples SM clock + throttle reasons via NVML while the timed region runs."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml as nv
            nv.nvmlInit()
            self.nv = nv
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and vis.split(",")[index].isdigit() else index
            self.h = nv.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def sample(self):
        if not self.nv:
            return
        nv = self.nv
        try:
            self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
            r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
                     0x4: "sw_power_cap", 0x80: "hw_power_brake", 0x2: "applications_clocks_setting",
                     0x100: "display_clock_setting", 0x10: "sync_boost"}
            for bit, name in names.items():
                if r & bit:
                    self.reasons.add(name)
        except Exception:
            pass

    def start(self):
        def loop():
            while not self._stop.is_set():
                self.sample()
                time.sleep(0.002)
        self._t = threading.Thread(target=loop, daemon=True)
        self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def _env_rank():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def _workload_name(cfg):
    return (f"{cfg.name}: {cfg.slots} slots, {'shared' if cfg.shared_table else 'per-slot'} {cfg.rows} x {cfg.dim} "
            f"{cfg.dtype} table, batch {cfg.batch}, Zipf(1.05) ids, mean {cfg.mean_len} ids/seq, "
            f"{'/'.join(sorted(set(cfg.pooltypes)))} pool"
            f"{' + CVM' if cfg.use_cvm else (' + CVM(strip)' if cfg.use_cvm is False else '')}")


def _config_block(cfg, world, parallelism):
    """The `config` object: the same keys in both arms, so the driver can compare them."""
    return {"workload": _workload_name(cfg), "global_batch": cfg.batch * world, "parallelism": parallelism}


def _pinned(cfg):
    """How far the reference pins this workload's results."""
    if cfg.dtype != "f32":
        return "restatement (the reference registers pool / cvm for float and double only: 16-bit pooling is defined here)"
    return "reference (refer.h kernels compiled from the reference tree + its own test formulas)"


# --------------------------------------------------------------------------------------
# CPU leg (cpu_baseline of our arm, and the whole of --impl reference).  The only place in
# this file that TIMES anything under oracle/ -- the reference's CPU arithmetic for this path.
# --------------------------------------------------------------------------------------
def host_table_and_batches(cfg, synth, n_batches, max_host_gb=24.0):
    """Host tables generated on the host cores (benchtools/synth_host.c, bit-identical to the
    device generator), with rows folded while they would not fit, and host batches whose ids
    index them.  No GPU involved."""
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:
        avail = 32 << 30
    limit = min(max_host_gb * (1 << 30), avail * 0.5)
    rows = cfg.rows
    while cfg.n_tables() * rows * cfg.dim * cfg.elem_size > limit:
        rows //= 2
    tabs = [synth.fill_table_host(rows, cfg.dim, cfg.dtype, cfg.seed + 1000 * t, cvm_cols=cfg.use_cvm is not None)
            for t in range(cfg.n_tables())]
    batches = []
    for b in range(n_batches):
        bt = synth.draw_batch(cfg, step=b)
        if rows != cfg.rows:
            bt.ids = [i % rows for i in bt.ids]
        batches.append(bt)
    return tabs, batches, rows


def cpu_leg(cfg, table_host, batches, budget_s=12.0, threads=None):
    """Times the reference CPU arithmetic on whole batches of the workload, ~budget_s of work in
    all.  Variants: `refer_fused` = the reference's refer::EmbSeqPool
    compiled from /root/reference (oracle/_ref) + pad / CVM glue, when the workload is the fp32
    SUM + CVM shape it covers; `port_fused` = the oracle's fused gather + pool + CVM (-O3 -mavx2,
    mirrors jit EmbSeqPool + fusion_seqpool_cvm_concat); `port_unfused` = gather -> pool -> cvm
    with the [K, D] and [B, D] intermediates (mirrors the three separate CPU kernels).  Each at
    all host threads and at one.  `value` = the fastest all-thread variant (kind "reference" when
    that is the reference's own kernel, else "port")."""
    from oracle import oracle as O
    threads = threads or (os.cpu_count() or 1)
    R = O.ref_lib()
    use_ref = (R is not None and cfg.dtype == "f32" and cfg.use_cvm is True and cfg.shared_table
               and all(p == "SUM" for p in cfg.pooltypes))
    S, Bsz, D = cfg.slots, cfg.batch, cfg.dim
    out = np.empty((Bsz, S * cfg.d_out), dtype=table_host[0].dtype)
    n_rows = table_host[0].shape[0]
    cvm_mode = O.CVM_NONE if cfg.use_cvm is None else (O.CVM_USE if cfg.use_cvm else O.CVM_STRIP)

    def run(kind, batch, nthreads):
        if kind == "refer_fused":
            t = table_host[0]
            for s in range(S):
                rc = R.ref_fused_lod_f32(t.ctypes.data, n_rows, D, batch.ids[s].ctypes.data, batch.lods[s].ctypes.data,
                                         Bsz, 0.0, 1, out.ctypes.data + s * D * 4, S * D, nthreads)
                assert rc == 0
        else:
            O.fused_forward(table_host, batch.ids, batch.lods, cfg.pooltypes, cvm_mode=cvm_mode,
                            bf16=cfg.dtype == "bf16", concat=True, threads=nthreads, fast=True,
                            unfused=kind == "port_unfused")

    def timed(kind, nthreads, budget):
        run(kind, batches[0], nthreads)  # warm
        n, t0 = 0, time.perf_counter()
        while True:
            run(kind, batches[n % len(batches)], nthreads)
            n += 1
            el = time.perf_counter() - t0
            if el >= budget or n >= 20000:
                return {"value": S * Bsz * n / el, "batches": n, "seconds": round(el, 3), "threads": nthreads}

    kinds = (["refer_fused"] if use_ref else []) + ["port_fused", "port_unfused"]
    per = budget_s / (2.0 * len(kinds))
    variants = {}
    for k in kinds:
        variants[k] = {"all_threads": timed(k, threads, per * 1.2), "one_thread": timed(k, 1, per * 0.8)}
    best = max(kinds, key=lambda k: variants[k]["all_threads"]["value"])
    headline = best  # the faster of the reference's own kernels and the AVX2 port is the baseline quoted
    v = variants[headline]["all_threads"]
    return {"value": v["value"], "unit": UNIT, "cores": threads,
            "kind": "reference" if headline == "refer_fused" else "port",
            "sample": (f"{v['batches']} full batches ({S} slots x {Bsz} seqs) in {v['seconds']:.2f}s at {threads} threads; "
                       + ("refer::EmbSeqPool from the reference tree (oracle/_ref) + pad/CVM glue"
                          if headline == "refer_fused" else f"oracle port {headline} (oracle/ctr_oracle.c, -O3 -mavx2)")
                       + (f"; table folded to {n_rows} rows" if n_rows != cfg.rows else "")),
            "value_1thread": variants[headline]["one_thread"]["value"],
            "fastest_variant": best, "variants": variants}


def run_reference(args):
    """The reference's own CPU implementation of the path on this box's host cores.  No GPU, no
    libctrfe.so: the table is generated on the host."""
    rank, _local_rank, _world = _env_rank()
    if rank != 0:
        return 0
    from benchtools import synth
    from oracle import oracle as O
    O.build()
    cfg = synth.configs()[args.workload]
    tabs, batches, rows = host_table_and_batches(cfg, synth, 4)
    # a step = one full batch; the run is bounded by a time budget scaled to steps + warmup
    base = cpu_leg(cfg, tabs, batches, budget_s=min(90.0, max(12.0, 0.3 * (args.steps + args.warmup))))
    value = base["value"]
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * cfg.slots * cfg.batch / value,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": cfg.dtype,
            "data": "synthetic",
            "config": _config_block(cfg, 1, f"host threads x{base['cores']} (the reference's CPU kernels; rank 0 only)"),
            "cpu_baseline": base,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), file=_JSON_OUT or sys.stdout, flush=True)
    return 0


# --------------------------------------------------------------------------------------
# parity: a sample of one timed batch against the CPU oracle (the checker, never the thing timed)
# --------------------------------------------------------------------------------------
def parity_check(cfg, batch, got, n_tab, sample=48, rng_seed=1):
    """got: host array [Bsz, S, d_out] (uint16 bits for bf16) of the batch's pooled rows.  For a
    sample of sequences of every slot the oracle pools exactly the rows they touch, regenerated
    on the host from the table's (seed, row, column) hash.  Tolerances as in tests/: pooled
    columns bit-exact; the two CVM log columns rtol 2e-7 / 1e-6 (+1e-6 abs) in fp32."""
    from benchtools import synth
    from oracle import oracle as O
    Bsz, S = got.shape[0], got.shape[1]
    rng = np.random.default_rng(rng_seed)
    pick = np.arange(Bsz) if sample >= Bsz else np.sort(rng.choice(Bsz, sample, replace=False))
    cvm_mode = O.CVM_NONE if cfg.use_cvm is None else (O.CVM_USE if cfg.use_cvm else O.CVM_STRIP)
    exact, max_rel, checked, ok = True, 0.0, 0, True
    for s in range(S):
        lod = batch.lods[s].astype(np.int64)
        seg_ids = [batch.ids[s][lod[i]:lod[i + 1]] for i in pick]
        flat = np.concatenate(seg_ids) if seg_ids else np.empty(0, np.int64)
        uniq, inv = np.unique(flat, return_inverse=True)
        compact = synth.table_rows_numpy(cfg.seed + (1000 * s if n_tab > 1 else 0), uniq, cfg.dim, cfg.dtype,
                                         cvm_cols=cfg.use_cvm is not None)
        if uniq.size == 0:
            compact = np.zeros((1, cfg.dim), compact.dtype)
        sub_lod = np.concatenate([[0], np.cumsum([len(x) for x in seg_ids])]).astype(np.uint64)
        want = O.fused_forward([compact], [inv.astype(np.int64)], [sub_lod], [cfg.pooltypes[s]], cvm_mode=cvm_mode,
                               bf16=cfg.dtype == "bf16")[0]
        g = got[pick, s, :]
        checked += len(pick)
        if cfg.dtype == "f32":
            c0 = 2 if cfg.use_cvm is True else 0
            if not np.array_equal(g[:, c0:], want[:, c0:]):
                exact = False
            with np.errstate(divide="ignore", invalid="ignore"):
                rel = np.abs(g.astype(np.float64) - want) / np.maximum(np.abs(want), 1e-6)
            max_rel = max(max_rel, float(np.nanmax(rel)) if rel.size else 0.0)
            if cfg.use_cvm is True:
                ok &= bool(np.allclose(g[:, 0], want[:, 0], rtol=2e-7, atol=0) and
                           np.allclose(g[:, 1], want[:, 1], rtol=1e-6, atol=1e-6))
            ok &= bool(np.array_equal(g[:, c0:], want[:, c0:]))
        else:
            same = np.array_equal(g.view(np.uint16), want.view(np.uint16))
            exact &= same
            if not same:
                gf = O.bf16_bits_to_f32(g) if cfg.dtype == "bf16" else g.astype(np.float32)
                wf = O.bf16_bits_to_f32(want) if cfg.dtype == "bf16" else want.astype(np.float32)
                rel = np.abs(gf - wf) / np.maximum(np.abs(wf), 1e-3)
                max_rel = max(max_rel, float(rel.max()))
                ok &= bool(max_rel <= 1e-2)
    return {"checked": int(checked), "slots": int(S), "max_rel_err": max_rel, "bit_exact": bool(exact), "ok": bool(ok),
            "bit_exact_means": ("pooled columns identical to the oracle's bits; the two CVM log columns within 1 ulp / 1e-6"
                                if cfg.use_cvm is True and cfg.dtype == "f32" else "every output bit identical to the oracle's"),
            "pinned": _pinned(cfg)}


def _out_to_host(torch, out, Bsz, S, d_out):
    if out.dtype == torch.bfloat16:
        return out.view(torch.int16).cpu().numpy().view(np.uint16).reshape(Bsz, S, d_out)
    return out.cpu().numpy().reshape(Bsz, S, d_out)


# --------------------------------------------------------------------------------------
# device-resident measurement of one workload on this rank
# --------------------------------------------------------------------------------------
def measure_device(torch, ops, synth, cfg, W, Bsz, *, rank, world, steps, warmup, regions, dist=None, overlap=True,
                   n_tab=1, batch_seed0=0, sample=48, do_8d=True, traffic_key=None, sampler=None, hot_cache=None):
    """Plans for n_batches resident batches (ids, LoD, outputs in HBM), `regions` timed regions of
    `steps` launches each (one ctrfe_plan_launch_many call per region), the same launches one at a
    time, the SURVEY 8(d) step (host LoD uploaded per step), and a parity sample of the last timed
    batch.  Returns a dict; `ms_region_max` is the max over ranks of the per-rank median region."""
    from paddle_infer_b200 import _lib as B
    dev = torch.device("cuda", torch.cuda.current_device())
    S, D, es = cfg.slots, cfg.dim, cfg.elem_size
    n_batches = 4 if Bsz * S * D * es > (256 << 20) else N_BATCHES
    host_batches = [synth.draw_batch(cfg, step=batch_seed0 + b, batch=Bsz) for b in range(n_batches)]
    plans, plans_h, k_tots = [], [], []
    for hb in host_batches:
        ids_d = [torch.from_numpy(i).to(dev) for i in hb.ids]
        lods_d = [torch.from_numpy(l.astype(np.int64)).to(dev) for l in hb.lods]
        plans.append(ops.ctr_fused_emb_seqpool_cvm(W, ids_d, lods_d, cfg.pooltypes, use_cvm=cfg.use_cvm, concat=True,
                                                   plan_only=True, hot_cache=hot_cache))
        if do_8d:  # the same batch with its LoD offsets on the host (uploaded by the library)
            plans_h.append(ops.ctr_fused_emb_seqpool_cvm(W, ids_d, hb.lods, cfg.pooltypes, use_cvm=cfg.use_cvm, concat=True,
                                                         plan_only=True, out=plans[-1].outs, hot_cache=hot_cache))
        k_tots.append(hb.k_tot)
    torch.cuda.synchronize()
    ctx = plans[0].ctx
    stream = torch.cuda.current_stream(dev)
    K, Wm = steps, max(warmup, 3)
    flags = B.LAUNCH_OVERLAP_PREVIOUS if overlap else 0

    def handles(ps, n, first=0):
        return (C.c_void_p * n)(*[ps[(first + i) % len(ps)].handle.value for i in range(n)])

    def launch(arr, n, fl):
        B.check(B.lib().ctrfe_plan_launch_many(arr, n, stream.cuda_stream, fl))

    launch(handles(plans, Wm), Wm, flags)
    ctx.check_ids()  # also a sync
    arrK = handles(plans, K, Wm)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(regions)]
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    if sampler is not None:
        sampler.start()
    for e0, e1 in evs:
        e0.record(stream)
        launch(arrK, K, flags)
        e1.record(stream)
    if sampler is not None:
        sampler.sample()
    torch.cuda.synchronize()
    if sampler is not None:
        sampler.stop()
    if dist is not None:
        dist.barrier()
    per_region = np.array([e0.elapsed_time(e1) for e0, e1 in evs])
    ms_region = float(np.median(per_region))
    t = torch.tensor([ms_region], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_region_max = float(t.item())
    ctx.check_ids()
    last = (Wm + K - 1) % n_batches  # the batch of the last timed launch

    # the same launches one at a time (an event between every two, no overlap)
    Ki = min(K, 200)
    iso = [torch.cuda.Event(enable_timing=True) for _ in range(Ki + 1)]
    iso[0].record(stream)
    for i in range(Ki):
        plans[(Wm + i) % n_batches].launch(False)
        iso[i + 1].record(stream)
    torch.cuda.synchronize()
    per = np.array([iso[i].elapsed_time(iso[i + 1]) for i in range(Ki)])

    out = {"ms_region": ms_region, "ms_region_max": ms_region_max, "ms_per_step": ms_region_max / K,
           "regions": regions, "region_ms_min_max": [float(per_region.min()), float(per_region.max())],
           "kernel_ms_isolated_mean": float(per.mean()), "kernel_ms_isolated_median": float(np.median(per)),
           "n_batches": n_batches, "launches_per_step": plans[0].num_kernels(), "kernels": plans[0].kernel_names()}
    k_used = [k_tots[(Wm + i) % n_batches] for i in range(K)]
    out["k_mean"] = float(np.mean(k_used))
    out["alg_bytes"] = float(np.mean([k * (8 + D * es) + S * (Bsz + 1) * 8 + S * Bsz * cfg.d_out * es for k in k_used]))

    if do_8d:
        # "ids and LoD already in pinned host memory; the one async LoD upload per step is
        # included".  Isolated: upload + launch, an event either side, one step at a time.  Back to back:
        # the uploads ride the library's copy stream beside the previous batch's kernel.
        fl8 = flags | B.LAUNCH_UPLOAD_LOD
        launch(handles(plans_h, min(Wm, n_batches)), min(Wm, n_batches), fl8)
        torch.cuda.synchronize()
        K8 = min(K, 100)
        e8 = [torch.cuda.Event(enable_timing=True) for _ in range(2 * K8)]
        for i in range(K8):
            one = handles(plans_h, 1, Wm + i)
            e8[2 * i].record(stream)
            launch(one, 1, B.LAUNCH_UPLOAD_LOD)
            e8[2 * i + 1].record(stream)
            torch.cuda.synchronize()
        iso8 = np.array([e8[2 * i].elapsed_time(e8[2 * i + 1]) for i in range(K8)])
        arr8 = handles(plans_h, K, Wm)
        b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        b0.record(stream)
        launch(arr8, K, fl8)
        b1.record(stream)
        torch.cuda.synchronize()
        lod_bytes = int(sum((((l.size + 2046) // 2048 * 4 + 15) // 16) * 16 + (((l.size - 1) * 2 + 15) // 16) * 16
                            for l in host_batches[0].lods))
        out["survey_8d"] = {
            "isolated_ms_median": float(np.median(iso8)), "isolated_value": S * Bsz * world / (float(np.median(iso8)) * 1e-3),
            "back_to_back_ms_per_step": b0.elapsed_time(b1) / K, "back_to_back_value": S * Bsz * world * K / (b0.elapsed_time(b1) * 1e-3),
            "lod_upload_bytes_per_step": lod_bytes,
            "what": "per step: the batch's LoD offsets (validated on the host beforehand; pinned image of 16-bit sequence "
                    "lengths, expanded to offsets by a small kernel on the copy stream) -> ONE cudaMemcpyAsync -> the "
                    "plan's launch; ids resident in HBM.  isolated = CUDA events around "
                    "upload + launch, one step at a time, median; back to back = K steps in one ctrfe_plan_launch_many "
                    "call, uploads on the library's copy stream beside the previous batch's kernel"}
        ctx.check_ids()

    # parity of the last timed batch (its plan's output block is still what that launch wrote)
    got = _out_to_host(torch, plans[last].outs, Bsz, S, cfg.d_out)
    out["parity"] = parity_check(cfg, host_batches[last], got, n_tab, sample=sample)
    peak, peak_src = _peaks()
    kern_ms = ms_region / K
    achieved = out["alg_bytes"] / (kern_ms * 1e-3) / 1e9
    iso_ach = out["alg_bytes"] / (out["kernel_ms_isolated_median"] * 1e-3) / 1e9
    out["roofline"] = {
        "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
        "traffic": _traffic_from_profiles(traffic_key) if traffic_key else None, "peak_source": peak_src,
        "kernel": out["kernels"][0] if len(out["kernels"]) == 1 else out["kernels"],
        "frac_uses": "ms per step of the timed regions (back-to-back launches, each allowed to start while the previous "
                     "one drains)" if overlap else "ms per step of the timed regions (stream-ordered launches)",
        "algorithmic_bytes_per_launch": out["alg_bytes"], "kernel_ms_mean": kern_ms,
        "kernel_ms_isolated_mean": out["kernel_ms_isolated_mean"], "kernel_ms_isolated_median": out["kernel_ms_isolated_median"],
        "achieved_isolated": iso_ach, "frac_isolated": iso_ach / peak, "frac_of_nominal_8000": achieved / 8000.0,
        "launch_overlap": overlap}
    tr = out["roofline"]["traffic"]
    if tr:  # what actually crossed the DRAM pins (ncu, cold-cache launch) against the same clock
        out["roofline"]["dram_gbs"] = tr / (kern_ms * 1e-3) / 1e9
        out["roofline"]["dram_frac"] = tr / (kern_ms * 1e-3) / 1e9 / peak
        out["roofline"]["traffic_over_algorithmic"] = tr / out["alg_bytes"]
        out["roofline"]["traffic_note"] = ("dram__bytes_read + dram__bytes_write per step of the committed ncu capture "
                                           "(profiles/r2_kernels_ncu.csv, cold cache); below the algorithmic bytes because "
                                           "the Zipf head is served by L1 / L2, so frac can exceed what DRAM alone could give")
    for p in plans_h:
        p.close()
    out["_plans"], out["_host_batches"], out["_ctx"] = plans, host_batches, ctx
    return out


def _free_measure(m):
    """Drop the plans (and with them the batches' device tensors) of a measurement."""
    for p in m.pop("_plans", []):
        p.close()
    m.pop("_host_batches", None)
    m.pop("_ctx", None)


def make_tables(torch, synth, cfg, dev):
    tdt = {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}[cfg.dtype]
    tables = []
    for t in range(cfg.n_tables()):
        tab = torch.empty((cfg.rows, cfg.dim), dtype=tdt, device=dev)
        synth.fill_table(tab, cfg.seed + 1000 * t, cvm_cols=cfg.use_cvm is not None)
        tables.append(tab)
    torch.cuda.synchronize()
    return tables


def config_record(torch, ops, synth, name, dev, args, traffic_key=None, hot_rows=0):
    """One BASELINE config at full size on this GPU: value, per-step times, roofline, parity.
    hot_rows > 0 (single shared table): the same run again with that many hot-row replicas in a
    compact region (shard.HotRowCache beside the unsharded table), reported under `with_hot_rows`."""
    cfg = synth.configs()[name]
    need = cfg.n_tables() * cfg.rows * cfg.dim * cfg.elem_size + (6 << 30)
    torch.cuda.empty_cache()
    free, _ = torch.cuda.mem_get_info()
    if free < need:
        return {"workload": _workload_name(cfg), "skipped": f"needs {need / 1e9:.0f} GB of HBM, {free / 1e9:.0f} GB free"}
    tables = make_tables(torch, synth, cfg, dev)
    W = tables if len(tables) > 1 else tables[0]
    steps = max(20, min(args.steps, int(0.25 / max(cfg.algorithmic_bytes(int(cfg.batch * cfg.slots * cfg.mean_len)) / 4e12, 1e-6))))
    m = measure_device(torch, ops, synth, cfg, W, cfg.batch, rank=0, world=1, steps=steps, warmup=max(args.warmup, 3),
                       regions=max(3, args.regions // 3), n_tab=len(tables), traffic_key=traffic_key or name)
    hot_rec = None
    if hot_rows and len(tables) == 1:
        from paddle_infer_b200 import shard as shard_mod
        sample = [synth.draw_batch(cfg, step=500 + b) for b in range(max(1, args.hot_sample_batches // 8))]
        sample_ids = torch.cat([torch.from_numpy(i).to(dev) for sb in sample for i in sb.ids])
        hot_ids = shard_mod.pick_hot_ids(sample_ids, hot_rows)
        del sample, sample_ids
        hc = shard_mod.HotRowCache(tables[0], hot_ids)
        mh = measure_device(torch, ops, synth, cfg, W, cfg.batch, rank=0, world=1, steps=steps, warmup=max(args.warmup, 3),
                            regions=max(3, args.regions // 3), n_tab=1, do_8d=False, hot_cache=hc)
        hit = float(np.mean([np.isin(np.concatenate(hb.ids), hot_ids.cpu().numpy()).mean() for hb in mh["_host_batches"]]))
        hot_rec = {"value": cfg.slots * cfg.batch / (mh["ms_per_step"] * 1e-3), "ms_per_step": mh["ms_per_step"],
                   "ms_per_step_isolated_median": mh["kernel_ms_isolated_median"], "kernels": mh["kernels"],
                   "frac": mh["roofline"]["frac"], "parity": mh["parity"],
                   "hot_row_cache": {"rows": int(hot_ids.numel()), "bytes": int(hc.nbytes), "hit_rate_on_timed_batches": hit,
                                     "built_from": "sample batches disjoint from the timed ones"},
                   "why": "the table is beyond the TLB's reach: every cold row costs a page walk; the replicas of the hottest "
                          "rows sit in a compact region the TLB and L2 hold (bit-identical results)"}
        _free_measure(mh)
        del mh, hc
    rec = {"workload": _workload_name(cfg), "value": cfg.slots * cfg.batch / (m["ms_per_step"] * 1e-3), "unit": UNIT,
           "dtype": cfg.dtype, "steps": steps, "regions": m["regions"], "ms_per_step": m["ms_per_step"],
           "ms_per_step_isolated_median": m["kernel_ms_isolated_median"], "launches_per_step": m["launches_per_step"],
           "roofline": m["roofline"], "parity": m["parity"], "survey_8d": m.get("survey_8d"),
           "table_gb": cfg.n_tables() * cfg.rows * cfg.dim * cfg.elem_size / 1e9}
    if hot_rec is not None:
        rec["with_hot_rows"] = hot_rec
    _free_measure(m)
    del tables, W, m
    torch.cuda.empty_cache()
    return rec


def _e2e_variant(torch, ops, cfg, W, m, Bsz, world, dist, args, dev, *, mode, ids="int64", inflight=1, groups=3, ramp=False,
                 direct_in=False, direct_out=False, Ke=40, sync_ranks=True):
    """One timed loop of host-buffer requests: every request reads that step's ids + LoD offsets from
    pinned host memory and lands the pooled rows in pinned host memory.  `inflight` requests are in
    flight at a time (1: a request is collected before the next is submitted)."""
    from collections import deque
    S = cfg.slots
    host_batches, ctx = m["_host_batches"], m["_ctx"]
    n_batches = len(host_batches)
    stream = torch.cuda.current_stream(dev)
    max_ids = max(int(i.size) for hb in host_batches for i in hb.ids)
    id_np = np.int32 if ids == "int32" else np.int64
    F = max(1, int(inflight))
    if mode == "pipe":
        units = [ops.HostPipe(W, S, Bsz, max_ids, cfg.pooltypes, use_cvm=cfg.use_cvm, ctx=ctx, n_groups=groups, direct_out=direct_out,
                              direct_in=direct_in, ramp=ramp, ids_dtype=ids) for _ in range(F)]
        # a request's ids sit in ONE pinned feed buffer, slot after slot (per-slot views of it), its LoD
        # offsets in another; the pointer tables are built once per buffer set, as a serving loop would
        preps, keep = [], []
        for hb in host_batches:
            feed = torch.from_numpy(np.concatenate(hb.ids).astype(id_np)).pin_memory()
            offs = np.concatenate([[0], np.cumsum([i.size for i in hb.ids])])
            lod_feed = torch.from_numpy(np.stack([l.astype(np.int64) for l in hb.lods])).pin_memory()
            keep.append((feed, lod_feed))
            preps.append(units[0].prepare([feed[int(offs[s]):int(offs[s + 1])] for s in range(S)], [lod_feed[s] for s in range(S)]))

        def submit(i):
            u = units[i % F]
            u.run_prepared(preps[i % n_batches], wait=None)
            return u

        def collect(u):
            u.wait()

        def h2d_of(i):
            return preps[i % n_batches][4]

        def result_of(i):
            return [units[i % F].host_out[s] for s in range(S)]
    else:
        units = [ops.HostFrontEnd(W, S, Bsz, max_ids, cfg.pooltypes, use_cvm=cfg.use_cvm, ctx=ctx, mode=mode) for _ in range(F)]
        pinned = [[torch.from_numpy(i.astype(id_np)).pin_memory() for i in hb.ids] for hb in host_batches]
        if mode == "zero_copy":
            lods_p = [[torch.from_numpy(l.astype(np.int64)).pin_memory() for l in hb.lods] for hb in host_batches]
            calls = [[units[j].bind(pinned[b], lods_p[b]) for j in range(F)] for b in range(n_batches)]

            def submit(i):
                c = calls[i % n_batches][i % F]
                return (c, c.submit(overlap_previous=F > 1))

            def collect(h):
                h[0].result(h[1])

            def h2d_of(i):
                return calls[i % n_batches][0].h2d_bytes
        else:
            def submit(i):
                units[0].run(pinned[i % n_batches], host_batches[i % n_batches].lods)
                return None

            def collect(h):
                pass

            def h2d_of(i):
                b = i % n_batches
                return sum(t.numel() * t.element_size() for t in pinned[b]) + sum(l.size * 8 for l in host_batches[b].lods)

        def result_of(i):
            return units[i % F]._result()

    def loop(n):
        pending = deque()
        for i in range(n):
            if len(pending) >= F:
                collect(pending.popleft())
            pending.append(submit(i))
        while pending:
            collect(pending.popleft())

    loop(max(3, 2 * F))
    if dist is not None and sync_ranks:
        dist.barrier()
    torch.cuda.synchronize()
    t_host0 = time.perf_counter()
    loop(Ke)
    torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - t_host0) * 1e3   # requests are collected on the host: host time is the truth
    te = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
    if dist is not None and sync_ranks:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    h2d = sum(h2d_of(i) for i in range(Ke))
    d2h = units[0].d2h_bytes * Ke
    # the last request's rows against the oracle: the host-buffer path is checked too
    res = result_of(Ke - 1)
    got = np.stack([(r.view(torch.int16).numpy().view(np.uint16) if r.dtype == torch.bfloat16 else r.numpy()) for r in res], axis=1)
    par = parity_check(cfg, host_batches[(Ke - 1) % n_batches], got, 1 if cfg.shared_table else S, sample=16)
    ms = float(te.item()) / Ke
    if mode == "pipe":
        path = (f"HostPipe (ctrfe_host_pipe_run, {groups} slot group(s){', ramped' if ramp else ''}): "
                + ("the kernels read the pinned ids / 64-bit offsets themselves" if direct_in else
                   "ids and 64-bit offsets go up on one copy engine (cudaMemcpyAsync from pinned feed buffers)")
                + ", the groups are pooled, "
                + ("the kernels write the rows into pinned host memory" if direct_out else "the rows come down on the other copy engine")
                + "; LoD pairs checked by the kernels")
    elif mode == "zero_copy":
        path = ("HostFrontEnd zero_copy (buffers bound once): the kernel reads pinned host ids / LoD and writes the pinned per-slot "
                "host outputs over PCIe inside the one launch; LoD pairs checked by the kernel")
    else:
        path = "HostFrontEnd copy: H2D ids + host-validated LoD upload, launch, D2H output"
    path += (f"; {F} requests in flight (request i + 1 is submitted before request i is collected; every request still "
             "reads its inputs from and lands its rows in pinned host memory)" if F > 1 else
             "; one request at a time (collected before the next is submitted)")
    rec = {"value": S * Bsz * world / (ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d // Ke, "d2h_bytes_per_step": d2h // Ke,
           "steps": Ke, "ms_per_step": ms, "parity": par, "pcie_gbs_per_gpu": (h2d + d2h) / Ke / (ms * 1e-3) / 1e9,
           "mode": mode, "ids_dtype": ids, "requests_in_flight": F, "path": path}
    for u in units:
        if hasattr(u, "close"):
            u.close()
    return rec


def e2e_leg(torch, ops, cfg, W, m, Bsz, world, dist, args, dev):
    """The same metric through the host-buffer call (`e2e`), plus -- N = 1, full run -- the other forms
    of that call measured the same way (`e2e.variants`)."""
    Ke = max(3, min(args.steps, args.e2e_steps))
    common = dict(Ke=Ke, groups=args.e2e_groups, ramp=args.e2e_ramp, direct_in=args.e2e_direct_in, direct_out=args.e2e_direct_out)
    mode, ids, inflight = args.e2e_mode, args.e2e_ids, args.e2e_inflight
    if mode == "auto":
        mode, ids, inflight, common["groups"] = E2E_AUTO[min(world, 8)] if world in E2E_AUTO else E2E_AUTO[8]
    rec = _e2e_variant(torch, ops, cfg, W, m, Bsz, world, dist, args, dev, mode=mode, ids=ids, inflight=inflight, **common)
    if world == 1 and (not args.quick or args.e2e_variants):
        variants = {}
        for name, kw in E2E_VARIANTS.items():
            try:
                v = _e2e_variant(torch, ops, cfg, W, m, Bsz, world, dist, args, dev, Ke=Ke, sync_ranks=False, **kw)
                variants[name] = {k: v[k] for k in ("ms_per_step", "value", "h2d_bytes_per_step", "d2h_bytes_per_step", "mode",
                                                    "ids_dtype", "requests_in_flight")}
                variants[name]["parity_ok"] = v["parity"]["ok"]
            except Exception as ex:  # a secondary figure must not cost the run its headline
                variants[name] = {"error": repr(ex)[:200]}
        # the unbound shape of the call: fresh pageable numpy arrays every request, pointer tables built
        # per call (ZeroCopyTensor::CopyFromCpu takes pageable user memory too)
        try:
            S, hbs = cfg.slots, m["_host_batches"]
            pipe = ops.HostPipe(W, S, Bsz, max(int(i.size) for hb in hbs for i in hb.ids), cfg.pooltypes, use_cvm=cfg.use_cvm,
                                ctx=m["_ctx"], n_groups=3)
            nu = max(3, Ke // 4)
            pipe.run([np.array(x) for x in hbs[0].ids], hbs[0].lods)
            t0u = time.perf_counter()
            for i in range(nu):
                hbu = hbs[i % len(hbs)]
                pipe.run([np.array(x) for x in hbu.ids], hbu.lods)
            variants["pipe_unbound_pageable"] = {"ms_per_step": (time.perf_counter() - t0u) * 1e3 / nu, "mode": "pipe",
                                                 "note": "fresh pageable numpy arrays and pointer tables every request"}
            pipe.close()
        except Exception as ex:
            variants["pipe_unbound_pageable"] = {"error": repr(ex)[:200]}
        rec["variants"] = variants
    return rec


def sharded_cfg4_block(torch, ops, synth, dist, rank, world, dev, args):
    """BASELINE config 4 on `world` GPUs (strong scaling): one logical 1e9 x 32 fp32 table row-sharded
    over the ranks (row r on GPU r % world), the global batch split over the ranks, remote rows
    fetched by the pooling kernel itself through NVLink peer mappings.  Measured with hot-row
    replicas of two sizes and without any; the one-GPU figure (whole table on rank 0) is measured
    in the same run, so the speed-up needs no other record."""
    from paddle_infer_b200 import shard as shard_mod
    cfg = synth.configs()["cfg4"]
    S, D = cfg.slots, cfg.dim
    steps = max(20, min(args.steps, 200))
    block = {"workload": _workload_name(cfg), "n_gpus": world, "scaling": "strong", "unit": UNIT}
    # ---- the denominator: the whole table on one GPU, same global batch
    one = None
    if rank == 0:
        one = config_record(torch, ops, synth, "cfg4", dev, args, hot_rows=min(h for h in args.hot_rows_list if h) if any(args.hot_rows_list) else 0)
    dist.barrier()
    one_l = [one]
    dist.broadcast_object_list(one_l, src=0)
    one = one_l[0]
    block["one_gpu"] = {k: one.get(k) for k in ("value", "ms_per_step", "ms_per_step_isolated_median", "skipped")}
    one_hot = one.get("with_hot_rows")
    if one_hot:  # the same replicas beside the whole table on one GPU: the like-for-like denominator of the hot variants
        block["one_gpu"]["with_hot_rows"] = {k: one_hot.get(k) for k in ("value", "ms_per_step", "hot_row_cache")}
    # ---- the sharded table
    tdt = torch.float32
    Bsz = cfg.batch // world
    table = shard_mod.ShardedTable(cfg.rows, D, tdt, sharing=os.environ.get("CTRFE_SHARING", "vmm"))
    synth.fill_table(table.local, cfg.seed, row0=rank, row_stride=world, cvm_cols=True)
    torch.cuda.synchronize()
    dist.barrier()
    # hot set: the most frequent ids of sample traffic drawn from the same distribution but DISJOINT from
    # the timed batches (the same amount of sample traffic per rank whatever the world size)
    n_sample = max(1, int(round(args.hot_sample_batches * (cfg.batch / 8) / Bsz)))
    sample = [synth.draw_batch(cfg, step=rank * 1000 + 500 + b, batch=Bsz) for b in range(n_sample)]
    sample_ids = torch.cat([torch.from_numpy(i).to(dev) for sb in sample for i in sb.ids])
    hot_all = shard_mod.pick_hot_ids(sample_ids, max(args.hot_rows_list))
    del sample, sample_ids
    variants = {}
    for n_hot in args.hot_rows_list:
        key = f"hot_rows_{n_hot}" if n_hot else "no_replicas"
        table.hot = shard_mod.HotRowCache(table, hot_all[:n_hot].contiguous()) if n_hot else None
        dist.barrier()
        m = measure_device(torch, ops, synth, cfg, table, Bsz, rank=rank, world=world, steps=steps,
                           warmup=max(args.warmup, 3), regions=max(3, args.regions // 3), dist=dist, do_8d=False,
                           batch_seed0=rank * 1000, sample=24)
        okt = torch.tensor([1.0 if m["parity"]["ok"] else 0.0, 1.0 if m["parity"]["bit_exact"] else 0.0,
                            -m["parity"]["max_rel_err"]], dtype=torch.float64, device=dev)
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
        hit = 0.0
        if n_hot:
            hot_np = table.hot.hot_ids.cpu().numpy()
            hit = float(np.mean([np.isin(np.concatenate(hb.ids), hot_np).mean() for hb in m["_host_batches"]]))
        nv_bytes = m["k_mean"] * (1.0 - hit) * (world - 1) / world * D * cfg.elem_size
        ms = m["ms_per_step"]
        value = S * cfg.batch / (ms * 1e-3)
        rec = {"value": value, "ms_per_step": ms, "ms_per_step_isolated_median": m["kernel_ms_isolated_median"],
               "steps": steps, "regions": m["regions"], "kernels": m["kernels"],
               "speedup_vs_one_gpu": (value / one["value"]) if one.get("value") else None,
               "speedup_vs_one_gpu_with_the_same_replicas": (value / one_hot["value"]) if (one_hot and n_hot and
                                                             n_hot == one_hot["hot_row_cache"]["rows"]) else None,
               "parity": {"checked_per_rank": m["parity"]["checked"], "ok": bool(okt[0].item() > 0.5),
                          "bit_exact": bool(okt[1].item() > 0.5), "max_rel_err": float(-okt[2].item()),
                          "pinned": m["parity"]["pinned"], "note": "every rank checks a sample of ITS last timed batch "
                          "against the oracle; the worst rank is reported"},
               "nvlink": {"remote_row_bytes_per_step_per_gpu": nv_bytes, "requested_gbs": nv_bytes / (ms * 1e-3) / 1e9,
                          "link_peak_gbs": NVLINK_PEAK_GBS, "frac_of_link_peak": nv_bytes / (ms * 1e-3) / 1e9 / NVLINK_PEAK_GBS,
                          "peak_source": "B200_PROFILING.md measured peer copy, per direction per GPU",
                          "note": "bytes the kernel asks the peer mappings for / time; repeated hot rows can hit L1"},
               "hot_row_cache": ({"rows": int(n_hot), "bytes_per_gpu": int(table.hot.nbytes), "hit_rate_on_timed_batches": hit,
                                  "built_from": f"{n_sample} sample batches of {Bsz} per rank, disjoint from the timed batches"}
                                 if n_hot else None)}
        variants[key] = rec
        _free_measure(m)
        del m
        table.hot = None
        torch.cuda.empty_cache()
    # ---- the comparator: bucket ids by owner -> NCCL all-to-all -> owner gather -> rows back -> pool
    # (shard.RoutedForward, the exchange the north-star names), with and without per-step dedup
    n_rb = 4
    rb = [synth.draw_batch(cfg, step=rank * 1000 + b, batch=Bsz) for b in range(n_rb)]
    rb_ids = [[torch.from_numpy(i).to(dev) for i in hb.ids] for hb in rb]
    for dedup in (False, True):
        router = shard_mod.RoutedForward(table, dedup=dedup)
        outs = [torch.empty((Bsz, S * cfg.d_out), dtype=tdt, device=dev) for _ in range(n_rb)]
        for b in range(n_rb):
            router(rb_ids[b], rb[b].lods, cfg.pooltypes, use_cvm=cfg.use_cvm, concat=True, out=outs[b])
        torch.cuda.synchronize()
        dist.barrier()
        Kr = 12
        r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        r0.record()
        for i in range(Kr):
            b = i % n_rb
            router(rb_ids[b], rb[b].lods, cfg.pooltypes, use_cvm=cfg.use_cvm, concat=True, out=outs[b])
        r1.record()
        torch.cuda.synchronize()
        tr = torch.tensor([r0.elapsed_time(r1) / Kr], dtype=torch.float64, device=dev)
        dist.all_reduce(tr, op=dist.ReduceOp.MAX)
        # the same batch through the peer-load kernel: the two paths must agree to the bit
        direct = ops.ctr_fused_emb_seqpool_cvm(table, rb_ids[0], rb[0].lods, cfg.pooltypes, use_cvm=cfg.use_cvm, concat=True)
        same = torch.tensor([1.0 if torch.equal(direct, outs[0]) else 0.0], dtype=torch.float64, device=dev)
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        par = parity_check(cfg, rb[0], _out_to_host(torch, outs[0], Bsz, S, cfg.d_out), 1, sample=24)
        uniq = float(np.mean([np.unique(np.concatenate(hb.ids)).size / max(hb.k_tot, 1) for hb in rb]))
        msr = float(tr.item())
        variants["nccl_router_dedup" if dedup else "nccl_router"] = {
            "value": S * cfg.batch / (msr * 1e-3), "ms_per_step": msr, "steps": Kr,
            "speedup_vs_one_gpu": (S * cfg.batch / (msr * 1e-3) / one["value"]) if one.get("value") else None,
            "equal_to_peer_load_kernel": bool(same.item() > 0.5), "parity": {"ok": par["ok"], "bit_exact": par["bit_exact"]},
            "distinct_ids_fraction": uniq if dedup else None,
            "what": "ids bucketed by owner shard (device sort + scan), counts and ids exchanged with NCCL all-to-all, rows "
                    "gathered by the owner, returned with NCCL all-to-all, pooled by the requester in sequence order"
                    + ("; every distinct id crosses the link once (torch.unique)" if dedup else "")}
        del outs
    del rb_ids, rb
    torch.cuda.empty_cache()
    block["variants"] = variants
    peer = {k: v for k, v in variants.items() if not k.startswith("nccl_router")}
    best = max(peer, key=lambda k: peer[k]["value"])
    block.update({"value": variants[best]["value"], "ms_per_step": variants[best]["ms_per_step"], "best_variant": best,
                  "speedup_vs_one_gpu": variants[best]["speedup_vs_one_gpu"],
                  "speedup_vs_one_gpu_with_the_same_replicas": variants[best].get("speedup_vs_one_gpu_with_the_same_replicas"),
                  "speedup_note": "speedup_vs_one_gpu: against one GPU holding the whole 128 GB table without replicas (TLB-bound); "
                                  "..._with_the_same_replicas: against one GPU with the same hot-row replicas beside the table",
                  "parallelism": f"table row-sharded x{world} (row r on GPU r % {world}), global batch {cfg.batch} split x{world}, "
                                 "remote rows loaded by the pooling kernel through NVLink peer mappings; no data-path collective "
                                 "(the NCCL all-to-all router is measured beside it as variants.nccl_router*)"})
    table.close()
    return block


def run_ours(args):
    import torch
    import __graft_entry__ as g
    rank, local_rank, world = _env_rank()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU: paddle_infer_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    from paddle_infer_b200 import affinity
    prev_affinity = None if os.environ.get("CTRFE_NO_BIND") == "1" else affinity.bind_to_gpu(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    if rank == 0:
        g.build()
    if world > 1:
        dist.barrier()
    from benchtools import synth
    from paddle_infer_b200 import ops

    cfg = synth.configs()[args.workload]
    if args.rows:  # diagnostic override: same workload shape on a smaller / larger table
        cfg.rows = args.rows
        cfg.name += f"[rows={args.rows}]"
    S, Bsz = cfg.slots, cfg.batch
    tables = make_tables(torch, synth, cfg, dev)
    W = tables if len(tables) > 1 else tables[0]
    sampler = ClockSampler(local_rank)
    m = measure_device(torch, ops, synth, cfg, W, Bsz, rank=rank, world=world, steps=args.steps, warmup=args.warmup,
                       regions=args.regions, dist=dist, overlap=not args.no_overlap, n_tab=len(tables),
                       batch_seed0=rank * 1000, traffic_key=args.workload, sampler=sampler)
    K = args.steps
    value = S * Bsz * K * world / (m["ms_region_max"] * 1e-3)
    e2e = e2e_leg(torch, ops, cfg, W, m, Bsz, world, dist, args, dev)
    e2e["host_thread_bound_to_gpu_local_cpus"] = prev_affinity is not None
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": max(args.warmup, 3),
        "ms_per_step": m["ms_region_max"] / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": cfg.dtype, "data": "synthetic",
        "config": dict(_config_block(cfg, world, f"replica x{world} (every rank a full table and its own batches; "
                                                 "no data-path collective)"),
                       l2_hygiene=f"table {cfg.n_tables() * cfg.rows * cfg.dim * cfg.elem_size / 1e9:.1f} GB >> L2; the timed loop "
                                  f"rotates {m['n_batches']} distinct resident batches (ids+LoD+out "
                                  f"{m['n_batches'] * m['alg_bytes'] / 1e6:.0f} MB > 126 MB L2); hot table rows stay cached as in "
                                  "steady-state serving",
                       timing=f"{m['regions']} regions of {K} launches, each region one ctrfe_plan_launch_many call between two "
                              "CUDA events; value = the median region, max over ranks; "
                              + ("each launch may start while the previous one drains (CTRFE_LAUNCH_OVERLAP_PREVIOUS: the "
                                 "batches are independent)" if not args.no_overlap else "stream-ordered launches"),
                       lod="value: ids, LoD and outputs resident in HBM; survey_8d: LoD uploaded per step; e2e: host buffers"),
        "clocks": sampler.summary(),
        "gpu_launches": K * m["launches_per_step"] * m["regions"],
        "region_ms_min_max": m["region_ms_min_max"],
        "roofline": m["roofline"], "parity": m["parity"], "survey_8d": m.get("survey_8d"), "e2e": e2e,
    }
    _free_measure(m)
    del m, W, tables
    torch.cuda.empty_cache()
    affinity.restore(prev_affinity)  # everything below may use every host core
    if not args.quick:
        if world == 1:
            others = [n for n in ("cfg1", "cfg2", "cfg3", "cfg4", "cfg5") if n != args.workload]
            line["configs"] = {n: config_record(torch, ops, synth, n, dev, args, hot_rows=2_000_000 if n == "cfg4" else 0)
                               for n in others}
            c4 = line["configs"].get("cfg4")
            if c4 is not None:
                line["sharded_cfg4"] = {"workload": c4["workload"], "n_gpus": 1, "scaling": "strong", "unit": UNIT,
                                        "value": c4.get("value"), "ms_per_step": c4.get("ms_per_step"),
                                        "speedup_vs_one_gpu": 1.0 if c4.get("value") else None, "parity": c4.get("parity"),
                                        "parallelism": "the whole 128 GB table on one GPU (the strong-scaling denominator); "
                                                       "see configs.cfg4 for the roofline"}
        else:
            line["sharded_cfg4"] = sharded_cfg4_block(torch, ops, synth, dist, rank, world, dev, args)
    if rank == 0 and world == 1 and not args.no_cpu_baseline and not args.quick:
        tabs, hb4, _rows = host_table_and_batches(cfg, synth, 4)
        line["cpu_baseline"] = cpu_leg(cfg, tabs, hb4, budget_s=args.cpu_budget)
    elif rank == 0:
        line["cpu_baseline"] = None
    if rank == 0:
        print(json.dumps(line), file=_JSON_OUT or sys.stdout, flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


_JSON_OUT = None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=["cfg1", "cfg2", "cfg3", "cfg4", "cfg5"])
    ap.add_argument("--regions", type=int, default=15, help="timed regions of --steps launches; the median region is reported")
    ap.add_argument("--e2e-steps", type=int, default=40)
    ap.add_argument("--e2e-mode", default="auto", choices=["auto", "pipe", "zero_copy", "copy"],
                    help="auto: E2E_AUTO[n_gpus] (zero_copy, 32-bit ids, 2 requests in flight on 1-2 GPUs; the copy-engine pipe on "
                         "more); anything else takes --e2e-ids / --e2e-inflight / --e2e-groups as given")
    ap.add_argument("--e2e-groups", type=int, default=3, help="slot groups of the host-buffer pipe")
    ap.add_argument("--e2e-direct-out", action="store_true", help="host pipe: kernels write the rows into host memory themselves")
    ap.add_argument("--e2e-variants", action="store_true", help="measure e2e.variants even with --quick")
    ap.add_argument("--e2e-ramp", action="store_true", help="host pipe: group sizes grow 1 : 2 : 4 : ...")
    ap.add_argument("--e2e-ids", default="int64", choices=["int64", "int32"], help="host pipe: width of the ids in the feed buffers")
    ap.add_argument("--e2e-inflight", type=int, default=1, help="host pipe: requests in flight (one pipe each)")
    ap.add_argument("--e2e-direct-in", action="store_true",
                    help="host pipe: kernels read ids / LoD from pinned host memory themselves (no upload); rows come down on a copy engine")
    ap.add_argument("--no-overlap", action="store_true",
                    help="timed loop without CTRFE_LAUNCH_OVERLAP_PREVIOUS (each launch waits for the one before)")
    ap.add_argument("--cpu-budget", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--quick", action="store_true", help="headline workload only: no configs / sharded_cfg4 / cpu_baseline blocks")
    ap.add_argument("--rows", type=int, default=0, help="diagnostic: override the table rows of the workload")
    ap.add_argument("--hot-sample-batches", type=int, default=48,
                    help="sample traffic for the hot set, in batches of 1/8 of the global batch, per rank")
    ap.add_argument("--hot-rows", default="2000000,0,16000000",
                    help="sharded_cfg4: hot-row replica counts to measure (0 = none), comma separated")
    args = ap.parse_args()
    args.hot_rows_list = [int(x) for x in str(args.hot_rows).split(",") if x != ""]
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3  # timing rule: at least 3 warm-up steps
    # stdout carries the ONE JSON line and nothing else: libraries that print there (NCCL's version banner,
    # compiler chatter from build()) are sent to stderr for the length of the run
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return run_reference(args) if args.impl == "reference" else run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
