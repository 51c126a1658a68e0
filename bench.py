#!/usr/bin/env python
"""bench.py -- pooled rows/sec of the fused gather + sequence_pool + CVM front end.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2] [--impl reference]

One "step" = one pass of the hot path over one synthetic batch (all slots, ONE kernel
launch).  Default workload: BASELINE.json configs[1] (DeepFM: 39 slots, one shared
1e8 x 16 fp32 table, batch 8192, Zipf ids, SUM + CVM).  N > 1 (torchrun): every rank
holds a replica of the table and its own batches -- weak scaling, no data-path
collective (the row-sharded mode is measured by --workload cfg4).

Prints ONE JSON line (contract in the task statement): value = whole-job pooled rows/s
with inputs resident in HBM; e2e = same metric through the host-buffer call (H2D of ids
+ LoD, D2H of the pooled block inside the timed region); roofline for the fused kernel
from CUDA events; cpu_baseline = the reference CPU arithmetic timed on this box's cores.
"""
from __future__ import annotations

import argparse
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
N_BATCHES = 16  # distinct pre-generated batches the timed loop rotates through


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def _random_row_ceiling(row_bytes):
    """Measured ceiling for uniform-random row reads of this size (profiles/random_row_ceiling.json)."""
    p = os.path.join(ROOT, "profiles", "random_row_ceiling.json")
    if not os.path.exists(p):
        return None
    with open(p) as f:
        tab = json.load(f)["gbs_by_row_bytes"]
    keys = sorted(int(k) for k in tab)
    k = max([k for k in keys if k <= row_bytes] or [keys[0]])
    return float(tab[str(k)])


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
            # torch's device index follows CUDA_VISIBLE_DEVICES; map through the UUID-less simple case
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


# --------------------------------------------------------------------------------------
# CPU leg (cpu_baseline of our arm, and the whole of --impl reference)
# --------------------------------------------------------------------------------------
def cpu_leg(cfg, table_host, batches, budget_s=10.0, threads=None):
    """Times the reference CPU arithmetic on whole batches of the workload until ~budget_s of
    work is done.  Uses oracle/_ref (the reference's refer::EmbSeqPool compiled from
    /root/reference) when present and the workload is the fp32 SUM+CVM shape it covers,
    else the oracle port.  Returns a cpu_baseline dict."""
    import ctypes as C
    from oracle import oracle as O
    threads = threads or (os.cpu_count() or 1)
    R = O.ref_lib()
    use_ref = (R is not None and cfg.dtype == "f32" and cfg.use_cvm is True and cfg.shared_table
               and all(p == "SUM" for p in cfg.pooltypes))
    S, Bsz, D = cfg.slots, cfg.batch, cfg.dim
    out = np.empty((Bsz, S * cfg.d_out), dtype=table_host[0].dtype)
    n_rows = table_host[0].shape[0]

    def run(batch, nthreads, fast=True):
        if use_ref:
            t = table_host[0]
            for s in range(S):
                rc = R.ref_fused_lod_f32(t.ctypes.data, n_rows, D, batch.ids[s].ctypes.data, batch.lods[s].ctypes.data,
                                         Bsz, 0.0, 1, out.ctypes.data + s * D * 4, S * D, nthreads)
                assert rc == 0
        else:
            cvm_mode = O.CVM_NONE if cfg.use_cvm is None else (O.CVM_USE if cfg.use_cvm else O.CVM_STRIP)
            O.fused_forward(table_host, batch.ids, batch.lods, cfg.pooltypes, cvm_mode=cvm_mode,
                            bf16=cfg.dtype == "bf16", concat=True, threads=nthreads, fast=fast)

    def timed(nthreads, budget):
        run(batches[0], nthreads)  # warm
        n, t0 = 0, time.perf_counter()
        while True:
            run(batches[n % len(batches)], nthreads)
            n += 1
            el = time.perf_counter() - t0
            if el >= budget or n >= 20000:
                return S * Bsz * n / el, n, el

    v_all, n_all, t_all = timed(threads, budget_s * 0.6)
    v_one, n_one, t_one = timed(1, budget_s * 0.4)
    return {"value": v_all, "unit": UNIT, "cores": threads,
            "kind": "reference" if use_ref else "port",
            "sample": (f"{n_all} full batches ({S} slots x {Bsz} seqs) in {t_all:.2f}s at {threads} threads; "
                       + ("refer::EmbSeqPool from the reference tree (oracle/_ref) + pad/CVM glue"
                          if use_ref else "oracle port (oracle/ctr_oracle.c, -O3 -mavx2)")
                       + (f"; table folded to {n_rows} rows" if n_rows != cfg.rows else "")),
            "value_1thread": v_one, "sample_1thread": f"{n_one} batches in {t_one:.2f}s"}


def host_table_and_batches(cfg, synth, n_batches, torch=None, dev_tables=None, max_host_gb=24.0):
    """Host copies of the table(s) (D2H from the device when it exists, else regenerated with
    rows folded so that it stays small) and host batches whose ids index them."""
    bytes_needed = cfg.n_tables() * cfg.rows * cfg.dim * cfg.elem_size
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:
        avail = 32 << 30
    limit = min(max_host_gb * (1 << 30), avail * 0.5)
    if dev_tables is not None and bytes_needed <= limit:
        tabs = []
        for t in dev_tables:
            h = t.view(torch.int16).cpu().numpy().view(np.uint16) if t.dtype == torch.bfloat16 else t.cpu().numpy()
            tabs.append(h)
        rows = cfg.rows
    else:
        rows = cfg.rows
        while cfg.n_tables() * rows * cfg.dim * cfg.elem_size > min(limit, 2 << 30):
            rows //= 2
        tabs = [synth.table_rows_numpy(cfg.seed + 1000 * t, np.arange(rows), cfg.dim, cfg.dtype,
                                       cvm_cols=cfg.use_cvm is not None) for t in range(cfg.n_tables())]
    batches = []
    for b in range(n_batches):
        bt = synth.draw_batch(cfg, step=b)
        if rows != cfg.rows:
            bt.ids = [i % rows for i in bt.ids]
        batches.append(bt)
    return tabs, batches, rows


def run_reference(args):
    rank, local_rank, world = _env_rank()
    if rank != 0:
        return 0
    from benchtools import synth
    from oracle import oracle as O
    O.build()
    cfg = synth.configs()[args.workload]
    torch = None
    dev_tables = None
    try:
        import torch as _torch
        if _torch.cuda.is_available():
            torch = _torch
            dev = torch.device("cuda", local_rank)
            tdt = {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}[cfg.dtype]
            import __graft_entry__ as g
            g.build()
            dev_tables = []
            for t in range(cfg.n_tables()):
                tab = torch.empty((cfg.rows, cfg.dim), dtype=tdt, device=dev)
                synth.fill_table(tab, cfg.seed + 1000 * t, cvm_cols=cfg.use_cvm is not None)
                dev_tables.append(tab)
            torch.cuda.synchronize()
    except Exception:
        dev_tables = None
    tabs, batches, rows = host_table_and_batches(cfg, synth, 4, torch, dev_tables)
    del dev_tables
    # K steps of a bounded sample each; one step = one full batch here
    base = cpu_leg(cfg, tabs, batches, budget_s=min(60.0, max(5.0, 0.2 * (args.steps + args.warmup))))
    value = base["value"]
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * cfg.slots * cfg.batch / value,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": cfg.dtype,
            "data": "synthetic", "config": {"workload": _workload_name(cfg), "host_threads": base["cores"]},
            "cpu_baseline": base,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)
    return 0


def run_ours(args):
    import torch
    import __graft_entry__ as g
    rank, local_rank, world = _env_rank()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU: paddle_infer_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    # pinned host buffers of the e2e leg should sit on the GPU's own NUMA node
    from paddle_infer_b200 import affinity
    prev_affinity = None if os.environ.get("CTRFE_NO_BIND") == "1" else affinity.bind_to_gpu(local_rank)
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
    S, D = cfg.slots, cfg.dim
    tdt = {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}[cfg.dtype]
    # cfg4 on N > 1 GPUs: ONE logical table row-sharded over the ranks (row r on rank r % N),
    # the global batch split over the ranks (strong scaling); remote rows are fetched by the
    # pooling kernel itself over NVLink.  Everything else: a full table replica per rank.
    sharded = args.workload == "cfg4" and world > 1
    # diagnostic (one GPU): the sharded kernel over G shards that all live on this GPU, on 1/G
    # of the batch -- what one rank of a G-GPU run does, minus NVLink
    local_G = args.local_shards if world == 1 and args.workload == "cfg4" else 0
    Bsz = cfg.batch // world if sharded else (cfg.batch // local_G if local_G else cfg.batch)
    n_batches = 4 if cfg.batch * cfg.slots * cfg.dim * cfg.elem_size > (256 << 20) else N_BATCHES
    shard_table = None
    tables = []
    if sharded:
        from paddle_infer_b200 import shard as shard_mod
        shard_table = shard_mod.ShardedTable(cfg.rows, D, tdt, sharing=os.environ.get("CTRFE_SHARING", "vmm"))
        synth.fill_table(shard_table.local, cfg.seed, row0=rank, row_stride=world, cvm_cols=cfg.use_cvm is not None)
        torch.cuda.synchronize()
        dist.barrier()
        W = shard_table
    elif local_G:
        from paddle_infer_b200 import shard as shard_mod
        parts = []
        for k in range(local_G):
            part = torch.empty((shard_mod.shard_num_rows(cfg.rows, local_G, k), D), dtype=tdt, device=dev)
            synth.fill_table(part, cfg.seed, row0=k, row_stride=local_G, cvm_cols=cfg.use_cvm is not None)
            parts.append(part)
        W = shard_table = shard_mod.LocalShards(parts)
        cfg.name += f"[diagnostic: {local_G} shards on one GPU, 1/{local_G} of the batch]"
    else:
        for t in range(cfg.n_tables()):
            tab = torch.empty((cfg.rows, D), dtype=tdt, device=dev)
            synth.fill_table(tab, cfg.seed + 1000 * t, cvm_cols=cfg.use_cvm is not None)
            tables.append(tab)
        W = tables if len(tables) > 1 else tables[0]

    # sharded mode: replicate the hottest rows locally (shard.HotRowCache).  The hot set comes
    # from sample traffic drawn from the same distribution but DISJOINT from the timed batches.
    hot_info = None
    if (sharded or local_G) and args.hot_rows > 0:
        # the same amount of sample traffic per rank whatever the world size
        n_sample = max(1, int(round(args.hot_sample_batches * (cfg.batch / 8) / Bsz)))
        sample = [synth.draw_batch(cfg, step=rank * 1000 + 500 + b, batch=Bsz) for b in range(n_sample)]
        sample_ids = torch.cat([torch.from_numpy(i).to(dev) for sb in sample for i in sb.ids])
        hot_ids = shard_mod.pick_hot_ids(sample_ids, args.hot_rows)
        shard_table.hot = shard_mod.HotRowCache(shard_table, hot_ids)
        hot_info = {"rows": int(hot_ids.numel()), "bytes": int(shard_table.hot.nbytes),
                    "built_from": f"{n_sample} sample batches of {Bsz} per rank, disjoint from the timed batches"}
        del sample, sample_ids
        if world > 1:
            dist.barrier()

    # ---- batches resident in HBM (ids, LoD, output blocks), one plan each ----------------
    host_batches = [synth.draw_batch(cfg, step=rank * 1000 + b, batch=Bsz) for b in range(n_batches)]
    plans, k_tots = [], []
    for hb in host_batches:
        ids_d = [torch.from_numpy(i).to(dev) for i in hb.ids]
        lods_d = [torch.from_numpy(l.astype(np.int64)).to(dev) for l in hb.lods]
        plans.append(ops.ctr_fused_emb_seqpool_cvm(W, ids_d, lods_d, cfg.pooltypes, use_cvm=cfg.use_cvm, concat=True,
                                                   plan_only=True))
        k_tots.append(hb.k_tot)
    torch.cuda.synchronize()
    ctx = plans[0].ctx
    stream = torch.cuda.current_stream(dev)

    K, Wm = args.steps, max(args.warmup, 0)
    overlap = not args.no_overlap
    for i in range(Wm):
        plans[i % n_batches].launch(overlap)
    ctx.check_ids()  # also a sync

    # The timed region: K launches back to back over distinct resident batches.  With overlap
    # (default) each launch may start while the previous one drains (ctrfe_plan_launch_ex,
    # CTRFE_LAUNCH_OVERLAP_PREVIOUS: the batches are independent) -- what a serving loop does.
    evs = [torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)]
    sampler = ClockSampler(local_rank)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.start()
    evs[0].record(stream)
    for i in range(K):
        plans[(Wm + i) % n_batches].launch(overlap)
    evs[1].record(stream)
    sampler.sample()  # the queue is still draining: a sample under load even for tiny K
    torch.cuda.synchronize()
    sampler.stop()
    if world > 1:
        dist.barrier()
    elapsed_ms = evs[0].elapsed_time(evs[-1])
    ctx.check_ids()
    kern_ms = elapsed_ms / K
    # the same launches one at a time (an event between every two, no overlap): the duration of
    # an isolated launch, reported beside the throughput figure
    Ki = min(K, 400)
    iso = [torch.cuda.Event(enable_timing=True) for _ in range(Ki + 1)]
    iso[0].record(stream)
    for i in range(Ki):
        plans[(Wm + i) % n_batches].launch(False)
        iso[i + 1].record(stream)
    torch.cuda.synchronize()
    per = np.array([iso[i].elapsed_time(iso[i + 1]) for i in range(Ki)])
    kern_ms_iso, kern_ms_med = float(per.mean()), float(np.median(per))
    k_used = [k_tots[(Wm + i) % n_batches] for i in range(K)]
    es = cfg.elem_size
    # SURVEY 8(d) per-unit figures x the units one launch (this rank's batch) processes
    alg_bytes = float(np.mean([k * (8 + D * es) + S * (Bsz + 1) * 8 + S * Bsz * cfg.d_out * es for k in k_used]))

    t = torch.tensor([elapsed_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    elapsed_max = float(t.item())
    value = S * Bsz * K * world / (elapsed_max * 1e-3)

    # ---- e2e: host buffers in, host rows out, through the public host-buffer call ----------
    # Every step reads that step's ids + LoD offsets from pinned host memory and lands the
    # pooled rows in pinned host memory; the call returns after the stream has drained.
    Ke = max(3, min(K, args.e2e_steps))
    max_ids = max(int(i.size) for hb in host_batches for i in hb.ids)
    fe = ops.HostFrontEnd(W, S, Bsz, max_ids, cfg.pooltypes, use_cvm=cfg.use_cvm, ctx=ctx, mode=args.e2e_mode)
    pinned = [[torch.from_numpy(i).pin_memory() for i in hb.ids] for hb in host_batches]
    if args.e2e_mode == "zero_copy":
        e2e_lods = [[torch.from_numpy(l.astype(np.int64)).pin_memory() for l in hb.lods] for hb in host_batches]
        calls = [fe.bind(pinned[b], e2e_lods[b]) for b in range(n_batches)]
        step_fns = [c.run for c in calls]
        h2d_of = [c.h2d_bytes for c in calls]
    else:
        step_fns = [(lambda b=b: fe.run(pinned[b], host_batches[b].lods)) for b in range(n_batches)]
        h2d_of = [sum(i.numel() * 8 for i in pinned[b]) + sum(l.size * 8 for l in host_batches[b].lods)
                  for b in range(n_batches)]
    for i in range(3):
        step_fns[i % n_batches]()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_host0 = time.perf_counter()
    e0.record(stream)
    h2d = d2h = 0
    for i in range(Ke):
        step_fns[i % n_batches]()
        h2d += h2d_of[i % n_batches]
        d2h += fe.d2h_bytes
    e1.record(stream)
    torch.cuda.synchronize()
    t_host = (time.perf_counter() - t_host0) * 1e3
    e2e_ms = max(e0.elapsed_time(e1), t_host)  # the call is synchronous: host time is the truth when larger
    te = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = S * Bsz * Ke * world / (float(te.item()) * 1e-3)

    # the same host-buffer call with two requests in flight (zero_copy only): each request has its
    # own front end (own pinned output); request i+1 is submitted before request i is waited for
    e2e_pipe = None
    if args.e2e_mode == "zero_copy":
        fes = [fe] + [ops.HostFrontEnd(W, S, Bsz, max_ids, cfg.pooltypes, use_cvm=cfg.use_cvm, ctx=ctx, mode="zero_copy")
                      for _ in range(2)]
        ring = [fes[j].bind(pinned[b], e2e_lods[b]) for b in range(n_batches) for j in range(3)]  # [batch][front end]
        def pick(i):
            return ring[(i % n_batches) * 3 + (i % 3)]
        for i in range(3):
            pick(i).result(pick(i).submit())
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        pending = []
        for i in range(Ke):
            c = pick(i)
            pending.append((c, c.submit(overlap_previous=True)))
            if len(pending) > 2:
                c0, ev0 = pending.pop(0)
                c0.result(ev0)
        for c0, ev0 in pending:
            c0.result(ev0)
        tp = (time.perf_counter() - t0) * 1e3
        tpt = torch.tensor([tp], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tpt, op=dist.ReduceOp.MAX)
        e2e_pipe = {"value": S * Bsz * Ke * world / (float(tpt.item()) * 1e-3), "ms_per_step": float(tpt.item()) / Ke,
                    "requests_in_flight": 3,
                    "note": "BoundHostCall.submit()/result(): a request is submitted before the two before it are waited for; "
                            "every request still reads its inputs from and writes its rows to pinned host memory"}

    peak, peak_src = _peaks()
    achieved = alg_bytes / (kern_ms * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
        "ms_per_step": elapsed_max / K, "higher_is_better": True, "scaling": "strong" if sharded else "weak",
        "vs_baseline": None,
        "dtype": cfg.dtype, "data": "synthetic",
        "config": {"workload": _workload_name(cfg), "global_batch": Bsz * world,
                   "parallelism": (f"table row-sharded x{world} (row r on GPU r % {world}), batch split x{world}, remote rows "
                                   "fetched by the pooling kernel over NVLink peer mappings" if sharded
                                   else f"replica x{world}"),
                   "l2_hygiene": f"table {cfg.n_tables() * cfg.rows * D * cfg.elem_size / 1e9:.1f} GB >> L2; timed loop "
                                 f"rotates {n_batches} distinct resident batches (ids+LoD+out "
                                 f"{n_batches * alg_bytes / 1e6:.0f} MB > 126 MB L2); hot table rows stay cached as in "
                                 "steady-state serving",
                   "lod": "resident in HBM (value); host LoD uploaded per call (e2e)",
                   "launches": ("back to back over independent batches, each allowed to start while the previous one drains "
                                "(ctrfe_plan_launch_ex, CTRFE_LAUNCH_OVERLAP_PREVIOUS)" if overlap else
                                "back to back, each waiting for the previous one")},
        "clocks": sampler.summary(),
        "gpu_launches": K * plans[0].num_kernels(),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": _traffic_from_profiles(args.workload), "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms_mean": kern_ms, "kernel_ms_isolated_mean": kern_ms_iso, "kernel_ms_isolated_median": kern_ms_med,
                     "launch_overlap": overlap,
                     "frac_of_nominal_8000": achieved / 8000.0, "kernel": "ctrfe::fused_pool_fast_kernel",
                     "random_row_ceiling_gbs": _random_row_ceiling(D * es),
                     "frac_of_random_row_ceiling": (achieved / _random_row_ceiling(D * es)) if _random_row_ceiling(D * es) else None,
                     "random_row_ceiling_note": "what a stripped uniform-random gather of rows this size sustains on this GPU "
                                                "(benchtools/gather_ceiling.cu, profiles/r1_gather_ceiling.txt); Zipf ids can "
                                                "exceed it through L2 hits"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d // Ke, "d2h_bytes_per_step": d2h // Ke,
                "steps": Ke, "ms_per_step": float(te.item()) / Ke,
                "path": ("HostFrontEnd zero_copy (buffers bound once): the kernel reads pinned host ids/LoD and writes "
                         "the pinned per-slot host outputs over PCIe inside the one launch, then the stream is synchronised" if args.e2e_mode == "zero_copy" else
                         "HostFrontEnd copy: H2D ids + LoD upload, launch, D2H output")},
    }
    if sharded:
        # rows whose owner is another GPU: the bytes the kernel asks the NVLink mappings for per
        # launch (an upper bound on link traffic -- peer rows are cached in L1, and the Zipf head
        # repeats within a launch)
        hit = 0.0
        if hot_info is not None:  # look-ups served by the local replicas never reach the link
            hot_np = shard_table.hot.hot_ids.cpu().numpy()
            hit = float(np.mean([np.isin(np.concatenate(hb.ids), hot_np).mean() for hb in host_batches]))
            hot_info["hit_rate_on_timed_batches"] = hit
            line["config"]["hot_row_cache"] = hot_info
        nv_bytes = float(np.mean(k_used)) * (1.0 - hit) * (world - 1) / world * D * es
        nv_gbs = nv_bytes / (kern_ms * 1e-3) / 1e9
        line["nvlink"] = {"remote_row_bytes_per_launch_per_gpu": nv_bytes, "requested_gbs": nv_gbs,
                          "link_peak_gbs": 770.0,
                          "peak_source": "B200_PROFILING.md measured peer copy, per direction per GPU",
                          "note": "requested bytes / time; above the link peak means L1 hits on repeated hot rows"}
    if local_G and hot_info is not None:
        line["config"]["hot_row_cache"] = hot_info
    line["e2e"]["host_thread_bound_to_gpu_local_cpus"] = prev_affinity is not None
    if e2e_pipe is not None:
        line["e2e"]["pipelined"] = e2e_pipe
    affinity.restore(prev_affinity)  # the CPU leg below gets every host core
    if rank == 0 and world == 1 and not args.no_cpu_baseline and not local_G:
        tabs, hb4, rows = host_table_and_batches(cfg, synth, 4, torch, tables)
        line["cpu_baseline"] = cpu_leg(cfg, tabs, hb4, budget_s=args.cpu_budget)
    elif rank == 0:
        line["cpu_baseline"] = None
    if rank == 0:
        print(json.dumps(line), flush=True)
    if shard_table is not None:
        del plans, fe
        shard_table.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=["cfg1", "cfg2", "cfg3", "cfg4", "cfg5"])
    ap.add_argument("--e2e-steps", type=int, default=40)
    ap.add_argument("--e2e-mode", default="zero_copy", choices=["zero_copy", "copy"])
    ap.add_argument("--no-overlap", action="store_true",
                    help="timed loop without CTRFE_LAUNCH_OVERLAP_PREVIOUS (each launch waits for the one before)")
    ap.add_argument("--cpu-budget", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--rows", type=int, default=0, help="diagnostic: override the table rows of the workload")
    ap.add_argument("--hot-sample-batches", type=int, default=96,
                    help="sample traffic for the hot set, in batches of 1/8 of the global batch, per rank")
    ap.add_argument("--local-shards", type=int, default=0,
                    help="diagnostic, one GPU, cfg4: run the sharded kernel over this many shards on the same GPU")
    ap.add_argument("--hot-rows", type=int, default=16_000_000,
                    help="sharded mode: rows replicated locally by the hot-row cache (0 = none)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3  # timing rule: at least 3 warm-up steps
    return run_reference(args) if args.impl == "reference" else run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
