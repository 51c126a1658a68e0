"""Probe: host-buffer step with copy-engine H2D of ids / LoD in slot chunks, kernel writing the
pooled rows straight to pinned host memory.  Compares with the all-zero-copy bound call.
    python benchtools/e2e_staged_probe.py [chunks ...]"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402

g.build()
from benchtools import synth  # noqa: E402
from paddle_infer_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
cfg = synth.configs()["cfg2"]
S, D, Bsz = cfg.slots, cfg.dim, cfg.batch
tab = torch.empty((cfg.rows, D), dtype=torch.float32, device=dev)
synth.fill_table(tab, cfg.seed, cvm_cols=True)
NB = 4
batches = [synth.draw_batch(cfg, step=b) for b in range(NB)]
ctx = ops.default_context(dev)
main = torch.cuda.current_stream(dev)
copy = torch.cuda.Stream(dev)


def setup(hb, chunks):
    sizes = [i.size for i in hb.ids]
    offs = np.concatenate([[0], np.cumsum(sizes)])
    ids_host = torch.empty(int(offs[-1]), dtype=torch.int64).pin_memory()
    lod_host = torch.empty(S * (Bsz + 1), dtype=torch.int64).pin_memory()
    for s in range(S):
        ids_host[offs[s]:offs[s + 1]] = torch.from_numpy(hb.ids[s])
        lod_host[s * (Bsz + 1):(s + 1) * (Bsz + 1)] = torch.from_numpy(hb.lods[s].astype(np.int64))
    ids_dev = torch.empty_like(ids_host, device=dev)
    lod_dev = torch.empty_like(lod_host, device=dev)
    out_host = torch.empty((S, Bsz, D), dtype=torch.float32).pin_memory()
    bounds = [round(c * S / chunks) for c in range(chunks + 1)]
    parts = []
    for c in range(chunks):
        s0, s1 = bounds[c], bounds[c + 1]
        ids_v = [ids_dev[offs[s]:offs[s + 1]] for s in range(s0, s1)]
        lod_v = [lod_dev[s * (Bsz + 1):(s + 1) * (Bsz + 1)] for s in range(s0, s1)]
        # plan over device ids / LoD whose outputs are the pinned host rows
        import ctypes as C
        from paddle_infer_b200 import _lib as B
        slots = (B.Slot * (s1 - s0))()
        for k, s in enumerate(range(s0, s1)):
            slots[k] = B.Slot(tab.data_ptr(), tab.shape[0], ids_v[k].data_ptr(), lod_v[k].data_ptr(), Bsz,
                              out_host[s].data_ptr(), D, B.POOLTYPES["SUM"], B.I64)
        desc = B.FusedDesc(s1 - s0, B.F32, slots, D, -1, 0.0, B.CVM_USE, 2, B.SCALE_DIVIDE, B.INIT_ZERO, B.LOD_DEVICE, 0)
        h = C.c_void_p()
        B.check(B.lib().ctrfe_plan_create(ctx.handle, C.byref(desc), main.cuda_stream, C.byref(h)))
        plan = ops.FusedPlan(ctx, h, None, (slots, ids_v, lod_v))
        parts.append((plan, (int(offs[s0]), int(offs[s1])), (s0 * (Bsz + 1), s1 * (Bsz + 1)), torch.cuda.Event()))
    return dict(ids_host=ids_host, lod_host=lod_host, ids_dev=ids_dev, lod_dev=lod_dev, out=out_host, parts=parts)


def run(st):
    for plan, (i0, i1), (l0, l1), ev in st["parts"]:
        with torch.cuda.stream(copy):
            st["ids_dev"][i0:i1].copy_(st["ids_host"][i0:i1], non_blocking=True)
            st["lod_dev"][l0:l1].copy_(st["lod_host"][l0:l1], non_blocking=True)
            ev.record(copy)
        main.wait_event(ev)
        plan.launch(overlap_previous=False)
    main.synchronize()


def bench(fn, n=200):
    for _ in range(5):
        fn(0)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(n):
        fn(i)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3


# reference: zero-copy bound calls
max_ids = max(int(i.size) for hb in batches for i in hb.ids)
fe = ops.HostFrontEnd(tab, S, Bsz, max_ids, "SUM", use_cvm=True, ctx=ctx, mode="zero_copy")
pin_ids = [[torch.from_numpy(i).pin_memory() for i in hb.ids] for hb in batches]
pin_lod = [[torch.from_numpy(l.astype(np.int64)).pin_memory() for l in hb.lods] for hb in batches]
calls = [fe.bind(pin_ids[b], pin_lod[b]) for b in range(NB)]
print("zero_copy bound call: %.3f ms/step" % bench(lambda i: calls[i % NB].run()))
ref = [r.clone() for r in calls[0].run()]
for chunks in [int(a) for a in sys.argv[1:]] or [1, 2, 3, 4, 6]:
    states = [setup(hb, chunks) for hb in batches]
    ms = bench(lambda i: run(states[i % NB]))
    run(states[0])
    same = all(torch.equal(states[0]["out"][s], ref[s]) for s in range(S))
    print("staged H2D, %d chunk(s): %.3f ms/step, same bits: %s" % (chunks, ms, same))
