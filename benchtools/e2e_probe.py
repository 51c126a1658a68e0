import sys, time, ctypes as C
sys.path.insert(0, "/root/repo")
import numpy as np, torch
import __graft_entry__ as g; g.build()
from benchtools import synth
from paddle_infer_b200 import ops, _lib as B
dev = torch.device("cuda:0")
cfg = synth.configs()["cfg2"]
S, Bsz, D = cfg.slots, cfg.batch, cfg.dim
tab = torch.empty((cfg.rows, D), dtype=torch.float32, device=dev); synth.fill_table(tab, cfg.seed)
hb = [synth.draw_batch(cfg, step=b) for b in range(4)]
ctx = ops.default_context(dev)
stream = torch.cuda.current_stream(dev)
def mk(b, ids_host, lod_host, out_host, concat=True):
    ids_p = [torch.from_numpy(i).pin_memory() for i in b.ids]; ids_d = [i.to(dev) for i in ids_p]
    lod_p = [torch.from_numpy(l.astype(np.int64)).pin_memory() for l in b.lods]; lod_d = [l.to(dev) for l in lod_p]
    if concat:
        out_p = torch.empty((Bsz, S*D), dtype=torch.float32, pin_memory=True); out_d = torch.empty((Bsz, S*D), device=dev)
        ld = S*D
    else:
        out_p = torch.empty((S, Bsz, D), dtype=torch.float32, pin_memory=True); out_d = torch.empty((S, Bsz, D), device=dev)
        ld = D
    slots = (B.Slot * S)()
    for s in range(S):
        o = (out_p if out_host else out_d)
        optr = o.data_ptr() + (s*D*4 if concat else s*Bsz*D*4)
        slots[s] = B.Slot(tab.data_ptr(), tab.shape[0], (ids_p if ids_host else ids_d)[s].data_ptr(),
                          (lod_p if lod_host else lod_d)[s].data_ptr(), Bsz, optr, ld, B.POOL_SUM, B.I64)
    desc = B.FusedDesc(S, B.F32, slots, D, -1, 0.0, B.CVM_USE, 2, 0, 0, B.LOD_DEVICE, 0)
    return desc, (ids_p, ids_d, lod_p, lod_d, out_p, out_d, slots)
def timeit(label, **kw):
    ds = [mk(b, **kw) for b in hb]
    for d, _ in ds: B.check(B.lib().ctrfe_fused_forward(ctx.handle, C.byref(d), stream.cuda_stream))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 40; e0.record(stream)
    for i in range(n):
        B.check(B.lib().ctrfe_fused_forward(ctx.handle, C.byref(ds[i % 4][0]), stream.cuda_stream))
    e1.record(stream); torch.cuda.synchronize()
    print(f"{label:50s} {e0.elapsed_time(e1)/n*1000:8.1f} us/launch")
timeit("all device", ids_host=False, lod_host=False, out_host=False)
timeit("lod host", ids_host=False, lod_host=True, out_host=False)
timeit("ids host", ids_host=True, lod_host=False, out_host=False)
timeit("ids+lod host", ids_host=True, lod_host=True, out_host=False)
timeit("out host (concat)", ids_host=False, lod_host=False, out_host=True)
timeit("out host (per-slot)", ids_host=False, lod_host=False, out_host=True, concat=False)
timeit("all host (concat)", ids_host=True, lod_host=True, out_host=True)
timeit("all host (per-slot)", ids_host=True, lod_host=True, out_host=True, concat=False)
# plain copies for reference
x = torch.empty(20447232//4, dtype=torch.float32, device=dev); xp = torch.empty_like(x, device="cpu").pin_memory()
for _ in range(3): xp.copy_(x, non_blocking=True)
torch.cuda.synchronize(); t=time.perf_counter()
for _ in range(20): xp.copy_(x, non_blocking=True)
torch.cuda.synchronize(); print("D2H 20.4MB memcpy", (time.perf_counter()-t)/20*1e6, "us")
y = torch.empty(12575765//8, dtype=torch.int64, device=dev); yp = torch.empty_like(y, device="cpu").pin_memory()
torch.cuda.synchronize(); t=time.perf_counter()
for _ in range(20): y.copy_(yp, non_blocking=True)
torch.cuda.synchronize(); print("H2D 12.6MB memcpy", (time.perf_counter()-t)/20*1e6, "us")
