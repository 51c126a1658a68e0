"""ctrfe_fused_forward_static: the fused front end for callers that cannot allocate.

Inside Paddle's executor the framework's allocator has normally taken nearly all of HBM, so
a plugin's own cudaMalloc may fail or fragment what is left, and cudaMalloc / cudaFree
synchronise the whole device, stalling the other predictor streams of a serving process.
This entry point needs no context and keeps its scratch in a fixed array inside the library.
The test drives it through the C ABI and requires bit-identical output to the ordinary
ctrfe_fused_forward on two calls in a row (the second one smaller, so stale scratch from the
first would show), then checks the stated size limit is enforced.  It lives in its own file,
collected last, and must pass both alone in a fresh process and after the other tests.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _call_both(B, ops, dev, rng, S, Bsz, N, D, max_len, pooltype, use_cvm):
    table = rng.uniform(-1, 1, (N, D)).astype(np.float32)
    table[:, 0] = rng.integers(0, 16, N)
    table[:, 1] = np.minimum(table[:, 0], rng.integers(0, 16, N))
    lods, ids = [], []
    for _ in range(S):
        l = rng.integers(0, max_len + 1, Bsz)
        lod = np.concatenate([[0], np.cumsum(l)]).astype(np.uint64)
        lods.append(lod)
        ids.append(rng.integers(0, N, int(lod[-1])).astype(np.int64))
    tt = torch.from_numpy(table).to(dev)
    ids_t = [torch.from_numpy(i).to(dev) for i in ids]
    # the ordinary entry point (context-owned, allocated scratch)
    want = ops.ctr_fused_emb_seqpool_cvm(tt, ids_t, lods, pooltype, use_cvm=use_cvm, concat=True)
    # the static-scratch entry point, straight through the C ABI
    d_out = D if use_cvm is not False else D - 2
    got = torch.full((Bsz, S * d_out), float("nan"), dtype=torch.float32, device=dev)
    slots = (B.Slot * S)()
    for s in range(S):
        slots[s] = B.Slot(tt.data_ptr(), N, ids_t[s].data_ptr(), lods[s].ctypes.data, Bsz,
                          got.data_ptr() + s * d_out * 4, S * d_out, B.POOLTYPES[pooltype], B.I64, ids_t[s].numel())
    cvm_mode = B.CVM_NONE if use_cvm is None else (B.CVM_USE if use_cvm else B.CVM_STRIP)
    desc = B.FusedDesc(S, B.F32, slots, D, -1, 0.0, cvm_mode, 2, B.SCALE_DIVIDE, B.INIT_ZERO, B.LOD_HOST, 0)
    key = torch.zeros(1, dtype=torch.int64).pin_memory()
    stream = torch.cuda.current_stream(dev).cuda_stream
    rc = B.lib().ctrfe_fused_forward_static(C.byref(desc), stream, key.data_ptr())
    assert rc == B.OK, B.lib().ctrfe_last_error()
    torch.cuda.synchronize()
    assert key.item() == -1, "no id was out of range, the flag must read all ones"
    assert torch.equal(got, want)


def test_static_entry_matches_the_ordinary_one_twice_in_a_row():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import __graft_entry__ as g
    g.build()
    from paddle_infer_b200 import _lib as B
    from paddle_infer_b200 import ops
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(20261106)
    _call_both(B, ops, dev, rng, S=7, Bsz=600, N=5000, D=16, max_len=9, pooltype="SUM", use_cvm=True)
    # second call: smaller batch, other pool and width -- nothing may leak from the first
    _call_both(B, ops, dev, rng, S=3, Bsz=37, N=300, D=8, max_len=5, pooltype="MAX", use_cvm=None)
    _call_both(B, ops, dev, rng, S=2, Bsz=64, N=300, D=11, max_len=4, pooltype="AVERAGE", use_cvm=False)

    # a bad id is reported through the host word and the row reads as zeros
    table = torch.ones(10, 8, device=dev)
    ids = torch.tensor([1, 99, 2], dtype=torch.int64, device=dev)
    lod = np.array([0, 3], np.uint64)
    out = torch.empty(1, 8, device=dev)
    slots = (B.Slot * 1)(B.Slot(table.data_ptr(), 10, ids.data_ptr(), lod.ctypes.data, 1, out.data_ptr(), 8,
                                B.POOL_SUM, B.I64, 3))
    desc = B.FusedDesc(1, B.F32, slots, 8, -1, 0.0, B.CVM_NONE, 2, B.SCALE_DIVIDE, B.INIT_ZERO, B.LOD_HOST, 0)
    key = torch.zeros(1, dtype=torch.int64).pin_memory()
    assert B.lib().ctrfe_fused_forward_static(C.byref(desc), torch.cuda.current_stream(dev).cuda_stream,
                                              key.data_ptr()) == B.OK
    torch.cuda.synchronize()
    assert key.item() == (0 << 44) | 1 and out.eq(2).all()

    # the stated limit: more than 4 MiB of LoD offsets is refused with a clear message
    n_big = (4 << 20) // 8 + 8
    big_lod = np.zeros(n_big + 1, np.uint64)
    big_out = torch.empty(n_big, 8, device=dev)
    slots = (B.Slot * 1)(B.Slot(table.data_ptr(), 10, ids.data_ptr(), big_lod.ctypes.data, n_big, big_out.data_ptr(), 8,
                                B.POOL_SUM, B.I64, 3))
    desc = B.FusedDesc(1, B.F32, slots, 8, -1, 0.0, B.CVM_NONE, 2, B.SCALE_DIVIDE, B.INIT_ZERO, B.LOD_HOST, 0)
    rc = B.lib().ctrfe_fused_forward_static(C.byref(desc), torch.cuda.current_stream(dev).cuda_stream, None)
    assert rc == B.ERR_ARG and b"static scratch" in B.lib().ctrfe_last_error()
    # as is a call with more than 64 slots
    slots65 = (B.Slot * 65)()
    desc = B.FusedDesc(65, B.F32, slots65, 8, -1, 0.0, B.CVM_NONE, 2, B.SCALE_DIVIDE, B.INIT_ZERO, B.LOD_HOST, 0)
    assert B.lib().ctrfe_fused_forward_static(C.byref(desc), None, None) == B.ERR_ARG
