"""Guards against silent wrong answers, exercised on the GPU through the C ABI: LoD checks the
kernels apply to device-resident offsets (framework CheckLoD, lod_tensor.cc:116-142), the
id-count bound, bad-id attribution with several plans in flight, duplicate feasign keys, the
fp64 cvm the reference registers (cvm_op.cu:192-194) and the many-plans launch call."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def env():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import __graft_entry__ as g
    g.build()
    from oracle import oracle as O
    from paddle_infer_b200 import _lib as B
    from paddle_infer_b200 import ops
    return ops, O, B, torch.device("cuda:0")


def _case(rng, n_seg, N, D, max_len=6):
    table = rng.uniform(-1, 1, (N, D)).astype(np.float32)
    table[:, 0] = rng.integers(0, 16, N)
    table[:, 1] = np.minimum(table[:, 0], rng.integers(0, 16, N))
    l = rng.integers(0, max_len + 1, n_seg)
    lod = np.concatenate([[0], np.cumsum(l)]).astype(np.uint64)
    ids = rng.integers(0, N, int(lod[-1])).astype(np.int64)
    return table, ids, lod


@pytest.mark.parametrize("D", [16, 11])  # the specialised kernel and the general one
def test_device_lod_is_checked_by_the_kernel(env, D):
    ops, O, B, dev = env
    rng = np.random.default_rng(7 + D)
    table, ids, lod = _case(rng, 300, 1000, D)
    tt, it = torch.from_numpy(table).to(dev), torch.from_numpy(ids).to(dev)
    good = ops.ctr_fused_emb_seqpool_cvm(tt, [it], [lod], "SUM", use_cvm=None, pad_value=0.25)[0].cpu().numpy()
    # (1) a decreasing pair in the middle: that sequence (and the one after it, whose start now
    # exceeds its end... or not) is pooled as empty, everything else is untouched
    bad = lod.astype(np.int64).copy()
    k = 150
    bad[k] = bad[k + 1] + 5  # sequence k-1 grows past k's end, sequence k: start > end
    lod_dev = torch.from_numpy(bad).to(dev)
    out = ops.ctr_fused_emb_seqpool_cvm(tt, [it], [lod_dev], "SUM", use_cvm=None, pad_value=0.25, check=False)[0]
    with pytest.raises(ops.InvalidArgument, match="LoD offsets of sequence 150"):
        ops.default_context(dev).check_ids()
    o = out.cpu().numpy()
    np.testing.assert_array_equal(o[:k - 1], good[:k - 1])
    np.testing.assert_array_equal(o[k + 1:], good[k + 1:])
    assert (o[k] == 0.25).all()  # pooled as an empty sequence
    ops.default_context(dev).check_ids()  # the record was cleared
    # (2) the last offset beyond the id tensor: no read past the end, reported
    over = lod.astype(np.int64).copy()
    over[-1] += 1000
    out = ops.ctr_fused_emb_seqpool_cvm(tt, [it], [torch.from_numpy(over).to(dev)], "SUM", use_cvm=None, pad_value=0.25,
                                        check=False)[0]
    with pytest.raises(ops.InvalidArgument, match="sequence 299"):
        ops.default_context(dev).check_ids()
    np.testing.assert_array_equal(out.cpu().numpy()[:299], good[:299])
    # (3) an offset with a non-zero high word
    big = lod.astype(np.int64).copy()
    big[10] += 1 << 32
    ops.ctr_fused_emb_seqpool_cvm(tt, [it], [torch.from_numpy(big).to(dev)], "SUM", use_cvm=None, check=False)
    with pytest.raises(ops.InvalidArgument):
        ops.default_context(dev).check_ids()
    # host LoD: the same faults are refused before anything is launched
    for l in (bad, over, big):
        with pytest.raises(B.CtrfeError):
            ops.ctr_fused_emb_seqpool_cvm(tt, [it], [l.astype(np.uint64)], "SUM", use_cvm=None)


def test_bad_id_is_attributed_to_the_plan_that_met_it(env):
    ops, O, B, dev = env
    rng = np.random.default_rng(11)
    N, D = 500, 16
    table, ids_a, lod_a = _case(rng, 200, N, D)
    _, ids_b, lod_b = _case(rng, 200, N, D)
    ids_a = ids_a.copy()
    pos = int(lod_a[100]) if lod_a[101] > lod_a[100] else int(lod_a[np.nonzero(np.diff(lod_a.astype(np.int64)))[0][0]])
    ids_a[pos] = N + 77
    tt = torch.from_numpy(table).to(dev)
    pa = ops.ctr_fused_emb_seqpool_cvm(tt, [torch.from_numpy(ids_a).to(dev)], [lod_a], "SUM", plan_only=True)
    pb = ops.ctr_fused_emb_seqpool_cvm(tt, [torch.from_numpy(ids_b).to(dev)], [lod_b], "SUM", plan_only=True)
    pa.launch()
    pb.launch()  # the LAST launched plan is the clean one
    with pytest.raises(ValueError) as e:
        pa.ctx.check_ids()
    assert e.value.slot == 0 and e.value.position == pos and e.value.id_value == N + 77
    pa.close(), pb.close()


def test_duplicate_feasign_keys_resolve_to_the_first_occurrence(env):
    ops, O, B, dev = env
    n = 50_000
    rng = np.random.default_rng(3)
    keys = rng.integers(1, 1 << 62, n).astype(np.int64)
    keys[n // 2:] = keys[: n - n // 2]  # every key twice
    for _ in range(3):
        km = ops.KeyMap(torch.from_numpy(keys).to(dev))
        W = torch.arange(n, dtype=torch.float32, device=dev).view(n, 1).repeat(1, 4).contiguous()
        out = ops.pull_box_sparse(km, W, [torch.from_numpy(keys).to(dev)])[0].cpu().numpy()
        want = np.concatenate([np.arange(n // 2), np.arange(n - n // 2)]).astype(np.float32)
        np.testing.assert_array_equal(out[:, 0], want)


def test_cvm_fp64_matches_numpy(env):
    """cvm on double rows: y0 = log(x0 + 1), y1 = log(x1 + 1) - y0 in fp64 (cvm_op.h:25-41 with
    T = double; the numpy statement of test_cvm_op.py:31-36)."""
    ops, O, B, dev = env
    rng = np.random.default_rng(5)
    x = rng.uniform(0, 1, (37, 11))
    for use_cvm in (True, False):
        y = ops.cvm(torch.from_numpy(x).to(dev), None, use_cvm).cpu().numpy()
        if use_cvm:
            want = x.copy()
            want[:, 0] = np.log(x[:, 0] + 1)
            want[:, 1] = np.log(x[:, 1] + 1) - want[:, 0]
            np.testing.assert_allclose(y, want, rtol=1e-14, atol=1e-15)
            np.testing.assert_array_equal(y[:, 2:], x[:, 2:])
        else:
            np.testing.assert_array_equal(y, x[:, 2:])


def test_launch_many_equals_single_launches(env):
    ops, O, B, dev = env
    rng = np.random.default_rng(9)
    table, _, _ = _case(rng, 1, 2000, 16)
    tt = torch.from_numpy(table).to(dev)
    plans, singles = [], []
    for b in range(5):
        _, ids, lod = _case(rng, 257 + b, 2000, 16)
        it = torch.from_numpy(ids).to(dev)
        plans.append(ops.ctr_fused_emb_seqpool_cvm(tt, [it, it], [lod, lod], ["SUM", "SQRT"], plan_only=True))
        singles.append(ops.ctr_fused_emb_seqpool_cvm(tt, [it, it], [lod, lod], ["SUM", "SQRT"]))
    for flags in (0, B.LAUNCH_OVERLAP_PREVIOUS):
        for p in plans:
            for o in p.outs:
                o.fill_(float("nan"))
        ops.launch_many(plans, overlap_previous=bool(flags))
        torch.cuda.synchronize()
        for p, want in zip(plans, singles):
            for g, w in zip(p.outs, want):
                assert torch.equal(g, w)
    assert plans[0].kernel_names() and "fused_pool" in plans[0].kernel_names()[0]
    for p in plans:
        p.close()


@pytest.mark.parametrize("n_groups", [1, 3, 5])
def test_host_pipe_equals_the_device_call(env, n_groups):
    """ctrfe_host_pipe_*: host ids / LoD in, host rows out, slots pipelined in groups -- the same
    bits as the device-resident call, on several runs in a row with different batches (the
    pipe's buffers are reused), for fp32 SUM + CVM and for bf16 mixed pools."""
    ops, O, B, dev = env
    rng = np.random.default_rng(40 + n_groups)
    for dtype, D, pools, use_cvm in (("f32", 16, "SUM", True), ("bf16", 64, ["SUM", "AVERAGE", "SQRT", "SUM", "SQRT"], None)):
        S, Bsz, N = 5, 300, 4000
        table = rng.uniform(-1, 1, (N, D)).astype(np.float32)
        table[:, 0] = rng.integers(0, 16, N)
        table[:, 1] = np.minimum(table[:, 0], rng.integers(0, 16, N))
        tt = torch.from_numpy(table).to(dev)
        if dtype == "bf16":
            tt = tt.to(torch.bfloat16)
        pipe = ops.HostPipe(tt, S, Bsz, 4000, pools, use_cvm=use_cvm, n_groups=n_groups)
        for _ in range(3):
            lods, ids = [], []
            for _s in range(S):
                l = rng.integers(0, 9, Bsz)
                lod = np.concatenate([[0], np.cumsum(l)]).astype(np.uint64)
                lods.append(lod)
                ids.append(torch.from_numpy(rng.integers(0, N, int(lod[-1])).astype(np.int64)).pin_memory())
            got = pipe.run(ids, lods)
            want = ops.ctr_fused_emb_seqpool_cvm(tt, [i.to(dev) for i in ids], lods, pools, use_cvm=use_cvm)
            for s in range(S):
                assert torch.equal(got[s], want[s].cpu()), (dtype, s)
        # a LoD that names more ids than were passed is refused on the host
        bad = [l.copy() for l in lods]
        bad[2][-1] += 5
        with pytest.raises(B.CtrfeError, match="exceeds the number of ids"):
            pipe.run(ids, bad)
        # an out-of-range id is reported by check_ids with the slot it came from
        ids[3] = ids[3].clone()
        if ids[3].numel():
            ids[3][0] = N + 3
            pipe.run(ids, lods)
            with pytest.raises(ValueError) as e:
                pipe.ctx.check_ids()
            assert e.value.id_value == N + 3 and e.value.slot == 3   # the caller's slot, not its place in the group
        pipe.close()


@pytest.mark.parametrize("n_groups,direct_out", [(1, False), (4, False), (3, True)])
def test_host_pipe_direct_input(env, n_groups, direct_out):
    """CTRFE_PIPE_DIRECT_IN: the kernels read ids and 64-bit offsets from the caller's pinned host
    memory themselves -- same bits as the device-resident call, from one feed buffer (arrays that
    follow each other) and from separate pinned arrays; pageable memory is refused; a bad id is
    fetched from the host array it sits in."""
    ops, O, B, dev = env
    rng = np.random.default_rng(70 + n_groups)
    S, Bsz, N, D = 6, 257, 5000, 16
    table = rng.uniform(-1, 1, (N, D)).astype(np.float32)
    table[:, 0] = rng.integers(0, 16, N)
    table[:, 1] = np.minimum(table[:, 0], rng.integers(0, 16, N))
    tt = torch.from_numpy(table).to(dev)
    pipe = ops.HostPipe(tt, S, Bsz, 4000, "SUM", use_cvm=True, n_groups=n_groups, direct_in=True, direct_out=direct_out)
    for rep in range(3):
        lens = [rng.integers(0, 9, Bsz) if s != 2 else np.zeros(Bsz, np.int64) for s in range(S)]   # slot 2: no ids at all
        lods = [np.concatenate([[0], np.cumsum(l)]).astype(np.int64) for l in lens]
        ids_np = [rng.integers(0, N, int(l[-1])).astype(np.int64) for l in lods]
        if rep % 2 == 0:   # one feed buffer: the slots' arrays follow each other
            feed = torch.from_numpy(np.concatenate(ids_np)).pin_memory()
            offs = np.concatenate([[0], np.cumsum([i.size for i in ids_np])])
            ids = [feed[int(offs[s]):int(offs[s + 1])] for s in range(S)]
            lod_feed = torch.from_numpy(np.stack(lods)).pin_memory()
            lod_t = [lod_feed[s] for s in range(S)]
        else:
            ids = [torch.from_numpy(i).pin_memory() for i in ids_np]
            lod_t = [torch.from_numpy(l).pin_memory() for l in lods]
        got = pipe.run(ids, lod_t)
        want = ops.ctr_fused_emb_seqpool_cvm(tt, [torch.from_numpy(i).to(dev) for i in ids_np],
                                             [l.astype(np.uint64) for l in lods], "SUM", use_cvm=True)
        for s in range(S):
            assert torch.equal(got[s], want[s].cpu()), (rep, s)
    with pytest.raises(B.CtrfeError, match="pinned host memory"):
        pipe.run([torch.from_numpy(i) for i in ids_np], lod_t)          # pageable ids
    with pytest.raises(B.CtrfeError, match="pinned host memory"):
        pipe.run(ids, [l.astype(np.uint64) for l in lods])               # pageable offsets
    ids[4][1] = N + 9
    pipe.run(ids, lod_t)
    with pytest.raises(ValueError) as e:
        pipe.ctx.check_ids()
    assert e.value.id_value == N + 9 and e.value.slot == 4
    pipe.close()


@pytest.mark.parametrize("ids_dtype", ["int64", "int32"])
def test_two_host_pipes_in_flight(env, ids_dtype):
    """Two pipes used in turn on one stream (run with wait=None, wait() before the rows are read), ramped
    group sizes, 32- and 64-bit ids: every request's rows equal the device-resident call's."""
    ops, O, B, dev = env
    rng = np.random.default_rng(91)
    S, Bsz, N, D = 11, 200, 3000, 16
    table = rng.uniform(-1, 1, (N, D)).astype(np.float32)
    table[:, 0] = rng.integers(0, 16, N)
    table[:, 1] = np.minimum(table[:, 0], rng.integers(0, 16, N))
    tt = torch.from_numpy(table).to(dev)
    pipes = [ops.HostPipe(tt, S, Bsz, 3000, "SUM", use_cvm=True, n_groups=4, ramp=True, ids_dtype=ids_dtype) for _ in range(2)]
    npdt = np.int32 if ids_dtype == "int32" else np.int64
    reqs = []
    for _ in range(7):
        lods = [np.concatenate([[0], np.cumsum(rng.integers(0, 8, Bsz))]).astype(np.uint64) for _s in range(S)]
        ids = [torch.from_numpy(rng.integers(0, N, int(l[-1])).astype(npdt)).pin_memory() for l in lods]
        reqs.append((ids, lods))
    wants = [ops.ctr_fused_emb_seqpool_cvm(tt, [i.to(dev).long() for i in ids], lods, "SUM", use_cvm=True) for ids, lods in reqs]
    pending = [None, None]
    for r in range(len(reqs) + 2):
        pp = pipes[r % 2]
        if pending[r % 2] is not None:
            got = pp.wait()
            for s in range(S):
                assert torch.equal(got[s], wants[pending[r % 2]][s].cpu()), (pending[r % 2], s)
            pending[r % 2] = None
        if r < len(reqs):
            pp.run(reqs[r][0], reqs[r][1], wait=None)
            pending[r % 2] = r
    for pp in pipes:
        pp.close()


def test_int16_ids_and_sparse_grad(env):
    """lookup_table_v2 takes int16 ids as the reference GPU kernel does (embedding_kernel.cu:107-117), also
    over an fp64 table; the is_sparse gradient is the reference's SelectedRows: rows = the ids as int64, one per
    id and unmerged, value = Out@GRAD (embedding_grad_kernel.cu:141-214)."""
    ops, O, B, dev = env
    rng = np.random.default_rng(17)
    N, D, K = 300, 12, 1000
    for np_dt, t_dt in ((np.float32, torch.float32), (np.float64, torch.float64), (np.float16, torch.float16)):
        table = rng.uniform(-1, 1, (N, D)).astype(np_dt)
        for id_dt in (np.int16, np.int32, np.int64):
            ids = rng.integers(0, N, K).astype(id_dt)
            ids[5] = 7
            got = ops.lookup_table_v2(torch.from_numpy(table).to(dev), torch.from_numpy(ids).to(dev), padding_idx=7)
            want = table[ids.astype(np.int64)]
            want[ids == 7] = 0
            assert got.dtype == t_dt and np.array_equal(got.cpu().numpy(), want), (np_dt, id_dt)
    bad = rng.integers(0, N, 50).astype(np.int16)
    bad[9] = -3
    with pytest.raises(ValueError) as e:
        ops.lookup_table_v2(torch.from_numpy(table).to(dev), torch.from_numpy(bad).to(dev))
    assert e.value.id_value == -3 and e.value.position == 9
    for id_dt in (np.int16, np.int32, np.int64):
        ids = rng.integers(0, N, (K // 4, 4)).astype(id_dt)
        dout = rng.uniform(-1, 1, (K // 4, 4, D)).astype(np.float32)
        sr = ops.lookup_table_v2_grad_sparse(torch.from_numpy(dout).to(dev), torch.from_numpy(ids).to(dev), N)
        assert sr.height == N and sr.rows.dtype == torch.int64
        assert np.array_equal(sr.rows.cpu().numpy(), ids.reshape(-1).astype(np.int64))
        assert np.array_equal(sr.value.cpu().numpy(), dout.reshape(-1, D))
        # merged into a dense table it equals the dense gradient
        dense = np.zeros((N, D), np.float64)
        np.add.at(dense, ids.reshape(-1).astype(np.int64), dout.reshape(-1, D).astype(np.float64))
        dw = ops.lookup_table_v2_grad(torch.from_numpy(dout).to(dev), torch.from_numpy(ids.astype(np.int64)).to(dev), N)
        np.testing.assert_allclose(dw.cpu().numpy(), dense, rtol=1e-5, atol=1e-5)


def test_host_layer_runs_without_torch(env):
    """The package core needs numpy and the library only: a child process that never imports torch
    runs the fused front end, the three ops on their own and the host pipe on device.DeviceArray /
    device.PinnedArray memory and matches the oracle."""
    import os
    import subprocess
    import sys
    import textwrap
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = textwrap.dedent('''
        import sys
        import numpy as np
        sys.path.insert(0, %r)
        from oracle import oracle as O
        from paddle_infer_b200 import ops
        from paddle_infer_b200.device import DeviceArray, PinnedArray
        rng = np.random.default_rng(5)
        N, D, S, Bsz = 3000, 16, 4, 200
        table = rng.uniform(-1, 1, (N, D)).astype(np.float32)
        table[:, 0] = rng.integers(0, 16, N); table[:, 1] = np.minimum(table[:, 0], rng.integers(0, 16, N))
        lods, ids = [], []
        for s in range(S):
            l = rng.integers(0, 8, Bsz)
            lods.append(np.concatenate([[0], np.cumsum(l)]).astype(np.uint64))
            ids.append(rng.integers(0, N, int(lods[-1][-1])).astype(np.int64))
        W = DeviceArray.from_numpy(table)
        ids_d = [DeviceArray.from_numpy(i) for i in ids]
        want = O.fused_forward([table], ids, lods, ["SUM"] * S, cvm_mode=O.CVM_USE, concat=True)
        got = ops.ctr_fused_emb_seqpool_cvm(W, ids_d, lods, "SUM", use_cvm=True, concat=True).numpy()
        g, w = got.reshape(Bsz, S, D), want.reshape(Bsz, S, D)
        assert np.array_equal(g[:, :, 2:], w[:, :, 2:]) and np.allclose(g[:, :, :2], w[:, :, :2], rtol=1e-6, atol=1e-6)
        rows = ops.lookup_table_v2(W, ids_d[0]).numpy()
        assert np.array_equal(rows, table[ids[0]])
        pooled = ops.sequence_pool(DeviceArray.from_numpy(rows), [list(lods[0])], "MAX", is_test=True).Out.numpy()
        assert np.array_equal(pooled, O.sequence_pool(rows, lods[0], "MAX"))
        y = ops.cvm(DeviceArray.from_numpy(np.abs(pooled)), None, True).numpy()
        assert np.allclose(y, O.cvm(np.abs(pooled), True), rtol=1e-6, atol=1e-6)
        # host buffers in, host buffers out, pinned memory from the library itself
        pipe = ops.HostPipe(W, S, Bsz, 4000, "SUM", use_cvm=True, n_groups=2)
        pin = [PinnedArray((i.size,), np.int64) for i in ids]
        for p, i in zip(pin, ids):
            p.array[:] = i
        outs = pipe.run([p.array for p in pin], lods)
        for s in range(S):
            assert np.array_equal(outs[s][:, 2:], w[:, s, 2:])
        assert "torch" not in sys.modules, "the host layer pulled torch in"
        print("NO_TORCH_OK")
    ''') % root
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600, cwd=root)
    assert r.returncode == 0 and "NO_TORCH_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]


def test_hot_rows_beside_an_unsharded_table_change_nothing(env):
    """shard.HotRowCache beside ONE unsharded table (a compact, TLB-friendly replica region for tables
    beyond the TLB's reach): the same bits as without it, for hot ids, cold ids and a mix, fp32 and
    bf16; slots reading different tables are refused."""
    ops, O, B, dev = env
    from paddle_infer_b200 import shard
    rng = np.random.default_rng(123)
    N, S, Bsz = 50_000, 4, 600
    for dtype, D, use_cvm in ((torch.float32, 32, True), (torch.bfloat16, 64, None)):
        table = torch.from_numpy(rng.uniform(-1, 1, (N, D)).astype(np.float32)).to(dev)
        table[:, 0] = torch.randint(0, 16, (N,), device=dev).float()
        table[:, 1] = torch.minimum(table[:, 0], torch.randint(0, 16, (N,), device=dev).float())
        table = table.to(dtype)
        lods, ids = [], []
        for s in range(S):
            l = rng.integers(0, 9, Bsz)
            lods.append(np.concatenate([[0], np.cumsum(l)]).astype(np.uint64))
            raw = rng.zipf(1.3, int(lods[-1][-1])) % N if s % 2 == 0 else rng.integers(0, N, int(lods[-1][-1]))
            ids.append(torch.from_numpy(raw.astype(np.int64)).to(dev))
        hot = shard.HotRowCache(table, shard.pick_hot_ids(torch.cat(ids), 3000))
        plain = ops.ctr_fused_emb_seqpool_cvm(table, ids, lods, ["SUM", "AVERAGE", "SQRT", "SUM"], use_cvm=use_cvm, concat=True)
        cached = ops.ctr_fused_emb_seqpool_cvm(table, ids, lods, ["SUM", "AVERAGE", "SQRT", "SUM"], use_cvm=use_cvm, concat=True,
                                               hot_cache=hot)
        assert torch.equal(plain, cached)
        p = ops.ctr_fused_emb_seqpool_cvm(table, ids, lods, "SUM", use_cvm=use_cvm, concat=True, plan_only=True, hot_cache=hot)
        assert "hot=1" in p.kernel_names()[0]
        p.close()
        other = table.clone()
        with pytest.raises(B.CtrfeError, match="same table"):
            ops.ctr_fused_emb_seqpool_cvm([table, other, table, table], ids, lods, "SUM", use_cvm=use_cvm, hot_cache=hot)


def test_plan_lod_image_compact_and_wide_forms(env):
    """A plan's host LoD travels as 16-bit lengths (expanded on the device) when every sequence is
    shorter than 65536, else as 32-bit offsets; ctrfe_plan_update_lod may flip between the forms and
    CTRFE_LAUNCH_UPLOAD_LOD re-uploads whichever is current.  Same bits as the one-shot call."""
    ops, O, B, dev = env
    rng = np.random.default_rng(77)
    N, D = 5000, 16
    table = torch.from_numpy(rng.uniform(-1, 1, (N, D)).astype(np.float32)).to(dev)
    n_seg = 2500  # more than one 2048-sequence step of the expansion kernel

    def make(long_seq):
        l = rng.integers(0, 7, n_seg)
        if long_seq:
            l[1234] = 70000
        lod = np.concatenate([[3], 3 + np.cumsum(l)]).astype(np.uint64)  # a LoD need not start at 0 here
        return lod

    lod_a, lod_b, lod_c = make(False), make(True), make(False)
    n_ids = int(max(lod_a[-1], lod_b[-1], lod_c[-1]))
    ids = torch.from_numpy(rng.integers(0, N, n_ids).astype(np.int64)).to(dev)
    import os
    os.environ["CTRFE_LOD_COMPACT_MIN"] = "0"  # small images take the compact form too (default: from 256 KB)
    try:
        plan = ops.ctr_fused_emb_seqpool_cvm(table, [ids, ids], [lod_a, lod_a], ["SUM", "SQRT"], use_cvm=None, plan_only=True)
    finally:
        del os.environ["CTRFE_LOD_COMPACT_MIN"]
    for lod in (lod_a, lod_b, lod_c, lod_b, lod_a):
        want = ops.ctr_fused_emb_seqpool_cvm(table, [ids, ids], [lod, lod], ["SUM", "SQRT"], use_cvm=None)
        plan.update_lod([lod, lod])
        for o in plan.outs:
            o.fill_(float("nan"))
        ops.launch_many([plan], upload_lod=True)
        torch.cuda.synchronize()
        for g, w in zip(plan.outs, want):
            assert torch.equal(g, w)
    bad = lod_a.copy()
    bad[100] = bad[101] + 1
    with pytest.raises(B.CtrfeError, match="non-decreasing"):
        plan.update_lod([lod_a, bad])
    over = lod_a.copy()
    over[-1] = n_ids + 1
    with pytest.raises(B.CtrfeError, match="exceeds the number of ids"):
        plan.update_lod([over, lod_a])
    plan.close()
