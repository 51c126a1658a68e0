"""GPU parity tests proper: the sm_100a kernels, called through the C ABI (via the ctypes
host mirror), against the CPU oracle on identical inputs and against the committed golden
fixtures.  Bars (BASELINE.json north_star):
  * gathered rows and MAX/LAST/FIRST pooling: bit-exact;
  * SUM/AVERAGE/SQRT in fp32: bit-exact (summation order is preserved), asserted as such;
  * CVM column 0: <= 1 ulp of the oracle's double-log (rtol 2e-7); column 1 (a cancellation):
    rtol 1e-6 + atol 1e-6 (the reference's own tests use atol 1e-5, op_test.py:1776);
  * fp16/bf16: rtol 1e-2 required, bit-exact observed (single rounding, same order).
"""
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
    # the native library must be the thing that runs
    assert B.lib().ctrfe_version() >= 100
    return ops, O, B, torch.device("cuda:0")


def _t(a, dev, bf16=False):
    t = torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    return t.view(torch.bfloat16) if bf16 else t


def _np(t):
    if t.dtype == torch.bfloat16:
        return t.view(torch.int16).cpu().numpy().view(np.uint16)
    return t.cpu().numpy()


def _assert_cvm_close(got, want, use_cvm, exact_tail=True):
    got, want = np.asarray(got, np.float32), np.asarray(want, np.float32)
    if use_cvm:
        np.testing.assert_allclose(got[..., 0], want[..., 0], rtol=2e-7, atol=0)
        np.testing.assert_allclose(got[..., 1], want[..., 1], rtol=1e-6, atol=1e-6)
        np.testing.assert_array_equal(got[..., 2:], want[..., 2:])
    else:
        np.testing.assert_array_equal(got, want)


# ------------------------------------------------------------------ golden fixtures on the GPU
def test_golden_lookup_table_v2(env, golden):
    ops, O, B, dev = env
    arrays, meta = golden["lookup_table_v2"]
    for name, m in meta.items():
        table, ids = arrays[f"{name}/table"], arrays[f"{name}/ids"]
        out = ops.lookup_table_v2(_t(table, dev), _t(ids, dev), m["padding_idx"])
        assert tuple(out.shape) == ids.shape + (31,)
        np.testing.assert_array_equal(out.cpu().numpy(), arrays[f"{name}/out"], err_msg=name)


def test_golden_sequence_pool(env, golden):
    ops, O, B, dev = env
    arrays, meta = golden["sequence_pool"]
    for cls, m in meta.items():
        x, want = arrays[f"{cls}/x"], arrays[f"{cls}/out"]
        a = m["attrs"]
        r = ops.sequence_pool(_t(x, dev), m["offsets"], a["pooltype"], a["pad_value"], is_test=a.get("is_test", False))
        got = r.Out.cpu().numpy()
        assert got.shape == want.shape, cls
        # vs the reference's numpy formula at the reference's tolerance
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-5, err_msg=cls)
        # vs the oracle: bit-exact
        np.testing.assert_array_equal(got, O.sequence_pool(x, m["offsets"][-1], a["pooltype"], a["pad_value"]), err_msg=cls)
        # two-level LoD: Out carries level 0 (sequence_pool_op.h:62-70)
        assert r.lod == ([m["offsets"][0]] if len(m["offsets"]) == 2 else [])
        if a["pooltype"] == "MAX" and not a.get("is_test", False):
            _, idx = O.sequence_pool(x, m["offsets"][-1], "MAX", a["pad_value"], is_test=False)
            np.testing.assert_array_equal(r.MaxIndex.cpu().numpy(), idx, err_msg=cls)
        else:
            assert r.MaxIndex is None


def test_golden_cvm(env, golden):
    ops, O, B, dev = env
    arrays, meta = golden["cvm"]
    for name, m in meta.items():
        x, want = arrays[f"{name}/x"], arrays[f"{name}/y"]
        got = ops.cvm(_t(x, dev), None, m["use_cvm"]).cpu().numpy()
        assert got.shape == want.shape
        _assert_cvm_close(got, want, m["use_cvm"])
        _assert_cvm_close(got, O.cvm(x, m["use_cvm"]), m["use_cvm"])


def test_golden_fusion_seqpool_cvm_concat(env, golden):
    ops, O, B, dev = env
    arrays, meta = golden["fusion_seqpool_cvm_concat"]
    for key, m in meta.items():
        S = m["n_slots"]
        from paddle_infer_b200.lod import lengths_to_offsets
        lods = [lengths_to_offsets(l) for l in m["lods_lengths"]]
        xs = [arrays[f"{key}/x{i}"] for i in range(S)]
        got = ops.fusion_seqpool_cvm_concat([_t(x, dev) for x in xs], lods, None, m["pooltype"]).cpu().numpy()
        want = arrays[f"{key}/out"]
        assert got.shape == want.shape
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-5, err_msg=key)
        ids = [np.arange(x.shape[0], dtype=np.int64) for x in xs]
        orc = O.fused_forward(xs, ids, [l[0] for l in lods], [m["pooltype"]] * S, cvm_mode=O.CVM_USE, concat=True,
                              scale_mode=O.SCALE_RECIPROCAL)
        w = m["w"]
        _assert_cvm_close(got.reshape(-1, S, w), orc.reshape(-1, S, w), True)


# ------------------------------------------------------------------ the fused path, swept
def _random_case(rng, S, Bsz, N, D, max_len, dtype, empty_frac=0.15):
    if dtype == "f32":
        table = rng.uniform(-1, 1, (N, D)).astype(np.float32)
    elif dtype == "f16":
        table = rng.uniform(-1, 1, (N, D)).astype(np.float16)
    else:
        from oracle import oracle as O
        table = O.f32_to_bf16_bits(rng.uniform(-1, 1, (N, D)).astype(np.float32))
    if D >= 2:  # show / click columns: small non-negative integers, click <= show
        show = rng.integers(0, 16, N)
        click = rng.integers(0, show + 1)
        if dtype == "bf16":
            from oracle import oracle as O
            table[:, 0] = O.f32_to_bf16_bits(show.astype(np.float32))
            table[:, 1] = O.f32_to_bf16_bits(click.astype(np.float32))
        else:
            table[:, 0], table[:, 1] = show, click
    lods, ids = [], []
    for _ in range(S):
        l = rng.integers(0, max_len + 1, Bsz)
        l[rng.random(Bsz) < empty_frac] = 0
        lod = np.concatenate([[0], np.cumsum(l)]).astype(np.uint64)
        lods.append(lod)
        ids.append(rng.integers(0, N, int(lod[-1])).astype(np.int64))
    return table, ids, lods


POOLS = ["SUM", "AVERAGE", "SQRT", "MAX", "LAST", "FIRST"]


@pytest.mark.parametrize("dtype", ["f32", "f16", "bf16"])
@pytest.mark.parametrize("D", [8, 16, 64, 128, 11, 3, 200, 1000])
def test_fused_matches_oracle(env, dtype, D):
    """All pooltypes x cvm modes; D covers 128-bit lanes (8..128), the scalar fallback
    (11, 3: the reference fixtures' widths), multi-tile rows (200, 1000)."""
    ops, O, B, dev = env
    rng = np.random.default_rng({"f32": 1, "f16": 2, "bf16": 3}[dtype] * 10007 + D)
    S, Bsz, N = 6, 97, 500
    table, ids, lods = _random_case(rng, S, Bsz, N, D, 9, dtype)
    bf16 = dtype == "bf16"
    tt = _t(table, dev, bf16)
    ids_t = [_t(i, dev) for i in ids]
    pad_idx = int(ids[0][0])
    for use_cvm, cvm_mode in ((None, O.CVM_NONE), (True, O.CVM_USE), (False, O.CVM_STRIP)):
        got = ops.ctr_fused_emb_seqpool_cvm(tt, ids_t, lods, POOLS, pad_value=0.1, padding_idx=pad_idx,
                                            use_cvm=use_cvm)
        want = O.fused_forward([table], ids, lods, POOLS, pad_value=0.1, padding_idx=pad_idx, cvm_mode=cvm_mode,
                               bf16=bf16)
        for s in range(S):
            g, w = _np(got[s]), want[s]
            assert g.shape == w.shape
            if dtype == "f32":
                _assert_cvm_close(g, w, use_cvm is True)
            else:
                gf = O.bf16_bits_to_f32(g) if bf16 else g.astype(np.float32)
                wf = O.bf16_bits_to_f32(w) if bf16 else w.astype(np.float32)
                np.testing.assert_allclose(gf, wf, rtol=1e-2, atol=1e-3)       # the stated bar
                tail = slice(2, None) if use_cvm is True else slice(None)
                np.testing.assert_array_equal(g[:, tail], w[:, tail])            # observed: bit-exact
                if POOLS[s] in ("MAX", "LAST", "FIRST"):
                    np.testing.assert_array_equal(g[:, tail], w[:, tail])


def test_fused_concat_int32_ids_and_per_slot_tables(env):
    ops, O, B, dev = env
    rng = np.random.default_rng(123)
    S, Bsz, N, D = 5, 64, 300, 16
    tabs, ids, lods = [], [], []
    for s in range(S):
        t, i, l = _random_case(rng, 1, Bsz, N + s, D, 7, "f32")
        tabs.append(t), ids.append(i[0].astype(np.int32)), lods.append(l[0])
    got = ops.ctr_fused_emb_seqpool_cvm([_t(t, dev) for t in tabs], [_t(i, dev) for i in ids], lods,
                                        ["SUM", "SQRT", "AVERAGE", "MAX", "SUM"], use_cvm=True, concat=True)
    want = O.fused_forward(tabs, ids, lods, ["SUM", "SQRT", "AVERAGE", "MAX", "SUM"], cvm_mode=O.CVM_USE, concat=True)
    _assert_cvm_close(got.cpu().numpy().reshape(Bsz, S, D), want.reshape(Bsz, S, D), True)


def test_fused_modes_reciprocal_firstrow_padinit(env):
    ops, O, B, dev = env
    rng = np.random.default_rng(77)
    table, ids, lods = _random_case(rng, 3, 50, 200, 16, 12, "f32", empty_frac=0.0)
    tt, ids_t = _t(table, dev), [_t(i, dev) for i in ids]
    for kw_gpu, kw_orc in (
        (dict(scale_mode=B.SCALE_RECIPROCAL), dict(scale_mode=O.SCALE_RECIPROCAL)),
        (dict(init_mode=B.INIT_PAD, pad_value=0.25), dict(init_mode=O.INIT_PAD, pad_value=0.25)),
    ):
        got = ops.ctr_fused_emb_seqpool_cvm(tt, ids_t, lods, ["SUM", "AVERAGE", "SQRT"], use_cvm=None, **kw_gpu)
        want = O.fused_forward([table], ids, lods, ["SUM", "AVERAGE", "SQRT"], **kw_orc)
        for g, w in zip(got, want):
            np.testing.assert_array_equal(g.cpu().numpy(), w)
    # first-row init (EmbSeqPool order, refer.h:513-528) on non-empty sequences, incl. a -0.0 row
    table2 = table.copy()
    table2[5] = -0.0
    lens = rng.integers(1, 6, 40)
    lod = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    idv = rng.integers(0, 200, int(lod[-1])).astype(np.int64)
    idv[lod[:-1].astype(np.int64)] = 5  # every sequence starts with the -0.0 row
    out = ops.fused_embedding_seq_pool(_t(table2, dev), _t(idv, dev).reshape(-1, 1), [lod])
    want = O.fused_forward([table2], [idv], [lod], ["SUM"], init_mode=O.INIT_FIRST_ROW)[0]
    np.testing.assert_array_equal(out.cpu().numpy().view(np.uint32), want.view(np.uint32))
    R = O.ref_lib()
    if R is not None:  # and the reference's own EmbSeqPool agrees to the bit
        ref = np.empty_like(want)
        assert R.ref_embseqpool_lod_f32(table2.ctypes.data, 200, 16, idv.ctypes.data, lod.ctypes.data, 40,
                                        ref.ctypes.data, 16) == 0
        np.testing.assert_array_equal(out.cpu().numpy().view(np.uint32), ref.view(np.uint32))


@pytest.mark.parametrize("tbl_w,idx_w", [(16, 2), (8, 3), (64, 5), (11, 4)])
def test_fused_embedding_seq_pool_index_width(env, tbl_w, idx_w):
    """Ids whose last dimension is wider than 1 (jit/refer/refer.h:477-528, the grid of
    jit/test.cc:669-720): column j of Ids pools into columns [j*D, (j+1)*D) of Out.  Checked
    per sequence against the reference's OWN refer::EmbSeqPool compiled into oracle/_ref, to the
    bit, and against a plain numpy statement when that library is absent."""
    ops, O, B, dev = env
    rng = np.random.default_rng(tbl_w * 10 + idx_w)
    N = 300
    table = rng.uniform(-1, 1, (N, tbl_w)).astype(np.float32)
    lens = rng.integers(1, 9, 37)
    lod = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    ids = rng.integers(0, N, (int(lod[-1]), idx_w)).astype(np.int64)
    out = ops.fused_embedding_seq_pool(_t(table, dev), _t(ids, dev), [lod]).cpu().numpy()
    assert out.shape == (37, tbl_w * idx_w)
    R = O.ref_lib()
    for i in range(37):
        sub = np.ascontiguousarray(ids[int(lod[i]):int(lod[i + 1])])
        if R is not None:
            ref = np.empty(tbl_w * idx_w, np.float32)
            assert R.ref_embseqpool_f32(table.ctypes.data, N, tbl_w, sub.ctypes.data, sub.shape[0], idx_w, ref.ctypes.data) == 0
        else:
            ref = np.empty(tbl_w * idx_w, np.float32)
            for w in range(idx_w):
                acc = table[sub[0, w]].copy()
                for h in range(1, sub.shape[0]):
                    acc = acc + table[sub[h, w]]
                ref[w * tbl_w:(w + 1) * tbl_w] = acc
        np.testing.assert_array_equal(out[i].view(np.uint32), ref.view(np.uint32))
    bad = ids.copy()
    bad[5, idx_w - 1] = N + 9
    with pytest.raises(ValueError) as e:
        ops.fused_embedding_seq_pool(_t(table, dev), _t(bad, dev), [lod])
    assert e.value.slot == idx_w - 1 and e.value.position == 5 and e.value.id_value == N + 9


def test_fused_seqpool_cvm_paddlebox(env):
    ops, O, B, dev = env
    rng = np.random.default_rng(31)
    xs, lods = [], []
    for s in range(4):
        lens = rng.integers(0, 6, 33)
        lod = [0] + list(np.cumsum(lens))
        x = rng.uniform(0, 2, (int(lod[-1]), 12)).astype(np.float32)
        xs.append(x), lods.append([lod])
    for use_cvm in (True, False):
        got = ops.fused_seqpool_cvm([_t(x, dev) for x in xs], lods, None, "SUM", 0.5, use_cvm, 2)
        for x, l, g in zip(xs, lods, got):
            want = O.fused_seqpool_cvm(x, l[0], 0.5, use_cvm, 2, log_mode=O.LOG_DOUBLE)
            _assert_cvm_close(g.cpu().numpy(), want, use_cvm)
            # and within 1e-6 of the float-log variant the reference GPU kernel computes
            wf = O.fused_seqpool_cvm(x, l[0], 0.5, use_cvm, 2, log_mode=O.LOG_FLOAT)
            np.testing.assert_allclose(g.cpu().numpy(), wf, rtol=1e-6, atol=1e-6)


# ------------------------------------------------------------------ edge cases
def test_edge_cases(env):
    ops, O, B, dev = env
    rng = np.random.default_rng(5)
    table = rng.uniform(-1, 1, (64, 16)).astype(np.float32)
    tt = _t(table, dev)
    # all sequences empty
    lod = np.zeros(10, np.uint64)
    out = ops.ctr_fused_emb_seqpool_cvm(tt, [torch.empty(0, dtype=torch.int64, device=dev)], [lod], "SUM",
                                        pad_value=0.3, use_cvm=None)[0]
    assert (out.cpu().numpy() == np.float32(0.3)).all()
    # one long sequence (many id batches), one sequence of one row, trailing empties
    k = 3000
    ids = rng.integers(0, 64, k + 1).astype(np.int64)
    lod = np.array([0, k, k + 1, k + 1, k + 1], np.uint64)
    for pool in POOLS:
        got = ops.ctr_fused_emb_seqpool_cvm(tt, [_t(ids, dev)], [lod], pool, use_cvm=None)[0].cpu().numpy()
        want = O.fused_forward([table], [ids], [lod], [pool])[0]
        np.testing.assert_array_equal(got, want, err_msg=pool)
    # NaN / inf / -0.0 / denormal rows: MAX keeps the reference's "first row, then strictly greater"
    t2 = table.copy()
    t2[0, :] = np.nan
    t2[1, :] = -np.inf
    t2[2, :] = -0.0
    t2[3, :] = 0.0
    t2[4, :] = 1e-40  # denormal: flushed in arithmetic, compared as zero (FTZ/DAZ)
    ids = np.array([0, 7, 1, 1, 2, 3, 3, 2, 8, 0, 4, 3, 4, 4], np.int64)
    lod = np.array([0, 2, 4, 6, 8, 10, 12, 14], np.uint64)
    for pool in ("MAX", "SUM", "FIRST", "LAST"):
        got = ops.ctr_fused_emb_seqpool_cvm(_t(t2, dev), [_t(ids, dev)], [lod], pool, use_cvm=None)[0].cpu().numpy()
        want = O.fused_forward([t2], [ids], [lod], [pool])[0]
        if pool == "SUM":
            # arithmetic on a NaN yields the device's canonical NaN (0x7fffffff) where x86 keeps the
            # quiet input payload: equal as values (NaN == NaN here), not as bits; -0.0 vs 0.0 is checked
            np.testing.assert_array_equal(got, want, err_msg=pool)
            np.testing.assert_array_equal(np.signbit(got), np.signbit(want) & ~np.isnan(want) | np.signbit(got) & np.isnan(want))
        else:  # selection pools move bits
            np.testing.assert_array_equal(got.view(np.uint32), want.view(np.uint32), err_msg=pool)


def test_bad_ids_raise_like_the_reference(env):
    ops, O, B, dev = env
    table = torch.ones(32, 8, device=dev)
    for bad in (32, -1, 1 << 40):
        ids = torch.tensor([1, 2, bad, 3], dtype=torch.int64, device=dev)
        with pytest.raises(B.BadIdError) as e:
            ops.lookup_table_v2(table, ids)
        assert e.value.position == 2 and e.value.id_value == bad
        assert "expected >= 0 and <" in str(e.value)
        with pytest.raises(ValueError) as e:
            ops.ctr_fused_emb_seqpool_cvm(table, [ids[:2], ids], [[0, 1, 2], [0, 2, 4]], "SUM", use_cvm=None)
        assert e.value.slot == 1 and e.value.position == 2
    # padding_idx is honoured before the range check, and the flag is clear again
    ids = torch.tensor([1, 99], dtype=torch.int64, device=dev)
    out = ops.lookup_table_v2(table, ids, padding_idx=99)
    assert out[0].eq(1).all() and out[1].eq(0).all()
    # host-side argument checks mirror the reference's enforce messages
    with pytest.raises(ops.InvalidArgument, match="does not contain LoD"):
        ops.sequence_pool(table, [], "SUM")
    with pytest.raises(ops.InvalidArgument, match="no more than 2"):
        ops.sequence_pool(table, [[0, 1], [0, 1], [0, 32]], "SUM")
    with pytest.raises(ops.InvalidArgument, match="illegal"):
        ops.sequence_pool(table, [[0, 3], [0, 32]], "SUM")
    with pytest.raises(ops.InvalidArgument, match="rank should be 2"):
        ops.cvm(torch.ones(2, 3, 4, device=dev))
    with pytest.raises(Exception, match="non-decreasing"):
        ops.sequence_pool(table, [[0, 5, 3, 32]], "SUM")


def test_unfused_chain_equals_fused(env):
    """lookup_table_v2 -> sequence_pool -> cvm as three launches equals the fused launch (fp32)."""
    ops, O, B, dev = env
    rng = np.random.default_rng(2)
    table, ids, lods = _random_case(rng, 2, 300, 1000, 16, 8, "f32")
    tt = _t(table, dev)
    for s, pool in enumerate(("SUM", "MAX")):
        rows = ops.lookup_table_v2(tt, _t(ids[s], dev))
        np.testing.assert_array_equal(rows.cpu().numpy(), table[ids[s]])
        pooled = ops.sequence_pool(rows, [list(lods[s])], pool, 0.0, is_test=True).Out
        y = ops.cvm(pooled, None, True)
        fused = ops.ctr_fused_emb_seqpool_cvm(tt, [_t(ids[s], dev)], [lods[s]], pool, use_cvm=True)[0]
        np.testing.assert_array_equal(y.cpu().numpy(), fused.cpu().numpy())


def test_plan_reuse_and_device_lod(env):
    ops, O, B, dev = env
    rng = np.random.default_rng(9)
    table, ids, lods = _random_case(rng, 4, 128, 2000, 16, 8, "f32")
    tt, ids_t = _t(table, dev), [_t(i, dev) for i in ids]
    lods_dev = [torch.from_numpy(l.astype(np.int64)).to(dev) for l in lods]
    plan = ops.ctr_fused_emb_seqpool_cvm(tt, ids_t, lods_dev, "SUM", use_cvm=True, concat=True, plan_only=True)
    assert plan.num_kernels() == 1
    want = O.fused_forward([table], ids, lods, ["SUM"] * 4, cvm_mode=O.CVM_USE, concat=True)
    for _ in range(3):
        plan.outs.zero_()
        plan.launch()
        torch.cuda.synchronize()
        _assert_cvm_close(plan.outs.cpu().numpy().reshape(128, 4, 16), want.reshape(128, 4, 16), True)
    plan.ctx.check_ids()
    plan.close()


# ------------------------------------------------------------------ BASELINE configs
def _run_config(env, cfg, batch=None, rows=None, slots=None, sample=None):
    """Run one synthetic config on the GPU; check a sample of sequences against the oracle
    fed with exactly the table rows the sample touches (regenerated on the host)."""
    ops, O, B, dev = env
    from benchtools import synth
    b = synth.draw_batch(cfg, 0, batch, slots, rows)
    N = cfg.rows if rows is None else rows
    S = len(b.ids)
    Bsz = b.lods[0].size - 1
    tdt = {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}[cfg.dtype]
    n_tab = 1 if cfg.shared_table else S
    tabs = []
    for t in range(n_tab):
        tab = torch.empty((N, cfg.dim), dtype=tdt, device=dev)
        synth.fill_table(tab, cfg.seed + 1000 * t, cvm_cols=cfg.use_cvm is not None)
        tabs.append(tab)
    ids_t = [torch.from_numpy(i).to(dev) for i in b.ids]
    out = ops.ctr_fused_emb_seqpool_cvm(tabs if n_tab > 1 else tabs[0], ids_t, b.lods, cfg.pooltypes[:S],
                                        use_cvm=cfg.use_cvm, concat=True)
    torch.cuda.synchronize()
    got = _np(out).reshape(Bsz, S, cfg.d_out)
    # oracle on a sample of sequences, with a compact table of just the rows they need
    rng = np.random.default_rng(1)
    pick = np.arange(Bsz) if sample is None or sample >= Bsz else np.sort(rng.choice(Bsz, sample, replace=False))
    cvm_mode = O.CVM_NONE if cfg.use_cvm is None else (O.CVM_USE if cfg.use_cvm else O.CVM_STRIP)
    for s in range(S):
        lod = b.lods[s].astype(np.int64)
        seg_ids = [b.ids[s][lod[i]:lod[i + 1]] for i in pick]
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
        if cfg.dtype == "f32":
            _assert_cvm_close(g, want, cfg.use_cvm is True)
        else:
            np.testing.assert_array_equal(g, want)
    return got, b


def test_config1_full(env):
    from benchtools import synth
    _run_config(env, synth.configs()["cfg1"])


def test_config2_full_size_sampled(env):
    """DeepFM: 39 slots x 8192, one shared 1e8 x 16 fp32 table (6.4 GB), Zipf ids."""
    from benchtools import synth
    cfg = synth.configs()["cfg2"]
    got, b = _run_config(env, cfg, sample=512)
    # size-independent property on the full output: column 2.. of an empty sequence is pad_value (0)
    for s in (0, 17, 38):
        empty = np.diff(b.lods[s].astype(np.int64)) == 0
        assert empty.any() and (got[empty, s, 2:] == 0).all()


def test_config3_reduced_rows(env):
    """DLRM: 26 tables, bf16 D=64, mixed SUM/AVERAGE/SQRT, no CVM; rows cut to 1e6 per table
    and batch to 4096 so 26 tables fit beside the other tests (full size runs in bench.py)."""
    from benchtools import synth
    _run_config(env, synth.configs()["cfg3"], batch=4096, rows=1_000_000, sample=256)


def test_config4_reduced_rows_one_gpu(env):
    """The sharded config's shape on one GPU: fp32 D=32 (128-byte rows), SUM + CVM; rows cut to
    2e6, batch to 4096.  Takes the balanced-load kernel."""
    from benchtools import synth
    _run_config(env, synth.configs()["cfg4"], batch=4096, rows=2_000_000, sample=512)


def _full_size(env, name, sample):
    """One BASELINE config at its stated table and batch size on one GPU.  The tables are
    generated on the device; the oracle gets exactly the rows a sample of sequences touches,
    regenerated on the host from the same (seed, row, column) hash -- so the 64-bit address
    arithmetic and the 32-bit row ids are exercised where only a > 100 GB table exercises them."""
    from benchtools import synth
    cfg = synth.configs()[name]
    need = cfg.n_tables() * cfg.rows * cfg.dim * cfg.elem_size + (4 << 30)
    torch.cuda.empty_cache()
    free, _total = torch.cuda.mem_get_info()
    if free < need:
        pytest.skip(f"{name} at full size needs {need / 1e9:.0f} GB of HBM, {free / 1e9:.0f} GB free")
    try:
        got, b = _run_config(env, cfg, sample=sample)
        assert got.shape == (cfg.batch, cfg.slots, cfg.d_out)
        # the rows with the highest ids in the batch sit at the far end of the table
        top = max(int(i.max()) for i in b.ids if i.size)
        assert top > 0.99 * cfg.rows
    finally:
        torch.cuda.empty_cache()


def test_config3_full_size(env):
    """DLRM at BASELINE size: 26 tables of 4e7 x 64 bf16 (133 GB), batch 65536, SUM/AVERAGE/SQRT."""
    _full_size(env, "cfg3", 256)


def test_config4_full_size_one_gpu(env):
    """The sharded config's table whole on one GPU: 1e9 x 32 fp32 (128 GB), batch 131072, SUM + CVM."""
    _full_size(env, "cfg4", 256)


def test_config5_full_size(env):
    """Long sequences at BASELINE size: 8 tables of 1e7 x 128 fp16, batch 4096, mean 200 ids, MAX / SQRT."""
    _full_size(env, "cfg5", 128)


@pytest.mark.parametrize("dtype,D", [("f32", 32), ("f16", 64), ("bf16", 64), ("f32", 16)])
def test_specialised_kernels_ragged_long_and_bad(env, dtype, D):
    """The specialised kernels (128-byte rows: balanced loads through shared memory; 64-byte
    rows: warp-level loop) on shapes the configs do not have: segment counts that do not fill
    a warp or a CTA, runs of ids longer than one batch (a 300-row and a 70-row segment among
    short ones), all-empty slots, int32 ids, and a bad id in the middle of a long run."""
    ops, O, B, dev = env
    rng = np.random.default_rng(D * 131 + len(dtype))
    N, S = 3000, 5
    table, _, _ = _random_case(rng, 1, 1, N, D, 1, dtype)
    bf16 = dtype == "bf16"
    tt = _t(table, dev, bf16)
    lods, ids = [], []
    for s, n_seg in enumerate([1, 7, 29, 333, 64]):
        l = rng.integers(0, 9, n_seg)
        if n_seg > 20:
            l[3], l[n_seg - 2] = 300, 70
        if s == 4:
            l[:] = 0
        lod = np.concatenate([[0], np.cumsum(l)]).astype(np.uint64)
        lods.append(lod)
        ids.append(rng.integers(0, N, int(lod[-1])).astype(np.int64))
    pools = ["SUM", "AVERAGE", "SQRT", "SUM", "SUM"]
    for use_cvm, cvm_mode in ((True, O.CVM_USE), (None, O.CVM_NONE)):
        for as_i32 in (False, True):
            ids_t = [_t(i.astype(np.int32) if as_i32 else i, dev) for i in ids]
            got = ops.ctr_fused_emb_seqpool_cvm(tt, ids_t, lods, pools, pad_value=0.5, use_cvm=use_cvm)
            want = O.fused_forward([table], ids, lods, pools, pad_value=0.5, cvm_mode=cvm_mode, bf16=bf16)
            for s in range(S):
                g, w = _np(got[s]), want[s]
                assert g.shape == w.shape
                if dtype == "f32":
                    _assert_cvm_close(g, w, use_cvm is True)
                else:
                    np.testing.assert_array_equal(g, w)
    # MAX over the same shapes (rows of 128 bytes and wider take the balanced-load kernel's
    # selecting form), with special values in the table: NaN, infinities, signed zeros
    tspec = table.copy()
    if dtype == "f32":
        tspec[5, :4] = [np.nan, np.inf, -np.inf, -0.0]
        tspec[6, :4] = [0.0, -np.inf, np.nan, 0.0]
    elif dtype == "f16":
        tspec[5, :4] = np.array([np.nan, np.inf, -np.inf, -0.0], np.float16)
        tspec[6, :4] = np.array([0.0, -np.inf, np.nan, 0.0], np.float16)
    else:
        tspec[5, :4] = O.f32_to_bf16_bits(np.array([np.nan, np.inf, -np.inf, -0.0], np.float32))
        tspec[6, :4] = O.f32_to_bf16_bits(np.array([0.0, -np.inf, np.nan, 0.0], np.float32))
    ids_m = [i.copy() for i in ids]
    for s in range(S):
        if ids_m[s].size > 6:
            ids_m[s][::7] = 5
            ids_m[s][3::11] = 6
    got = ops.ctr_fused_emb_seqpool_cvm(_t(tspec, dev, bf16), [_t(i, dev) for i in ids_m], lods, "MAX", pad_value=0.5,
                                        use_cvm=None)
    want = O.fused_forward([tspec], ids_m, lods, ["MAX"] * S, pad_value=0.5, bf16=bf16)
    for s in range(S):
        g, w = _np(got[s]), want[s]
        if dtype == "f32":
            np.testing.assert_array_equal(g.view(np.uint32), w.view(np.uint32))
        else:
            np.testing.assert_array_equal(g.view(np.uint16), w.view(np.uint16))
    bad = [i.copy() for i in ids]
    pos = int(lods[3][3]) + 150   # inside the 300-row segment
    bad[3][pos] = N + 5
    with pytest.raises(ValueError) as e:
        ops.ctr_fused_emb_seqpool_cvm(tt, [_t(i, dev) for i in bad], lods, pools, use_cvm=None)
    assert e.value.slot == 3 and e.value.position == pos and e.value.id_value == N + 5


def test_config5_long_sequences(env):
    """8 slots x avg 200 ids, D=128 fp16, MAX and SQRT; rows cut to 2e6 per table."""
    from benchtools import synth
    _run_config(env, synth.configs()["cfg5"], batch=512, rows=2_000_000, sample=128)


def test_linearity_property_full_batch(env):
    """SUM pooling is linear in the table: pool(2*T) == 2*pool(T) exactly (power-of-two scale),
    checked on every output of a cfg2-shaped batch without the oracle."""
    ops, O, B, dev = env
    from benchtools import synth
    cfg = synth.configs()["cfg2"]
    b = synth.draw_batch(cfg, 3, batch=8192, slots=8, rows=1_000_000)
    tab = torch.empty((1_000_000, 16), dtype=torch.float32, device=dev)
    synth.fill_table(tab, cfg.seed)
    ids_t = [torch.from_numpy(i).to(dev) for i in b.ids]
    a = ops.ctr_fused_emb_seqpool_cvm(tab, ids_t, b.lods, "SUM", use_cvm=None, concat=True)
    c = ops.ctr_fused_emb_seqpool_cvm(tab * 2, ids_t, b.lods, "SUM", use_cvm=None, concat=True)
    assert torch.equal(a * 2, c)
    # checksum of checksums: sum over all pooled rows == sum over all gathered rows (fp64)
    total_rows = sum(tab[i].double().sum().item() for i in ids_t)
    assert abs(a.double().sum().item() - total_rows) <= 1e-6 * max(1.0, abs(total_rows))


def test_host_front_end_modes(env):
    """The host-buffer doorway (what bench.py's e2e leg calls): zero-copy (the kernel reads the
    pinned ids / LoD and writes the pinned output itself), the same with buffers bound once, and
    the explicit-copy mode all equal the device-resident call."""
    ops, O, B, dev = env
    rng = np.random.default_rng(44)
    S, Bsz, N, D = 5, 200, 3000, 16
    table, ids, lods = _random_case(rng, S, Bsz, N, D, 9, "f32")
    tt = _t(table, dev)
    want = O.fused_forward([table], ids, lods, ["SUM"] * S, cvm_mode=O.CVM_USE)
    ids_p = [torch.from_numpy(i).pin_memory() for i in ids]
    lods_p = [torch.from_numpy(l.astype(np.int64)).pin_memory() for l in lods]
    max_ids = max(i.size for i in ids)
    fe = ops.HostFrontEnd(tt, S, Bsz, max_ids, "SUM", use_cvm=True, mode="zero_copy")
    got = fe.run(ids_p, lods_p, check=True)
    for s in range(S):
        assert not got[s].is_cuda
        _assert_cvm_close(got[s].numpy(), want[s], True)
    fe.host_out.zero_()
    call = fe.bind(ids_p, lods_p)
    got = call.run()
    for s in range(S):
        _assert_cvm_close(got[s].numpy(), want[s], True)
    assert call.h2d_bytes == sum(i.size * 8 for i in ids) + S * (Bsz + 1) * 8
    assert call.d2h_bytes == S * Bsz * D * 4
    # new ids in the same bound buffers: the next run() sees them
    ids_p[0].copy_(torch.from_numpy(np.roll(ids[0], 1)))
    want0 = O.fused_forward([table], [np.roll(ids[0], 1)], [lods[0]], ["SUM"], cvm_mode=O.CVM_USE)[0]
    _assert_cvm_close(call.run()[0].numpy(), want0, True)
    ids_p[0].copy_(torch.from_numpy(ids[0]))
    # asynchronous form: submit() now, result() later, two requests in flight on two front ends
    fe_b = ops.HostFrontEnd(tt, S, Bsz, max_ids, "SUM", use_cvm=True, mode="zero_copy")
    call_b = fe_b.bind(ids_p, lods_p)
    fe.host_out.zero_()
    ev_a = call.submit()
    ev_b = call_b.submit(overlap_previous=True)
    got_a, got_b = call.result(ev_a), call_b.result(ev_b)
    for s in range(S):
        _assert_cvm_close(got_a[s].numpy(), want[s], True)
        assert torch.equal(got_a[s], got_b[s])
    # explicit-copy mode, concat layout
    fe2 = ops.HostFrontEnd(tt, S, Bsz, max_ids, "SUM", use_cvm=True, mode="copy", concat=True)
    got2 = fe2.run(ids_p, lods, check=True).numpy().reshape(Bsz, S, D)
    for s in range(S):
        _assert_cvm_close(got2[:, s, :], want[s], True)
    with pytest.raises(ops.InvalidArgument, match="pinned"):
        fe.run([torch.from_numpy(i) for i in ids], lods_p)


# ------------------------------------------------------------------ feasign key map (pull_box_sparse front)
def test_feasign_keymap_pull_and_fused(env):
    """uint64 feasign keys resolved inside the kernel: the standalone pull (values by key) is a
    bit-copy, and keys -> pool -> CVM in one launch equals the oracle on the pulled values."""
    ops, O, B, dev = env
    rng = np.random.default_rng(2026)
    n, D, S, Bsz = 200_000, 16, 5, 300
    keys = np.unique(rng.integers(1, 2**64 - 1, n + 1000, dtype=np.uint64))[:n]
    rng.shuffle(keys)
    assert keys.size == n and (keys >> np.uint64(63)).any()   # some keys use the top bit
    table = rng.uniform(-1, 1, (n, D)).astype(np.float32)
    show = rng.integers(0, 16, n)
    table[:, 0], table[:, 1] = show, rng.integers(0, show + 1)
    km = ops.KeyMap(_t(keys.view(np.int64), dev))
    assert km.n_keys == n and (1 << km.log2_slots) >= 2 * n
    tt = _t(table, dev)
    lods, qs = [], []
    for s in range(S):
        l = rng.integers(0, 9, Bsz)
        lod = np.concatenate([[0], np.cumsum(l)]).astype(np.uint64)
        q = keys[rng.integers(0, n, int(lod[-1]))].copy()
        q[rng.random(q.size) < 0.05] = 0                                   # padding feasign
        miss = rng.random(q.size) < 0.05
        q[miss] = rng.integers(1, 2**64 - 1, int(miss.sum()), dtype=np.uint64)  # (almost surely) unknown keys
        lods.append(lod)
        qs.append(q)
    q_t = [_t(q.view(np.int64), dev) for q in qs]
    # the op on its own
    pulled = ops.pull_box_sparse(km, tt, q_t)
    for s in range(S):
        np.testing.assert_array_equal(pulled[s].cpu().numpy().view(np.uint32), O.pull_sparse(keys, table, qs[s]).view(np.uint32))
    # fused: keys -> rows -> pool -> CVM, nothing in between
    padded = np.concatenate([table, np.zeros((1, D), np.float32)])
    rows = [O.pull_sparse_rows(keys, q, zero_row=n) for q in qs]
    pools = ["SUM", "AVERAGE", "MAX", "SQRT", "LAST"]
    for use_cvm, cvm_mode in ((True, O.CVM_USE), (None, O.CVM_NONE)):
        got = ops.ctr_fused_emb_seqpool_cvm(tt, q_t, lods, pools, use_cvm=use_cvm, keymap=km)
        want = O.fused_forward([padded], rows, lods, pools, cvm_mode=cvm_mode)
        for s in range(S):
            _assert_cvm_close(got[s].cpu().numpy(), want[s], use_cvm is True)
    # same values through PaddleBox's two-op form: pull, then fused_seqpool_cvm over the values
    two_step = ops.fused_seqpool_cvm(pulled, [[l.tolist()] for l in lods], None, "SUM", 0.0, True)
    one_step = ops.ctr_fused_emb_seqpool_cvm(tt, q_t, lods, "SUM", use_cvm=True, keymap=km)
    for s in range(S):
        assert torch.equal(two_step[s], one_step[s])
    # int32 ids cannot be feasign keys; sharded tables cannot take a key map
    with pytest.raises((ops.InvalidArgument, ValueError, RuntimeError)):
        ops.ctr_fused_emb_seqpool_cvm(tt, [q.to(torch.int32) for q in q_t], lods, "SUM", keymap=km)


def test_overlapped_plan_launches_equal_plain_ones(env):
    """ctrfe_plan_launch_ex(CTRFE_LAUNCH_OVERLAP_PREVIOUS): a burst of launches over independent
    batches, each allowed to start while the previous one drains, writes the same bits as
    ordinary stream-ordered launches."""
    ops, O, B, dev = env
    rng = np.random.default_rng(77)
    S, Bsz, N, D = 8, 3000, 20000, 16
    table, _, _ = _random_case(rng, 1, 1, N, D, 1, "f32")
    tt = _t(table, dev)
    plans, wants = [], []
    for b in range(6):
        _, ids, lods = _random_case(rng, S, Bsz, N, D, 9, "f32")
        ids_t = [_t(i, dev) for i in ids]
        lods_t = [_t(l.astype(np.int64), dev) for l in lods]
        plans.append(ops.ctr_fused_emb_seqpool_cvm(tt, ids_t, lods_t, "SUM", use_cvm=True, concat=True, plan_only=True))
        wants.append((ids, lods))
    for p in plans:
        p.launch()
    torch.cuda.synchronize()
    plain = [p.outs.clone() for p in plans]
    for p in plans:
        p.outs.zero_()
    for rep in range(3):
        for p in plans:
            p.launch(overlap_previous=True)
    plans[0].ctx.check_ids()
    for p, ref in zip(plans, plain):
        assert torch.equal(p.outs, ref)
    ids, lods = wants[0]
    want = O.fused_forward([table], ids, lods, ["SUM"] * S, cvm_mode=O.CVM_USE, concat=True)
    _assert_cvm_close(plans[0].outs.cpu().numpy().reshape(Bsz, S, D), want.reshape(Bsz, S, D), True)
    with pytest.raises(B.CtrfeError):
        B.check(B.lib().ctrfe_plan_launch_ex(plans[0].handle, plans[0].ctx.stream(), 2))
    # a plan with two launches (accumulating and selecting pools -> specialised + general kernel),
    # overlapped with itself and with its neighbours
    mixed = []
    for b in range(3):
        _, ids, lods = _random_case(rng, 4, 2000, N, D, 30, "f32")
        p = ops.ctr_fused_emb_seqpool_cvm(tt, [_t(i, dev) for i in ids], [_t(l.astype(np.int64), dev) for l in lods],
                                          ["SUM", "MAX", "SQRT", "LAST"], use_cvm=None, concat=True, plan_only=True)
        assert p.num_kernels() == 2
        mixed.append((p, O.fused_forward([table], ids, lods, ["SUM", "MAX", "SQRT", "LAST"], concat=True)))
    for rep in range(3):
        for p, _ in mixed:
            p.launch(overlap_previous=True)
    mixed[0][0].ctx.check_ids()
    for p, want in mixed:
        np.testing.assert_array_equal(p.outs.cpu().numpy(), want)


# ------------------------------------------------------------------ backward ops
def test_backward_ops_match_the_oracle(env):
    ops, O, B, dev = env
    rng = np.random.default_rng(404)
    for w, n_seg in ((5, 9), (16, 400), (100, 77)):
        l = rng.integers(0, 8, n_seg)
        l[::5] = 0
        lod = np.concatenate([[0], np.cumsum(l)]).astype(np.uint64)
        K = int(lod[-1])
        x = rng.normal(size=(K, w)).astype(np.float32)
        dout = rng.normal(size=(n_seg, w)).astype(np.float32)
        fwd = ops.sequence_pool(_t(x, dev), [lod.tolist()], "MAX", is_test=False)
        _, idx = O.sequence_pool(x, lod, "MAX", is_test=False)
        for pt in ("SUM", "AVERAGE", "SQRT", "LAST", "FIRST", "MAX"):
            got = ops.sequence_pool_grad(_t(dout, dev), [lod.tolist()], pt, MaxIndex=fwd.MaxIndex if pt == "MAX" else None)
            want = O.sequence_pool_grad(dout, lod, pt, max_index=idx if pt == "MAX" else None)
            np.testing.assert_array_equal(got.cpu().numpy(), want, err_msg=f"{pt} w={w}")
        # X longer than its LoD covers: the uncovered rows get zeros
        got = ops.sequence_pool_grad(_t(dout, dev), [lod.tolist()], "SUM", x_rows=K + 3)
        np.testing.assert_array_equal(got.cpu().numpy(), O.sequence_pool_grad(dout, lod, "SUM", k_rows=K + 3))
        # cvm_grad on the pooled rows (no LoD) and on LoD rows
        cv = rng.uniform(0, 9, (n_seg, 2)).astype(np.float32)
        if w > 2:
            for use_cvm in (True, False):
                dy = rng.normal(size=(n_seg, w if use_cvm else w - 2)).astype(np.float32)
                got = ops.cvm_grad(_t(dy, dev), _t(cv, dev), use_cvm)
                np.testing.assert_array_equal(got.cpu().numpy(), O.cvm_grad(dy, cv, use_cvm))
            dyk = rng.normal(size=(K, w)).astype(np.float32)
            got = ops.cvm_grad(_t(dyk, dev), _t(cv, dev), True, lod=[lod.tolist()])
            np.testing.assert_array_equal(got.cpu().numpy(), O.cvm_grad(dyk, cv, True, lod=lod))
        # embedding grad (atomics: equal up to rounding)
        N = 50
        ids = rng.integers(0, N, K).astype(np.int64)
        pad = int(ids[0]) if K else -1
        for id_dt in (np.int64, np.int32):
            got = ops.lookup_table_v2_grad(_t(x, dev), _t(ids.astype(id_dt), dev), N, padding_idx=pad)
            np.testing.assert_allclose(got.cpu().numpy(), O.embedding_grad(x, ids, N, padding_idx=pad), rtol=1e-5, atol=1e-5)
    bad = np.array([1, 2, 60, 3], np.int64)
    with pytest.raises(ValueError) as e:
        ops.lookup_table_v2_grad(_t(np.ones((4, 3), np.float32), dev), _t(bad, dev), 50)
    assert e.value.position == 2 and e.value.id_value == 60


def test_more_slots_than_one_launch_carries(env):
    """150 slots of mixed pooltypes: descriptors travel in the kernel parameters 64 at a time, so
    this is five launches (3 accumulating + 2 selecting... whatever the split, more than two),
    which overlap each other; every slot must still match the oracle, with concat and without."""
    ops, O, B, dev = env
    rng = np.random.default_rng(150)
    S, Bsz, N, D = 150, 40, 700, 16
    table, ids, lods = _random_case(rng, S, Bsz, N, D, 6, "f32")
    tt = _t(table, dev)
    pools = [POOLS[s % len(POOLS)] for s in range(S)]
    ids_t = [_t(i, dev) for i in ids]
    want = O.fused_forward([table], ids, lods, pools, cvm_mode=O.CVM_USE)
    plan = ops.ctr_fused_emb_seqpool_cvm(tt, ids_t, [_t(l.astype(np.int64), dev) for l in lods], pools, use_cvm=True,
                                         concat=True, plan_only=True)
    assert plan.num_kernels() >= 3
    for overlap in (False, True, True):
        plan.outs.zero_()
        plan.launch(overlap_previous=overlap)
        plan.ctx.check_ids()
        got = plan.outs.cpu().numpy().reshape(Bsz, S, D)
        for s in range(S):
            _assert_cvm_close(got[:, s, :], want[s], True)
    outs = ops.ctr_fused_emb_seqpool_cvm(tt, ids_t, lods, pools, use_cvm=True)
    for s in range(S):
        _assert_cvm_close(outs[s].cpu().numpy(), want[s], True)


@pytest.mark.parametrize("switch", ["CTRFE_NO_FAST=1"])
def test_every_kernel_variant_passes_the_shape_sweep(env, switch):
    """The library picks its kernel per process (environment switches read once): re-run the
    ragged / long / special-value sweep and the reduced cfg3 / cfg4 / cfg5 runs in a child process
    with each switch, so that the general kernel on everything (CTRFE_NO_FAST=1) meets the same
    bit-exact bar as the specialised kernels the configs normally take."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    k, v = switch.split("=")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(root, "tests", "test_gpu_parity.py"), "-x", "-q", "-m", "gpu",
                        "-k", "specialised or reduced_rows or long_sequences"],
                       capture_output=True, text=True, timeout=900, cwd=root, env=dict(os.environ, **{k: v}))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]


@pytest.mark.parametrize("lod", [[0, 10], [0, 2, 7, 10]])
def test_reference_gtest_sequence_pooling_grad_sum_on_gpu(env, lod):
    """TEST(SequencePoolingGrad, CUDA_SUM) of the reference (sequence_pooling_test.cc:133-145):
    same fixture, same EXPECT_EQ bar, through ctrfe_sequence_pool_grad."""
    ops, O, B, dev = env
    w = 128
    dout = np.arange((len(lod) - 1) * w, dtype=np.float32).reshape(-1, w)
    dx = ops.sequence_pool_grad(_t(dout, dev), [lod], "SUM").cpu().numpy()
    assert dx.shape == (lod[-1], w)
    for i in range(len(lod) - 1):
        for r in range(lod[i], lod[i + 1]):
            np.testing.assert_array_equal(dx[r], dout[i])


def test_analyzer_seq_pool1_shape(env):
    """The reference's end-to-end CTR inference test model in shape
    (inference/tests/api/analyzer_seq_pool1_tester_helper.h:34,66-71,136-145): 154 slots of
    pre-embedded LoD float rows, width 11 (2 CVM columns + 9), SUM pool -> cvm -> concat, which
    seqpool_cvm_concat_fuse_pass turns into one fusion_seqpool_cvm_concat op.  The model file is
    a download the reference does not ship, so values are random; the fused op must equal the
    oracle slot by slot."""
    ops, O, B, dev = env
    rng = np.random.default_rng(154)
    S, Bsz, w = 154, 16, 11
    xs, lods = [], []
    for s in range(S):
        l = rng.integers(0, 5, Bsz)
        lod = np.concatenate([[0], np.cumsum(l)]).astype(np.uint64)
        x = rng.uniform(0.1, 1, (int(lod[-1]), w)).astype(np.float32)
        x[:, 0] = rng.integers(1, 9, x.shape[0])
        x[:, 1] = rng.integers(0, 2, x.shape[0])
        xs.append(x)
        lods.append(lod)
    got = ops.fusion_seqpool_cvm_concat([_t(x, dev) for x in xs], [[l.tolist()] for l in lods], None, "SUM").cpu().numpy()
    assert got.shape == (Bsz, S * w)
    ids = [np.arange(x.shape[0], dtype=np.int64) for x in xs]
    want = O.fused_forward(xs, ids, lods, ["SUM"] * S, cvm_mode=O.CVM_USE, concat=True, scale_mode=O.SCALE_RECIPROCAL)
    _assert_cvm_close(got.reshape(Bsz, S, w), want.reshape(Bsz, S, w), True)
