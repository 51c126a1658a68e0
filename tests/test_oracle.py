"""Pins the CPU oracle (oracle/ctr_oracle.c) before anything trusts it:
  * against the golden fixtures produced by the reference's own numpy test formulas
    (tests/golden/make_golden.py), at the reference's own tolerance (op_test.py:1436-1447,
    atol = rtol = 1e-5) -- and much tighter where the arithmetic allows;
  * against the reference's jit::refer::SeqPool / EmbSeqPool compiled into oracle/_ref
    (bit-exact), when that library is present.
"""
import ctypes as C

import numpy as np
import pytest

from oracle import oracle as O


def _last_level(meta):
    return meta["offsets"][-1]


# ---------------------------------------------------------------- sequence_pool goldens
def test_sequence_pool_matches_reference_formulas(golden):
    arrays, meta = golden["sequence_pool"]
    assert len(meta) >= 40
    for cls, m in meta.items():
        x, want = arrays[f"{cls}/x"], arrays[f"{cls}/out"]
        a = m["attrs"]
        got = O.sequence_pool(x, _last_level(m), a["pooltype"], a["pad_value"], is_test=True)
        assert got.shape == want.shape, cls
        if a["pooltype"] in ("MAX", "LAST", "FIRST"):
            np.testing.assert_array_equal(got, want, err_msg=cls)  # selection: bit-exact
        else:
            np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-5, err_msg=cls)
            # numpy sums pairwise in float32; a sequential fp32 sum of <= 13 terms in
            # (0.1, 1) differs by a few ulp at most
            np.testing.assert_allclose(got, want, rtol=4e-7, atol=0, err_msg=cls)


def test_sequence_pool_max_index(golden):
    arrays, meta = golden["sequence_pool"]
    for cls in ("TestSeqMaxPool", "TestSeqMaxPoolLen0", "TestSeqMaxPool2DLen0LoDLevel2"):
        m = meta[cls]
        x = arrays[f"{cls}/x"]
        out, idx = O.sequence_pool(x, _last_level(m), "MAX", m["attrs"]["pad_value"], is_test=False)
        lod = _last_level(m)
        xf = x.reshape(x.shape[0], -1)
        of, idf = out.reshape(out.shape[0], -1), idx.reshape(out.shape[0], -1)
        for i in range(len(lod) - 1):
            if lod[i] == lod[i + 1]:
                assert (idf[i] == -1).all()
            else:
                assert ((idf[i] >= lod[i]) & (idf[i] < lod[i + 1])).all()
                np.testing.assert_array_equal(xf[idf[i], np.arange(xf.shape[1])], of[i])


def test_sequence_pool_double(golden):
    arrays, meta = golden["sequence_pool"]
    for cls in ("TestSeqAvgPoolLen0", "TestSeqSqrtPool2DLen0", "TestSeqSumPool"):
        m = meta[cls]
        x = arrays[f"{cls}/x"].astype(np.float64)
        got = O.sequence_pool(x, _last_level(m), m["attrs"]["pooltype"], m["attrs"]["pad_value"])
        np.testing.assert_allclose(got, arrays[f"{cls}/out"], rtol=1e-6, atol=1e-6)


# ---------------------------------------------------------------- cvm goldens
def test_cvm_matches_reference_formula(golden):
    arrays, meta = golden["cvm"]
    for name, m in meta.items():
        x, want = arrays[f"{name}/x"], arrays[f"{name}/y"]
        for log_mode in (O.LOG_DOUBLE, O.LOG_FLOAT):
            got = O.cvm(x, m["use_cvm"], log_mode)
            assert got.shape == want.shape
            if not m["use_cvm"]:
                np.testing.assert_array_equal(got, want)
                continue
            np.testing.assert_array_equal(got[:, 2:], want[:, 2:])
            np.testing.assert_allclose(got[:, 0], want[:, 0], rtol=2e-7, atol=0)   # <= 1 ulp
            np.testing.assert_allclose(got[:, 1], want[:, 1], rtol=1e-6, atol=1e-6)
        # the double-log mode reproduces test_cvm_op.py's math.log to the bit on column 0
        got = O.cvm(x, True, O.LOG_DOUBLE) if m["use_cvm"] else None
        if got is not None:
            np.testing.assert_array_equal(got[:, 0], want[:, 0])


# ---------------------------------------------------------------- fused CPU op goldens
@pytest.mark.parametrize("scale_mode", [O.SCALE_DIVIDE, O.SCALE_RECIPROCAL])
def test_fusion_seqpool_cvm_concat(golden, scale_mode):
    arrays, meta = golden["fusion_seqpool_cvm_concat"]
    for key, m in meta.items():
        S, w = m["n_slots"], m["w"]
        lods = [np.concatenate([[0], np.cumsum(l[0])]) for l in m["lods_lengths"]]
        xs = [arrays[f"{key}/x{i}"] for i in range(S)]
        # no table here: the fused CPU op takes pre-embedded X.  Express it through the
        # gather path with an identity table (row k of slot i = X_i[k]).
        ids = [np.arange(x.shape[0], dtype=np.int64) for x in xs]
        got = O.fused_forward(xs, ids, lods, [m["pooltype"]] * S, cvm_mode=O.CVM_USE, concat=True,
                              scale_mode=scale_mode)
        want = arrays[f"{key}/out"]
        assert got.shape == want.shape == (len(lods[0]) - 1, S * w)
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-5, err_msg=key)


# ---------------------------------------------------------------- lookup_table_v2 goldens
def test_lookup_table_v2(golden):
    arrays, meta = golden["lookup_table_v2"]
    for name, m in meta.items():
        got = O.embedding(arrays[f"{name}/table"], arrays[f"{name}/ids"], m["padding_idx"])
        np.testing.assert_array_equal(got, arrays[f"{name}/out"], err_msg=name)


def test_embedding_bounds_check():
    table = np.zeros((5, 3), np.float32)
    for bad in (5, -1, 1 << 40):
        with pytest.raises(O.BadId):
            O.embedding(table, np.array([0, bad, 1], np.int64))
    # padding_idx is tested before the range check (embedding_kernel.cc:56-57)
    out = O.embedding(table + 1, np.array([0, 7], np.int64), padding_idx=7)
    assert (out[1] == 0).all() and (out[0] == 1).all()


# ---------------------------------------------------------------- the real reference code
needs_ref = pytest.mark.skipif(O.ref_lib() is None, reason="oracle/_ref not built (no /root/reference)")


@needs_ref
def test_oracle_equals_reference_refer_seqpool():
    """jit/test.cc:628-666 grid: w, h in {1..31 subset, 100, 2000} x {Sum, Avg, Sqrt}."""
    R = O.ref_lib()
    rng = np.random.default_rng(7)
    types = {"SUM": R.ref_seqpool_type_sum(), "AVERAGE": R.ref_seqpool_type_avg(), "SQRT": R.ref_seqpool_type_sqrt()}
    for w in (1, 2, 3, 7, 8, 15, 16, 30, 31, 100, 2000):
        for h in (1, 2, 3, 8, 13, 16, 31, 100, 2000):
            x = rng.uniform(-1, 1, (h, w)).astype(np.float32)
            for name, t in types.items():
                want = np.empty(w, np.float32)
                assert R.ref_seqpool_f32(x.ctypes.data, want.ctypes.data, h, w, t) == 0
                got = O.sequence_pool(x, [0, h], name, scale_mode=O.SCALE_RECIPROCAL)[0]
                np.testing.assert_array_equal(got, want, err_msg=f"{name} h={h} w={w}")
                # the sequence_pool op (true division) agrees within 1e-6 relative
                div = O.sequence_pool(x, [0, h], name, scale_mode=O.SCALE_DIVIDE)[0]
                np.testing.assert_allclose(div, want, rtol=1e-6, atol=1e-12)


@needs_ref
def test_oracle_equals_reference_refer_seqpool_f64():
    R = O.ref_lib()
    rng = np.random.default_rng(8)
    x = rng.uniform(-1, 1, (37, 23))
    for name, t in (("SUM", R.ref_seqpool_type_sum()), ("AVERAGE", R.ref_seqpool_type_avg()),
                    ("SQRT", R.ref_seqpool_type_sqrt())):
        want = np.empty(23)
        assert R.ref_seqpool_f64(x.ctypes.data, want.ctypes.data, 37, 23, t) == 0
        got = O.sequence_pool(x, [0, 37], name, scale_mode=O.SCALE_RECIPROCAL)[0]
        np.testing.assert_array_equal(got, want)


@needs_ref
def test_oracle_equals_reference_refer_embseqpool():
    """jit/test.cc:669-720 grid: tbl_h = 1e4, tbl_w in a few widths, idx_h in {1,2,9,13,16}."""
    R = O.ref_lib()
    rng = np.random.default_rng(9)
    for tbl_w in (1, 2, 8, 16, 31, 64):
        table = rng.uniform(-1, 1, (10000, tbl_w)).astype(np.float32)
        for idx_h in (1, 2, 9, 13, 16):
            idx = rng.integers(0, 10000, idx_h).astype(np.int64)
            want = np.empty(tbl_w, np.float32)
            assert R.ref_embseqpool_f32(table.ctypes.data, 10000, tbl_w, idx.ctypes.data, idx_h, 1,
                                        want.ctypes.data) == 0
            got = O.fused_forward([table], [idx], [[0, idx_h]], ["SUM"], init_mode=O.INIT_FIRST_ROW)[0][0]
            np.testing.assert_array_equal(got, want)
            # zero-initialised sum (refer::SeqPool order) gives the same values
            got0 = O.fused_forward([table], [idx], [[0, idx_h]], ["SUM"])[0][0]
            np.testing.assert_array_equal(got0, want)
    # the reference rejects out-of-range ids (refer.h:493-510); so does the oracle
    bad = np.array([3, 10000], np.int64)
    out = np.empty(8, np.float32)
    t8 = np.zeros((10000, 8), np.float32)
    assert R.ref_embseqpool_f32(t8.ctypes.data, 10000, 8, bad.ctypes.data, 2, 1, out.ctypes.data) != 0
    with pytest.raises(O.BadId):
        O.fused_forward([t8], [bad], [[0, 2]], ["SUM"])


# ---------------------------------------------------------------- internal consistency
def test_fused_equals_unfused_chain_fp32():
    rng = np.random.default_rng(11)
    S, B, D, N = 5, 40, 16, 1000
    table = rng.uniform(-1, 1, (N, D)).astype(np.float32)
    table[:, 0] = rng.integers(0, 16, N)
    table[:, 1] = np.minimum(table[:, 0], rng.integers(0, 16, N))
    lens = [rng.integers(0, 9, B) for _ in range(S)]
    lods = [np.concatenate([[0], np.cumsum(l)]) for l in lens]
    ids = [rng.integers(0, N, int(l[-1])).astype(np.int64) for l in lods]
    pools = ["SUM", "AVERAGE", "SQRT", "MAX", "LAST"]
    for cvm_mode in (O.CVM_NONE, O.CVM_USE, O.CVM_STRIP):
        fused = O.fused_forward([table], ids, lods, pools, pad_value=0.1, padding_idx=int(ids[0][0]),
                                cvm_mode=cvm_mode)
        unf = O.fused_forward([table], ids, lods, pools, pad_value=0.1, padding_idx=int(ids[0][0]),
                              cvm_mode=cvm_mode, unfused=True)
        for s in range(S):
            # step by step with the single-op entry points
            rows = O.embedding(table, ids[s], padding_idx=int(ids[0][0]))
            pooled = O.sequence_pool(rows, lods[s], pools[s], 0.1)
            want = pooled if cvm_mode == O.CVM_NONE else O.cvm(pooled, cvm_mode == O.CVM_USE)
            np.testing.assert_array_equal(fused[s], want)
            np.testing.assert_array_equal(unf[s], want)
    # threads do not change results
    a = O.fused_forward([table], ids, lods, pools, cvm_mode=O.CVM_USE, concat=True, threads=4)
    b = O.fused_forward([table], ids, lods, pools, cvm_mode=O.CVM_USE, concat=True, threads=1)
    np.testing.assert_array_equal(a, b)


def test_half_conversions_match_numpy():
    L = O.lib()
    allh = np.arange(0, 1 << 16, dtype=np.uint16)
    want = allh.view(np.float16).astype(np.float32)
    got = np.array([L.orc_half_to_float(int(h)) for h in allh[::7]], dtype=np.float32)
    w7 = want[::7]
    finite = ~np.isnan(w7)  # signalling-NaN payloads are quietened by the hardware convert numpy uses
    np.testing.assert_array_equal(got.view(np.uint32)[finite], w7.view(np.uint32)[finite])
    assert np.isnan(got[~finite]).all()
    rng = np.random.default_rng(3)
    f = np.concatenate([rng.standard_normal(3000).astype(np.float32) * s for s in (1e-8, 1e-5, 1e-3, 1, 1e3, 1e5)])
    f = np.concatenate([f, np.array([0.0, -0.0, np.inf, -np.inf, 65504, 65519.99, 65520, 6e-8, 2.98e-8, 2.99e-8],
                                    np.float32)])
    with np.errstate(over="ignore"):
        wanth = f.astype(np.float16).view(np.uint16)
    goth = np.array([L.orc_float_to_half(float(v)) for v in f], dtype=np.uint16)
    np.testing.assert_array_equal(goth, wanth)
    # bf16: vectorised helper == C scalar; round trip of exact values
    gotb = np.array([L.orc_float_to_bf16(float(v)) for v in f], dtype=np.uint16)
    np.testing.assert_array_equal(O.f32_to_bf16_bits(f), gotb)
    np.testing.assert_array_equal(O.f32_to_bf16_bits(O.bf16_bits_to_f32(gotb)), gotb)


def test_flush_to_zero_like_the_executor():
    """naive_executor.cc:66 runs kernels under FTZ+DAZ: a denormal summand counts as 0."""
    x = np.array([[1e-40], [1e-40]], np.float32)
    assert O.sequence_pool(x, [0, 2], "SUM")[0, 0] == 0.0
    # MAX compares under DAZ too: a denormal is not greater than 0
    x = np.array([[0.0], [1e-40]], np.float32)
    assert O.sequence_pool(x, [0, 2], "MAX")[0, 0] == 0.0


def test_fused_seqpool_cvm_paddlebox_mode():
    """fused_seqpool_cvm_op.cu:33-102: accumulator starts at pad_value, also for empty segments."""
    rng = np.random.default_rng(5)
    x = rng.uniform(0, 3, (20, 6)).astype(np.float32)
    lod = [0, 3, 3, 10, 20]
    out = O.fused_seqpool_cvm(x, lod, pad_value=0.25, use_cvm=True)
    acc = np.stack([0.25 + x[a:b].astype(np.float64).sum(0) for a, b in zip(lod[:-1], lod[1:])])
    np.testing.assert_allclose(out[:, 2:], acc[:, 2:], rtol=1e-6)
    np.testing.assert_allclose(out[:, 0], np.log(acc[:, 0] + 1), rtol=1e-6)
    np.testing.assert_allclose(out[:, 1], np.log(acc[:, 1] + 1) - np.log(acc[:, 0] + 1), rtol=1e-5, atol=1e-6)
    out2 = O.fused_seqpool_cvm(x, lod, pad_value=0.25, use_cvm=False, cvm_offset=2)
    np.testing.assert_allclose(out2, acc[:, 2:], rtol=1e-6)
    # equals the generic fused path with INIT_PAD over an identity table
    g = O.fused_forward([x], [np.arange(20, dtype=np.int64)], [lod], ["SUM"], pad_value=0.25,
                        init_mode=O.INIT_PAD, cvm_mode=O.CVM_STRIP)[0]
    np.testing.assert_array_equal(g, out2)


def test_pull_sparse_restatement():
    """pull_box_sparse semantics: values by key, zeros for key 0 and for unknown keys."""
    keys = np.array([7, 2**63 + 5, 11, 3], dtype=np.uint64)
    vals = np.arange(12, dtype=np.float32).reshape(4, 3) + 1
    q = np.array([3, 0, 2**63 + 5, 999, 7, 7], dtype=np.uint64)
    got = O.pull_sparse(keys, vals, q)
    np.testing.assert_array_equal(got, np.stack([vals[3], np.zeros(3), vals[1], np.zeros(3), vals[0], vals[0]]))
    # int64 views of the same bits (how Paddle holds feasigns) behave the same
    np.testing.assert_array_equal(O.pull_sparse(keys.view(np.int64), vals, q.view(np.int64)), got)
    assert O.pull_sparse(np.zeros(0, np.uint64), np.zeros((0, 3), np.float32), q).shape == (6, 3)


def test_backward_restatements_against_the_reference_formulas():
    """sequence_pool_grad / cvm_grad / lookup_table_v2_grad of the oracle vs the formulas of
    sequence_pooling.cc:438-496, cvm_op.h:43-58 and embedding_grad_kernel.cc:63-88 written out
    in numpy."""
    rng = np.random.default_rng(9)
    lod = [0, 3, 3, 7, 8, 8]
    w = 5
    dout = rng.normal(size=(5, w)).astype(np.float32)
    x = rng.normal(size=(8, w)).astype(np.float32)
    _, idx = O.sequence_pool(x, lod, "MAX", is_test=False)
    for pt in ("SUM", "AVERAGE", "SQRT", "LAST", "FIRST", "MAX"):
        want = np.zeros((8, w), np.float32)
        for i in range(5):
            a, e = lod[i], lod[i + 1]
            if a == e:
                continue
            h = np.float32(e - a)
            if pt == "SUM":
                want[a:e] = dout[i]
            elif pt == "AVERAGE":
                want[a:e] = dout[i] / h
            elif pt == "SQRT":
                want[a:e] = dout[i] / np.sqrt(h)
            elif pt == "LAST":
                want[e - 1] = dout[i]
            elif pt == "FIRST":
                want[a] = dout[i]
            else:
                for c in range(w):
                    want[idx[i, c], c] = dout[i, c]
        got = O.sequence_pool_grad(dout, lod, pt, max_index=idx if pt == "MAX" else None)
        np.testing.assert_array_equal(got, want, err_msg=pt)
    # rows past the last offset stay zero
    assert (O.sequence_pool_grad(dout, lod, "SUM", k_rows=10)[8:] == 0).all()
    # cvm_grad: without LoD, and with X carrying LoD (one CVM row per sequence)
    dy = rng.normal(size=(4, 6)).astype(np.float32)
    cv = rng.uniform(0, 5, (4, 2)).astype(np.float32)
    np.testing.assert_array_equal(O.cvm_grad(dy, cv, True), np.concatenate([cv, dy[:, 2:]], axis=1))
    np.testing.assert_array_equal(O.cvm_grad(dy, cv, False), np.concatenate([cv, dy], axis=1))
    np.testing.assert_array_equal(O.cvm_grad(dy, cv[:2], True, lod=[0, 1, 4]),
                                  np.concatenate([cv[[0, 1, 1, 1]], dy[:, 2:]], axis=1))
    # embedding grad: in-order accumulation, padding rows skipped, bad ids refused
    ids = np.array([3, 1, 3, 0, 2, 3], np.int64)
    g = rng.normal(size=(6, 4)).astype(np.float32)
    want = np.zeros((5, 4), np.float32)
    for i, r in enumerate(ids):
        if r != 2:
            want[r] += g[i]
    np.testing.assert_array_equal(O.embedding_grad(g, ids, 5, padding_idx=2), want)
    with pytest.raises(IndexError):
        O.embedding_grad(g, np.array([0, 1, 5, 0, 0, 0]), 5)


@pytest.mark.parametrize("lod", [[0, 10], [0, 2, 7, 10]])
def test_reference_gtest_sequence_pooling_grad_sum(lod):
    """The reference's own functor test, replayed on the oracle
    (operators/math/sequence_pooling_test.cc:20-145): Out@GRAD[i, m] = i * 128 + m, width 128;
    every row of sequence i of X@GRAD must equal row i of Out@GRAD, exactly."""
    w = 128
    dout = np.arange((len(lod) - 1) * w, dtype=np.float32).reshape(-1, w)
    dx = O.sequence_pool_grad(dout, lod, "SUM")
    assert dx.shape == (lod[-1], w)
    for i in range(len(lod) - 1):
        for r in range(lod[i], lod[i + 1]):
            np.testing.assert_array_equal(dx[r], dout[i])
