"""Row-sharded table mode (BASELINE config 4): placement arithmetic and handle exchange on CPU
with a world_size-2 gloo group; the sharded kernel path on one GPU (several shards in one
process); the real multi-process NVLink path when the box has >= 2 GPUs."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ------------------------------------------------------------------ host logic (no GPU)
def test_placement_arithmetic():
    from paddle_infer_b200 import shard
    for n_rows in (0, 1, 7, 8, 9, 1000, 1_000_003):
        for G in (2, 4, 8):
            counts = [shard.shard_num_rows(n_rows, G, g) for g in range(G)]
            assert sum(counts) == n_rows
            ids = np.arange(n_rows, dtype=np.int64)
            own, loc = shard.owner_of(ids, G), shard.local_row_of(ids, G)
            for g in range(G):
                mine = ids[own == g]
                assert mine.size == counts[g]
                np.testing.assert_array_equal(mine, shard.global_rows_of_shard(n_rows, G, g))
                np.testing.assert_array_equal(loc[own == g], np.arange(counts[g]))   # dense, in order
    with pytest.raises(ValueError):
        shard.check_n_shards(3)
    with pytest.raises(ValueError):
        shard.shard_num_rows(10, 4, 4)


def _gloo_worker(rank, world, port, q):
    """One simulated GPU: builds its shard of the logical table, exchanges handles and shards
    over gloo, and checks that assembling rows by (owner, local row) reproduces the table and
    that the pooled result of its local batch equals the unsharded oracle."""
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        sys.path.insert(0, ROOT)
        import torch.distributed as dist
        from benchtools import synth
        from oracle import oracle as O
        from paddle_infer_b200 import shard
        dist.init_process_group("gloo", rank=rank, world_size=world)
        N, D, seed = 5003, 8, 77
        mine = shard.global_rows_of_shard(N, world, rank)
        my_shard = synth.table_rows_numpy(seed, mine, D, "f32")
        assert my_shard.shape[0] == shard.shard_num_rows(N, world, rank)
        # the 64-byte handle exchange used at set-up (dummy handles here: no GPU)
        handles = shard.exchange_handles(bytes([rank]) * 64)
        assert handles == [bytes([r]) * 64 for r in range(world)]
        # stand-in for the NVLink mapping: every rank gets to see every shard
        shards = [None] * world
        dist.all_gather_object(shards, my_shard)
        # this rank's slice of a data-parallel batch
        rng = np.random.default_rng(1000 + rank)
        lens = rng.integers(0, 7, 50)
        lod = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
        ids = rng.integers(0, N, int(lod[-1])).astype(np.int64)
        # fetch each row from its owner at its local row: what the kernel's address math does
        own, loc = shard.owner_of(ids, world), shard.local_row_of(ids, world)
        rows = np.stack([shards[o][l] for o, l in zip(own, loc)]) if ids.size else np.zeros((0, D), np.float32)
        full = synth.table_rows_numpy(seed, np.arange(N), D, "f32")
        np.testing.assert_array_equal(rows, full[ids])
        got = O.fused_forward([rows], [np.arange(ids.size, dtype=np.int64)], [lod], ["SUM"], cvm_mode=O.CVM_USE)[0]
        want = O.fused_forward([full], [ids], [lod], ["SUM"], cvm_mode=O.CVM_USE)[0]
        np.testing.assert_array_equal(got, want)
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, "FAIL: " + traceback.format_exc()))
        raise e


def test_two_rank_gloo_sharding():
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res


def _gloo_router_worker(rank, world, port, q, dedup):
    """The all-to-all router (shard.RoutedForward) over gloo: bucketing by owner, the three counted
    exchanges and the inverse permutation are the product's code; the owner-side gather and the
    pooling are numpy / oracle stand-ins, since there is no GPU here."""
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        sys.path.insert(0, ROOT)
        import types
        import torch
        import torch.distributed as dist
        from benchtools import synth
        from oracle import oracle as O
        from paddle_infer_b200 import shard
        dist.init_process_group("gloo", rank=rank, world_size=world)
        N, D, seed, S = 3001, 8, 21, 3
        mine = shard.global_rows_of_shard(N, world, rank)
        local = torch.from_numpy(synth.table_rows_numpy(seed, mine, D, "f32"))
        table = types.SimpleNamespace(n_shards=world, rank=rank, group=dist.group.WORLD, local=local, dim=D,
                                      dtype=torch.float32)

        def pool(rows_in, back_slots, lods, pooltype, **kw):
            outs = O.fused_forward([rows_in.numpy()] * len(back_slots), [b.numpy() for b in back_slots], lods,
                                   [pooltype] * len(back_slots), cvm_mode=O.CVM_USE)
            return np.concatenate(outs, axis=1)

        router = shard.RoutedForward(table, dedup=dedup, gather=lambda w, r: w[r], pool=pool)
        rng = np.random.default_rng(500 + rank)          # every rank its own batch, ragged, one slot empty
        lods, Ids = [], []
        for s_ in range(S):
            lens = rng.integers(0, 6, 40) if s_ != 1 or rank != 0 else np.zeros(40, np.int64)
            lods.append(np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64))
            Ids.append(torch.from_numpy(rng.integers(0, 64 if s_ == 0 else N, int(lods[-1][-1])).astype(np.int64)))
        got = router(Ids, lods, "SUM")
        full = synth.table_rows_numpy(seed, np.arange(N), D, "f32")
        want = np.concatenate(O.fused_forward([full] * S, [i.numpy() for i in Ids], lods, ["SUM"] * S,
                                              cvm_mode=O.CVM_USE), axis=1)
        np.testing.assert_array_equal(got, want)
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, "FAIL: " + traceback.format_exc()))
        raise e


@pytest.mark.parametrize("dedup", [False, True])
def test_two_rank_gloo_router(dedup):
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_router_worker, args=(r, 2, port, q, dedup)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res


# ------------------------------------------------------------------ the kernel path, one GPU
@pytest.mark.gpu
@pytest.mark.parametrize("G", [2, 4, 8])
def test_sharded_kernel_on_one_gpu(G):
    """Several shards in one process: the sharded addressing (owner = id % G, local row = id // G)
    gives bit-identical results to the single-table kernel and to the oracle."""
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import __graft_entry__ as g
    g.build()
    from benchtools import synth
    from oracle import oracle as O
    from paddle_infer_b200 import ops, shard
    dev = torch.device("cuda:0")
    N, D, seed, S, Bsz = 40_001, 32, 5, 4, 300
    full = synth.table_rows_numpy(seed, np.arange(N), D, "f32")
    shards = [torch.from_numpy(full[shard.global_rows_of_shard(N, G, k)]).to(dev) for k in range(G)]
    table = shard.LocalShards(shards)
    assert table.n_rows == N
    rng = np.random.default_rng(G)
    lods, ids = [], []
    for _ in range(S):
        l = rng.integers(0, 9, Bsz)
        lod = np.concatenate([[0], np.cumsum(l)]).astype(np.uint64)
        lods.append(lod)
        ids.append(rng.integers(0, N, int(lod[-1])).astype(np.int64))
    ids_t = [torch.from_numpy(i).to(dev) for i in ids]
    pools = ["SUM", "AVERAGE", "MAX", "SQRT"]
    got = ops.ctr_fused_emb_seqpool_cvm(table, ids_t, lods, pools, use_cvm=None)
    ref = ops.ctr_fused_emb_seqpool_cvm(torch.from_numpy(full).to(dev), ids_t, lods, pools, use_cvm=None)
    want = O.fused_forward([full], ids, lods, pools)
    for s in range(S):
        assert torch.equal(got[s], ref[s])
        np.testing.assert_array_equal(got[s].cpu().numpy(), want[s])
    # bad ids are still caught against the LOGICAL row count
    bad = torch.tensor([1, N, 2], dtype=torch.int64, device=dev)
    with pytest.raises(ValueError):
        ops.ctr_fused_emb_seqpool_cvm(table, [bad], [[0, 3]], "SUM", use_cvm=None)


@pytest.mark.gpu
def test_hot_row_cache_changes_nothing_but_the_source_of_the_rows():
    """Replicas of the most frequent rows (shard.HotRowCache): same bits out, with and without."""
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import __graft_entry__ as g
    g.build()
    from benchtools import synth
    from oracle import oracle as O
    from paddle_infer_b200 import ops, shard
    dev = torch.device("cuda:0")
    G, N, D, seed, S, Bsz = 4, 50_021, 32, 11, 3, 400
    full = synth.table_rows_numpy(seed, np.arange(N), D, "f32")
    full[7] = -0.0  # a row of negative zeros must be replicated with its sign
    table = shard.LocalShards([torch.from_numpy(full[shard.global_rows_of_shard(N, G, k)]).to(dev) for k in range(G)])
    rng = np.random.default_rng(5)
    lods, ids = [], []
    for _ in range(S):
        l = rng.integers(0, 9, Bsz)
        lod = np.concatenate([[0], np.cumsum(l)]).astype(np.uint64)
        lods.append(lod)
        ids.append(synth.zipf_ids(rng, int(lod[-1]), N))   # a real head: most look-ups hit few rows
    ids[0][:5] = 7
    ids_t = [torch.from_numpy(i).to(dev) for i in ids]
    plain = ops.ctr_fused_emb_seqpool_cvm(table, ids_t, lods, "SUM", use_cvm=True)
    hot_ids = shard.pick_hot_ids(torch.cat(ids_t), 500)
    assert hot_ids.numel() == 500 and 7 in hot_ids.tolist()
    table.hot = shard.HotRowCache(table, hot_ids)
    # every replica sits at its key's hash slot and is an exact copy of the table row
    keys = table.hot.keys.cpu().numpy().view(np.uint32)
    used = np.nonzero(keys != 0xffffffff)[0]
    assert 490 <= used.size <= 500 and set(keys[used].tolist()) <= set(hot_ids.tolist())
    np.testing.assert_array_equal(table.hot.rows.cpu().numpy().view(np.uint32)[used], full[keys[used]].view(np.uint32))
    hit = np.isin(np.concatenate(ids), hot_ids.cpu().numpy()).mean()
    assert hit > 0.4
    cached = ops.ctr_fused_emb_seqpool_cvm(table, ids_t, lods, "SUM", use_cvm=True)
    want = O.fused_forward([full], ids, lods, ["SUM"] * S, cvm_mode=O.CVM_USE)
    for s in range(S):
        assert torch.equal(plain[s], cached[s])
        np.testing.assert_array_equal(cached[s].cpu().numpy()[:, 2:], want[s][:, 2:])
    # AVERAGE without CVM, and a cache so small that its hash is nearly full
    table.hot = shard.HotRowCache(table, hot_ids[:10])
    a = ops.ctr_fused_emb_seqpool_cvm(table, ids_t, lods, "AVERAGE", use_cvm=None)
    table.hot = None
    b = ops.ctr_fused_emb_seqpool_cvm(table, ids_t, lods, "AVERAGE", use_cvm=None)
    for s in range(S):
        assert torch.equal(a[s], b[s])


# ------------------------------------------------------------------ the NVLink path, >= 2 GPUs
@pytest.mark.gpu
def test_sharded_across_processes():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    n = 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", "29611", os.path.join(ROOT, "tests", "dist_sharded_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "sharded ok" in r.stdout
