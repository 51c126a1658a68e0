"""Launched under torchrun (one process per GPU) by tests/test_sharded.py and by hand:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29611 tests/dist_sharded_check.py

Every rank owns rows r % G == rank of a logical table, maps its peers' shards over NVLink
(CUDA IPC), pools its own slice of the batch with remote rows fetched by the kernel itself,
and checks the result bit-for-bit against the CPU oracle on the unsharded table."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import __graft_entry__ as g
    if rank == 0:
        g.build()
    dist.barrier()
    from benchtools import synth
    from oracle import oracle as O
    from paddle_infer_b200 import ops, shard

    N, D, seed, S, Bsz = 200_003, 32, 9, 6, 512
    table = shard.ShardedTable(N, D, torch.float32, sharing=os.environ.get("CTRFE_SHARING", "vmm"))
    assert table.local_rows == shard.shard_num_rows(N, world, rank)
    synth_rows = shard.global_rows_of_shard(N, world, rank)
    table.local.copy_(torch.from_numpy(synth.table_rows_numpy(seed, synth_rows, D, "f32")))
    torch.cuda.synchronize()
    dist.barrier()  # every shard is filled before anyone reads

    full = synth.table_rows_numpy(seed, np.arange(N), D, "f32")
    rng = np.random.default_rng(100 + rank)
    lods, ids = [], []
    for _ in range(S):
        l = rng.integers(0, 9, Bsz)
        lod = np.concatenate([[0], np.cumsum(l)]).astype(np.uint64)
        lods.append(lod)
        ids.append(rng.integers(0, N, int(lod[-1])).astype(np.int64))
    ids_t = [torch.from_numpy(i).to(dev) for i in ids]
    pools = ["SUM", "AVERAGE", "MAX", "SQRT", "SUM", "LAST"]
    got = ops.ctr_fused_emb_seqpool_cvm(table, ids_t, lods, pools, use_cvm=None)
    want = O.fused_forward([full], ids, lods, pools)
    for s in range(S):
        np.testing.assert_array_equal(got[s].cpu().numpy(), want[s])
    # with CVM, concatenated
    got = ops.ctr_fused_emb_seqpool_cvm(table, ids_t, lods, "SUM", use_cvm=True, concat=True).cpu().numpy()
    want = O.fused_forward([full], ids, lods, ["SUM"] * S, cvm_mode=O.CVM_USE, concat=True)
    np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-6)
    np.testing.assert_array_equal(got.reshape(Bsz, S, D)[:, :, 2:], want.reshape(Bsz, S, D)[:, :, 2:])
    # hot-row replicas: same bits, fewer NVLink look-ups
    plain = ops.ctr_fused_emb_seqpool_cvm(table, ids_t, lods, "SUM", use_cvm=True, concat=True)
    table.hot = shard.HotRowCache(table, shard.pick_hot_ids(torch.cat(ids_t), 4096))
    cached = ops.ctr_fused_emb_seqpool_cvm(table, ids_t, lods, "SUM", use_cvm=True, concat=True)
    assert torch.equal(plain, cached)
    table.hot = None
    # the bucket / NCCL all-to-all router (the measured comparator): same bits, with and without dedup
    for dedup in (False, True):
        routed = shard.RoutedForward(table, dedup=dedup)(ids_t, lods, "SUM", use_cvm=True, concat=True)
        assert torch.equal(plain, routed), f"router (dedup={dedup}) differs from the peer-load kernel"
    frac_remote = float(np.mean(np.concatenate(ids) % world != rank))
    dist.barrier()
    table.close()
    if rank == 0:
        print(f"sharded ok: {world} GPUs, {frac_remote:.0%} of rows fetched over NVLink, bit-exact vs the oracle; "
              "hot-row replicas and the NCCL all-to-all router (with and without dedup) give the same bits")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
