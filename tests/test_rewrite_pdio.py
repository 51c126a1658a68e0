"""the program rewriter (ProgramDesc in, ProgramDesc out) and the
`.pdiparams` tensor-stream reader, without Paddle installed."""
import io
import os

import numpy as np
import pytest

from paddle_infer_b200 import pdio
from paddle_infer_b200 import rewrite as R


# ------------------------------------------------------------------ a DeepFM-shaped front end
def _front_end_program(n_slots=4, shared_table=True, pooltype="SUM", extra_consumer=False):
    prog = R.ProgramDesc()
    blk = prog.blocks.add()
    blk.idx, blk.parent_idx = 0, -1
    blk.vars.append(b"\x0a\x03emb")  # an opaque VarDesc: must survive untouched
    prog.version = b"\x08\x00"
    ys = []
    for i in range(n_slots):
        w = "emb" if shared_table else f"emb_{i}"
        blk.ops.add().CopyFrom(R.make_op("lookup_table_v2", {"W": [w], "Ids": [f"ids_{i}"]}, {"Out": [f"rows_{i}"]},
                                         [("padding_idx", R.LONG, -1)]))
        blk.ops.add().CopyFrom(R.make_op("sequence_pool", {"X": [f"rows_{i}"]},
                                         {"Out": [f"pool_{i}"], "MaxIndex": [f"maxidx_{i}"]},
                                         [("pooltype", R.STRING, pooltype), ("pad_value", R.FLOAT, 0.0),
                                          ("is_test", R.BOOLEAN, True)]))
        blk.ops.add().CopyFrom(R.make_op("cvm", {"X": [f"pool_{i}"], "CVM": ["cvm_in"]}, {"Y": [f"y_{i}"]},
                                         [("use_cvm", R.BOOLEAN, True)]))
        ys.append(f"y_{i}")
    if extra_consumer:  # a second reader of an intermediate: that chain must not be fused
        blk.ops.add().CopyFrom(R.make_op("scale", {"X": ["pool_0"]}, {"Out": ["pool_0_scaled"]}))
    blk.ops.add().CopyFrom(R.make_op("concat", {"X": ys}, {"Out": ["front"]}, [("axis", R.INT, 1)]))
    blk.ops.add().CopyFrom(R.make_op("fc", {"Input": ["front"], "W": ["fc_w"]}, {"Out": ["logit"]}))
    return prog


def test_fuse_front_end_structure():
    prog = _front_end_program(4)
    raw = prog.SerializeToString()
    out, stats = R.rewrite_program_bytes(raw)
    assert stats == {"fused_chains": 4, "fused_concats": 0, "retargeted_ops": 0}
    new = R.ProgramDesc()
    new.ParseFromString(out)
    ops = [o.type for o in new.blocks[0].ops]
    assert ops == [R.FUSED_OP, "concat", "fc"]
    f = new.blocks[0].ops[0]
    assert R.op_input(f, "W") == ["emb"] and R.op_input(f, "Ids") == [f"ids_{i}" for i in range(4)]
    assert R.op_output(f, "Out") == [f"y_{i}" for i in range(4)]           # the concat keeps its inputs
    assert (R.op_attr(f, "pooltype"), R.op_attr(f, "use_cvm"), R.op_attr(f, "padding_idx")) == ("SUM", True, -1)
    assert list(new.blocks[0].vars) == [b"\x0a\x03emb"] and new.version == b"\x08\x00"   # carried through
    # the input program itself was not modified
    again = R.ProgramDesc()
    again.ParseFromString(raw)
    assert [o.type for o in again.blocks[0].ops][:3] == ["lookup_table_v2", "sequence_pool", "cvm"]


def test_fuse_groups_by_table_and_respects_other_consumers():
    prog = _front_end_program(3, shared_table=False)
    assert R.fuse_front_end(prog) == 3
    assert [o.type for o in prog.blocks[0].ops] == [R.FUSED_OP] * 3 + ["concat", "fc"]   # one op per table
    prog = _front_end_program(3, extra_consumer=True)
    assert R.fuse_front_end(prog) == 2                                     # slot 0's pooled rows have two readers
    types = [o.type for o in prog.blocks[0].ops]
    assert types.count(R.FUSED_OP) == 1 and types[:3] == ["lookup_table_v2", "sequence_pool", "cvm"]
    # what is left is retargeted 1:1 -- except the gather, whose Out feeds an op that reads its LoD: a
    # custom op cannot pass LoD on (custom_operator.cc:293), so it stays with the built-in kernel
    assert R.retarget_ops(prog) == 2
    assert [o.type for o in prog.blocks[0].ops][:3] == ["lookup_table_v2", "b200_sequence_pool", "b200_cvm"]


def test_retarget_only():
    prog = _front_end_program(2)
    out, stats = R.rewrite_program_bytes(prog.SerializeToString(), fuse=False)
    assert stats == {"fused_chains": 0, "fused_concats": 0, "retargeted_ops": 4}   # the two gathers keep their LoD-sharing kernel
    new = R.ProgramDesc()
    new.ParseFromString(out)
    op = new.blocks[0].ops[1]
    assert op.type == "b200_sequence_pool" and R.op_attr(op, "pooltype") == "SUM" and R.op_input(op, "X") == ["rows_0"]


def _pass_tester_program(second_pool_reads=None):
    """The graph of TEST(SeqPoolCVMConcatFusePass, basic)
    (framework/ir/seqpool_cvm_concat_fuse_pass_tester.cc:86-163): three sequence_pool(SUM)
    a,b,c -> (MaxIndex d,f,h; Out e,g,i) -> cvm(use_cvm, CVM=n) -> j,k,l -> concat(axis 1) -> m."""
    prog = R.ProgramDesc()
    blk = prog.blocks.add()
    blk.idx, blk.parent_idx = 0, -1
    for x, idx, out in (("a", "d", "e"), ("b", "f", "g"), ("c", "h", "i")):
        blk.ops.add().CopyFrom(R.make_op("sequence_pool", {"X": [x]}, {"MaxIndex": [idx], "Out": [out]},
                                         [("pooltype", R.STRING, "SUM")]))
    for x, y in (("e", "j"), ("g", "k"), ("i", "l")):
        blk.ops.add().CopyFrom(R.make_op("cvm", {"X": [x], "CVM": ["n"]}, {"Y": [y]}, [("use_cvm", R.BOOLEAN, True)]))
    blk.ops.add().CopyFrom(R.make_op("concat", {"X": ["j", "k", "l"]}, {"Out": ["m"]}, [("axis", R.INT, 1)]))
    if second_pool_reads:
        blk.ops.add().CopyFrom(R.make_op("scale", {"X": [second_pool_reads]}, {"Out": ["z"]}))
    return prog


def test_reference_fuse_pass_pattern():
    """Same expectations as the reference's pass tester: seven ops become one
    fusion op (its 16 removed + 1 added graph nodes are 6 ops + concat and their 9 intermediate
    variables), inputs in concat order, CVM carried, axis kept."""
    prog = _pass_tester_program()
    assert R.fuse_seqpool_cvm_concat(prog) == 1
    ops = prog.blocks[0].ops
    assert [o.type for o in ops] == [R.FUSION_CONCAT_OP]
    f = ops[0]
    assert R.op_input(f, "X") == ["a", "b", "c"] and R.op_input(f, "CVM") == ["n"] and R.op_output(f, "Out") == ["m"]
    assert (R.op_attr(f, "pooltype"), R.op_attr(f, "use_cvm"), R.op_attr(f, "axis")) == ("SUM", True, 1)
    # an intermediate with a second reader keeps the whole concat as it was (then retargeted 1:1)
    prog = _pass_tester_program(second_pool_reads="g")
    assert R.fuse_seqpool_cvm_concat(prog) == 0
    out, stats = R.rewrite_program_bytes(prog.SerializeToString())
    assert stats == {"fused_chains": 0, "fused_concats": 0, "retargeted_ops": 6}
    # through the byte-level entry
    out, stats = R.rewrite_program_bytes(_pass_tester_program().SerializeToString())
    assert stats == {"fused_chains": 0, "fused_concats": 1, "retargeted_ops": 0}


# ------------------------------------------------------------------ the tensor stream
def test_tensor_stream_round_trip(tmp_path):
    rng = np.random.default_rng(0)
    table = rng.standard_normal((1000, 16)).astype(np.float32)
    other = rng.integers(0, 9, (7, 3)).astype(np.int64)
    half = rng.standard_normal((5, 8)).astype(np.float16)
    path = os.path.join(tmp_path, "model.pdiparams")
    with open(path, "wb") as f:
        pdio.write_tensor(f, other)
        pdio.write_tensor(f, table, lod=[[0, 400, 1000]])
        pdio.write_tensor(f, half)
    recs = list(pdio.iter_records(path))
    assert [(r.dtype, r.dims) for r in recs] == [("int64", [7, 3]), ("float32", [1000, 16]), ("float16", [5, 8])]
    assert recs[1].lod == [[0, 400, 1000]] and recs[0].lod == []
    np.testing.assert_array_equal(pdio.read_tensor(path, recs[1]), table)
    np.testing.assert_array_equal(pdio.read_tensor(path, recs[0]), other)
    np.testing.assert_array_equal(pdio.read_tensor(path, recs[2]), half)
    # byte-level layout of a record, as SerializeToStream / TensorToStream write it
    buf = io.BytesIO()
    pdio.write_tensor(buf, np.arange(6, dtype=np.float32).reshape(2, 3))
    b = buf.getvalue()
    assert b[:4] == b"\x00\x00\x00\x00" and b[4:12] == (0).to_bytes(8, "little")      # version, 0 LoD levels
    assert b[12:16] == b"\x00\x00\x00\x00"                                              # tensor version
    assert int.from_bytes(b[16:20], "little") == 6 and b[20:26] == b"\x08\x05\x10\x02\x10\x03"  # FP32=5, dims 2,3
    assert len(b) == 26 + 24
    # packed dims (what a proto3-style writer would emit) parse too; negative dims survive
    assert pdio.parse_tensor_desc(b"\x08\x05\x12\x02\x02\x03") == (5, [2, 3])
    assert pdio.parse_tensor_desc(pdio.build_tensor_desc(5, [-1, 640])) == (5, [-1, 640])
    with pytest.raises(ValueError, match="not supported"):
        pdio.read_record_header(io.BytesIO(b"\x01\x00\x00\x00" + b"\x00" * 40))


@pytest.mark.gpu
def test_table_streams_into_device_memory_whole_and_sharded(tmp_path):
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    rng = np.random.default_rng(1)
    table = rng.standard_normal((10_007, 16)).astype(np.float32)
    path = os.path.join(tmp_path, "w.pdiparams")
    with open(path, "wb") as f:
        pdio.write_tensor(f, np.zeros((3, 3), np.float32))
        pdio.write_tensor(f, table)
    rec = list(pdio.iter_records(path))[1]
    dev = torch.device("cuda:0")
    whole = torch.empty((10_007, 16), device=dev)
    pdio.load_table_to_device(path, rec, whole, chunk_rows=1000)
    np.testing.assert_array_equal(whole.cpu().numpy(), table)
    for G in (2, 8):
        for g in range(G):
            rows = len(range(g, 10_007, G))
            part = torch.empty((rows, 16), device=dev)
            pdio.load_table_to_device(path, rec, part, row0=g, row_stride=G, chunk_rows=333)
            np.testing.assert_array_equal(part.cpu().numpy(), table[g::G])


@pytest.mark.gpu
def test_rewritten_program_computes_the_same_front_end():
    """Interpret the op list before and after the rewrite with this repo's operator mirrors."""
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import __graft_entry__ as g
    g.build()
    from paddle_infer_b200 import ops
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(3)
    S, Bsz, N, D = 4, 64, 500, 16
    table = rng.uniform(-1, 1, (N, D)).astype(np.float32)
    table[:, 0] = rng.integers(0, 9, N)
    table[:, 1] = np.minimum(table[:, 0], rng.integers(0, 9, N))
    scope, lods = {"emb": torch.from_numpy(table).to(dev)}, {}
    for i in range(S):
        l = rng.integers(0, 6, Bsz)
        lod = [0] + list(np.cumsum(l))
        scope[f"ids_{i}"] = torch.from_numpy(rng.integers(0, N, lod[-1]).astype(np.int64)).to(dev)
        lods[f"ids_{i}"] = [lod]

    def run(prog):
        sc, ld = dict(scope), dict(lods)
        for op in prog.blocks[0].ops:
            if op.type == "lookup_table_v2":
                out = R.op_output(op, "Out")[0]
                sc[out] = ops.lookup_table_v2(sc[R.op_input(op, "W")[0]], sc[R.op_input(op, "Ids")[0]],
                                              R.op_attr(op, "padding_idx"))
                ld[out] = ld[R.op_input(op, "Ids")[0]]                      # ShareLoD(Ids -> Out)
            elif op.type == "sequence_pool":
                x = R.op_input(op, "X")[0]
                sc[R.op_output(op, "Out")[0]] = ops.sequence_pool(sc[x], ld[x], R.op_attr(op, "pooltype"),
                                                                  R.op_attr(op, "pad_value"), True).Out
            elif op.type == "cvm":
                sc[R.op_output(op, "Y")[0]] = ops.cvm(sc[R.op_input(op, "X")[0]], None, R.op_attr(op, "use_cvm"))
            elif op.type == R.FUSED_OP:
                ids = R.op_input(op, "Ids")
                outs = ops.ctr_fused_emb_seqpool_cvm(sc[R.op_input(op, "W")[0]], [sc[i] for i in ids],
                                                     [ld[i][-1] for i in ids], R.op_attr(op, "pooltype"),
                                                     pad_value=R.op_attr(op, "pad_value"),
                                                     padding_idx=R.op_attr(op, "padding_idx"),
                                                     use_cvm=R.op_attr(op, "use_cvm"), cvm_offset=R.op_attr(op, "cvm_offset"))
                for name, t in zip(R.op_output(op, "Out"), outs):
                    sc[name] = t
            elif op.type == "concat":
                sc[R.op_output(op, "Out")[0]] = torch.cat([sc[n] for n in R.op_input(op, "X")], dim=R.op_attr(op, "axis"))
            elif op.type == "fc":
                break
        return sc["front"]

    prog = _front_end_program(S)
    before = run(prog)
    assert R.fuse_front_end(prog) == S
    after = run(prog)
    assert before.shape == (Bsz, S * D) and torch.equal(before, after)


def test_retarget_rules_for_lod_and_new_clones():
    """A lookup_table_v2 whose rows go to an op that does not read LoD is retargeted; the PaddleBox ops
    get their clones; pull_box_sparse only with the resident table's variables; a pool with a
    non-zero pad_value is not folded into the concat fusion; a variable written twice is not matched."""
    prog = R.ProgramDesc()
    blk = prog.blocks.add()
    blk.idx, blk.parent_idx = 0, -1
    blk.ops.add().CopyFrom(R.make_op("lookup_table_v2", {"W": ["emb"], "Ids": ["ids"]}, {"Out": ["rows"]},
                                     [("padding_idx", R.LONG, -1)]))
    blk.ops.add().CopyFrom(R.make_op("scale", {"X": ["rows"]}, {"Out": ["z"]}))
    blk.ops.add().CopyFrom(R.make_op("fused_seqpool_cvm", {"X": ["a", "b"], "CVM": ["n"]}, {"Out": ["p", "q"]},
                                     [("pooltype", R.STRING, "SUM"), ("pad_value", R.FLOAT, 0.0), ("use_cvm", R.BOOLEAN, True),
                                      ("cvm_offset", R.INT, 2)]))
    blk.ops.add().CopyFrom(R.make_op("fused_embedding_seq_pool", {"W": ["emb"], "Ids": ["ids2"]}, {"Out": ["o"]},
                                     [("combiner", R.STRING, "sum")]))
    blk.ops.add().CopyFrom(R.make_op("pull_box_sparse", {"Ids": ["k0", "k1"]}, {"Out": ["v0", "v1"]}, [("size", R.INT, 11)]))
    assert R.retarget_ops(prog) == 3
    assert [o.type for o in blk.ops] == ["b200_lookup_table_v2", "scale", "b200_fused_seqpool_cvm",
                                         "b200_fused_embedding_seq_pool", "pull_box_sparse"]
    assert R.retarget_ops(prog, pull_box_table={"W": "vals", "MapKeys": "mk", "MapRows": "mr"}) == 1
    p = blk.ops[4]
    assert p.type == "b200_pull_box_sparse" and R.op_input(p, "W") == ["vals"] and R.op_input(p, "MapKeys") == ["mk"]
    assert R.op_input(p, "Ids") == ["k0", "k1"] and R.op_output(p, "Out") == ["v0", "v1"] and R.op_attr(p, "size") == 11
    # pad_value != 0 blocks the concat fusion
    prog = _pass_tester_program()
    R.set_attr(prog.blocks[0].ops[1], "pad_value", R.FLOAT, 0.5)
    assert R.fuse_seqpool_cvm_concat(prog) == 0
    # a variable with two writers is not matched
    prog = _pass_tester_program()
    prog.blocks[0].ops.add().CopyFrom(R.make_op("scale", {"X": ["a"]}, {"Out": ["e"]}))
    assert R.fuse_seqpool_cvm_concat(prog) == 0
    # the fused front end carries check_ids
    prog = _front_end_program(2)
    assert R.fuse_front_end(prog, check_ids=False) == 2
    assert R.op_attr(prog.blocks[0].ops[0], "check_ids") is False
