"""Program rewriter: make a saved Paddle inference program run the B200 front end.

Paddle's custom-operator loader refuses to register an op under a built-in name
(paddle/fluid/framework/custom_operator.cc:722-725), so the replacement ops carry new names
(paddle_shim/ctrfe_paddle_ops.cc) and the PROGRAM is retargeted instead -- in memory, the
model file on disk is not touched.  Two rewrites over a `ProgramDesc`
(paddle/fluid/framework/framework.proto):

  retarget_ops   lookup_table_v2 -> b200_lookup_table_v2, sequence_pool -> b200_sequence_pool,
                 cvm -> b200_cvm: same inputs / outputs / attrs, 1:1.
  fuse_seqpool_cvm_concat  the reference's own seqpool_cvm_concat_fuse_pass pattern (no gather)
  fuse_front_end every chain  lookup_table_v2(W, Ids_i) -> sequence_pool(pooltype P) -> cvm
                 whose slots share W, P, pad_value and use_cvm becomes ONE
                 ctr_fused_emb_seqpool_cvm(W, Ids@VECTOR) -> Out@VECTOR op writing the cvm
                 outputs (so a following concat(axis=1) keeps working).  This is the pattern
                 seqpool_cvm_concat_fuse_pass matches on the CPU
                 (framework/ir/seqpool_cvm_concat_fuse_pass.cc:102-198), extended upstream by
                 the lookup; like that pass it only fires when each intermediate has a single
                 consumer.

The ProgramDesc messages are described here by hand (field numbers from framework.proto:46-78,
218-249) so that no Paddle install and no protoc are needed; variables (`BlockDesc.vars`) and
every field this module does not name are carried through as opaque bytes / unknown fields.
"""
from __future__ import annotations

from typing import Optional, Dict, List, Tuple

from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

# AttrType values (framework.proto:25-41)
INT, FLOAT, STRING, BOOLEAN, LONG = 0, 1, 2, 6, 9

_F = descriptor_pb2.FieldDescriptorProto


def _build_messages():
    fd = descriptor_pb2.FileDescriptorProto()
    fd.name = "ctrfe_framework_subset.proto"
    fd.package = "ctrfe.pd"
    fd.syntax = "proto2"

    def field(msg, name, num, ftype, label=_F.LABEL_OPTIONAL, type_name=None):
        f = msg.field.add()
        f.name, f.number, f.type, f.label = name, num, ftype, label
        if type_name:
            f.type_name = type_name
        return f

    op = fd.message_type.add()
    op.name = "OpDesc"
    attr = op.nested_type.add()
    attr.name = "Attr"
    field(attr, "name", 1, _F.TYPE_STRING, _F.LABEL_REQUIRED)
    field(attr, "type", 2, _F.TYPE_INT32, _F.LABEL_REQUIRED)  # AttrType enum, kept numeric
    field(attr, "i", 3, _F.TYPE_INT32)
    field(attr, "f", 4, _F.TYPE_FLOAT)
    field(attr, "s", 5, _F.TYPE_STRING)
    field(attr, "ints", 6, _F.TYPE_INT32, _F.LABEL_REPEATED)
    field(attr, "floats", 7, _F.TYPE_FLOAT, _F.LABEL_REPEATED)
    field(attr, "strings", 8, _F.TYPE_STRING, _F.LABEL_REPEATED)
    field(attr, "b", 10, _F.TYPE_BOOL)
    field(attr, "bools", 11, _F.TYPE_BOOL, _F.LABEL_REPEATED)
    field(attr, "block_idx", 12, _F.TYPE_INT32)
    field(attr, "l", 13, _F.TYPE_INT64)
    field(attr, "blocks_idx", 14, _F.TYPE_INT32, _F.LABEL_REPEATED)
    field(attr, "longs", 15, _F.TYPE_INT64, _F.LABEL_REPEATED)
    field(attr, "float64s", 16, _F.TYPE_DOUBLE, _F.LABEL_REPEATED)
    field(attr, "var_name", 17, _F.TYPE_STRING)
    field(attr, "vars_name", 18, _F.TYPE_STRING, _F.LABEL_REPEATED)
    field(attr, "float64", 19, _F.TYPE_DOUBLE)
    var = op.nested_type.add()
    var.name = "Var"
    field(var, "parameter", 1, _F.TYPE_STRING, _F.LABEL_REQUIRED)
    field(var, "arguments", 2, _F.TYPE_STRING, _F.LABEL_REPEATED)
    field(op, "inputs", 1, _F.TYPE_MESSAGE, _F.LABEL_REPEATED, ".ctrfe.pd.OpDesc.Var")
    field(op, "outputs", 2, _F.TYPE_MESSAGE, _F.LABEL_REPEATED, ".ctrfe.pd.OpDesc.Var")
    field(op, "type", 3, _F.TYPE_STRING, _F.LABEL_REQUIRED)
    field(op, "attrs", 4, _F.TYPE_MESSAGE, _F.LABEL_REPEATED, ".ctrfe.pd.OpDesc.Attr")
    field(op, "is_target", 5, _F.TYPE_BOOL)

    blk = fd.message_type.add()
    blk.name = "BlockDesc"
    field(blk, "idx", 1, _F.TYPE_INT32, _F.LABEL_REQUIRED)
    field(blk, "parent_idx", 2, _F.TYPE_INT32, _F.LABEL_REQUIRED)
    field(blk, "vars", 3, _F.TYPE_BYTES, _F.LABEL_REPEATED)  # VarDesc messages, carried opaquely
    field(blk, "ops", 4, _F.TYPE_MESSAGE, _F.LABEL_REPEATED, ".ctrfe.pd.OpDesc")
    field(blk, "forward_block_idx", 5, _F.TYPE_INT32)

    prog = fd.message_type.add()
    prog.name = "ProgramDesc"
    field(prog, "blocks", 1, _F.TYPE_MESSAGE, _F.LABEL_REPEATED, ".ctrfe.pd.BlockDesc")
    field(prog, "version", 4, _F.TYPE_BYTES)         # Version message, opaque
    field(prog, "op_version_map", 5, _F.TYPE_BYTES)  # OpVersionMap message, opaque

    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    get = getattr(message_factory, "GetMessageClass", None)
    if get is None:  # older protobuf
        factory = message_factory.MessageFactory(pool)
        get = factory.GetPrototype
    return (get(pool.FindMessageTypeByName("ctrfe.pd.ProgramDesc")),
            get(pool.FindMessageTypeByName("ctrfe.pd.BlockDesc")),
            get(pool.FindMessageTypeByName("ctrfe.pd.OpDesc")))


ProgramDesc, BlockDesc, OpDesc = _build_messages()

RETARGET = {"lookup_table_v2": "b200_lookup_table_v2", "sequence_pool": "b200_sequence_pool", "cvm": "b200_cvm",
            "fused_seqpool_cvm": "b200_fused_seqpool_cvm", "fused_embedding_seq_pool": "b200_fused_embedding_seq_pool"}
# Ops that read the LoD of their input.  The custom-op runner never copies an output's LoD back
# (framework/custom_operator.cc:293), so a retargeted lookup_table_v2 / cvm cannot honour the
# reference's ShareLoD (lookup_table_v2_op.cc:65, cvm_op.cc:46): such an op is left to the built-in
# kernel when one of these reads its output (the fused rewrites need no LoD to travel).
LOD_READERS = {"sequence_pool", "b200_sequence_pool", "fused_seqpool_cvm", "b200_fused_seqpool_cvm",
               "fusion_seqpool_cvm_concat", "b200_fusion_seqpool_cvm_concat", "fused_embedding_seq_pool",
               "b200_fused_embedding_seq_pool", "cvm_grad", "sequence_pool_grad"}
SHARES_LOD = {"lookup_table_v2": "Out", "cvm": "Y"}
FUSED_OP = "ctr_fused_emb_seqpool_cvm"


# ---- small accessors over OpDesc ------------------------------------------------------------------
def op_input(op, name: str) -> List[str]:
    for v in op.inputs:
        if v.parameter == name:
            return list(v.arguments)
    return []


def op_output(op, name: str) -> List[str]:
    for v in op.outputs:
        if v.parameter == name:
            return list(v.arguments)
    return []


def op_attr(op, name: str, default=None):
    for a in op.attrs:
        if a.name == name:
            return {INT: a.i, FLOAT: a.f, STRING: a.s, BOOLEAN: a.b, LONG: a.l}.get(a.type, default)
    return default


def set_attr(op, name: str, atype: int, value) -> None:
    a = op.attrs.add()
    a.name, a.type = name, atype
    if atype == INT:
        a.i = int(value)
    elif atype == FLOAT:
        a.f = float(value)
    elif atype == STRING:
        a.s = str(value)
    elif atype == BOOLEAN:
        a.b = bool(value)
    elif atype == LONG:
        a.l = int(value)
    else:
        raise ValueError("unsupported attr type")


def make_op(op_type: str, inputs: Dict[str, List[str]], outputs: Dict[str, List[str]], attrs=()):
    op = OpDesc()
    op.type = op_type
    for k, args in inputs.items():
        v = op.inputs.add()
        v.parameter = k
        v.arguments.extend(args)
    for k, args in outputs.items():
        v = op.outputs.add()
        v.parameter = k
        v.arguments.extend(args)
    for name, atype, value in attrs:
        set_attr(op, name, atype, value)
    return op


# ---- the rewrites ------------------------------------------------------------------------------------
def _reads_lod(op_type: str) -> bool:
    return op_type in LOD_READERS or op_type.startswith("sequence_")


def retarget_ops(program, pull_box_table: Optional[Dict[str, str]] = None) -> int:
    """1:1 renaming of the ops to their b200_* clones.  A lookup_table_v2 / cvm whose output feeds
    an op that reads LoD stays with the built-in kernel (see LOD_READERS).  pull_box_table =
    {"W": ..., "MapKeys": ..., "MapRows": ...} (variable names of the resident value table and its
    key map) also retargets pull_box_sparse, which needs those inputs.  Returns the number of ops
    changed."""
    n = 0
    for blk in program.blocks:
        use = _consumers(blk)
        for op in blk.ops:
            if op.type == "pull_box_sparse" and pull_box_table:
                ids, outs = op_input(op, "Ids"), op_output(op, "Out")
                new = make_op("b200_pull_box_sparse",
                              {"W": [pull_box_table["W"]], "MapKeys": [pull_box_table["MapKeys"]],
                               "MapRows": [pull_box_table["MapRows"]], "Ids": ids}, {"Out": outs},
                              [("size", INT, op_attr(op, "size", 1)), ("is_sparse", BOOLEAN, op_attr(op, "is_sparse", False)),
                               ("is_distributed", BOOLEAN, op_attr(op, "is_distributed", False))])
                op.CopyFrom(new)
                n += 1
                continue
            if op.type not in RETARGET:
                continue
            if op.type in SHARES_LOD:
                outs = op_output(op, SHARES_LOD[op.type])
                if any(_reads_lod(blk.ops[u].type) for o in outs for u in use.get(o, [])):
                    continue  # its output's LoD is needed downstream and a custom op cannot set it
            op.type = RETARGET[op.type]
            n += 1
    return n


def _consumers(blk) -> Dict[str, List[int]]:
    use: Dict[str, List[int]] = {}
    for i, op in enumerate(blk.ops):
        for v in op.inputs:
            for a in v.arguments:
                use.setdefault(a, []).append(i)
    return use


def _single_producers(blk) -> Dict[str, int]:
    """variable -> the op that writes it, for variables written exactly once in the block (a name
    written twice cannot be matched to one producer, so the rewrites leave it alone)."""
    count: Dict[str, int] = {}
    last: Dict[str, int] = {}
    for i, op in enumerate(blk.ops):
        for v in op.outputs:
            for a in v.arguments:
                count[a] = count.get(a, 0) + 1
                last[a] = i
    return {a: i for a, i in last.items() if count[a] == 1}


def fuse_front_end(program, check_ids: bool = True) -> int:
    """Replaces lookup_table_v2 -> sequence_pool -> cvm chains by ctr_fused_emb_seqpool_cvm ops,
    one per group of slots that share (W, pooltype, pad_value, padding_idx, use_cvm).  check_ids:
    the fused op synchronises and raises on an id outside the table, as the reference's CPU
    lookup_table_v2 does (False: a serving loop that checks elsewhere).  Returns the number of
    chains fused."""
    fused = 0
    for blk in program.blocks:
        use = _consumers(blk)
        producer = _single_producers(blk)
        groups: Dict[Tuple, List[Tuple[int, int, int, str, str]]] = {}
        for ci, cvm in enumerate(blk.ops):
            if cvm.type != "cvm":
                continue
            x = op_input(cvm, "X")
            if len(x) != 1 or x[0] not in producer:
                continue
            pi = producer[x[0]]
            pool = blk.ops[pi]
            if pool.type != "sequence_pool" or use.get(x[0], []) != [ci]:
                continue
            px = op_input(pool, "X")
            if len(px) != 1 or px[0] not in producer:
                continue
            li = producer[px[0]]
            look = blk.ops[li]
            if look.type != "lookup_table_v2" or use.get(px[0], []) != [pi]:
                continue
            pooltype = op_attr(pool, "pooltype", "AVERAGE")
            if pooltype not in ("SUM", "AVERAGE", "SQRT", "MAX", "LAST", "FIRST"):
                continue
            if pooltype == "MAX" and not op_attr(pool, "is_test", False) and use.get((op_output(pool, "MaxIndex") or [""])[0]):
                continue  # somebody reads MaxIndex: leave the chain alone
            w, ids = op_input(look, "W"), op_input(look, "Ids")
            if len(w) != 1 or len(ids) != 1:
                continue
            key = (w[0], pooltype, float(op_attr(pool, "pad_value", 0.0)), int(op_attr(look, "padding_idx", -1)),
                   bool(op_attr(cvm, "use_cvm", True)))
            groups.setdefault(key, []).append((li, pi, ci, ids[0], op_output(cvm, "Y")[0]))
        if not groups:
            continue
        dead = set()
        inserts: List[Tuple[int, object]] = []
        for (w, pooltype, pad_value, padding_idx, use_cvm), chains in groups.items():
            chains.sort(key=lambda c: c[2])
            # the fused op replaces the group's LAST cvm; if some consumer of an earlier slot's
            # output runs before that point the group is left alone
            pos = chains[-1][2]
            if any(u < pos for c in chains for u in use.get(c[4], [])):
                continue
            op = make_op(FUSED_OP, {"W": [w], "Ids": [c[3] for c in chains]}, {"Out": [c[4] for c in chains]},
                         [("pooltype", STRING, pooltype), ("pad_value", FLOAT, pad_value),
                          ("padding_idx", LONG, padding_idx), ("use_cvm", BOOLEAN, use_cvm), ("cvm_offset", INT, 2),
                          ("check_ids", BOOLEAN, check_ids)])
            # the fused op goes where the LAST of its cvm ops stood: every input exists by then
            inserts.append((chains[-1][2], op))
            for li, pi, ci, _, _ in chains:
                dead.update((li, pi, ci))
            fused += len(chains)
        new_ops = []
        at = {pos: op for pos, op in inserts}
        for i, op in enumerate(blk.ops):
            if i in at:
                new_ops.append(at[i])
            if i not in dead:
                cp = OpDesc()
                cp.CopyFrom(op)
                new_ops.append(cp)
        del blk.ops[:]
        for op in new_ops:
            blk.ops.add().CopyFrom(op)
    return fused


FUSION_CONCAT_OP = "b200_fusion_seqpool_cvm_concat"


def fuse_seqpool_cvm_concat(program) -> int:
    """The reference's seqpool_cvm_concat_fuse_pass on a ProgramDesc
    (framework/ir/seqpool_cvm_concat_fuse_pass.cc:102-198): a concat(axis=1) whose every input is
    the Y of a cvm(use_cvm=true) fed by the Out of a sequence_pool -- all with one pooltype out of
    SUM / AVERAGE / SQRT (the pass itself matches SUM only; the fused op takes all three,
    fusion_seqpool_cvm_concat_op.cc:132-157) and one CVM input -- becomes one
    b200_fusion_seqpool_cvm_concat(X = the pools' inputs, CVM) -> the concat's Out, when nobody
    else reads the intermediate variables.  Returns the number of concats fused."""
    fused = 0
    for blk in program.blocks:
        use = _consumers(blk)
        producer = _single_producers(blk)
        dead = set()
        inserts = {}
        for ki, cat in enumerate(blk.ops):
            if cat.type != "concat" or int(op_attr(cat, "axis", 0)) != 1:
                continue
            xs, pools, cvms, cvm_in, pooltype, ok = [], [], [], None, None, True
            for y in op_input(cat, "X"):
                ci = producer.get(y)
                if ci is None or blk.ops[ci].type != "cvm" or not op_attr(blk.ops[ci], "use_cvm", True) \
                        or use.get(y, []) != [ki]:
                    ok = False
                    break
                cvm = blk.ops[ci]
                x = op_input(cvm, "X")
                pi = producer.get(x[0]) if len(x) == 1 else None
                if pi is None or blk.ops[pi].type != "sequence_pool" or use.get(x[0], []) != [ci]:
                    ok = False
                    break
                pool = blk.ops[pi]
                pt = op_attr(pool, "pooltype", "AVERAGE")
                idx = op_output(pool, "MaxIndex")
                c_in = op_input(cvm, "CVM")
                if float(op_attr(pool, "pad_value", 0.0)) != 0.0:
                    ok = False  # the fused op has no pad_value: an empty sequence would give cvm(0) instead of cvm(pad_value)
                    break
                if pt not in ("SUM", "AVERAGE", "SQRT") or (pooltype not in (None, pt)) or (idx and use.get(idx[0])) \
                        or len(c_in) != 1 or (cvm_in not in (None, c_in[0])) or len(op_input(pool, "X")) != 1:
                    ok = False
                    break
                pooltype, cvm_in = pt, c_in[0]
                xs.append(op_input(pool, "X")[0])
                pools.append(pi)
                cvms.append(ci)
            if not ok or not xs:
                continue
            op = make_op(FUSION_CONCAT_OP, {"X": xs, "CVM": [cvm_in]}, {"Out": op_output(cat, "Out")},
                         [("pooltype", STRING, pooltype), ("use_cvm", BOOLEAN, True), ("axis", INT, 1)])
            inserts[ki] = op
            dead.update(pools + cvms + [ki])
            fused += 1
        if not inserts:
            continue
        new_ops = []
        for i, op in enumerate(blk.ops):
            if i in inserts:
                new_ops.append(inserts[i])
            if i not in dead:
                cp = OpDesc()
                cp.CopyFrom(op)
                new_ops.append(cp)
        del blk.ops[:]
        for op in new_ops:
            blk.ops.add().CopyFrom(op)
    return fused


def rewrite_program_bytes(pdmodel: bytes, fuse: bool = True, check_ids: bool = True,
                          pull_box_table: Optional[Dict[str, str]] = None) -> Tuple[bytes, Dict[str, int]]:
    """Parses a serialized ProgramDesc (`*.pdmodel`), applies the rewrites, serializes it again."""
    prog = ProgramDesc()
    prog.ParseFromString(pdmodel)
    stats = {"fused_chains": fuse_front_end(prog, check_ids) if fuse else 0,
             "fused_concats": fuse_seqpool_cvm_concat(prog) if fuse else 0}
    stats["retargeted_ops"] = retarget_ops(prog, pull_box_table)
    return prog.SerializeToString(), stats
