#!/usr/bin/env python
"""Generate tests/golden/*.npz from the REFERENCE's own test formulas.

Run once in the authoring container (where /root/reference exists); the outputs are
committed so that the GPU box (no /root/reference) can replay them.

    python tests/golden/make_golden.py

What it does: the reference's op tests hold no stored vectors -- each computes its
expectation with a small numpy function.  Those test modules import paddle (absent
here), so instead of importing them this script parses them with `ast`, lifts out
the pure-numpy pieces *verbatim*, and executes them:

  python/paddle/fluid/tests/unittests/sequence/test_sequence_pool.py
      compute_seqpool_sum / _avg / _sqrt           :26-54
      every TestSeq*Pool* class: set_lod() and compute() resolved through its bases
  python/paddle/fluid/tests/unittests/test_reorder_lod_tensor.py
      convert_to_offset                            :26-31
  python/paddle/fluid/tests/unittests/test_cvm_op.py
      cvm_compute                                  :22-36
  python/paddle/fluid/tests/unittests/test_fusion_seqpool_cvm_concat_op.py
      the lods/w fixtures                          :30,79-102 (composition :41-63)
  python/paddle/fluid/tests/unittests/test_lookup_table_v2_op.py
      table[ids] with the shapes of :54-57, :91-94 and padding of :110-129

Inputs are drawn here with a fixed seed using the value ranges those tests use
(uniform(0.1, 1); MAX adds +2.0 / +1.0 to one row per segment).
"""
from __future__ import annotations

import ast
import json
import os
import types
from math import log

import numpy as np

REF = os.environ.get("CTRFE_REFERENCE_ROOT", "/root/reference")
UT = os.path.join(REF, "python/paddle/fluid/tests/unittests")
HERE = os.path.dirname(os.path.abspath(__file__))
SEED = 20261106


def _module_ast(path):
    with open(path) as f:
        return ast.parse(f.read(), filename=path)


def _lift_functions(path, names, env):
    """exec the named top-level FunctionDefs of `path` into env, unchanged."""
    tree = _module_ast(path)
    found = {}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in names:
            mod = ast.Module(body=[node], type_ignores=[])
            exec(compile(mod, path, "exec"), env)
            found[node.name] = env[node.name]
    missing = set(names) - set(found)
    assert not missing, f"{path}: missing {missing}"
    return found


def _class_table(path):
    """class name -> (base names, {method name: FunctionDef})"""
    tree = _module_ast(path)
    table = {}
    for node in tree.body:
        if isinstance(node, ast.ClassDef):
            bases = [b.id for b in node.bases if isinstance(b, ast.Name)]
            methods = {n.name: n for n in node.body if isinstance(n, ast.FunctionDef)}
            table[node.name] = (bases, methods)
    return table


def _resolve(table, cls, method):
    while cls in table:
        bases, methods = table[cls]
        if method in methods:
            return methods[method], cls
        cls = bases[0] if bases else None
    return None, None


def _bind(fn_node, path, env):
    mod = ast.Module(body=[fn_node], type_ignores=[])
    ns = dict(env)
    exec(compile(mod, path, "exec"), ns)
    return ns[fn_node.name]


def _family(table, cls):
    """Which data-shape family a test class belongs to (by walking its bases)."""
    chain = []
    c = cls
    while c in table:
        chain.append(c)
        b = table[c][0]
        c = b[0] if b else None
    if "TestSeqMaxPool2D" in chain:
        return (3, 11), 1.0
    if "TestSeqAvgPool2D" in chain:
        return (3, 17), None
    if "TestSeqMaxPool" in chain:
        return (23,), 2.0
    return (23,), None


def gen_sequence_pool(rng):
    path = os.path.join(UT, "sequence/test_sequence_pool.py")
    env = {"np": np}
    _lift_functions(os.path.join(UT, "test_reorder_lod_tensor.py"), ["convert_to_offset"], env)
    _lift_functions(path, ["compute_seqpool_sum", "compute_seqpool_avg", "compute_seqpool_sqrt"], env)
    convert_to_offset = env["convert_to_offset"]
    table = _class_table(path)
    arrays, meta = {}, {}
    for cls in table:
        if not cls.startswith("TestSeq"):
            continue
        lod_node, _ = _resolve(table, cls, "set_lod")
        comp_node, comp_owner = _resolve(table, cls, "compute")
        if lod_node is None or comp_node is None:
            continue
        lod = _bind(lod_node, path, env)(None)  # lengths per level, verbatim from the test
        offset = convert_to_offset(lod)
        level = len(lod) - 1
        feat, bump = _family(table, cls)
        k = int(sum(lod[level]))
        x = rng.uniform(0.1, 1, (k,) + feat).astype("float32")
        if bump is not None:  # TestSeqMaxPool.set_data :166-169 / TestSeqMaxPool2D.set_data :386-390
            for i in range(len(offset[level]) - 1):
                l = offset[level][i + 1] - offset[level][i]
                if l > 0:
                    x[offset[level][i] + rng.integers(l)] += bump
        out = np.zeros((len(lod[level]),) + feat).astype("float32")
        holder = types.SimpleNamespace(attrs={})
        _bind(comp_node, path, env)(holder, x, offset, out)  # the reference's own numpy formula
        arrays[f"{cls}/x"] = x
        arrays[f"{cls}/out"] = out
        meta[cls] = {"lod_lengths": lod, "offsets": [[int(v) for v in lv] for lv in offset],
                     "attrs": holder.attrs, "compute_from": comp_owner, "feat": list(feat)}
    return arrays, meta


def gen_cvm(rng):
    env = {"np": np, "log": log}
    f = _lift_functions(os.path.join(UT, "test_cvm_op.py"), ["cvm_compute"], env)["cvm_compute"]
    arrays, meta = {}, {}
    # test_cvm_op.py:63-72 (LoD, use_cvm False), :104-114 (use_cvm True), :141-152 (use_cvm False)
    for name, batch, use_cvm in (("lod_nocvm", 8, False), ("dense_cvm", 2, True), ("dense_nocvm", 2, False),
                                 ("dense_cvm_b64", 64, True)):
        x = rng.random((batch, 11)).astype("float32")
        arrays[f"{name}/x"] = x
        arrays[f"{name}/y"] = f(x, 11, use_cvm)
        meta[name] = {"use_cvm": use_cvm, "item_width": 11}
    # counts-like inputs (show >= click, integers), the shape real CTR data has
    x = rng.random((32, 11)).astype("float32")
    show = rng.integers(0, 64, 32)
    x[:, 0] = show
    x[:, 1] = rng.integers(0, show + 1)
    arrays["counts_cvm/x"] = x
    arrays["counts_cvm/y"] = f(x, 11, True)
    meta["counts_cvm"] = {"use_cvm": True, "item_width": 11}
    return arrays, meta


def gen_fusion_seqpool_cvm_concat(rng):
    env = {"np": np, "log": log}
    _lift_functions(os.path.join(UT, "test_reorder_lod_tensor.py"), ["convert_to_offset"], env)
    pools = _lift_functions(os.path.join(UT, "sequence/test_sequence_pool.py"),
                            ["compute_seqpool_sum", "compute_seqpool_avg", "compute_seqpool_sqrt"], env)
    cvm_compute = _lift_functions(os.path.join(UT, "test_cvm_op.py"), ["cvm_compute"], env)["cvm_compute"]
    convert_to_offset = env["convert_to_offset"]
    # fixtures: test_fusion_seqpool_cvm_concat_op.py:30 (default), :82, :88, :94-95, :101-102
    cases = {
        "default": ([[[2, 3, 5]], [[1, 5, 2]]], 11),
        "case1": ([[[1]]], 11),
        "case2": ([[[1]], [[1]], [[1]]], 11),
        "case3": ([[[1, 3, 4, 6]]], 10),
        "case4": ([[[2, 13, 4]], [[1, 1, 1]], [[5, 3, 1]], [[9, 10, 3]]], 3),
    }
    fn = {"SUM": pools["compute_seqpool_sum"], "AVERAGE": pools["compute_seqpool_avg"],
          "SQRT": pools["compute_seqpool_sqrt"]}
    arrays, meta = {}, {}
    for cname, (lods, w) in cases.items():
        for pooltype in ("SUM", "AVERAGE", "SQRT"):
            bs = len(lods[0][0])
            outs = []
            key = f"{cname}_{pooltype}"
            for i, lod in enumerate(lods):
                x = rng.uniform(0.1, 1, [sum(lod[0]), w]).astype("float32")
                offset = convert_to_offset(lod)
                out = np.zeros((bs, w)).astype("float32")
                fn[pooltype](x, offset, out)          # :47-55
                out = cvm_compute(out, w, True)       # :49,52,55
                arrays[f"{key}/x{i}"] = x
                outs.append(out)
            arrays[f"{key}/out"] = np.concatenate(outs, axis=1)  # :63
            meta[key] = {"lods_lengths": lods, "w": w, "pooltype": pooltype, "n_slots": len(lods)}
    return arrays, meta


def gen_lookup_table_v2(rng):
    arrays, meta = {}, {}
    table = rng.random((17, 31)).astype("float64")           # :54
    ids = rng.integers(0, 17, 4).astype("int64")              # :55
    arrays["basic/table"], arrays["basic/ids"], arrays["basic/out"] = table, ids, table[ids]
    meta["basic"] = {"padding_idx": -1}
    ids3 = rng.integers(0, 17, (2, 4, 5)).astype("int32")     # :92
    arrays["tensor_ids/table"], arrays["tensor_ids/ids"] = table, ids3
    arrays["tensor_ids/out"] = table[ids3.flatten()].reshape((2, 4, 5, 31))  # :94
    meta["tensor_ids"] = {"padding_idx": -1}
    pad = int(rng.choice(ids, 1)[0])                          # :111-114
    out = table[ids].copy()
    out[ids == pad] = np.zeros(31)
    arrays["padding/table"], arrays["padding/ids"], arrays["padding/out"] = table, ids, out
    meta["padding"] = {"padding_idx": pad}
    pad3 = int(rng.choice(ids3.flatten(), 1)[0])              # :124-128
    out3 = table[ids3.flatten()].reshape((2, 4, 5, 31)).copy()
    out3[ids3 == pad3] = np.zeros(31)
    arrays["tensor_ids_padding/table"], arrays["tensor_ids_padding/ids"] = table, ids3
    arrays["tensor_ids_padding/out"] = out3
    meta["tensor_ids_padding"] = {"padding_idx": pad3}
    return arrays, meta


def main():
    rng = np.random.default_rng(SEED)
    for name, gen in (("sequence_pool", gen_sequence_pool), ("cvm", gen_cvm),
                      ("fusion_seqpool_cvm_concat", gen_fusion_seqpool_cvm_concat),
                      ("lookup_table_v2", gen_lookup_table_v2)):
        arrays, meta = gen(rng)
        np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **arrays)
        with open(os.path.join(HERE, f"{name}.json"), "w") as f:
            json.dump(meta, f, indent=1, sort_keys=True)
        print(f"{name}: {len(meta)} cases, {len(arrays)} arrays")


if __name__ == "__main__":
    main()
