"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol
include/ctrfe.h declares, mirrors the header's enums, and the host layer refuses to run
without a GPU (no CPU fallback).  No compute calls here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    import __graft_entry__ as g
    g.build()
    from paddle_infer_b200 import _lib
    return _lib


def _header():
    with open(os.path.join(ROOT, "include", "ctrfe.h")) as f:
        return f.read()


def test_library_exports_every_declared_symbol(built):
    declared = set(re.findall(r"^\s*(?:int|const char\*)\s+(ctrfe_\w+)\s*\(", _header(), re.M))
    assert declared, "no declarations parsed from include/ctrfe.h"
    assert declared == set(built.EXPORTS)
    L = built.lib()
    for name in declared:
        assert getattr(L, name) is not None
    assert L.ctrfe_version() == int(re.search(r"#define CTRFE_VERSION (\d+)", _header()).group(1))


def test_python_enums_match_header(built):
    h = _header()
    consts = {}
    for body in re.findall(r"enum\s*\{(.*?)\};", h, re.S):
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        for name, val in re.findall(r"(CTRFE_\w+)\s*=\s*(\d+)", body):
            consts[name] = int(val)
    B = built
    pairs = {"CTRFE_OK": B.OK, "CTRFE_ERR_BAD_ID": B.ERR_BAD_ID, "CTRFE_ERR_ARG": B.ERR_ARG,
             "CTRFE_ERR_CUDA": B.ERR_CUDA, "CTRFE_ERR_UNSUPPORTED": B.ERR_UNSUPPORTED,
             "CTRFE_F32": B.F32, "CTRFE_F16": B.F16, "CTRFE_BF16": B.BF16, "CTRFE_F64": B.F64,
             "CTRFE_I64": B.I64, "CTRFE_I32": B.I32,
             "CTRFE_POOL_AVERAGE": B.POOL_AVERAGE, "CTRFE_POOL_SUM": B.POOL_SUM, "CTRFE_POOL_SQRT": B.POOL_SQRT,
             "CTRFE_POOL_LAST": B.POOL_LAST, "CTRFE_POOL_FIRST": B.POOL_FIRST, "CTRFE_POOL_MAX": B.POOL_MAX,
             "CTRFE_CVM_NONE": B.CVM_NONE, "CTRFE_CVM_USE": B.CVM_USE, "CTRFE_CVM_STRIP": B.CVM_STRIP,
             "CTRFE_SCALE_DIVIDE": B.SCALE_DIVIDE, "CTRFE_SCALE_RECIPROCAL": B.SCALE_RECIPROCAL,
             "CTRFE_INIT_ZERO": B.INIT_ZERO, "CTRFE_INIT_PAD": B.INIT_PAD, "CTRFE_INIT_FIRST_ROW": B.INIT_FIRST_ROW,
             "CTRFE_LOD_HOST": B.LOD_HOST, "CTRFE_LOD_DEVICE": B.LOD_DEVICE, "CTRFE_I16": B.I16,
             "CTRFE_PIPE_DIRECT_OUT": B.PIPE_DIRECT_OUT, "CTRFE_PIPE_DIRECT_IN": B.PIPE_DIRECT_IN, "CTRFE_PIPE_RAMP": B.PIPE_RAMP,
             "CTRFE_LAUNCH_UPLOAD_LOD": B.LAUNCH_UPLOAD_LOD, "CTRFE_COPY_H2D": B.COPY_H2D, "CTRFE_COPY_D2H": B.COPY_D2H,
             "CTRFE_COPY_D2D": B.COPY_D2D}
    for k, v in pairs.items():
        assert consts[k] == v, k
    # the oracle shares the numeric values (it is driven with the same descriptors in tests)
    from oracle import oracle as O
    assert (O.F32, O.F16, O.BF16) == (B.F32, B.F16, B.BF16)
    assert O.POOLTYPES == B.POOLTYPES
    assert (O.CVM_NONE, O.CVM_USE, O.CVM_STRIP) == (B.CVM_NONE, B.CVM_USE, B.CVM_STRIP)


def test_struct_layouts_match_header(built):
    # ctrfe_slot: 7 x 8-byte fields + 2 x int32 + n_ids; ctrfe_fused_desc: see header
    assert C.sizeof(built.Slot) == 72 and built.Slot.n_ids.offset == 64
    assert C.sizeof(built.FusedDesc) == 88 and built.FusedDesc.hot_cache.offset == 72
    assert built.FusedDesc.keymap.offset == 80
    assert C.sizeof(built.KeyMap) == 32 and built.KeyMap.n_keys.offset == 24
    assert C.sizeof(built.HotCache) == 24 and built.HotCache.log2_slots.offset == 16
    assert built.FusedDesc.slots.offset == 8 and built.FusedDesc.dim.offset == 16
    assert built.FusedDesc.pad_value.offset == 32 and built.FusedDesc.lod_location.offset == 52
    assert built.FusedDesc.n_shards.offset == 56 and built.FusedDesc.shard_tables.offset == 64


def test_header_cites_reference_for_every_entry_point():
    h = _header()
    # split into comment+declaration chunks and require a reference file citation near each compute entry
    for name in ("ctrfe_fused_forward", "ctrfe_embedding", "ctrfe_sequence_pool", "ctrfe_cvm", "ctrfe_check_ids"):
        idx = h.index(f"int {name}(")
        comment = h[h.rfind("/*", 0, idx):idx]
        assert re.search(r"\.(cc|cu|h):\d+", comment), f"{name}: no reference file:line in its comment"


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "paddle_infer_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cc", ".h")):
                with open(os.path.join(dirpath, f)) as fh:
                    src = fh.read()
                assert "oracle" not in src.replace("no oracle", ""), f"{f} mentions the oracle"


def test_no_cpu_fallback_without_gpu(built):
    import torch
    from paddle_infer_b200 import ops
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError, match="no CPU"):
        ops.lookup_table_v2(torch.zeros(4, 4), torch.zeros(2, dtype=torch.int64))
    with pytest.raises(RuntimeError, match="no CPU"):
        ops.sequence_pool(torch.zeros(4, 4), [[0, 4]], "SUM")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.Context()
    # the C ABI itself reports a CUDA error instead of computing on the host
    h = C.c_void_p()
    rc = built.lib().ctrfe_ctx_create(0, C.byref(h))
    assert rc == built.ERR_CUDA and b"cuda" in built.lib().ctrfe_last_error().lower()


def test_lod_helpers():
    from paddle_infer_b200 import lod
    assert lod.lengths_to_offsets([[2, 0, 1, 2], [0, 4, 0, 7, 0]]) == [[0, 2, 2, 3, 5], [0, 0, 4, 4, 11, 11]]
    assert lod.offsets_to_lengths([[0, 2, 2, 3, 5], [0, 0, 4, 4, 11, 11]]) == [[2, 0, 1, 2], [0, 4, 0, 7, 0]]
    assert lod.check_lod([[0, 2, 2, 3, 5], [0, 0, 4, 4, 11, 11]], 11)
    assert not lod.check_lod([[0, 2, 2, 3, 5], [0, 0, 4, 4, 11, 11]], 12)
    assert not lod.check_lod([[0, 2, 1]])          # decreasing
    assert not lod.check_lod([[1, 2]])             # does not start at 0
    assert not lod.check_lod([[0, 3], [0, 1, 2]])  # level 0 end != level 1 sequence count
    assert lod.num_sequences([[0, 1], [0, 2, 5, 5]]) == 3
    with pytest.raises(ValueError):
        lod.lengths_to_offsets([[1, -1]])


def test_synth_generators_are_deterministic_and_shaped():
    from benchtools import synth
    cfgs = synth.configs()
    assert [cfgs[k].slots for k in ("cfg1", "cfg2", "cfg3", "cfg4", "cfg5")] == [26, 39, 26, 26, 8]
    b1 = synth.draw_batch(cfgs["cfg1"], step=0)
    b2 = synth.draw_batch(cfgs["cfg1"], step=0)
    b3 = synth.draw_batch(cfgs["cfg1"], step=1)
    assert all((x == y).all() for x, y in zip(b1.ids, b2.ids))
    assert any((x.size != y.size) or (x != y).any() for x, y in zip(b1.ids, b3.ids))
    assert len(b1.ids) == 26 and all(l.size == 513 for l in b1.lods)
    lens = np.diff(b1.lods[0].astype(np.int64))
    assert lens.min() == 0 and lens.max() <= 16 and 3.0 < lens.mean() < 4.5
    assert all(0 <= i.min() and i.max() < 1_000_000 for i in b1.ids)
    # Zipf head: rank 1 alone is ~9 % of the draws
    ids = np.concatenate(b1.ids)
    assert 0.05 < (ids == synth.hottest_id(1_000_000)).mean() < 0.15
    # SURVEY 8(d) byte formula on cfg1's nominal K
    assert cfgs["cfg1"].algorithmic_bytes(53248) == 53248 * 40 + 26 * 513 * 8 + 13312 * 32
    t = synth.table_rows_numpy(5, np.arange(100), 8, "f32")
    assert ((t[:, 0] >= 0) & (t[:, 0] < 16) & (t[:, 0] == np.floor(t[:, 0]))).all()
    assert ((t[:, 1] >= 0) & (t[:, 1] <= t[:, 0])).all()
    assert (np.abs(t[:, 2:]) <= 1).all()
    # long-sequence law
    l5 = synth.draw_lengths(np.random.default_rng(0), 4096, 200)
    assert 150 < l5.mean() < 230 and l5.max() <= 1024


def test_package_core_does_not_import_torch():
    """`import paddle_infer_b200.ops` needs numpy and the library only (torch is imported lazily, and
    only when a torch tensor is passed in)."""
    import subprocess
    import sys
    code = ("import sys; sys.path.insert(0, %r); import paddle_infer_b200.ops, paddle_infer_b200.device, paddle_infer_b200.lod; "
            "assert 'torch' not in sys.modules; print('ok')" % ROOT)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "ok" in r.stdout, r.stderr[-2000:]


def test_paddle_shim_compiles_against_the_reference_headers():
    """The PD_BUILD_OP shim is a syntax-only compile against the reference's own paddle/extension.h
    closure (paddle_shim/check_syntax.sh).  /root/reference exists on the authoring box only."""
    import shutil
    import subprocess
    ref = os.environ.get("CTRFE_REFERENCE_ROOT", "/root/reference")
    if not os.path.isdir(os.path.join(ref, "paddle")) or shutil.which("g++") is None:
        pytest.skip("reference tree or g++ not present")
    script = os.path.join(ROOT, "paddle_infer_b200", "paddle_shim", "check_syntax.sh")
    r = subprocess.run(["sh", script], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "syntax OK" in r.stdout


def test_reference_arm_prints_one_contract_line():
    """`bench.py --impl reference` (the CPU arm the driver times beside the GPU arm) runs without a GPU,
    loads neither CUDA nor libctrfe.so, and prints exactly one JSON line with the contract's keys.
    Run on cfg1 so it takes seconds; the driver runs it on the default workload."""
    import json
    import subprocess
    import sys
    code = ("import atexit, sys, runpy\n"
            "def _maps():\n"
            "    print('LOADED_CTRFE' if 'libctrfe' in open('/proc/self/maps').read() else 'NO_CTRFE', file=sys.stderr)\n"
            "atexit.register(_maps)\n"
            "sys.argv = ['bench.py', '--impl', 'reference', '--workload', 'cfg1', '--steps', '3', '--warmup', '1']\n"
            "runpy.run_path(%r, run_name='__main__')\n" % os.path.join(ROOT, "bench.py"))
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines
    line = json.loads(lines[0])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert k in line, k
    assert line["impl"] == "reference" and line["value"] > 0 and line["vs_baseline"] is None
    assert line["e2e"]["value"] == line["value"] and line["e2e"]["h2d_bytes_per_step"] == 0
    cb = line["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == line["value"] and cb["sample"]
    assert "workload" in line["config"] and "model" not in line["config"]
    assert "NO_CTRFE" in r.stderr
