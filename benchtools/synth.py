"""Synthetic CTR workloads (BASELINE.md section 5 / SURVEY.md section 8d).

* tables   generated from a counter-based hash of (seed, row, col): on the device by
           libsynth.so (fill_table), on the host by table_rows_numpy for any subset of rows,
           bit-identically -- a 100 GB table never crosses PCIe and the CPU oracle can be
           fed exactly the rows a sample touches.
* lengths  min(Poisson(L), 4L) with a deterministic 2 % of segments forced empty;
           L >= 100: round(LogNormal(ln L - 0.125, 0.5)) clipped to [0, 1024].
* ids      rank r ~ Zipf(alpha = 1.05) on [1, N] by inverse CDF: the exact discrete law for the
           first 1e4 ranks (P(r=1) = 0.079 at N = 1e8), the continuous power law beyond, then
           scattered: id = (r * 0x9E3779B97F4A7C15 mod 2^64) mod N.

ids/lengths are drawn on ONE side (numpy) and copied, so both sides see the same batch.
Bench/test tooling; nothing here is on the product path.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libsynth.so")
BASE_SEED = 20261106
GOLDEN = 0x9E3779B97F4A7C15
ALPHA = 1.05

_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)
_G = np.uint64(GOLDEN)


@dataclass
class Config:
    name: str
    slots: int
    batch: int
    rows: int              # rows per table
    dim: int
    dtype: str             # "f32" | "f16" | "bf16"
    mean_len: float
    pooltypes: List[str]
    use_cvm: Optional[bool]   # True / False / None (no cvm op)
    shared_table: bool = True
    cfg_index: int = 0
    note: str = ""

    @property
    def seed(self) -> int:
        return BASE_SEED + self.cfg_index

    @property
    def elem_size(self) -> int:
        return 4 if self.dtype == "f32" else 2

    @property
    def d_out(self) -> int:
        return self.dim - 2 if self.use_cvm is False else self.dim

    def n_tables(self) -> int:
        return 1 if self.shared_table else self.slots

    def algorithmic_bytes(self, k_tot: int) -> int:
        """K_tot*(8 + D*s) + S*(B+1)*8 + S*B*D_out*s."""
        s = self.elem_size
        return k_tot * (8 + self.dim * s) + self.slots * (self.batch + 1) * 8 + self.slots * self.batch * self.d_out * s


def configs() -> dict:
    """The five BASELINE.json configs (assumptions marked in BASELINE.md section 4)."""
    c = {}
    c["cfg1"] = Config("cfg1_wide_deep", 26, 512, 1_000_000, 8, "f32", 4, ["SUM"] * 26, True, True, 1)
    c["cfg2"] = Config("cfg2_deepfm", 39, 8192, 100_000_000, 16, "f32", 4, ["SUM"] * 39, True, True, 2)
    c["cfg3"] = Config("cfg3_dlrm", 26, 65536, 40_000_000, 64, "bf16", 4,
                       [("SUM", "AVERAGE", "SQRT")[i % 3] for i in range(26)], None, False, 3)
    c["cfg4"] = Config("cfg4_sharded", 26, 131072, 1_000_000_000, 32, "f32", 4, ["SUM"] * 26, True, True, 4)
    c["cfg5"] = Config("cfg5_longseq", 8, 4096, 10_000_000, 128, "f16", 200,
                       ["MAX"] * 4 + ["SQRT"] * 4, None, False, 5)
    return c


# ------------------------------------------------------------------------------------------
# table values
# ------------------------------------------------------------------------------------------
def _mix64(z: np.ndarray) -> np.ndarray:
    z = z ^ (z >> np.uint64(30))
    z = z * _M1
    z = z ^ (z >> np.uint64(27))
    z = z * _M2
    z = z ^ (z >> np.uint64(31))
    return z


def _unit24(seed: int, rows: np.ndarray, col: int) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = rows.astype(np.uint64) * _G + np.uint64(col) * _M1 + np.uint64(seed) * _M2
        z = _mix64(z)
    return (z >> np.uint64(40)).astype(np.float32) * np.float32(1.0 / 16777216.0)


def table_rows_numpy(seed: int, rows: np.ndarray, dim: int, dtype: str, cvm_cols: bool = True) -> np.ndarray:
    """Rows `rows` of the logical table, float32 / float16 / bf16-bits(uint16), bit-identical
    to what synth_fill_table writes on the device."""
    rows = np.asarray(rows, dtype=np.int64)
    out = np.empty((rows.size, dim), dtype=np.float32)
    show = None
    for c in range(dim):
        u = _unit24(seed, rows, c)
        if cvm_cols and c == 0:
            show = np.floor(u * np.float32(16.0))
            out[:, 0] = show
        elif cvm_cols and c == 1:
            out[:, 1] = np.floor(u * (show + np.float32(1.0)))
        else:
            out[:, c] = u * np.float32(2.0) - np.float32(1.0)
    if dtype == "f32":
        return out
    if dtype == "f16":
        return out.astype(np.float16)
    if dtype == "bf16":
        x = out.view(np.uint32)
        return ((x + (np.uint32(0x7FFF) + ((x >> np.uint32(16)) & np.uint32(1)))) >> np.uint32(16)).astype(np.uint16)
    raise ValueError(dtype)


_lib = None


def _synth_lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            raise ImportError(f"{_LIB} missing: run __graft_entry__.build()")
        _lib = C.CDLL(_LIB)
        _lib.synth_fill_table.argtypes = [C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_ulonglong, C.c_longlong,
                                          C.c_longlong, C.c_int, C.c_void_p]
    return _lib


def fill_table(t, seed: int, row0: int = 0, cvm_cols: bool = True, row_stride: int = 1) -> None:
    """Fill the CUDA tensor t [rows, dim] with logical rows row0, row0 + row_stride, ... (one
    shard of a row-sharded table when row_stride > 1)."""
    import torch
    code = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}[t.dtype]
    assert t.is_cuda and t.is_contiguous()
    with torch.cuda.device(t.device):
        rc = _synth_lib().synth_fill_table(t.data_ptr(), t.shape[0], t.shape[1], code, seed, row0, row_stride,
                                           int(cvm_cols),
                                           torch.cuda.current_stream(t.device).cuda_stream)
    if rc != 0:
        raise RuntimeError(f"synth_fill_table failed: cudaError {rc}")


_host_lib = None


def fill_table_host(rows: int, dim: int, dtype: str, seed: int, cvm_cols: bool = True, threads: int = 0) -> np.ndarray:
    """Logical rows 0..rows-1 generated on the host cores (benchtools/synth_host.c, OpenMP),
    bit-identical to fill_table / table_rows_numpy.  bf16 comes back as uint16 bit patterns."""
    global _host_lib
    if _host_lib is None:
        import subprocess
        src, out = os.path.join(_HERE, "synth_host.c"), os.path.join(_HERE, "libsynth_host.so")
        if not os.path.exists(out) or os.path.getmtime(out) < os.path.getmtime(src):
            subprocess.check_call(["gcc", "-O3", "-fopenmp", "-fPIC", "-shared", "-ffp-contract=off", "-o", out, src, "-lm"])
        _host_lib = C.CDLL(out)
        _host_lib.synth_fill_table_host.argtypes = [C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_ulonglong, C.c_longlong,
                                                    C.c_longlong, C.c_int, C.c_int]
    npdt = {"f32": np.float32, "f16": np.float16, "bf16": np.uint16}[dtype]
    t = np.empty((rows, dim), dtype=npdt)
    rc = _host_lib.synth_fill_table_host(t.ctypes.data, rows, dim, {"f32": 0, "f16": 1, "bf16": 2}[dtype], seed, 0, 1,
                                         int(cvm_cols), threads or (os.cpu_count() or 1))
    if rc != 0:
        raise RuntimeError("synth_fill_table_host failed")
    return t


# ------------------------------------------------------------------------------------------
# lengths and ids
# ------------------------------------------------------------------------------------------
def draw_lengths(rng: np.random.Generator, n: int, mean_len: float) -> np.ndarray:
    if mean_len >= 100:
        l = np.rint(rng.lognormal(np.log(mean_len) - 0.125, 0.5, n)).astype(np.int64)
        l = np.clip(l, 0, 1024)
    else:
        l = np.minimum(rng.poisson(mean_len, n), int(4 * mean_len)).astype(np.int64)
    l[np.arange(n) % 50 == 7] = 0  # a deterministic 2 % of segments are empty
    return l


ZIPF_EXACT_HEAD = 10_000  # ranks drawn from the exact discrete law; beyond: the continuous power law


_zipf_cache = {}


def _zipf_head(n_rows: int, alpha: float):
    """cdf of ranks 1..M under P(r) = r^-alpha / H(N, alpha), M = min(ZIPF_EXACT_HEAD, N); the mass
    beyond M is the midpoint-rule integral of x^-alpha over [M + 0.5, N + 0.5) (relative error of
    the normaliser < 1e-9 at M = 1e4)."""
    key = (n_rows, alpha)
    if key not in _zipf_cache:
        m = min(ZIPF_EXACT_HEAD, n_rows)
        w = np.arange(1, m + 1, dtype=np.float64) ** (-alpha)
        a = 1.0 - alpha
        tail = ((n_rows + 0.5) ** a - (m + 0.5) ** a) / a if n_rows > m else 0.0
        z = w.sum() + tail
        _zipf_cache[key] = (np.cumsum(w) / z, m)
    return _zipf_cache[key]


def zipf_ids(rng: np.random.Generator, k: int, n_rows: int, alpha: float = ALPHA) -> np.ndarray:
    cdf, m = _zipf_head(n_rows, alpha)
    u = rng.random(k)
    r = np.empty(k, dtype=np.uint64)
    head = u < cdf[-1]
    r[head] = np.searchsorted(cdf, u[head], side="right") + 1
    if n_rows > m:
        a = 1.0 - alpha
        ut = (u[~head] - cdf[-1]) / (1.0 - cdf[-1])
        lo, hi = (m + 0.5) ** a, (n_rows + 0.5) ** a
        x = (lo + ut * (hi - lo)) ** (1.0 / a)
        r[~head] = np.clip(np.rint(x), m + 1, n_rows).astype(np.uint64)
    else:
        r[~head] = m
    with np.errstate(over="ignore"):
        ids = (r * _G) % np.uint64(n_rows)
    return ids.astype(np.int64)


def zipf_rank1_probability(n_rows: int, alpha: float = ALPHA) -> float:
    return float(_zipf_head(n_rows, alpha)[0][0])


def hottest_id(n_rows: int) -> int:
    """The id Zipf rank 1 maps to (used as padding_idx in the correctness-only runs)."""
    return int((GOLDEN % (1 << 64)) % n_rows)


@dataclass
class Batch:
    """One synthetic batch on the host: per slot ids (int64) and LoD offsets (uint64)."""
    ids: List[np.ndarray]
    lods: List[np.ndarray]
    k_tot: int = field(default=0)


def draw_batch(cfg: Config, step: int = 0, batch: Optional[int] = None, slots: Optional[int] = None,
               rows: Optional[int] = None) -> Batch:
    B = cfg.batch if batch is None else batch
    S = cfg.slots if slots is None else slots
    N = cfg.rows if rows is None else rows
    rng = np.random.default_rng([cfg.seed, step])
    ids, lods, k_tot = [], [], 0
    for _ in range(S):
        l = draw_lengths(rng, B, cfg.mean_len)
        lod = np.zeros(B + 1, dtype=np.uint64)
        lod[1:] = np.cumsum(l)
        k = int(lod[-1])
        ids.append(zipf_ids(rng, k, N))
        lods.append(lod)
        k_tot += k
    return Batch(ids, lods, k_tot)
