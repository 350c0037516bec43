"""oracle/loader.py — ctypes access to the CPU checker.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
`Oracle`  = oracle/liboracle.so, our C restatement (paged_attn_oracle.c).
`Ref`     = oracle/_ref/libref_<isa>.so, the reference's own ggml CPU code + ref_harness.cpp (prebuilt in the
            build container; None on a machine where it was never built).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
F32, F16, Q4_0, Q8_0 = 0, 1, 2, 8


def row_bytes(t: int, n: int) -> int:
    return {F32: n * 4, F16: n * 2, Q8_0: n // 32 * 34, Q4_0: n // 32 * 18}[t]


class _AttnParams(C.Structure):
    _fields_ = [
        ("n_q", C.c_int), ("n_head", C.c_int), ("n_head_kv", C.c_int), ("head_dim", C.c_int),
        ("type_k", C.c_int), ("type_v", C.c_int),
        ("q", C.c_void_p), ("nbq1", C.c_size_t), ("nbq2", C.c_size_t),
        ("k", C.c_void_p), ("nbk1", C.c_size_t), ("nbk2", C.c_size_t),
        ("v", C.c_void_p), ("nbv1", C.c_size_t), ("nbv2", C.c_size_t),
        ("scale", C.c_float), ("max_bias", C.c_float), ("logit_softcap", C.c_float),
        ("dst", C.c_void_p),
        ("n_kv", C.c_int64), ("mask", C.c_void_p), ("nb_mask1", C.c_size_t),
        ("block_table", C.c_void_p), ("bt_stride", C.c_int64), ("seq_lens", C.c_void_p),
        ("n_seqs", C.c_int), ("block_size", C.c_int),
        ("q_seq", C.c_void_p), ("q_pos", C.c_void_p),
    ]


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Oracle:
    def __init__(self):
        path = os.path.join(_HERE, "liboracle.so")
        src = os.path.join(_HERE, "paged_attn_oracle.c")
        if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(src):
            subprocess.check_call(["make", "-s", "-C", _HERE, "oracle"])
        L = self.lib = C.CDLL(path)
        L.orc_from_float.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64]
        L.orc_to_float.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64]
        L.orc_kv_write.argtypes = [C.c_void_p, C.c_size_t, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                   C.c_size_t, C.c_size_t, C.c_int64]
        L.orc_block_copy.argtypes = [C.c_void_p, C.c_size_t, C.c_int64, C.c_void_p, C.c_int64]
        L.orc_attention.argtypes = [C.POINTER(_AttnParams)]
        L.orc_k_shift.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t, C.c_int64, C.c_void_p, C.c_void_p,
                                  C.c_int64, C.c_int, C.c_int, C.c_int] + [C.c_float]*6 + [C.c_void_p]
        L.orc_nmse.restype = C.c_double
        L.orc_nmse.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.orc_max_rel_err.restype = C.c_double
        L.orc_max_rel_err.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t]
        L.orc_fp32_to_fp16.restype = C.c_uint16
        L.orc_fp32_to_fp16.argtypes = [C.c_float]
        L.orc_fp16_to_fp32.restype = C.c_float
        L.orc_fp16_to_fp32.argtypes = [C.c_uint16]

    def from_float(self, t: int, x: np.ndarray) -> np.ndarray:
        """rows of x (last axis) -> bytes, shape [..., row_bytes]"""
        x = np.ascontiguousarray(x, dtype=np.float32)
        k = x.shape[-1]
        out = np.empty(x.shape[:-1] + (row_bytes(t, k),), dtype=np.uint8)
        assert t in (F32, F16) or k % 32 == 0
        # quant blocks are independent and rows are whole blocks, so one flat call converts every row
        assert self.lib.orc_from_float(t, _p(x.reshape(-1)), _p(out.reshape(-1)), x.size) == 0
        return out

    def to_float(self, t: int, b: np.ndarray, k: int) -> np.ndarray:
        b = np.ascontiguousarray(b, dtype=np.uint8)
        rb = row_bytes(t, k)
        rows = b.reshape(-1, rb)
        out = np.empty((rows.shape[0], k), dtype=np.float32)
        assert self.lib.orc_to_float(t, _p(rows), _p(out), out.size) == 0
        return out.reshape(b.shape[:-1] + (k,))

    def kv_write(self, src, slot_mapping, cache, t, n_head_kv, head_dim, nb_token, nb_head, n_cells):
        src = np.ascontiguousarray(src, dtype=np.float32)
        sm = None if slot_mapping is None else np.ascontiguousarray(slot_mapping, dtype=np.int32)
        assert cache.dtype == np.uint8 and cache.flags.c_contiguous
        rc = self.lib.orc_kv_write(_p(src), src.strides[0], src.shape[0], _p(sm), _p(cache), t, n_head_kv, head_dim,
                                   nb_token, nb_head, n_cells)
        assert rc == 0

    def k_shift(self, cache, t, n_head_kv, head_dim, shift, cells=None, *, n_rot=None, mode=0, n_ctx_orig=8192,
                freq_base=500000.0, freq_scale=1.0, ext_factor=0.0, attn_factor=1.0, beta_fast=32.0, beta_slow=1.0,
                freq_factors=None):
        """in place on a flat token-major cache (uint8 [n_cells * row_bytes])"""
        assert cache.dtype == np.uint8 and cache.flags.c_contiguous
        hb = row_bytes(t, head_dim)
        n_cells = cache.size // (hb * n_head_kv)
        sh = np.ascontiguousarray(shift, dtype=np.int32); assert sh.size == n_cells
        cl = None if cells is None else np.ascontiguousarray(cells, dtype=np.int32)
        ff = None if freq_factors is None else np.ascontiguousarray(freq_factors, dtype=np.float32)
        rc = self.lib.orc_k_shift(_p(cache), t, n_head_kv, head_dim, hb * n_head_kv, hb, n_cells, _p(sh), _p(cl),
                                  0 if cl is None else cl.size, n_rot or head_dim, mode, n_ctx_orig, freq_base, freq_scale,
                                  ext_factor, attn_factor, beta_fast, beta_slow, _p(ff))
        assert rc == 0, rc

    def block_copy(self, cache, block_bytes, n_blocks, pairs):
        pairs = np.ascontiguousarray(pairs, dtype=np.int32)
        assert self.lib.orc_block_copy(_p(cache), block_bytes, n_blocks, _p(pairs), pairs.shape[0]) == 0

    def attention(self, q, k, v, type_k, type_v, n_head_kv, scale, *, nbk=None, nbv=None, block_table=None,
                  seq_lens=None, block_size=16, q_seq=None, q_pos=None, mask=None, n_kv=0, max_bias=0.0,
                  softcap=0.0) -> np.ndarray:
        """q F32 [n_q, n_head, d]; k, v uint8 flat caches (token-major unless nbk/nbv = (nb1, nb2) given)."""
        q = np.ascontiguousarray(q, dtype=np.float32)
        n_q, n_head, d = q.shape
        P = _AttnParams()
        P.n_q, P.n_head, P.n_head_kv, P.head_dim = n_q, n_head, n_head_kv, d
        P.type_k, P.type_v = type_k, type_v
        P.q, P.nbq1, P.nbq2 = _p(q), n_head * d * 4, d * 4
        hk, hv = row_bytes(type_k, d), row_bytes(type_v, d)
        P.k = _p(k)
        P.nbk1, P.nbk2 = nbk if nbk else (hk * n_head_kv, hk)
        P.v = _p(v)
        P.nbv1, P.nbv2 = nbv if nbv else (hv * n_head_kv, hv)
        P.scale, P.max_bias, P.logit_softcap = scale, max_bias, softcap
        out = np.empty_like(q)
        P.dst = _p(out)
        keep = []
        if block_table is not None:
            bt = np.ascontiguousarray(block_table, dtype=np.int32)
            sl = np.ascontiguousarray(seq_lens, dtype=np.int32)
            P.block_table, P.bt_stride, P.seq_lens = _p(bt), bt.shape[1], _p(sl)
            P.n_seqs, P.block_size = bt.shape[0], block_size
            qs = None if q_seq is None else np.ascontiguousarray(q_seq, dtype=np.int32)
            qp = None if q_pos is None else np.ascontiguousarray(q_pos, dtype=np.int32)
            P.q_seq, P.q_pos = _p(qs), _p(qp)
            keep += [bt, sl, qs, qp]
        else:
            P.n_kv = n_kv
            if mask is not None:
                m = np.ascontiguousarray(mask, dtype=np.float16)
                P.mask, P.nb_mask1 = _p(m), m.strides[0]
                keep.append(m)
        rc = self.lib.orc_attention(C.byref(P))
        assert rc == 0, rc
        return out

    def nmse(self, ref, out) -> float:
        ref = np.ascontiguousarray(ref, dtype=np.float32).reshape(-1)
        out = np.ascontiguousarray(out, dtype=np.float32).reshape(-1)
        return float(self.lib.orc_nmse(_p(ref), _p(out), ref.size))

    def max_rel_err(self, ref, out) -> float:
        row = ref.shape[-1]
        ref = np.ascontiguousarray(ref, dtype=np.float32).reshape(-1)
        out = np.ascontiguousarray(out, dtype=np.float32).reshape(-1)
        return float(self.lib.orc_max_rel_err(_p(ref), _p(out), ref.size, row))


def _cpu_flags() -> set:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


class Ref:
    """the reference's ggml CPU backend (oracle/_ref); construct with Ref.load() -> Ref | None"""

    @staticmethod
    def load():
        flags = _cpu_flags()
        order = []
        if {"avx512f", "avx512bw", "avx512vl", "avx512dq", "avx512cd", "avx512_vnni"} <= flags:
            order.append("avx512")
        if {"avx2", "fma", "f16c", "bmi2"} <= flags:
            order.append("avx2")
        for isa in order:
            path = os.path.join(_HERE, "_ref", f"libref_{isa}.so")
            if os.path.exists(path):
                return Ref(path, isa)
        return None

    def __init__(self, path: str, isa: str):
        self.isa = isa
        L = self.lib = C.CDLL(path)
        L.ref_n_threads_default.restype = C.c_int
        L.ref_cpu_features.restype = C.c_char_p
        L.ref_quantize_row_ref.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64]
        L.ref_quantize_row_cpu.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64]
        L.ref_to_float.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64]
        L.ref_cpy_f32.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_int]
        L.ref_attention.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int, C.c_int,
                                    C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float,
                                    C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_double)]
        L.ref_k_shift.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int] + [C.c_float]*6 \
                                 + [C.c_void_p, C.c_int]
        self.n_threads = L.ref_n_threads_default()

    def features(self) -> str:
        return self.lib.ref_cpu_features().decode()

    def quantize(self, t: int, x: np.ndarray, simd: bool = False) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float32)
        k = x.shape[-1]
        out = np.empty(x.shape[:-1] + (row_bytes(t, k),), dtype=np.uint8)
        fn = self.lib.ref_quantize_row_cpu if simd else self.lib.ref_quantize_row_ref
        assert t == F16 or k % 32 == 0
        assert fn(t, _p(x.reshape(-1)), _p(out.reshape(-1)), x.size) == 0
        return out

    def to_float(self, t: int, b: np.ndarray, k: int) -> np.ndarray:
        b = np.ascontiguousarray(b, dtype=np.uint8)
        rows = b.reshape(-1, row_bytes(t, k))
        out = np.empty((rows.shape[0], k), dtype=np.float32)
        assert self.lib.ref_to_float(t, _p(rows), _p(out), out.size) == 0
        return out.reshape(b.shape[:-1] + (k,))

    def k_shift(self, cache, t, n_head_kv, head_dim, shift, *, n_rot=None, mode=0, n_ctx_orig=8192, freq_base=500000.0,
                freq_scale=1.0, ext_factor=0.0, attn_factor=1.0, beta_fast=32.0, beta_slow=1.0, freq_factors=None, n_threads=2):
        """the reference's K-shift graph (build_kv_self_shift) on the CPU backend, in place, every cell"""
        assert cache.dtype == np.uint8 and cache.flags.c_contiguous
        n_cells = cache.size // (row_bytes(t, head_dim) * n_head_kv)
        sh = np.ascontiguousarray(shift, dtype=np.int32); assert sh.size == n_cells
        ff = None if freq_factors is None else np.ascontiguousarray(freq_factors, dtype=np.float32)
        rc = self.lib.ref_k_shift(t, _p(cache), n_cells, n_head_kv, head_dim, _p(sh), n_rot or head_dim, mode, n_ctx_orig,
                                  freq_base, freq_scale, ext_factor, attn_factor, beta_fast, beta_slow, _p(ff), n_threads)
        assert rc == 0, rc

    def cpy_f32(self, t: int, src: np.ndarray, n_threads: int = 1) -> np.ndarray:
        src = np.ascontiguousarray(src, dtype=np.float32)
        n_rows, k = src.shape
        out = np.empty((n_rows, row_bytes(t, k)), dtype=np.uint8)
        assert self.lib.ref_cpy_f32(t, _p(src), k, n_rows, _p(out), n_threads) == 0
        return out

    def attention(self, mode: int, q, k, v, type_k, type_v, n_head_kv, n_kv, mask, scale, max_bias=0.0, softcap=0.0,
                  n_threads: int | None = None, n_iter: int = 1):
        """mode 0: GGML_OP_FLASH_ATTN_EXT; mode 1: F32 mul_mat/soft_max_ext formulation.
        q [n_q, n_head, d] F32; k, v flat token-major caches (uint8); mask F16 [pad64(n_q), n_kv] or None.
        -> (out [n_q, n_head, d], best microseconds)"""
        q = np.ascontiguousarray(q, dtype=np.float32)
        n_q, n_head, d = q.shape
        m = None
        if mask is not None:
            m = np.ascontiguousarray(mask, dtype=np.float16)
            pad = (n_q + 63) // 64 * 64
            assert m.shape == (pad, n_kv), (m.shape, pad, n_kv)
        out = np.empty_like(q)
        us = C.c_double(0.0)
        rc = self.lib.ref_attention(mode, n_q, n_head, n_head_kv, d, n_kv, type_k, type_v, _p(q), _p(k), _p(v), _p(m),
                                    scale, max_bias, softcap, _p(out), n_threads or self.n_threads, n_iter, C.byref(us))
        assert rc == 0
        return out, us.value
