"""oracle/kv_state_oracle.py — TEST INFRASTRUCTURE ONLY (see oracle/loader.py for who may import this).

Restatement of the reference's KV-cache session stream for a single sequence, over the reference's OWN cache layout:
a flat unified cache of `size` cells where the sequence occupies arbitrary cells (src/llama-kv-cache.cpp):

  state_write(io, seq_id)      :848-884   cell ranges that carry seq_id, in cell-index order; u32 cell_count
  state_write_meta             :905-923   per cell: i32 pos, u32 n_seq_id (0 for a single-sequence stream)
  state_write_data             :925-971   u32 v_trans, u32 n_layer; per layer i32 type, u64 row size, the rows of
                                          every range; first all K layers, then all V layers (v_trans == 0)
  state_read_meta / _data      :1008-1240 the checks and their messages

Pinned: tests/test_ref_kv_cache.py compares this writer byte for byte with the reference's own state_write (single-sequence
and whole-cache form) run through oracle/_ref/refkv, the reference's llama_kv_cache_unified compiled where it lies.
"""
from __future__ import annotations

import struct

import numpy as np


def unified_state_write(cells, seq_id, k_l, v_l, type_k, type_v):
    """cells: list of (pos, set_of_seq_ids) per cell index (pos < 0: empty); k_l / v_l: per layer uint8 arrays
    [size, size_row] (the flat k_l[il] / v_l[il] tensors seen as rows).  Returns the stream bytes."""
    size = len(cells)
    ranges, count, begin = [], 0, size
    for i, (pos, ids) in enumerate(cells):                       # :856-872
        if (seq_id == -1 and ids) or seq_id in ids:
            count += 1
            if begin == size:
                begin = i
        elif begin != size:
            ranges.append((begin, i))
            begin = size
    if begin != size:
        ranges.append((begin, size))
    out = [struct.pack("<I", count)]                             # :879
    for a, b in ranges:                                          # state_write_meta
        for i in range(a, b):
            pos, ids = cells[i]
            n_seq_id = len(ids) if seq_id == -1 else 0
            out.append(struct.pack("<iI", pos, n_seq_id))
            if n_seq_id:
                out.extend(struct.pack("<i", s) for s in sorted(ids))
    out.append(struct.pack("<II", 0, len(k_l)))                  # v_trans = 0 (flash attention), n_layer
    for t, layers in ((type_k, k_l), (type_v, v_l)):
        for rows in layers:
            rows = np.ascontiguousarray(rows, dtype=np.uint8)
            out.append(struct.pack("<iQ", t, rows.shape[1]))
            for a, b in ranges:
                out.append(rows[a:b].tobytes())
    return b"".join(out)


def unified_state_read_errors(data: bytes, n_layer, type_k, k_row, type_v, v_row):
    """the reference's verdict on a single-sequence stream for a cache of this shape: None if accepted, else the
    message of the failing check (state_read_meta :1020-1027, state_read_data :1105-1175)"""
    try:
        off = 0
        (count,) = struct.unpack_from("<I", data, off); off += 4
        for _ in range(count):
            _pos, n_seq_id = struct.unpack_from("<iI", data, off); off += 8
            if n_seq_id != 0:
                return "invalid seq_id-agnostic kv cell"
        v_trans, nl = struct.unpack_from("<II", data, off); off += 8
        if nl != n_layer:
            return "mismatched layer count"
        if v_trans != 0:
            return "incompatible V transposition"
        for what, t_want, row_want in (("key", type_k, k_row), ("value", type_v, v_row)):
            for _ in range(n_layer):
                t, row = struct.unpack_from("<iQ", data, off); off += 12
                if t != t_want:
                    return f"mismatched {what} type"
                if row != row_want:
                    return f"mismatched {what} row size"
                if off + count * row > len(data):
                    raise struct.error("short")
                off += count * row
    except struct.error:
        return "truncated"
    return None
