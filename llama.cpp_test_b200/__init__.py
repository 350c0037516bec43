"""llama.cpp_test_b200 — test/bench driver bindings for libb200attn.so.

The product is the C-ABI shared library (include/b200_paged_attn.h, b200_ggml_abi.h, b200_host.h) built from
csrc/ (hand-written sm_100a CUDA) and host/ (C++ block managers mirroring the reference's
BlockManager / BlockKVCache / llama_kv_cache interfaces).  This Python module only binds it with ctypes
so pytest and bench.py can drive it with torch-allocated device buffers; torch is plumbing (memory,
streams, NCCL), never the compute path.  There is NO fallback: if the library is missing, importing
raises; if no GPU is present, compute calls return B200_ERR_CUDA and the wrappers raise.

The directory name contains a dot, so import it with `load()` from __graft_entry__ / tests/conftest.py
(importlib by path) under the module name `llama_cpp_test_b200`.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200_LIB") or os.path.join(_HERE, "libb200attn.so")     # B200_LIB: A/B builds of the same ABI

TYPE_F32, TYPE_F16, TYPE_Q4_0, TYPE_Q8_0 = 0, 1, 2, 8
TYPE_NAMES = {TYPE_F32: "f32", TYPE_F16: "f16", TYPE_Q4_0: "q4_0", TYPE_Q8_0: "q8_0"}
KERNEL_AUTO, KERNEL_GENERIC, KERNEL_STREAMK, KERNEL_PREFILL, KERNEL_STREAMK_DQ = 0, 1, 2, 3, 4

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build it with `make -C {os.path.join(_HERE, 'csrc')}` "
        "(or __graft_entry__.build()); this package has no fallback path")

lib = C.CDLL(LIB_PATH)


class AttnDesc(C.Structure):
    """struct b200_attn_desc (include/b200_paged_attn.h)"""
    _fields_ = [
        ("n_q", C.c_int32), ("n_head", C.c_int32), ("n_head_kv", C.c_int32), ("head_dim", C.c_int32),
        ("type_k", C.c_int32), ("type_v", C.c_int32),
        ("q", C.c_void_p), ("nbq1", C.c_int64), ("nbq2", C.c_int64),
        ("k", C.c_void_p), ("nbk1", C.c_int64), ("nbk2", C.c_int64),
        ("v", C.c_void_p), ("nbv1", C.c_int64), ("nbv2", C.c_int64),
        ("dst", C.c_void_p),
        ("scale", C.c_float), ("max_bias", C.c_float), ("logit_softcap", C.c_float),
        ("n_kv", C.c_int64),
        ("mask", C.c_void_p), ("nb_mask1", C.c_int64),
        ("block_table", C.c_void_p), ("bt_stride", C.c_int64),
        ("seq_lens", C.c_void_p),
        ("n_seqs", C.c_int32), ("block_size", C.c_int32),
        ("q_seq", C.c_void_p), ("q_pos", C.c_void_p),
        ("max_seq_len", C.c_int64), ("total_tokens", C.c_int64),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t),
        ("kernel", C.c_int32), ("flags", C.c_int32),
    ]


class KVWriteDesc(C.Structure):
    """struct b200_kv_write_desc"""
    _fields_ = [
        ("k_src", C.c_void_p), ("v_src", C.c_void_p), ("src_nb1", C.c_int64),
        ("n_tokens", C.c_int64),
        ("slot_mapping", C.c_void_p),
        ("k_cache", C.c_void_p), ("v_cache", C.c_void_p),
        ("type_k", C.c_int32), ("type_v", C.c_int32), ("n_head_kv", C.c_int32), ("head_dim", C.c_int32),
        ("nbk_token", C.c_int64), ("nbk_head", C.c_int64), ("nbv_token", C.c_int64), ("nbv_head", C.c_int64),
        ("n_cells", C.c_int64),
    ]


class KShiftDesc(C.Structure):
    """struct b200_k_shift_desc"""
    _fields_ = [
        ("caches", C.POINTER(C.c_void_p)), ("n_caches", C.c_int32),
        ("type_k", C.c_int32), ("n_head_kv", C.c_int32), ("head_dim", C.c_int32),
        ("nb_token", C.c_int64), ("nb_head", C.c_int64), ("n_cells", C.c_int64),
        ("shift", C.c_void_p),
        ("cells", C.c_void_p), ("n_rows", C.c_int64),
        ("n_rot", C.c_int32), ("mode", C.c_int32), ("n_ctx_orig", C.c_int32),
        ("freq_base", C.c_float), ("freq_scale", C.c_float), ("ext_factor", C.c_float), ("attn_factor", C.c_float),
        ("beta_fast", C.c_float), ("beta_slow", C.c_float),
        ("freq_factors", C.c_void_p),
    ]


class GgmlTensor(C.Structure):
    """struct b200_ggml_tensor == the reference's struct ggml_tensor (include/b200_ggml_abi.h)"""
    pass


GgmlTensor._fields_ = [
    ("type", C.c_int32), ("buffer", C.c_void_p),
    ("ne", C.c_int64 * 4), ("nb", C.c_size_t * 4),
    ("op", C.c_int32), ("op_params", C.c_int32 * 16), ("flags", C.c_int32),
    ("src", C.POINTER(GgmlTensor) * 10),
    ("view_src", C.POINTER(GgmlTensor)), ("view_offs", C.c_size_t),
    ("data", C.c_void_p), ("name", C.c_char * 64), ("extra", C.c_void_p), ("padding", C.c_char * 8),
]

GGML_OP_CPY, GGML_OP_FLASH_ATTN_EXT, GGML_TYPE_I32 = 32, 62, 26

# every symbol the headers declare (tests/test_abi.py checks the .so exports exactly these)
EXPORTS = {
    "b200_paged_attn.h": [
        "b200_paged_attention_workspace", "b200_paged_attention", "b200_supports_attention", "b200_kv_write",
        "b200_block_copy", "b200_kv_rows_gather", "b200_kv_rows_scatter", "b200_k_shift", "b200_decode_step", "b200_store_attention",
        "b200_peer_alloc", "b200_peer_free", "b200_peer_export", "b200_peer_open", "b200_peer_close", "b200_peer_publish",
        "b200_peer_wait", "b200_peer_exchange", "b200_peer_exchange_ll", "b200_peer_read_state",
        "b200_static_workspace_bytes", "b200_paged_attention_static_ws", "b200_last_error", "b200_version", "b200_device_sm_count", "b200_launch_count"],
    "b200_ggml_abi.h": ["b200_flash_attn_ext_workspace", "b200_flash_attn_ext", "b200_cpy", "b200_supports_op"],
    "b200_host.h": [
        "b200_bm_create", "b200_bm_destroy", "b200_bm_allocate", "b200_bm_free", "b200_bm_evict", "b200_bm_status",
        "b200_fork_trace", "b200_pc_create", "b200_pc_destroy", "b200_pc_append", "b200_pc_prepare_batch", "b200_pc_undo_batch", "b200_pc_seq_cp", "b200_pc_seq_rm",
        "b200_pc_seq_keep", "b200_pc_clear", "b200_pc_seq_len", "b200_pc_n_free", "b200_pc_ref_count", "b200_pc_export",
        "b200_pc_slots_of", "b200_pc_defrag", "b200_pc_restore_seq", "b200_pc_context_shift", "b200_kv_state_seq_size", "b200_kv_state_seq_write",
        "b200_kv_state_seq_parse", "b200_kv_state_seq_positions", "b200_pc_used_cells", "b200_kv_state_all_write", "b200_kv_state_all_restore",
        "b200_kvc_create", "b200_kvc_destroy", "b200_kvc_clear", "b200_kvc_defrag", "b200_kvc_seq_rm", "b200_kvc_seq_cp", "b200_kvc_seq_keep",
        "b200_kvc_seq_add", "b200_kvc_seq_div", "b200_kvc_seq_pos_max", "b200_kvc_n_tokens", "b200_kvc_used_cells", "b200_kvc_unsupported",
        "b200_kvc_slots_of", "b200_kvc_find_slot", "b200_kvc_update", "b200_kvc_update_groups", "b200_shard_sequences", "b200_head_shard"],
}

lib.b200_paged_attention_workspace.restype = C.c_size_t
lib.b200_paged_attention_workspace.argtypes = [C.POINTER(AttnDesc)]
lib.b200_paged_attention.restype = C.c_int
lib.b200_paged_attention.argtypes = [C.POINTER(AttnDesc), C.c_void_p]
lib.b200_paged_attention_static_ws.restype = C.c_int
lib.b200_paged_attention_static_ws.argtypes = [C.POINTER(AttnDesc), C.c_void_p]
lib.b200_static_workspace_bytes.restype = C.c_size_t
lib.b200_supports_attention.restype = C.c_int
lib.b200_supports_attention.argtypes = [C.POINTER(AttnDesc)]
lib.b200_kv_write.restype = C.c_int
lib.b200_kv_write.argtypes = [C.POINTER(KVWriteDesc), C.c_void_p]
lib.b200_block_copy.restype = C.c_int
lib.b200_block_copy.argtypes = [C.POINTER(C.c_void_p), C.c_int32, C.c_int64, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p]
lib.b200_kv_rows_gather.restype = C.c_int
lib.b200_kv_rows_gather.argtypes = [C.POINTER(C.c_void_p), C.c_int32, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
lib.b200_kv_rows_scatter.restype = C.c_int
lib.b200_kv_rows_scatter.argtypes = lib.b200_kv_rows_gather.argtypes
lib.b200_k_shift.restype = C.c_int
lib.b200_k_shift.argtypes = [C.POINTER(KShiftDesc), C.c_void_p]
lib.b200_peer_alloc.restype = C.c_void_p
lib.b200_peer_alloc.argtypes = [C.c_size_t]
lib.b200_peer_free.argtypes = [C.c_void_p]
lib.b200_peer_export.argtypes = [C.c_void_p, C.c_char_p]
lib.b200_peer_open.restype = C.c_void_p
lib.b200_peer_open.argtypes = [C.c_char_p]
lib.b200_peer_close.argtypes = [C.c_void_p]
lib.b200_peer_publish.restype = C.c_int
lib.b200_peer_publish.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
lib.b200_peer_exchange.restype = C.c_int
lib.b200_peer_exchange.argtypes = lib.b200_peer_publish.argtypes
lib.b200_peer_exchange_ll.restype = C.c_int
lib.b200_peer_exchange_ll.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_void_p), C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
lib.b200_peer_wait.restype = C.c_int
lib.b200_peer_wait.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]
lib.b200_peer_read_state.restype = C.c_int
lib.b200_peer_read_state.argtypes = [C.c_void_p, C.POINTER(C.c_uint32)]
lib.b200_decode_step.restype = C.c_int
lib.b200_decode_step.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]
lib.b200_store_attention.restype = C.c_int
lib.b200_store_attention.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
lib.b200_last_error.restype = C.c_char_p
lib.b200_version.restype = C.c_char_p
lib.b200_device_sm_count.restype = C.c_int
lib.b200_launch_count.restype = C.c_int64
lib.b200_flash_attn_ext_workspace.restype = C.c_size_t
lib.b200_flash_attn_ext_workspace.argtypes = [C.POINTER(GgmlTensor)]
lib.b200_flash_attn_ext.restype = C.c_int
lib.b200_flash_attn_ext.argtypes = [C.POINTER(GgmlTensor), C.c_void_p, C.c_size_t, C.c_void_p]
lib.b200_cpy.restype = C.c_int
lib.b200_cpy.argtypes = [C.POINTER(GgmlTensor), C.POINTER(GgmlTensor), C.c_void_p]
lib.b200_supports_op.restype = C.c_int
lib.b200_supports_op.argtypes = [C.POINTER(GgmlTensor)]

lib.b200_bm_create.restype = C.c_void_p
lib.b200_bm_create.argtypes = [C.c_int64]
lib.b200_bm_destroy.argtypes = [C.c_void_p]
lib.b200_bm_allocate.restype = C.c_int32
lib.b200_bm_allocate.argtypes = [C.c_void_p]
lib.b200_bm_free.argtypes = [C.c_void_p, C.c_int32]
lib.b200_bm_evict.restype = C.c_int32
lib.b200_bm_evict.argtypes = [C.c_void_p]
lib.b200_bm_status.restype = C.c_int64
lib.b200_bm_status.argtypes = [C.c_void_p, C.POINTER(C.c_uint8), C.c_int64]
lib.b200_fork_trace.argtypes = [C.c_int]
lib.b200_pc_create.restype = C.c_void_p
lib.b200_pc_create.argtypes = [C.c_int64, C.c_int32]
lib.b200_pc_destroy.argtypes = [C.c_void_p]
lib.b200_pc_append.restype = C.c_int64
lib.b200_pc_append.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_int64]
lib.b200_pc_seq_cp.restype = C.c_int32
lib.b200_pc_seq_cp.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int64, C.c_int64]
lib.b200_pc_seq_rm.restype = C.c_int32
lib.b200_pc_seq_rm.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.c_int64]
lib.b200_pc_seq_keep.argtypes = [C.c_void_p, C.c_int32]
lib.b200_pc_clear.argtypes = [C.c_void_p]
lib.b200_pc_seq_len.restype = C.c_int64
lib.b200_pc_seq_len.argtypes = [C.c_void_p, C.c_int32]
lib.b200_pc_n_free.restype = C.c_int64
lib.b200_pc_n_free.argtypes = [C.c_void_p]
lib.b200_pc_ref_count.restype = C.c_int32
lib.b200_pc_ref_count.argtypes = [C.c_void_p, C.c_int64]
lib.b200_pc_prepare_batch.restype = C.c_int32
lib.b200_pc_prepare_batch.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                      C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int64),
                                      C.POINTER(C.c_int32), C.c_int64, C.POINTER(C.c_int64)]
lib.b200_pc_undo_batch.restype = None
lib.b200_pc_undo_batch.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_int64]
lib.b200_pc_export.restype = C.c_int32
lib.b200_pc_export.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.c_int64, C.POINTER(C.c_int32), C.c_int64, C.POINTER(C.c_int32)]
lib.b200_pc_slots_of.restype = C.c_int64
lib.b200_pc_slots_of.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.c_int64]
lib.b200_pc_defrag.restype = C.c_int64
lib.b200_pc_defrag.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.c_int64, C.c_int64]
lib.b200_pc_restore_seq.restype = C.c_int64
lib.b200_pc_restore_seq.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.POINTER(C.c_int32)]
lib.b200_pc_context_shift.restype = C.c_int32
lib.b200_pc_context_shift.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.c_int64, C.POINTER(C.c_int32), C.c_int64,
                                      C.POINTER(C.c_int32), C.c_int64, C.POINTER(C.c_int32), C.c_int64, C.POINTER(C.c_int64)]
lib.b200_kv_state_seq_size.restype = C.c_int64
lib.b200_kv_state_seq_size.argtypes = [C.c_int64, C.c_int32, C.c_int64, C.c_int64]
lib.b200_kv_state_seq_write.restype = C.c_int64
lib.b200_kv_state_seq_write.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int64, C.c_int32, C.c_int64,
                                        C.c_void_p, C.c_void_p]
lib.b200_kv_state_seq_parse.restype = C.c_int32
lib.b200_kv_state_seq_parse.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int64, C.c_int32, C.c_int64,
                                        C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
lib.b200_kv_state_seq_positions.restype = C.c_int64
lib.b200_kv_state_seq_positions.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_int32), C.c_int64]
lib.b200_pc_used_cells.restype = C.c_int64
lib.b200_pc_used_cells.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.c_int64]
lib.b200_kv_state_all_write.restype = C.c_int64
lib.b200_kv_state_all_write.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_int32, C.c_int64, C.c_int32, C.c_int64,
                                        C.c_void_p, C.c_void_p]
lib.b200_kv_state_all_restore.restype = C.c_int64
lib.b200_kv_state_all_restore.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int64, C.c_int32, C.c_int64,
                                          C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_int64,
                                          C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
lib.b200_kvc_create.restype = C.c_void_p
lib.b200_kvc_create.argtypes = [C.c_int64, C.c_int32]
lib.b200_kvc_destroy.restype = None
lib.b200_kvc_destroy.argtypes = [C.c_void_p]
for _n in ("clear", "defrag"):
    getattr(lib, "b200_kvc_" + _n).restype = None
    getattr(lib, "b200_kvc_" + _n).argtypes = [C.c_void_p]
lib.b200_kvc_seq_rm.restype = C.c_int32
lib.b200_kvc_seq_rm.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32]
lib.b200_kvc_seq_cp.restype = None
lib.b200_kvc_seq_cp.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32]
lib.b200_kvc_seq_keep.restype = None
lib.b200_kvc_seq_keep.argtypes = [C.c_void_p, C.c_int32]
lib.b200_kvc_seq_add.restype = None
lib.b200_kvc_seq_add.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32]
lib.b200_kvc_seq_div.restype = None
lib.b200_kvc_seq_div.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32]
lib.b200_kvc_seq_pos_max.restype = C.c_int32
lib.b200_kvc_seq_pos_max.argtypes = [C.c_void_p, C.c_int32]
for _n in ("n_tokens", "used_cells", "unsupported"):
    getattr(lib, "b200_kvc_" + _n).restype = C.c_int32
    getattr(lib, "b200_kvc_" + _n).argtypes = [C.c_void_p]
lib.b200_kvc_slots_of.restype = C.c_int64
lib.b200_kvc_slots_of.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.c_int64]
lib.b200_kvc_find_slot.restype = C.c_int32
lib.b200_kvc_find_slot.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                   C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int64),
                                   C.POINTER(C.c_int32), C.c_int64, C.POINTER(C.c_int64)]
lib.b200_kvc_update.restype = C.c_int32
lib.b200_kvc_update.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
lib.b200_kvc_update_groups.restype = C.c_int64
lib.b200_kvc_update_groups.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.c_int64]
lib.b200_shard_sequences.restype = C.c_int32
lib.b200_shard_sequences.argtypes = [C.POINTER(C.c_int32), C.c_int64, C.c_int32, C.POINTER(C.c_int32)]
lib.b200_head_shard.restype = C.c_int32
lib.b200_head_shard.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_int32)]
STATE_ERRORS = {-1: "truncated stream", -2: "invalid seq_id-agnostic kv cell", -3: "positions are not 0..n-1",
                -4: "mismatched layer count", -5: "incompatible V transposition", -6: "mismatched key/value type",
                -7: "mismatched key/value row size", -8: "not enough cells in kv cache", -9: "invalid seq_id"}


class B200Error(RuntimeError):
    pass


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise B200Error(f"{what}: status {rc}: {lib.b200_last_error().decode(errors='replace')}")


def row_bytes(type_id: int, n: int) -> int:
    """ggml_row_size for the KV types (ggml/src/ggml.c:1169)"""
    if type_id == TYPE_F32:
        return n * 4
    if type_id == TYPE_F16:
        return n * 2
    if type_id == TYPE_Q8_0:
        return n // 32 * 34
    if type_id == TYPE_Q4_0:
        return n // 32 * 18
    raise ValueError(f"unsupported type {type_id}")


from .driver import *  # noqa: E402,F401,F403  (torch-tensor helpers and host-class wrappers)
from .sharding import *  # noqa: E402,F401,F403  (multi-GPU partitioning + the one all-gather)
