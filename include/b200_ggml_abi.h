/*
 * b200_ggml_abi.h — entry points that take the reference's own op node (struct ggml_tensor *), so a
 * ggml backend can route GGML_OP_FLASH_ATTN_EXT / GGML_OP_CPY nodes here without any translation.
 *
 * They replace, with the same argument meaning:
 *   b200_flash_attn_ext  <- void ggml_cuda_flash_attn_ext(ggml_backend_cuda_context &, ggml_tensor * dst)
 *                           ggml/src/ggml-cuda/fattn.cu:244 (called from ggml-cuda.cu:2298)
 *   b200_cpy             <- void ggml_cuda_cpy(ggml_backend_cuda_context &, const ggml_tensor * src0, ggml_tensor * src1)
 *                           ggml/src/ggml-cuda/cpy.cu:510 (called from ggml-cuda.cu:2126)
 *   b200_supports_op     <- ggml_backend_cuda_device_supports_op, cases GGML_OP_FLASH_ATTN_EXT
 *                           (ggml-cuda.cu:3227-3245) and GGML_OP_CPY (:3080-3133)
 *
 * The node layout is the reference snapshot's `struct ggml_tensor` (ggml/include/ggml.h:578-610).  This
 * header does NOT include ggml.h (it is not available on the GPU box); it declares a layout-identical
 * mirror, and tests/test_abi.py compiles a probe against the real ggml.h whenever the reference tree is
 * present to prove sizeof/offsetof and the enum values below are identical.
 *
 * Paged extension (not in ggml; GGML_MAX_SRC is 10, ggml.h:220, so the node has room):
 *   dst->src[4] I32 [bt_stride, n_seqs]  block table          (absent => dense-mask mode, reference semantics)
 *   dst->src[5] I32 [n_seqs]             seq_lens
 *   dst->src[6] I32 [n_q]                q_seq   (optional)
 *   dst->src[7] I32 [n_q]                q_pos   (optional)
 *   dst->op_params[4] = block_size (int32), op_params[5..6] = max_seq_len (int64, 0 = unknown)
 *   a paged node carries no mask (rejected: the block table only says "causal prefix of the row's own sequence"); its
 *   src[3] may instead hold the fused KV store of the step, which replaces build_attn's two ggml_cpy nodes
 *   (src/llama-graph.cpp:1375-1397) and runs before the attention reads the cache:
 *   dst->src[3] I32 [n_tokens]           slot mapping: destination cell of each new row (out of range: skipped)
 *   dst->src[8] F32 [n_head_kv*d, n_tokens]  this step's K rows (k_cur), contiguous   } either may be absent
 *   dst->src[9] F32 [n_head_kv*d, n_tokens]  this step's V rows (v_cur), contiguous   }
 */
#ifndef B200_GGML_ABI_H
#define B200_GGML_ABI_H

#include "b200_paged_attn.h"

#ifdef __cplusplus
extern "C" {
#endif

#define B200_GGML_MAX_DIMS      4    /* GGML_MAX_DIMS      ggml.h:218 */
#define B200_GGML_MAX_SRC       10   /* GGML_MAX_SRC       ggml.h:220 */
#define B200_GGML_MAX_OP_PARAMS 64   /* GGML_MAX_OP_PARAMS ggml.h:222 */
#define B200_GGML_MAX_NAME      64   /* GGML_MAX_NAME      ggml.h:225 */

#define B200_GGML_OP_DUP            1    /* enum ggml_op, ggml.h:448-525 of the reference snapshot */
#define B200_GGML_OP_CPY            32
#define B200_GGML_OP_CONT           33
#define B200_GGML_OP_FLASH_ATTN_EXT 62
#define B200_GGML_TYPE_I32          26
#define B200_GGML_PREC_F32          1

/* layout-identical to struct ggml_tensor (336 bytes on LP64) */
struct b200_ggml_tensor {
    int32_t  type;                                   /* enum ggml_type            @0   */
    void    *buffer;                                 /* ggml_backend_buffer *     @8   */
    int64_t  ne[B200_GGML_MAX_DIMS];                 /*                           @16  */
    size_t   nb[B200_GGML_MAX_DIMS];                 /*                           @48  */
    int32_t  op;                                     /* enum ggml_op              @80  */
    int32_t  op_params[B200_GGML_MAX_OP_PARAMS/4];   /*                           @84  */
    int32_t  flags;                                  /*                           @148 */
    struct b200_ggml_tensor *src[B200_GGML_MAX_SRC]; /*                           @152 */
    struct b200_ggml_tensor *view_src;               /*                           @232 */
    size_t   view_offs;                              /*                           @240 */
    void    *data;                                   /*                           @248 */
    char     name[B200_GGML_MAX_NAME];               /*                           @256 */
    void    *extra;                                  /*                           @320 */
    char     padding[8];                             /*                           @328 */
};

/* `dst` is the GGML_OP_FLASH_ATTN_EXT node; all ->data pointers are device pointers (as in the CUDA backend).
 * workspace: device scratch of at least b200_flash_attn_ext_workspace(dst) bytes (the backend's pool). */
B200_API size_t b200_flash_attn_ext_workspace(const struct b200_ggml_tensor *dst);
B200_API int    b200_flash_attn_ext(const struct b200_ggml_tensor *dst, void *workspace, size_t workspace_bytes, void *stream);

/* GGML_OP_CPY: src0 -> src1 (src1 is the destination view, as in ggml_cuda_cpy).  The two uses on the hot path — F32 ->
 * {F32,F16,Q8_0,Q4_0} contiguous rows (KV store) and same-type contiguous copies (block moves) — have their own kernels; beyond
 * them, row by row: F32 / F16 sources with any strides (permuted, not contiguous along dim 0) to F32 / F16 / Q8_0 / Q4_0, and
 * Q8_0 / Q4_0 rows to F32 (the cast of build_rope_shift), for rows of a multiple of 4 (32 with a quantised side) up to 8192
 * elements, into a destination that is contiguous or has the source's shape with contiguous rows.  Everything else
 * (other types, odd row lengths) is declined by b200_supports_op and stays with the backend's own kernel. */
B200_API int    b200_cpy(const struct b200_ggml_tensor *src0, struct b200_ggml_tensor *src1, void *stream);

/* 1 if b200_flash_attn_ext / b200_cpy will take this node, else 0 (the backend keeps its own kernel). */
B200_API int    b200_supports_op(const struct b200_ggml_tensor *op);

#ifdef __cplusplus
}
#endif
#endif
