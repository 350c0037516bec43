// ggml_entry.cu — entry points taking the reference's op node (include/b200_ggml_abi.h).
// Lowers a GGML_OP_FLASH_ATTN_EXT / GGML_OP_CPY node to the plain descriptors of b200_paged_attn.h,
// reading the node exactly as launch_fattn does (ggml/src/ggml-cuda/fattn-common.cuh:683-700, :782-815)
// and as ggml_cuda_cpy does (ggml/src/ggml-cuda/cpy.cu:510-584).
#include "common.cuh"
#include "b200_ggml_abi.h"

#include <cstring>

namespace b200 {

static_assert(sizeof(b200_ggml_tensor) == 336, "ggml_tensor mirror out of sync with ggml/include/ggml.h:578-610");

int launch_kv_write(const b200_kv_write_desc &d, cudaStream_t stream);
int launch_cpy_rows(const void *src, void *dst, int type_src, int type_dst, const int64_t ne[4], const int64_t snb[4],
                    const int64_t dnb[4], cudaStream_t stream);

static int64_t nelements(const b200_ggml_tensor *t) { return t->ne[0]*t->ne[1]*t->ne[2]*t->ne[3]; }

static bool is_contiguous(const b200_ggml_tensor *t) {
    if (!type_ok(t->type)) return false;
    const int64_t blck = (t->type == B200_TYPE_Q8_0 || t->type == B200_TYPE_Q4_0) ? QK : 1;
    if (t->ne[0] % blck) return false;
    size_t nb = (size_t) row_bytes(t->type, blck);
    if (t->nb[0] != nb) return false;
    nb = (size_t) row_bytes(t->type, t->ne[0]);
    for (int i = 1; i < B200_GGML_MAX_DIMS; ++i) {
        if (t->ne[i] != 1 && t->nb[i] != nb) return false;
        nb *= (size_t) t->ne[i];
    }
    return true;
}

static bool node_to_desc(const b200_ggml_tensor *dst, b200_attn_desc &d, b200_kv_write_desc *store = nullptr) {
    if (store) memset(store, 0, sizeof(*store));
    if (!dst || dst->op != B200_GGML_OP_FLASH_ATTN_EXT) { set_error("flash_attn_ext: not a FLASH_ATTN_EXT node"); return false; }
    const b200_ggml_tensor *Q = dst->src[0], *K = dst->src[1], *V = dst->src[2], *M = dst->src[3];
    if (!Q || !K || !V) { set_error("flash_attn_ext: missing q/k/v source"); return false; }
    // asserts of launch_fattn (fattn-common.cuh:691-700) and of the CPU op (ggml-cpu.c:12242-12261)
    if (Q->type != B200_TYPE_F32 || dst->type != B200_TYPE_F32) { set_error("flash_attn_ext: q and dst must be F32"); return false; }
    if (Q->ne[3] != 1 || K->ne[3] != 1 || V->ne[3] != 1) { set_error("flash_attn_ext: batch dim (ne3) must be 1"); return false; }
    if (Q->nb[0] != 4 || K->ne[0] != Q->ne[0] || V->ne[0] != Q->ne[0]) { set_error("flash_attn_ext: head size mismatch / q rows not contiguous"); return false; }
    if (M && !dst->src[4] && (M->type != B200_TYPE_F16 || M->ne[1] < Q->ne[1] || M->ne[0] != K->ne[1])) { set_error("flash_attn_ext: mask must be F16 [n_kv, >= n_q]"); return false; }
    if (dst->ne[0] != Q->ne[0] || dst->ne[1] != Q->ne[2] || dst->ne[2] != Q->ne[1] || !is_contiguous(dst)) { set_error("flash_attn_ext: dst must be contiguous [d, n_head, n_q]"); return false; }
    memset(&d, 0, sizeof(d));
    d.head_dim = (int32_t) Q->ne[0]; d.n_q = (int32_t) Q->ne[1]; d.n_head = (int32_t) Q->ne[2];
    d.n_head_kv = (int32_t) K->ne[2];
    d.type_k = K->type; d.type_v = V->type;
    d.q = Q->data; d.nbq1 = (int64_t) Q->nb[1]; d.nbq2 = (int64_t) Q->nb[2];
    d.k = K->data; d.nbk1 = (int64_t) K->nb[1]; d.nbk2 = (int64_t) K->nb[2];
    d.v = V->data; d.nbv1 = (int64_t) V->nb[1]; d.nbv2 = (int64_t) V->nb[2];
    d.dst = dst->data;
    memcpy(&d.scale,         &dst->op_params[0], 4);
    memcpy(&d.max_bias,      &dst->op_params[1], 4);
    memcpy(&d.logit_softcap, &dst->op_params[2], 4);
    d.n_kv = K->ne[1];
    const b200_ggml_tensor *BT = dst->src[4];
    if (BT) {                                                     // paged extension
        const b200_ggml_tensor *SL = dst->src[5], *QS = dst->src[6], *QP = dst->src[7];
        if (BT->type != B200_GGML_TYPE_I32 || !SL || SL->type != B200_GGML_TYPE_I32) { set_error("flash_attn_ext: block_table/seq_lens must be I32"); return false; }
        if ((QS && QS->type != B200_GGML_TYPE_I32) || (QP && QP->type != B200_GGML_TYPE_I32)) { set_error("flash_attn_ext: q_seq/q_pos must be I32"); return false; }
        // a block table says "row t sees positions [0, min(seq_len, q_pos + 1)) of its sequence" and nothing else: a node that also
        // carries a mask (non-causal, sliding window, ALiBi) would silently lose it.  In a paged node src[3] is instead the
        // I32 slot mapping of the fused KV store: src[8] / src[9] = this step's new K / V rows (F32 [n_head_kv*d, n_tokens],
        // contiguous), written to cell slot_mapping[i] of the caches src[1] / src[2] view BEFORE the attention reads them —
        // build_attn's two ggml_cpy nodes (src/llama-graph.cpp:1375-1397) and the op in one node, as b200_decode_step does
        if (M && M->type != B200_GGML_TYPE_I32) { set_error("flash_attn_ext: a paged node (src[4] block table) must not carry a mask (src[3])"); return false; }
        const b200_ggml_tensor *KN = dst->src[8], *VN = dst->src[9];
        if ((KN || VN) && !M) { set_error("flash_attn_ext: new K/V rows (src[8], src[9]) need the I32 slot mapping in src[3]"); return false; }
        if (M) {
            if (!KN && !VN) { set_error("flash_attn_ext: a slot mapping (src[3]) without new K/V rows (src[8], src[9])"); return false; }
            const int64_t n_tok = M->ne[0], row = (int64_t) d.n_head_kv*d.head_dim;
            for (const b200_ggml_tensor *X : {KN, VN}) {
                if (X && (X->type != B200_TYPE_F32 || !is_contiguous(X) || nelements(X) != n_tok*row)) {
                    set_error("flash_attn_ext: new K/V rows must be contiguous F32 with n_tokens x n_head_kv x head_dim elements"); return false;
                }
            }
            if (store) {
                store->k_src = KN ? static_cast<const float *>(KN->data) : nullptr;
                store->v_src = VN ? static_cast<const float *>(VN->data) : nullptr;
                store->src_nb1 = row*4;
                store->n_tokens = n_tok;
                store->slot_mapping = static_cast<const int32_t *>(M->data);
                store->k_cache = const_cast<void *>(d.k); store->v_cache = const_cast<void *>(d.v);
                store->type_k = d.type_k; store->type_v = d.type_v; store->n_head_kv = d.n_head_kv; store->head_dim = d.head_dim;
                store->nbk_token = d.nbk1; store->nbk_head = d.nbk2; store->nbv_token = d.nbv1; store->nbv_head = d.nbv2;
                store->n_cells = d.n_kv;
            }
        }
        d.block_table = static_cast<const int32_t *>(BT->data);
        d.bt_stride   = (int64_t) (BT->nb[1]/4);
        d.n_seqs      = (int32_t) BT->ne[1];
        d.seq_lens    = static_cast<const int32_t *>(SL->data);
        d.q_seq       = QS ? static_cast<const int32_t *>(QS->data) : nullptr;
        d.q_pos       = QP ? static_cast<const int32_t *>(QP->data) : nullptr;
        d.block_size  = dst->op_params[4];
        memcpy(&d.max_seq_len, &dst->op_params[5], 8);
    } else if (M) {
        d.mask = M->data; d.nb_mask1 = (int64_t) M->nb[1];
    }
    return true;
}

// element-wise strided copy between F32 / F16 tensors of any strides (test_cpy "not contiguous", ggml_cont of a permuted
// tensor): element i of the source in its logical order -> element i of the destination in its own
struct cpy_elem_args {
    const char *src; char *dst; int type_src, type_dst;
    int64_t n, sne[4], snb[4], dne[4], dnb[4];
};
__global__ void __launch_bounds__(256) cpy_elem_kernel(const __grid_constant__ cpy_elem_args a) {
    for (int64_t i = (int64_t) blockIdx.x*256 + threadIdx.x; i < a.n; i += (int64_t) gridDim.x*256) {
        int64_t r = i, so = 0, d_ = i, doff = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) { so += (r % a.sne[k])*a.snb[k]; r /= a.sne[k]; doff += (d_ % a.dne[k])*a.dnb[k]; d_ /= a.dne[k]; }
        const float v = a.type_src == B200_TYPE_F32 ? *reinterpret_cast<const float *>(a.src + so)
                                                    : __half2float(*reinterpret_cast<const __half *>(a.src + so));
        if (a.type_dst == B200_TYPE_F32) *reinterpret_cast<float *>(a.dst + doff) = v;
        else                             *reinterpret_cast<__half *>(a.dst + doff) = __float2half_rn(v);
    }
}
static bool cpy_elem_ok(const b200_ggml_tensor *s0, const b200_ggml_tensor *s1) {
    const auto fp = [](int t) { return t == B200_TYPE_F32 || t == B200_TYPE_F16; };
    return fp(s0->type) && fp(s1->type) && nelements(s0) == nelements(s1) && nelements(s0) > 0;
}

// the row-wise general copy (cpy_rows_kernel, kv_write.cu): which (src0, src1) pairs it takes
static bool cpy_rows_ok(const b200_ggml_tensor *s0, const b200_ggml_tensor *s1) {
    const auto kv_type = [](int t) { return t == B200_TYPE_F32 || t == B200_TYPE_F16 || t == B200_TYPE_Q8_0 || t == B200_TYPE_Q4_0; };
    const auto quant = [](int t) { return t == B200_TYPE_Q8_0 || t == B200_TYPE_Q4_0; };
    if (!kv_type(s0->type) || !kv_type(s1->type)) return false;
    if (nelements(s0) != nelements(s1) || nelements(s0) == 0) return false;
    const int64_t ne0 = s0->ne[0];
    if (ne0 % 4 || ne0 > 8192 || ((quant(s0->type) || quant(s1->type)) && ne0 % QK)) return false;
    if (quant(s0->type) && (s1->type != B200_TYPE_F32 || s0->nb[0] != (size_t) row_bytes(s0->type, QK))) return false;   // cpy.cu takes Q -> F32 only
    if (is_contiguous(s1)) return true;                            // row r of the source = the r-th run of ne0 elements
    if (s1->nb[0] != (size_t) (quant(s1->type) ? row_bytes(s1->type, QK) : row_bytes(s1->type, 1))) return false;
    for (int i = 0; i < B200_GGML_MAX_DIMS; ++i) if (s0->ne[i] != s1->ne[i]) return false;    // same shape, own strides
    return true;
}

__global__ void __launch_bounds__(256) copy16_kernel(const uint4 *src, uint4 *dst, int64_t n16) {
    for (int64_t i = (int64_t) blockIdx.x*256 + threadIdx.x; i < n16; i += (int64_t) gridDim.x*256) dst[i] = src[i];
}
__global__ void __launch_bounds__(256) copy1_kernel(const char *src, char *dst, int64_t n) {
    for (int64_t i = (int64_t) blockIdx.x*256 + threadIdx.x; i < n; i += (int64_t) gridDim.x*256) dst[i] = src[i];
}

} // namespace b200

using namespace b200;

extern "C" {

B200_API size_t b200_flash_attn_ext_workspace(const struct b200_ggml_tensor *dst) {
    b200_attn_desc d;
    if (!node_to_desc(dst, d)) return 0;
    return b200_paged_attention_workspace(&d);
}

B200_API int b200_flash_attn_ext(const struct b200_ggml_tensor *dst, void *workspace, size_t workspace_bytes, void *stream) {
    b200_attn_desc d;
    b200_kv_write_desc store;
    if (!node_to_desc(dst, d, &store)) return B200_ERR_INVALID;
    d.workspace = workspace; d.workspace_bytes = workspace_bytes;
    if (store.slot_mapping) {                                     // the fused KV store of a paged node, first
        if (!b200_supports_attention(&d)) { set_error("flash_attn_ext: unsupported shape/type"); return B200_ERR_UNSUPPORTED; }   // nothing is written for a node that cannot run
        return b200_store_attention(&store, &d, stream);
    }
    return b200_paged_attention(&d, stream);
}

B200_API int b200_cpy(const struct b200_ggml_tensor *src0, struct b200_ggml_tensor *src1, void *stream) {
    if (!src0 || !src1) { set_error("cpy: NULL tensor"); return B200_ERR_INVALID; }
    if (nelements(src0) != nelements(src1)) { set_error("cpy: element counts differ"); return B200_ERR_INVALID; }   // ggml.c:2953
    const int64_t n = nelements(src0);
    if (n == 0) return B200_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool fast = is_contiguous(src0) && is_contiguous(src1) && (src0->type == src1->type || src0->type == B200_TYPE_F32);
    if (!fast) {                                                   // strided / permuted source, F16 or quantised source: row by row
        if (cpy_elem_ok(src0, src1)) {
            cpy_elem_args e;
            e.src = static_cast<const char *>(src0->data); e.dst = static_cast<char *>(src1->data);
            e.type_src = src0->type; e.type_dst = src1->type; e.n = n;
            for (int i = 0; i < 4; ++i) { e.sne[i] = src0->ne[i]; e.snb[i] = (int64_t) src0->nb[i]; e.dne[i] = src1->ne[i]; e.dnb[i] = (int64_t) src1->nb[i]; }
            const int64_t blocks = (n + 255)/256;
            cpy_elem_kernel<<<(unsigned) (blocks < 148*16 ? blocks : 148*16), 256, 0, st>>>(e);
            count_launch();
            return check_launch("cpy_elem") ? B200_OK : B200_ERR_CUDA;
        }
        if (!cpy_rows_ok(src0, src1)) { set_error("cpy: this pair of tensors is not taken (types %d -> %d)", src0->type, src1->type); return B200_ERR_UNSUPPORTED; }
        int64_t ne[4], snb[4], dnb[4];
        for (int i = 0; i < 4; ++i) { ne[i] = src0->ne[i]; snb[i] = (int64_t) src0->nb[i]; dnb[i] = (int64_t) src1->nb[i]; }
        if (is_contiguous(src1)) {                                 // destination rows one after the other, whatever its own shape
            const int64_t rb = row_bytes(src1->type, ne[0]);
            dnb[1] = rb; dnb[2] = rb*ne[1]; dnb[3] = rb*ne[1]*ne[2];
        }
        return launch_cpy_rows(src0->data, src1->data, src0->type, src1->type, ne, snb, dnb, st);
    }
    if (src0->type == src1->type) {                                // block move: bytes
        const int64_t bytes = row_bytes(src0->type, n);
        const int grid = (int) ((bytes/16 + 255)/256 < 148*8 ? ((bytes/16 + 255)/256 > 0 ? (bytes/16 + 255)/256 : 1) : 148*8);
        if (((((uintptr_t) src0->data) | ((uintptr_t) src1->data) | (uintptr_t) bytes) & 15) == 0)
            copy16_kernel<<<grid, 256, 0, st>>>(static_cast<const uint4 *>(src0->data), static_cast<uint4 *>(src1->data), bytes/16);
        else
            copy1_kernel<<<grid, 256, 0, st>>>(static_cast<const char *>(src0->data), static_cast<char *>(src1->data), bytes);
        count_launch();
        return check_launch("cpy") ? B200_OK : B200_ERR_CUDA;
    }
    if (src0->type != B200_TYPE_F32) { set_error("cpy: source must be F32 or the destination type"); return B200_ERR_UNSUPPORTED; }
    // KV store: treat as rows of `row` elements written to consecutive cells of a one-head cache
    int64_t row = src0->ne[0];
    if (row % QK || row > 8192) { set_error("cpy: row of %lld elements not supported", (long long) row); return B200_ERR_UNSUPPORTED; }
    b200_kv_write_desc w;
    memset(&w, 0, sizeof(w));
    w.k_src = static_cast<const float *>(src0->data); w.src_nb1 = row*4;
    w.n_tokens = n/row;
    w.k_cache = src1->data; w.type_k = src1->type; w.type_v = src1->type;
    w.n_head_kv = 1; w.head_dim = (int32_t) row;
    w.nbk_token = row_bytes(src1->type, row); w.nbk_head = w.nbk_token;
    w.n_cells = w.n_tokens;
    return b200_kv_write(&w, stream);
}

B200_API int b200_supports_op(const struct b200_ggml_tensor *op) {
    if (!op) return 0;
    if (op->op == B200_GGML_OP_FLASH_ATTN_EXT) {
        b200_attn_desc d;
        if (!node_to_desc(op, d)) return 0;
        return b200_supports_attention(&d);
    }
    if (op->op == B200_GGML_OP_CPY || op->op == B200_GGML_OP_DUP || op->op == B200_GGML_OP_CONT) {
        const b200_ggml_tensor *s0 = op->src[0], *s1 = op->op == B200_GGML_OP_CPY ? op->src[1] : op;
        if (!s0 || !s1) return 0;
        if (is_contiguous(s0) && is_contiguous(s1)) {
            if (s0->type == s1->type) return 1;
            if (s0->type == B200_TYPE_F32) return s0->ne[0] % QK == 0 && s0->ne[0] <= 8192 ? 1 : 0;
        }
        return (cpy_elem_ok(s0, s1) || cpy_rows_ok(s0, s1)) ? 1 : 0;   // permuted / strided / F16 / quantised sources
    }
    return 0;
}

} // extern "C"
