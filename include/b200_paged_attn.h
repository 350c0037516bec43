/*
 * b200_paged_attn.h — C ABI of libb200attn.so: the B200-native (sm_100a) replacement for the
 * reference's paged-attention hot path.  Plain pointers and sizes only; every pointer named
 * "device" is a CUDA device pointer, `stream` is a cudaStream_t passed as void*.
 *
 * What each entry point replaces in the reference (paths relative to the reference root):
 *   b200_paged_attention   <- ggml_cuda_flash_attn_ext            ggml/src/ggml-cuda/fattn.cu:244
 *                             (launch_fattn, fattn-common.cuh:675-836; CPU twin
 *                              ggml_compute_forward_flash_attn_ext_f16, ggml-cpu/ggml-cpu.c:12219)
 *   b200_kv_write          <- ggml_cuda_cpy for the KV store       ggml/src/ggml-cuda/cpy.cu:510-584
 *                             (graph nodes built at src/llama-graph.cpp:1375-1397) and the fork's
 *                             BlockKVCache::writeToCache            
 *   b200_block_copy        <- per-move ggml_cpy / cudaMemcpyAsync  cpy.cu:544-547,
 *                             src/llama-context.cpp:653-711 (defrag moves) and the fork's
 *                             BlockKVCache::copyBlocks
 *   b200_flash_attn_ext / b200_cpy / b200_supports_op (b200_ggml_abi.h) take the ggml op node itself.
 *
 * All functions return B200_OK (0 == GGML_STATUS_SUCCESS) or a negative b200_status; they never
 * throw, never synchronise the stream and never allocate device memory, so they are CUDA-graph
 * capturable (SURVEY.md §8b "Threading").  There is no CPU fallback: without a CUDA device every
 * compute entry returns B200_ERR_CUDA.
 */
#ifndef B200_PAGED_ATTN_H
#define B200_PAGED_ATTN_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_API __attribute__((visibility("default")))

typedef enum {
    B200_OK              =  0,   /* GGML_STATUS_SUCCESS */
    B200_ERR_FAILED      = -1,   /* GGML_STATUS_FAILED */
    B200_ERR_ALLOC       = -2,   /* GGML_STATUS_ALLOC_FAILED: workspace too small */
    B200_ERR_UNSUPPORTED = -3,   /* shape / type outside b200_supports_* */
    B200_ERR_INVALID     = -4,   /* malformed descriptor (NULL pointer, misaligned, ...) */
    B200_ERR_CUDA        = -5    /* CUDA runtime error; see b200_last_error() */
} b200_status;

/* ggml type ids accepted for K/V (ggml/include/ggml.h:351-389) */
enum { B200_TYPE_F32 = 0, B200_TYPE_F16 = 1, B200_TYPE_Q4_0 = 2, B200_TYPE_Q8_0 = 8 };

/* kernel selection, for tests and ablation; AUTO in production */
enum { B200_KERNEL_AUTO = 0, B200_KERNEL_GENERIC = 1, B200_KERNEL_STREAMK = 2, B200_KERNEL_PREFILL = 3,
       B200_KERNEL_STREAMK_DQ = 4 /* REMOVED in round 2 (a tcgen05 formulation of Q8_0 decode, 20-25 % slower than the
                                     mma.sync kernel and never chosen by AUTO): asking for it is B200_ERR_UNSUPPORTED */ };

/*
 * Attention descriptor.  Tensor meaning follows ggml_flash_attn_ext (ggml/include/ggml.h:1794-1807):
 *   q   F32  [d, n_q, n_head]    element (i, t, h) at q + t*nbq1 + h*nbq2 (+ i*4)
 *   k,v type [d, n_cells, n_head_kv]  cell c, head g row at k + c*nbk1 + g*nbk2   (flat llama.cpp cache)
 *   dst F32  [d, n_head, n_q]    contiguous
 * Two addressing modes:
 *   dense  (block_table == NULL): cells [0, n_kv) all scanned, optional F16 mask [n_kv, >= n_q]
 *          with -inf hiding cells — identical semantics to the reference op.
 *   paged  (block_table != NULL): query row t belongs to sequence q_seq[t] (or t when NULL) and sees
 *          positions [0, min(seq_lens[s], q_pos[t]+1)); position p lives in cell
 *          block_table[s*bt_stride + p/block_size]*block_size + p%block_size.  No mask is read; with max_bias > 0
 *          the ALiBi bias the reference's F16 mask would carry, -|p - pos(query)|, is derived from p.
 * Shapes: head_dim 128 runs on the tensor-core kernels (stream-K decode for any GQA ratio up to 8 or a multiple of 8,
 * the tcgen05 tile kernel for prefill chunks); head_dim 64 / 80 / 96 / 112 / 256, other GQA ratios and ALiBi through a
 * block table run on the CUDA-core kernel (80 and 112: F16 / F32 rows only).  b200_supports_attention() says whether
 * a descriptor is taken at all.
 */
typedef struct b200_attn_desc {
    int32_t n_q, n_head, n_head_kv, head_dim;
    int32_t type_k, type_v;
    const void *q;   int64_t nbq1, nbq2;      /* device */
    const void *k;   int64_t nbk1, nbk2;      /* device */
    const void *v;   int64_t nbv1, nbv2;      /* device */
    void       *dst;                          /* device */
    float scale, max_bias, logit_softcap;

    int64_t     n_kv;                         /* dense: cells to scan; paged: cells in the pool (bounds) */
    const void *mask; int64_t nb_mask1;       /* device F16, dense mode only, may be NULL */

    const int32_t *block_table; int64_t bt_stride;   /* device */
    const int32_t *seq_lens;                          /* device [n_seqs] */
    int32_t        n_seqs, block_size;
    const int32_t *q_seq;                             /* device [n_q] or NULL */
    const int32_t *q_pos;                             /* device [n_q] or NULL */
    int64_t        max_seq_len;   /* host-known upper bound on seq_lens (0: bt_stride*block_size) */
    int64_t        total_tokens;  /* host-known sum of seq_lens, a scheduling hint only (0: unknown) */

    void   *workspace; size_t workspace_bytes;        /* device scratch, see b200_paged_attention_workspace */
    int32_t kernel;                                   /* B200_KERNEL_* */
    int32_t flags;                                    /* B200_ATTN_* */
} b200_attn_desc;

/*
 * B200_ATTN_REUSE_PLAN: the device-side schedule (tile counts / prefix sums / row tiles, computed from seq_lens,
 * q_seq, q_pos by a one-CTA plan kernel) of the PREVIOUS attention call on this stream with this workspace is still
 * valid — same n_q, head counts, kernel, and unchanged seq_lens / q_seq / q_pos contents — so the plan launch is
 * skipped.  That is the situation of layers 1..L-1 of one decode/prefill step (the reference rebuilds nothing per
 * layer either: one mask per graph, src/llama-graph.cpp:403-477).  b200_decode_step sets it by itself where the
 * descriptors of consecutive layers differ only in q / k / v / dst.  Kernels without a plan ignore it.
 */
enum { B200_ATTN_REUSE_PLAN = 1 };

B200_API size_t b200_paged_attention_workspace(const b200_attn_desc *d);
B200_API int    b200_paged_attention(const b200_attn_desc *d, void *stream);
B200_API int    b200_supports_attention(const b200_attn_desc *d);   /* 1 / 0 */

/*
 * Slot-mapped KV write with fused F32 -> {F16,Q8_0,Q4_0,F32} conversion.  K and V in one launch.
 *   k_src,v_src F32 [n_tokens][n_head_kv*head_dim], row strides src_nb1 bytes (either may be NULL)
 *   slot_mapping device int32 [n_tokens]: destination cell; negative or >= n_cells -> token skipped
 *   cache cell c, head g at cache + c*nb_token + g*nb_head
 * Quantisation is bit-exact with quantize_row_q8_0_ref / quantize_row_q4_0_ref
 * (ggml/src/ggml-quants.c:194-216, :31-66) and ggml_fp32_to_fp16_row (ggml/src/ggml.c:387).
 */
typedef struct b200_kv_write_desc {
    const float *k_src; const float *v_src; int64_t src_nb1;
    int64_t n_tokens;
    const int32_t *slot_mapping;
    void *k_cache; void *v_cache;
    int32_t type_k, type_v, n_head_kv, head_dim;
    int64_t nbk_token, nbk_head, nbv_token, nbv_head;
    int64_t n_cells;
} b200_kv_write_desc;

B200_API int b200_kv_write(const b200_kv_write_desc *d, void *stream);

/*
 * Batched block copy (copy-on-write, defrag moves, swap-in): for every cache in caches[0..n_caches)
 * (HOST array of device base pointers, e.g. k_l[il], v_l[il] of all layers) copy block pairs[i][0] to
 * pairs[i][1]; a block is block_bytes contiguous bytes (multiple of 16).  One launch.  No block may
 * be both a source and a destination in one call.
 */
B200_API int b200_block_copy(void *const *caches, int32_t n_caches, int64_t block_bytes, int64_t n_blocks,
                             const int32_t *pairs /* device [n_pairs][2] */, int64_t n_pairs, void *stream);

/*
 * Token-ordered rows <-> scattered cells, for KV session state on the paged layout.  The reference reads / writes
 * whole cell ranges of k_l[il] / v_l[il] with ggml_backend_tensor_get / _set (io.write_tensor,
 * src/llama-kv-cache.cpp:940-946; state_read_data :1147-1150); with blocks the rows of a sequence are scattered, so:
 *   gather : staging[c][i][0..row_bytes) = caches[c] + slots[i]*nb_token   for every cache c (all layers in one launch)
 *   scatter: the inverse
 * caches is a HOST array of device base pointers; slots is a DEVICE int32 [n_rows] (PagedKVCache::slots_of /
 * b200_pc_restore_seq); staging is DEVICE [n_caches][n_rows][row_bytes].  row_bytes, nb_token, pointers: multiples
 * of 16.  Rows whose slot is negative or >= n_cells are skipped.
 */
B200_API int b200_kv_rows_gather(void *const *caches, int32_t n_caches, int64_t nb_token, int64_t row_bytes, int64_t n_cells,
                                 const int32_t *slots, int64_t n_rows, void *staging, void *stream);
B200_API int b200_kv_rows_scatter(void *const *caches, int32_t n_caches, int64_t nb_token, int64_t row_bytes, int64_t n_cells,
                                  const int32_t *slots, int64_t n_rows, const void *staging, void *stream);

/*
 * K-shift: re-rotate cached K rows in place after their positions moved (context shift, llama_kv_self_seq_add /
 * seq_div).  Replaces the graph llama_context::build_kv_self_shift builds (src/llama-context.cpp:524-571; per layer
 * build_rope_shift :447-495: ggml_rope_ext_inplace on the F16 view [head_dim, n_head_kv, kv_size], and for a quantised
 * cache ggml_cast -> F32, rope, ggml_cpy back) with ONE launch over all layers; arithmetic is
 * ggml_compute_forward_rope_f32 / _f16 (ggml/src/ggml-cpu/ggml-cpu.c:10669-10853, :10855-11037; rope_yarn :10578-10593,
 * ggml_rope_cache_init :10595-10611) and the quantisers of b200_kv_write.
 *   caches        HOST array of n_caches device K bases (k_l[il]); cell c, head g at base + c*nb_token + g*nb_head
 *   shift         DEVICE int32, indexed by CELL: the rope position of that cell = the delta its position moved by
 *                 (llm_graph_input_k_shift::set_input, llama-context.cpp:507-520)
 *   cells         DEVICE int32 [n_rows] cells to rotate, or NULL: every cell in [0, n_cells) (n_rows ignored), which
 *                 is what the reference does (cells with shift 0 come back unchanged up to re-quantisation)
 *   n_rot, mode   ggml_rope_ext's n_dims and mode: 0 = GGML_ROPE_TYPE_NORM (pairs 2i, 2i+1), 2 = NEOX (pairs i, i + n_rot/2)
 *   freq_factors  DEVICE float [n_rot/2] or NULL (rope_freqs / long-short factors)
 * Cells outside [0, n_cells) are skipped.  F32 / F16 / Q8_0 / Q4_0 caches; head_dim and n_rot even, n_rot <= head_dim,
 * head_dim % 32 == 0 for the quantised types.
 */
typedef struct b200_k_shift_desc {
    void *const *caches; int32_t n_caches;
    int32_t type_k, n_head_kv, head_dim;
    int64_t nb_token, nb_head, n_cells;
    const int32_t *shift;
    const int32_t *cells; int64_t n_rows;
    int32_t n_rot, mode, n_ctx_orig;
    float freq_base, freq_scale, ext_factor, attn_factor, beta_fast, beta_slow;
    const float *freq_factors;
} b200_k_shift_desc;

B200_API int b200_k_shift(const b200_k_shift_desc *d, void *stream);

/*
 * One decode/prefill step over several layers in ONE call: for l in [0, n_layers): b200_kv_write(&writes[l]) (skipped
 * when writes == NULL) then b200_paged_attention(&attns[l]), all on `stream`.  This is the loop
 * llama_context::decode runs through ggml_backend_sched for the attention part of every layer
 * (src/llama-context.cpp:1291-1353); batching it removes the per-call host overhead of an FFI host.
 * Stops at the first failing layer and returns its status.
 *
 * Each layer is b200_store_attention(&writes[l], &attns[l]) below, which the results are identical to "store, then
 * attention" of: when the layer is a decode step on the stream-K kernel (n_q == writes[l].n_tokens <= n_seqs, ONE new
 * token per sequence) the attention kernel is started UNDER the store and waits for it only in front of the one tile per
 * sequence the store writes into.  The device checks first (plan kernel) that row r is the row of sequence r (q_seq NULL
 * or the identity) and that stored row r goes to the cell of that row's last visible position — what build_attn's
 * k_cur / v_cur copies do for a decode ubatch, src/llama-graph.cpp:1375-1397 — and falls back to waiting in front of
 * everything when any row does not.
 */
B200_API int b200_decode_step(const b200_kv_write_desc *writes, const b200_attn_desc *attns, int32_t n_layers, void *stream);
B200_API int b200_store_attention(const b200_kv_write_desc *store /* or NULL */, const b200_attn_desc *d, void *stream);

/*
 * KV-head sharding (one process per GPU, rank g computes the query heads of its kv heads): the one exchange step,
 * every rank's output slice to every rank, over NVLink peer memory with the library's own kernels instead of a
 * library collective (the payload is ~1 MiB per layer: the cost is latency).  The reference has no counterpart
 * (its multi-GPU path is cudaMemcpyPeerAsync row splits, ggml/src/ggml-cuda/ggml-cuda.cu:598,1592).
 *   b200_peer_alloc / _free      cudaMalloc'ed, zero-filled device memory that can be exported
 *   b200_peer_export / _open     CUDA IPC handle (64 bytes) of such memory / map a peer's handle into this process
 *   b200_peer_publish            gathered_p[slot][rank] = local for every peer p (own copy included), then
 *                                flag_p[rank] = epoch (system-scope release); epoch = ++state[0] on the device.
 *                                Two consecutive exchanges must name different slots (e.g. slot = layer index).
 *   b200_peer_wait               wait until flags[r] >= state[0] for every r (bounded spin; on timeout state[2] != 0)
 *   b200_peer_exchange           publish + wait in ONE launch (the CTA that raises the flags waits for the peers')
 *   b200_peer_exchange_ll        lowest latency: every 8-byte store carries 4 payload bytes + the epoch (no fence, no
 *                                separate flag); one launch pushes to every peer's LL buffer [n_slots][world][2*bytes]
 *                                and unpacks what the peers pushed into the LOCAL gathered buffer
 * gathered buffers are [n_slots][world][bytes]; flags uint32 [world]; state uint32 [4], all from b200_peer_alloc.
 * peer_bufs / peer_flags are HOST arrays of `world` device pointers (entry `rank` = this process's own memory).
 * Both launches are CUDA-graph capturable.
 */
B200_API void *b200_peer_alloc(size_t bytes);
B200_API int   b200_peer_free(void *p);
B200_API int   b200_peer_export(void *p, unsigned char handle[64]);
B200_API void *b200_peer_open(const unsigned char handle[64]);
B200_API int   b200_peer_close(void *p);
B200_API int   b200_peer_publish(const void *local, int64_t bytes, void *const *peer_bufs, void *const *peer_flags,
                                 void *state, int32_t slot, int32_t rank, int32_t world, void *stream);
B200_API int   b200_peer_wait(void *flags, void *state, int32_t world, void *stream);
B200_API int   b200_peer_exchange(const void *local, int64_t bytes, void *const *peer_bufs, void *const *peer_flags,
                                  void *state, int32_t slot, int32_t rank, int32_t world, void *stream);
B200_API int   b200_peer_exchange_ll(const void *local, int64_t bytes, void *const *peer_ll, void *gathered, void *state,
                                     int32_t slot, int32_t rank, int32_t world, void *stream);
B200_API int   b200_peer_read_state(const void *state, uint32_t out[4]);   /* synchronous D2H of the 4 state words */

/*
 * b200_paged_attention_static_ws — no counterpart in the reference.
 * b200_paged_attention's stream-K decode path for callers that cannot call cudaMalloc (code under CUDA-graph capture,
 * tight serving loops, harnesses that forbid it): d->workspace / d->workspace_bytes are IGNORED and the kernels' scratch
 * (device-side plan + partials of split rows) is a file-scope __device__ array of B200_STATIC_WS_BYTES that every GPU
 * loading the library holds until exit.  The call allocates nothing (no cudaMalloc / cudaFree / allocating library
 * routine).  Paged mode on the stream-K kernel only (d->kernel AUTO or STREAMK; whatever b200_supports_attention
 * accepts with kernel = B200_KERNEL_STREAMK); a call whose b200_paged_attention_workspace() exceeds
 * B200_STATIC_WS_BYTES returns B200_ERR_ALLOC with the two sizes in b200_last_error().  All calls on one device share
 * the one array (it keeps the last call's contents; nothing in it needs to start at zero): calls that overlap on the
 * GPU — two streams, two host threads — race, so order them on one stream or with events.
 */
#define B200_STATIC_WS_BYTES (16u << 20)     /* 16 MiB: config 2 (64 rows x 32 heads) needs 5.9 MiB, config 4 (32 x 64) 10.7 MiB */
B200_API size_t b200_static_workspace_bytes(void);
B200_API int    b200_paged_attention_static_ws(const b200_attn_desc *d, void *stream);

/* diagnostics */
B200_API const char *b200_last_error(void);
B200_API const char *b200_version(void);
B200_API int         b200_device_sm_count(void);
B200_API int64_t     b200_launch_count(void);      /* kernels launched by this library so far */

#ifdef __cplusplus
}
#endif
#endif
