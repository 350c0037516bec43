/*
 * b200_host.h — C ABI over the host-side block managers of libb200attn.so.
 *
 * C++ callers include the class headers directly (llama.cpp_test_b200/host/src/ headers: BlockManager,
 * BlockKVCache, KVCacheManager — the fork's own interfaces, reference tests/test_block_manager.cpp:3,
 * tests/test_block_kv_cache.cpp:17 — and PagedKVCache, the block-table form of llama_kv_cache_unified,
 * reference src/llama-kv-cache.h:63-178).  This header is the same surface for FFI callers (ctypes tests,
 * C hosts).  Integer results (block ids, slots, tables, ref-counts) are exact; nothing here touches the GPU.
 */
#ifndef B200_HOST_H
#define B200_HOST_H

#include "b200_paged_attn.h"

#ifdef __cplusplus
extern "C" {
#endif

/* fork BlockManager (first-fit lowest index, evict lowest allocated, -1 when impossible) */
typedef struct b200_block_manager b200_block_manager;
B200_API b200_block_manager *b200_bm_create(int64_t num_blocks);
B200_API void    b200_bm_destroy(b200_block_manager *bm);
B200_API int32_t b200_bm_allocate(b200_block_manager *bm);
B200_API void    b200_bm_free(b200_block_manager *bm, int32_t index);
B200_API int32_t b200_bm_evict(b200_block_manager *bm);
B200_API int64_t b200_bm_status(const b200_block_manager *bm, uint8_t *out, int64_t cap);   /* returns num_blocks */
B200_API void    b200_fork_trace(int enabled);           /* the fork's std::cerr trace lines on/off */

/* PagedKVCache */
typedef struct b200_paged_cache b200_paged_cache;
B200_API b200_paged_cache *b200_pc_create(int64_t n_blocks, int32_t block_size);
B200_API void    b200_pc_destroy(b200_paged_cache *pc);
/* returns number of CoW pairs written (>= 0), -1 if out of blocks, -2 if cow_cap too small.
 * slot_mapping receives n_new cells; cow_pairs receives [n][2] = (src_block, dst_block). */
B200_API int64_t b200_pc_append(b200_paged_cache *pc, int32_t seq, int64_t n_new, int32_t *slot_mapping,
                                int32_t *cow_pairs, int64_t cow_cap);
B200_API int32_t b200_pc_seq_cp(b200_paged_cache *pc, int32_t src, int32_t dst, int64_t p0, int64_t p1);   /* 1 ok / 0 refused */
B200_API int32_t b200_pc_seq_rm(b200_paged_cache *pc, int32_t seq, int64_t p0, int64_t p1);
B200_API void    b200_pc_seq_keep(b200_paged_cache *pc, int32_t seq);
B200_API void    b200_pc_clear(b200_paged_cache *pc);
B200_API int64_t b200_pc_seq_len(const b200_paged_cache *pc, int32_t seq);
B200_API int64_t b200_pc_n_free(const b200_paged_cache *pc);
B200_API int32_t b200_pc_ref_count(const b200_paged_cache *pc, int64_t block);
B200_API int32_t b200_pc_export(const b200_paged_cache *pc, const int32_t *seqs, int64_t n,
                                int32_t *block_table, int64_t stride, int32_t *seq_lens);   /* 1 ok / 0 stride too small */

#ifdef __cplusplus
}
#endif
#endif
