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
/* One ubatch of llama_context::decode's KV section in the paged form (src/llama-context.cpp:1291-1353 find_slot +
 * src/llama-graph.cpp:403-477 mask fill): token i of sequence seq[i] at position pos[i] (pos NULL: positions continue
 * each sequence) gets slot_mapping[i] (its cell, for b200_kv_write), q_seq[i] (row of its sequence in seqs[], the
 * order of first appearance: export those rows with b200_pc_export) and q_pos[i].  seqs[] needs room for one entry
 * per distinct sequence (n_tokens always suffices), cow_pairs for one (src, dst) pair per distinct sequence.
 * 0 ok; 1 not enough free blocks (llama_decode's "no KV slot"); -1 positions do not continue their sequence or bad
 * arguments; -3 cow_cap too small.  Nothing changes unless 0 is returned. */
B200_API int32_t b200_pc_prepare_batch(b200_paged_cache *pc, int64_t n_tokens, const int32_t *seq, const int32_t *pos,
                                       int32_t *slot_mapping, int32_t *q_seq, int32_t *q_pos, int32_t *seqs, int64_t *n_seqs,
                                       int32_t *cow_pairs, int64_t cow_cap, int64_t *n_cow);
/* llama_kv_slot_restorer::restore (src/llama-kv-cache.h:192-237): drop the tokens a successful b200_pc_prepare_batch placed
 * (seq[] as passed to it, q_pos[] and cow_pairs[n_cow][2] as it returned) when the ubatch's graph failed
 * (src/llama-context.cpp:1322-1343).  A block a copy-on-write of that batch gave a sequence is swapped back for the shared
 * original and released: the device copy may never have run.  cow_pairs may be NULL when n_cow is 0. */
B200_API void    b200_pc_undo_batch(b200_paged_cache *pc, int64_t n_tokens, const int32_t *seq, const int32_t *q_pos,
                                    const int32_t *cow_pairs, int64_t n_cow);
B200_API int32_t b200_pc_seq_cp(b200_paged_cache *pc, int32_t src, int32_t dst, int64_t p0, int64_t p1);   /* 1 ok / 0 refused */
B200_API int32_t b200_pc_seq_rm(b200_paged_cache *pc, int32_t seq, int64_t p0, int64_t p1);
B200_API void    b200_pc_seq_keep(b200_paged_cache *pc, int32_t seq);
B200_API void    b200_pc_clear(b200_paged_cache *pc);
B200_API int64_t b200_pc_seq_len(const b200_paged_cache *pc, int32_t seq);
B200_API int64_t b200_pc_n_free(const b200_paged_cache *pc);
B200_API int32_t b200_pc_ref_count(const b200_paged_cache *pc, int64_t block);
B200_API int32_t b200_pc_export(const b200_paged_cache *pc, const int32_t *seqs, int64_t n,
                                int32_t *block_table, int64_t stride, int32_t *seq_lens);   /* 1 ok / 0 stride too small */

/* cell index of every cached position of `seq` in position order; returns the count, -1 for an unknown sequence,
 * -(count)-1 when cap is too small */
B200_API int64_t b200_pc_slots_of(const b200_paged_cache *pc, int32_t seq, int32_t *cells, int64_t cap);
/* block compaction (reference defrag_prepare, src/llama-kv-cache.cpp:711-846, at block granularity): writes up to
 * min(cap, max_moves) (src_block, dst_block) pairs for b200_block_copy and rewrites the block tables; returns the
 * number of moves.  max_moves < 0: as many as cap allows. */
B200_API int64_t b200_pc_defrag(b200_paged_cache *pc, int32_t *moves /* [cap][2] */, int64_t cap, int64_t max_moves);

/*
 * Context shift (reference callers: llama_kv_self_seq_rm(seq, p0, p1) then llama_kv_self_seq_add(seq, p1, -1, -(p1 - p0)),
 * src/llama-kv-cache.cpp:173-242, :327-383): positions [p0, p1) of `seq` are dropped, the tokens behind them move down.
 * On return run, in this order: b200_block_copy(cow_pairs), b200_kv_rows_gather(src cells) + b200_kv_rows_scatter(dst
 * cells) over row_moves for K and V, b200_k_shift over shift_cells with shift[cell] = -(p1 - p0).
 *   cow_pairs  [cow_cap][2] (src_block, dst_block)      row_moves [move_cap][2] (src_cell, dst_cell), position order
 *   shift_cells [shift_cap] cells whose K rows need re-rotation      counts[3] receives {n_cow, n_moves, n_shift}
 * Returns 0; -1 unknown sequence / bad interval; -2 not enough free blocks; -3 a capacity is too small (counts holds
 * what is needed).  Nothing changes unless 0 is returned.
 */
B200_API int32_t b200_pc_context_shift(b200_paged_cache *pc, int32_t seq, int64_t p0, int64_t p1,
                                       int32_t *cow_pairs, int64_t cow_cap, int32_t *row_moves, int64_t move_cap,
                                       int32_t *shift_cells, int64_t shift_cap, int64_t counts[3]);

/*
 * KV session state of ONE sequence in the reference's stream format (llama_kv_cache_unified::state_write / state_read
 * with seq_id >= 0 and flash-attention caches, src/llama-kv-cache.cpp:848-1240; users: llama_state_seq_get_data /
 * llama_state_seq_set_data, examples/save-load-state).  Rows are token-ordered: gather them from the blocks with
 * b200_kv_rows_gather, restore them with b200_kv_rows_scatter (b200_paged_attn.h).
 *   k_rows / v_rows  host [n_layer][n_cells][size_row] bytes
 *   type_k / type_v  ggml_type ids (== B200_TYPE_*), size_row = ggml_row_size(type, n_embd_gqa)
 */
enum {
    B200_STATE_ERR_TRUNCATED    = -1,   /* stream shorter than its own header says */
    B200_STATE_ERR_SEQ_AGNOSTIC = -2,   /* a cell carries seq ids (whole-cache stream, not a sequence stream)   :1024 */
    B200_STATE_ERR_POSITIONS    = -3,   /* positions are not 0..n-1 (outside the block-table model)                  */
    B200_STATE_ERR_LAYERS       = -4,   /* mismatched layer count                                               :1112 */
    B200_STATE_ERR_VTRANS       = -5,   /* incompatible V transposition                                         :1120 */
    B200_STATE_ERR_TYPE         = -6,   /* mismatched key / value type                                     :1132,:1160 */
    B200_STATE_ERR_ROW_SIZE     = -7,   /* mismatched key / value row size                                 :1141,:1169 */
    B200_STATE_ERR_NO_ROOM      = -8,   /* not enough cells in kv cache (whole-cache restore)                    :1043 */
    B200_STATE_ERR_SEQ_ID       = -9    /* invalid seq_id in a cell (whole-cache stream)                         :1067 */
};
B200_API int64_t b200_kv_state_seq_size(int64_t n_cells, int32_t n_layer, int64_t k_size_row, int64_t v_size_row);
/* returns bytes written; -(bytes needed) when out is NULL or cap too small; -1 on bad arguments */
B200_API int64_t b200_kv_state_seq_write(uint8_t *out, int64_t cap, int64_t n_cells, int32_t n_layer,
                                         int32_t type_k, int64_t k_size_row, int32_t type_v, int64_t v_size_row,
                                         const uint8_t *k_rows, const uint8_t *v_rows);
/* validates a stream against the cache's shape; returns 0 or a B200_STATE_ERR_*; k_off/v_off [n_layer] receive the
 * byte offsets of each layer's rows inside `data` (either may be NULL).  The stream's rows must hold exactly the positions
 * 0..n-1, in any order (the reference writes cells in cell-index order): b200_kv_state_seq_positions gives the position of
 * every stream row, and the restore scatters row i to the cell of position pos[i]. */
B200_API int64_t b200_kv_state_seq_positions(const uint8_t *data, int64_t size, int32_t *pos_out, int64_t cap);   /* n_cells; -(n)-1: cap too small */
B200_API int32_t b200_kv_state_seq_parse(const uint8_t *data, int64_t size, int32_t n_layer,
                                         int32_t type_k, int64_t k_size_row, int32_t type_v, int64_t v_size_row,
                                         int64_t *n_cells, int64_t *k_off, int64_t *v_off);
/*
 * The whole-cache stream: llama_kv_cache_unified::state_write(io, -1) / state_read(io, -1) (llama_state_get_data /
 * llama_state_set_data; src/llama-kv-cache.cpp:848-903, meta with sequence ids :905-923 and :1038-1087).  Every cell some
 * sequence holds is written once, in ascending cell order, with its position and the ids of all sequences that see it.
 *   b200_pc_used_cells   those cells in stream order: gather their rows with b200_kv_rows_gather; returns the count,
 *                        -(count)-1 when cap is too small
 *   b200_kv_state_all_write   frames the gathered rows ([n_layer][n_cells][size_row]); returns bytes written, -(bytes
 *                        needed) when out is NULL or cap too small, -1 on bad arguments
 *   b200_kv_state_all_restore validates the stream (same codes as the sequence parser, plus SEQ_ID / NO_ROOM), then
 *                        clears the cache and rebuilds every sequence: src_rows[i] (row index inside each layer's rows
 *                        of the stream) must be copied to cell dst_cells[i] (b200_kv_rows_scatter).  Whole blocks of a
 *                        prefix that two sequences hold in the same stream cells stay shared (ref-counted); other shared
 *                        cells are written once per sequence.  Returns the number of pairs; a B200_STATE_ERR_*;
 *                        -(pairs needed)-100 when cap is too small (nothing changed).  Nothing changes on a parse
 *                        error; on NO_ROOM the cache is left cleared, as in the reference.
 */
B200_API int64_t b200_pc_used_cells(const b200_paged_cache *pc, int32_t *cells, int64_t cap);
B200_API int64_t b200_kv_state_all_write(uint8_t *out, int64_t cap, const b200_paged_cache *pc, int32_t n_layer,
                                         int32_t type_k, int64_t k_size_row, int32_t type_v, int64_t v_size_row,
                                         const uint8_t *k_rows, const uint8_t *v_rows);
B200_API int64_t b200_kv_state_all_restore(b200_paged_cache *pc, const uint8_t *data, int64_t size, int32_t n_layer,
                                           int32_t type_k, int64_t k_size_row, int32_t type_v, int64_t v_size_row,
                                           int32_t *src_rows, int32_t *dst_cells, int64_t cap,
                                           int64_t *n_cells, int64_t *k_off, int64_t *v_off);

/* state_read_meta for one sequence: drop `seq`, reserve n_cells tokens, slot_mapping[n_cells] = their cells.
 * Returns n_cells, or -1 when the pool has no room (the sequence is gone either way, as in the reference). */
B200_API int64_t b200_pc_restore_seq(b200_paged_cache *pc, int32_t seq, int64_t n_cells, int32_t *slot_mapping);

/*
 * The reference's llama_kv_cache method surface (struct llama_kv_cache : llama_memory_i, src/llama-kv-cache.h:17-26,
 * src/llama-memory.h:8-23; callers llama_kv_self_* in src/llama-context.cpp) on the block manager, for FFI hosts — the C++
 * class is llama_kv_cache_paged (llama.cpp_test_b200/host/src/llama_kv_cache_paged.h), same names and argument order.
 * seq_rm of a middle range stays open until the matching seq_add(seq, p1, -1, -(p1 - p0)) closes it into a context shift;
 * the device work of shifts and defrags is handed out by b200_kvc_update (= kv_self_update) in execution order, one GROUP
 * per context shift closed since the last update (b200_kvc_update_groups gives each group's end in the three lists): for
 * every group in turn b200_block_copy(its cow_pairs), rows gather/scatter over its row_moves, b200_k_shift(its shift_cells,
 * shift_delta) — a group is finished before the next starts, because a later shift may reuse blocks an earlier one freed
 * or read rows it wrote — and after the last group b200_block_copy(defrag_moves).  Requests the block tables cannot express (seq_div, a shift without a removal, a removal
 * never closed) set b200_kvc_unsupported and make update / find_slot refuse.
 */
typedef struct b200_kv_cache b200_kv_cache;
B200_API b200_kv_cache *b200_kvc_create(int64_t n_blocks, int32_t block_size);
B200_API void    b200_kvc_destroy(b200_kv_cache *kv);
B200_API void    b200_kvc_clear(b200_kv_cache *kv);
B200_API void    b200_kvc_defrag(b200_kv_cache *kv);
B200_API int32_t b200_kvc_seq_rm(b200_kv_cache *kv, int32_t seq, int32_t p0, int32_t p1);
B200_API void    b200_kvc_seq_cp(b200_kv_cache *kv, int32_t src, int32_t dst, int32_t p0, int32_t p1);
B200_API void    b200_kvc_seq_keep(b200_kv_cache *kv, int32_t seq);
B200_API void    b200_kvc_seq_add(b200_kv_cache *kv, int32_t seq, int32_t p0, int32_t p1, int32_t delta);
B200_API void    b200_kvc_seq_div(b200_kv_cache *kv, int32_t seq, int32_t p0, int32_t p1, int32_t d);
B200_API int32_t b200_kvc_seq_pos_max(b200_kv_cache *kv, int32_t seq);
B200_API int32_t b200_kvc_n_tokens(const b200_kv_cache *kv);
B200_API int32_t b200_kvc_used_cells(const b200_kv_cache *kv);
B200_API int32_t b200_kvc_unsupported(const b200_kv_cache *kv);
B200_API int64_t b200_kvc_slots_of(const b200_kv_cache *kv, int32_t seq, int32_t *cells, int64_t cap);
/* find_slot(ubatch): as b200_pc_prepare_batch; 1 also while deferred work is pending (decode runs kv_self_update first) */
B200_API int32_t b200_kvc_find_slot(b200_kv_cache *kv, int32_t n_tokens, const int32_t *seq, const int32_t *pos,
                                    int32_t *slot_mapping, int32_t *q_seq, int32_t *q_pos, int32_t *seqs, int64_t *n_seqs,
                                    int32_t *cow_pairs, int64_t cow_cap, int64_t *n_cow);
/* kv_self_update.  caps[4] / counts[5] = {cow pairs, row moves, shift cells, defrag moves, (counts only) groups}: call with
 * caps == NULL to learn the counts (returns -3 if there is work, 1 if none), then with arrays that fit.  1 done, 0
 * unsupported request. */
B200_API int32_t b200_kvc_update(b200_kv_cache *kv, int32_t *cow_pairs, int32_t *row_moves, int32_t *shift_cells, int32_t *shift_delta,
                                 int32_t *defrag_moves, const int64_t caps[4], int64_t counts[5]);
/* between the sizing call and the consuming call of b200_kvc_update: ends[g][3] = the cumulative {cow pairs, row moves,
 * shift cells} at the end of group g.  Returns the number of groups, or -(n_groups) - 1 when cap is too small. */
B200_API int64_t b200_kvc_update_groups(const b200_kv_cache *kv, int64_t *ends, int64_t cap);

/*
 * Multi-GPU partitioning of the path (one process per GPU; no reference counterpart — its multi-GPU path splits matmul rows,
 * ggml/src/ggml-cuda/ggml-cuda.cu:598,1592).  Deterministic integer policy, so every rank computes the same answer alone.
 *   b200_shard_sequences   partition A (configs 1-3, 5): sequences -> ranks balanced by context length (longest first to
 *                          the lightest rank); rank_of[i] = owner of sequence i.  No data-path collective afterwards.
 *   b200_head_shard        partition B (config 4, 70B GQA): rank g holds kv heads [out[0], out[0] + out[1]) of every block
 *                          and computes query heads [out[2], out[2] + out[3]); the block table is replicated; the one
 *                          exchange of the outputs is b200_peer_exchange_ll (b200_paged_attn.h).  -2 / -3: heads do not divide.
 * The peer buffers of that exchange are set up with b200_peer_alloc / _export (64-byte handle) / _open; carrying the handles
 * between the processes is the host's own IPC (the Python driver uses one torch.distributed.all_gather_object).
 */
B200_API int32_t b200_shard_sequences(const int32_t *ctx_lens, int64_t n_seqs, int32_t world, int32_t *rank_of);
B200_API int32_t b200_head_shard(int32_t rank, int32_t world, int32_t n_head, int32_t n_head_kv, int32_t out[4]);

#ifdef __cplusplus
}
#endif
#endif
