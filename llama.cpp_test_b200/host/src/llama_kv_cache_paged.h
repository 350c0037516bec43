// llama_kv_cache_paged.h — the reference's llama_kv_cache interface (src/llama-kv-cache.h:17-26 over llama_memory_i,
// src/llama-memory.h:8-23; the unified implementation's methods, llama-kv-cache.h:93-111) implemented on PagedKVCache.
//
// A maintainer derives it from the real llama_kv_cache (the method names, argument order and return types are the
// reference's; llama_pos / llama_seq_id are int32_t as in include/llama.h:60-63) and hands `update()`'s work lists to the
// C ABI the way llama_context::kv_self_update turns has_shift / do_defrag into graphs (src/llama-context.cpp:719-790).
//
// What maps and what does not:
//   clear / seq_cp / seq_keep / seq_pos_max / get_n_tokens / get_used_cells / defrag      exact (block granularity for defrag)
//   seq_rm(seq, p0, p1)   tail and whole-sequence removal: immediate.  A middle range is remembered: the reference's callers
//                         follow it with seq_add(seq, p1, end, p0 - p1) (context shift, examples/main/main.cpp) and the
//                         pair becomes ONE PagedKVCache::context_shift; a pending range that is not closed that way by
//                         the next update() / find_slot() fails there (positions with a hole cannot be a block table)
//   seq_add               only as the closing half of a context shift (delta == p0 - p1 of the pending range, range to
//                         the end of the sequence); anything else sets `unsupported` (update() then fails, the caller
//                         falls back as for get_can_shift() == false).  Several shifts may close between two update()
//                         calls (llama-server shifts several slots before one decode), of one sequence or of many: each
//                         is its own GROUP of device work and the groups are executed one after another (see pending_work)
//   seq_div               outside the block-table model: sets `unsupported`
//   find_slot(ubatch)     -> find_slot(seq, n_tokens): slot mapping + copy-on-write pairs (llama_decode's "no KV slot": false)
#pragma once
#include "paged_kv_cache.h"

#include <cstdint>
#include <limits>
#include <map>
#include <set>
#include <utility>
#include <vector>

typedef int32_t llama_pos;
typedef int32_t llama_seq_id;

class llama_kv_cache_paged {
public:
    // device work a call to update() hands out, in execution order.
    // One GROUP per context shift, in the order the shifts were closed: group g owns cow_pairs[groups[g-1].cow_end,
    // groups[g].cow_end), row_moves[.. move_end) and shift_cells[.. shift_end), and the consumer runs a group to the end
    // (copies, gather-all, scatter-all, rotation) before it starts the next one.  The order matters: a block one shift
    // frees may be the copy-on-write destination of the next while it is still the source of the first one's row moves,
    // and a sequence forked between two shifts shares blocks the first shift has yet to write.  (Running the flat lists
    // phase by phase is only right when there is a single group.)
    struct pending_work {
        struct group { size_t cow_end, move_end, shift_end; };
        std::vector<std::pair<int32_t, int32_t>> cow_pairs;    // b200_block_copy
        std::vector<std::pair<int32_t, int32_t>> row_moves;    // b200_kv_rows_gather (first) / _scatter (second), K and V
        std::vector<int32_t>                     shift_cells;  // b200_k_shift: cells ...
        std::vector<int32_t>                     shift_delta;  // ... and the delta of each (the reference's cell.delta)
        std::vector<group>                       groups;       // cumulative ends of each shift's share of the three lists above
        std::vector<std::pair<int32_t, int32_t>> defrag_moves; // b200_block_copy, after every group
        bool empty() const { return cow_pairs.empty() && row_moves.empty() && shift_cells.empty() && defrag_moves.empty(); }
    };

    llama_kv_cache_paged(int64_t n_blocks, int block_size) : cache_(n_blocks, block_size) {}

    // ---- llama_kv_cache
    int32_t  get_n_tokens()   const;                          // sum over sequences (shared prefixes count per sequence, as in the reference)
    uint32_t get_used_cells() const;                          // cells of allocated blocks
    bool     get_can_shift()  const { return true; }
    bool     get_can_edit()   const { return get_can_shift(); }

    // ---- llama_memory_i
    void clear();
    // llama_context::decode's closing check (src/llama-context.cpp:1468-1480): fragmentation = 1 - (used + padding)/n with
    // n = the padded extent of the used cells (cell_max, .cpp:679-689; get_padding 256 with flash attention, :674-677),
    // contexts below 2048 cells never defragment; above the threshold a defrag is queued for the next update().  In blocks:
    // n = cells up to the highest block in use, used = cells of the blocks in use.  Returns whether it queued one.
    bool defrag_if_fragmented(float defrag_thold, uint32_t padding = 256);
    void defrag() { do_defrag_ = true; }                      // deferred, like llama_kv_cache_unified::defrag (.cpp:149-153)
    bool seq_rm  (llama_seq_id seq_id,                              llama_pos p0, llama_pos p1);
    void seq_cp  (llama_seq_id seq_id_src, llama_seq_id seq_id_dst, llama_pos p0, llama_pos p1);
    void seq_keep(llama_seq_id seq_id);
    void seq_add (llama_seq_id seq_id,                              llama_pos p0, llama_pos p1, llama_pos delta);
    void seq_div (llama_seq_id seq_id,                              llama_pos p0, llama_pos p1, int d);
    llama_pos seq_pos_max(llama_seq_id seq_id);

    // ---- the decode loop
    // reserve n_tokens at the end of seq_id: their cells (slot mapping) and the copy-on-write pairs to run first.
    // false: no room (llama_decode returns 1) or an open middle removal (a request the block table cannot express)
    bool find_slot(llama_seq_id seq_id, int32_t n_tokens, std::vector<int32_t> & slot_mapping,
                   std::vector<std::pair<int32_t, int32_t>> & cow_pairs);
    // the reference's own shape, find_slot(const llama_ubatch &) (src/llama-kv-cache.cpp:453-672): every token of the ubatch,
    // sequences interleaved; also yields the op's per-row inputs.  Same statuses as PagedKVCache::prepare_batch (0 / 1 /
    // -1); 1 as well while deferred work is pending.
    int find_slot(int32_t n_tokens, const llama_seq_id * seq_id, const llama_pos * pos, PagedKVCache::BatchInputs & out);
    // kv_self_update: everything deferred since the last call.  false: a pending request could not be expressed
    // (`unsupported()` tells which kind) — the work lists are then empty and the cache is unchanged by this call.
    bool update(pending_work & out);
    bool has_pending() const { return !work_.empty() || do_defrag_ || !open_.empty(); }
    bool unsupported() const { return unsupported_; }

    PagedKVCache       & blocks()       { return cache_; }
    const PagedKVCache & blocks() const { return cache_; }

private:
    PagedKVCache cache_;
    pending_work work_;                                       // accumulated by closed context shifts
    std::map<llama_seq_id, std::pair<llama_pos, llama_pos>> open_;   // middle removals waiting for their seq_add
    bool do_defrag_ = false, unsupported_ = false;
};
