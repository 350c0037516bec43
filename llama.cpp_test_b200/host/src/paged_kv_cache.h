// paged_kv_cache.h — token-granular block manager for the device-resident paged KV cache.
//
// This is the bookkeeping half of the reference's llama_kv_cache_unified (src/llama-kv-cache.h:63-178:
// find_slot .cpp:453-672, seq_cp :244-293, seq_rm :173-242, seq_keep :295-325, clear) re-expressed over
// fixed-size blocks in the fork's vocabulary (BlockManager first-fit allocation), extended
// with what north_star asks for and the reference lacks: per-sequence block tables, reference counts and
// copy-on-write.  The reference shares prefix cells by inserting the new seq_id into each cell's set
// (seq_cp is metadata-only); here seq_cp shares whole blocks by bumping ref-counts, and the first append
// into a shared, partially filled tail block emits a (src,dst) copy pair for b200_block_copy.
//
// Cell numbering is the flat llama.cpp cache's: cell = block*block_size + offset, so slot_mapping entries
// index the same k_l/v_l tensors (src/llama-kv-cache.cpp:100-101) the ggml views address.
#pragma once
#include <cstdint>
#include <map>
#include <queue>
#include <utility>
#include <vector>

class PagedKVCache {
public:
    typedef int32_t seq_id;

    PagedKVCache(int64_t n_blocks, int block_size);

    int     block_size() const { return block_size_; }
    int64_t n_blocks()   const { return (int64_t) ref_.size(); }
    int64_t n_free()     const { return n_free_; }
    int     ref_count(int64_t block) const;

    // Reserve room for n_new tokens at the end of `seq` (created if absent).  Appends n_new cell indices to
    // slot_mapping and any copy-on-write (src_block, dst_block) pairs to cow_pairs.  All-or-nothing: on
    // exhaustion nothing changes and false is returned (llama_decode's "no KV slot", src/llama-context.cpp:1302).
    bool append(seq_id seq, int64_t n_new, std::vector<int32_t> & slot_mapping,
                std::vector<std::pair<int32_t, int32_t>> & cow_pairs);

    // One ubatch of llama_context::decode's KV section (src/llama-context.cpp:1291-1353: find_slot for the whole ubatch,
    // then llm_graph_input_attn_kv_unified::set_input, src/llama-graph.cpp:403-477, fills the [n_kv, n_tokens] mask with
    // 0 where the cell holds the token's sequence at a position <= the token's, -inf elsewhere) in the paged form:
    // token i of sequence seq[i] at position pos[i] gets its cell slot_mapping[i] and the query-row inputs q_seq[i] (the
    // row of its sequence in `seqs`, the order of first appearance = the row order of export_tables(seqs)) and
    // q_pos[i] = pos[i]; a row then sees positions [0, q_pos] of its own block-table row, which is exactly the set the
    // mask marks.  Sequences may interleave in the batch.  pos == nullptr: positions continue each sequence.
    // Returns 0; 1 when the blocks do not suffice (llama_decode's "could not find a KV slot", nothing changes);
    // -1 for positions that do not continue their sequence (a paged sequence has no holes), nothing changes.
    struct BatchInputs {
        std::vector<int32_t> slot_mapping, q_seq, q_pos;
        std::vector<seq_id>  seqs;
        std::vector<std::pair<int32_t, int32_t>> cow_pairs;
    };
    int prepare_batch(int64_t n_tokens, const seq_id * seq, const int32_t * pos, BatchInputs & out);
    // llama_kv_slot_restorer::restore (src/llama-kv-cache.h:192-237; llama_context::decode calls it when the graph of a ubatch
    // fails, src/llama-context.cpp:1322-1343): the tokens a prepare_batch placed are dropped again — every sequence of the
    // batch is cut back to the position of its first token in the batch, and a block that a copy-on-write of the batch gave a
    // sequence is swapped back for the shared original (the copy may never have run) and released.
    void undo_batch(const BatchInputs & in);

    // llama_kv_self_seq_cp semantics restricted to prefixes (p0 <= 0): dst becomes positions [0, p1) of src,
    // sharing blocks.  p1 < 0 means "to the end".  Returns false for p0 > 0 or unknown src.
    bool seq_cp(seq_id src, seq_id dst, int64_t p0, int64_t p1);
    // llama_kv_self_seq_rm: whole sequence (p0 <= 0, p1 < 0 or >= len) or tail truncation (p1 < 0 or >= len);
    // removing a middle range returns false like the recurrent cache does.  seq < 0 matches every sequence.
    bool seq_rm(seq_id seq, int64_t p0, int64_t p1);
    void seq_keep(seq_id seq);
    void clear();

    // Context shift: the reference callers' pair llama_kv_self_seq_rm(seq, p0, p1) + llama_kv_self_seq_add(seq, p1, -1,
    // -(p1 - p0)) (examples/main/main.cpp context shift; seq_rm src/llama-kv-cache.cpp:173-242, seq_add :327-383) as ONE
    // operation on the block table: positions [p0, p1) are dropped and the tokens behind them move down by p1 - p0.
    //   * p0 and p1 - p0 multiples of the block size: the dropped blocks leave the table, no row moves;
    //   * otherwise the surviving rows move: row_moves gets (src_cell, dst_cell) pairs in position order — gather all
    //     sources, then scatter (b200_kv_rows_gather / _scatter, K and V), after the copy-on-write copies;
    //   * every block that gets written (row moves, K rotation) and is shared is copied first: cow_pairs (src, dst) for
    //     b200_block_copy; its allocation happens before any block is freed, so no source is overwritten;
    //   * shift_cells gets the cell of every token whose position changed, in position order: their K rows need the
    //     rotation by -(p1 - p0) (b200_k_shift with shift[cell] = -(p1 - p0); the reference's cell delta, .h:31).
    // Returns false (nothing changes) for an unknown sequence, an empty / out-of-range interval, or too few free blocks.
    bool context_shift(seq_id seq, int64_t p0, int64_t p1, std::vector<std::pair<int32_t, int32_t>> & cow_pairs,
                       std::vector<std::pair<int32_t, int32_t>> & row_moves, std::vector<int32_t> & shift_cells);
    // upper bounds of what context_shift would append (for fixed-size callers); false when it would refuse
    bool context_shift_sizes(seq_id seq, int64_t p0, int64_t p1, int64_t & n_cow, int64_t & n_moves, int64_t & n_shift) const;

    bool    has_seq(seq_id seq) const { return seqs_.count(seq) != 0; }
    int64_t seq_len(seq_id seq) const;                 // tokens cached (seq_pos_max + 1), 0 if absent
    int64_t n_seqs() const { return (int64_t) seqs_.size(); }
    std::vector<seq_id> seq_ids() const { std::vector<seq_id> v; for (const auto & kv : seqs_) v.push_back(kv.first); return v; }
    const std::vector<int32_t> * blocks_of(seq_id seq) const;

    // Cell index of every cached position of `seq`, in position order (the gather list of a state save).
    bool slots_of(seq_id seq, std::vector<int32_t> & cells) const;

    // Every cell some sequence holds, in ascending cell order (the order llama_kv_cache_unified::state_write(io, -1) walks its
    // cells, src/llama-kv-cache.cpp:856-872), with its position and the sorted ids of the sequences that see it (a cell of a
    // shared block carries several; sharers agree on the position because K rows are rotated by it).
    struct UsedCell { int32_t cell; int32_t pos; std::vector<seq_id> ids; };
    void used_cells(std::vector<UsedCell> & out) const;

    // Compaction (the block-granular form of llama_kv_cache_unified::defrag_prepare, src/llama-kv-cache.cpp:711-846,
    // which moves cells from the end of the cache into holes at the front): used blocks with the highest indices move
    // into the lowest free indices until the used blocks are [0, n_used).  Appends (src_block, dst_block) pairs for
    // b200_block_copy — no block is both a source and a destination — and rewrites every block table that names a
    // moved block (shared blocks move once, ref-counts travel with them).  max_moves < 0: no limit
    // (the reference caps moves per graph, llama-kv-cache.cpp:724-729).  Returns the number of moves.
    int64_t defrag(std::vector<std::pair<int32_t, int32_t>> & moves, int64_t max_moves = -1);

    // Export rows of the device block table / seq_lens for the given sequences (row i <-> seqs[i]);
    // unused entries are filled with 0 (a valid block, never dereferenced past seq_len).
    // Returns false if a row would not fit in `stride` entries.
    bool export_tables(const seq_id * seqs, int64_t n, int32_t * block_table, int64_t stride, int32_t * seq_lens) const;

private:
    struct Seq { std::vector<int32_t> blocks; int64_t len = 0; };
    int32_t alloc_block();                  // lowest free index (BlockManager::allocateBlock policy), -1 if none
    void    unref(int32_t b);
    void    release(Seq & s, int64_t keep_tokens);

    int                  block_size_;
    std::vector<int32_t> ref_;
    std::priority_queue<int32_t, std::vector<int32_t>, std::greater<int32_t>> free_;
    int64_t              n_free_;
    std::map<seq_id, Seq> seqs_;
};
