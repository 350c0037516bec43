// kv_state.cpp — the reference's KV-cache session stream for ONE sequence, written from / read into the paged layout.
//
// Stream layout = llama_kv_cache_unified::state_write(io, seq_id >= 0) with flash-attention caches (v_trans = 0),
// src/llama-kv-cache.cpp:848-971:
//     u32 cell_count
//     cell_count x { i32 pos, u32 n_seq_id = 0 }                               state_write_meta :905-923
//     u32 v_trans = 0, u32 n_layer                                             state_write_data :925-971
//     n_layer x { i32 k_type (ggml_type), u64 k_size_row, cell_count x k_size_row bytes }
//     n_layer x { i32 v_type,             u64 v_size_row, cell_count x v_size_row bytes }
// The unified cache writes the rows of a cell range straight out of k_l[il] (io.write_tensor); here the rows of a
// sequence are scattered over blocks, so the device side gathers them into token order first (b200_kv_rows_gather,
// csrc/kv_write.cu) and this file only frames them.  Reading mirrors state_read_meta / state_read_data
// (:1008-1050, :1105-1240): every mismatch the reference rejects is rejected with its own code, nothing is modified
// on failure of the parse, and the restore itself is seq_rm + append (= find_slot) + b200_kv_rows_scatter.
#include "b200_host.h"
#include "capi_types.h"

#include <algorithm>
#include <cstring>
#include <map>
#include <vector>

namespace {

struct writer {
    uint8_t *p; int64_t cap, n = 0;
    void put(const void *src, int64_t bytes) { if (p && n + bytes <= cap) memcpy(p + n, src, (size_t) bytes); n += bytes; }
    template <typename T> void val(T v) { put(&v, (int64_t) sizeof(T)); }
};

struct reader {
    const uint8_t *p; int64_t size, n = 0; bool ok = true;
    template <typename T> T val() { T v{}; if (n + (int64_t) sizeof(T) > size) { ok = false; return v; } memcpy(&v, p + n, sizeof(T)); n += (int64_t) sizeof(T); return v; }
    bool skip(int64_t bytes) { if (bytes < 0 || n + bytes > size) { ok = false; return false; } n += bytes; return true; }
};

} // namespace

extern "C" {

B200_API int64_t b200_kv_state_seq_size(int64_t n_cells, int32_t n_layer, int64_t k_size_row, int64_t v_size_row) {
    if (n_cells < 0 || n_layer < 0 || k_size_row < 0 || v_size_row < 0) return -1;
    return 4 + n_cells*8 + 8 + (int64_t) n_layer*(2*(4 + 8) + n_cells*(k_size_row + v_size_row));
}

B200_API int64_t b200_kv_state_seq_write(uint8_t *out, int64_t cap, int64_t n_cells, int32_t n_layer,
                                         int32_t type_k, int64_t k_size_row, int32_t type_v, int64_t v_size_row,
                                         const uint8_t *k_rows, const uint8_t *v_rows) {
    const int64_t need = b200_kv_state_seq_size(n_cells, n_layer, k_size_row, v_size_row);
    if (need < 0 || n_cells > 0xffffffffLL || (n_cells > 0 && n_layer > 0 && (!k_rows || !v_rows))) return -1;
    if (!out || cap < need) return -need;                     // caller sizes the buffer from the magnitude
    writer w{out, cap};
    w.val<uint32_t>((uint32_t) n_cells);
    for (int64_t i = 0; i < n_cells; ++i) { w.val<int32_t>((int32_t) i); w.val<uint32_t>(0); }   // pos = token index
    w.val<uint32_t>(0);                                       // v_trans: flash-attention caches are not transposed
    w.val<uint32_t>((uint32_t) n_layer);
    for (int32_t il = 0; il < n_layer; ++il) {
        w.val<int32_t>(type_k); w.val<uint64_t>((uint64_t) k_size_row);
        w.put(k_rows + (int64_t) il*n_cells*k_size_row, n_cells*k_size_row);
    }
    for (int32_t il = 0; il < n_layer; ++il) {
        w.val<int32_t>(type_v); w.val<uint64_t>((uint64_t) v_size_row);
        w.put(v_rows + (int64_t) il*n_cells*v_size_row, n_cells*v_size_row);
    }
    return w.n;
}

B200_API int32_t b200_kv_state_seq_parse(const uint8_t *data, int64_t size, int32_t n_layer,
                                         int32_t type_k, int64_t k_size_row, int32_t type_v, int64_t v_size_row,
                                         int64_t *n_cells_out, int64_t *k_off, int64_t *v_off) {
    if (!data || size < 0 || n_layer < 0) return B200_STATE_ERR_TRUNCATED;
    reader r{data, size};
    const uint32_t n_cells = r.val<uint32_t>();
    // the paged cache stores position p of a sequence at table index p, so a restored sequence must hold exactly the
    // positions 0..n-1 (the unified cache accepts any positions; a shifted or holed sequence is outside the block-table
    // model) — in ANY stream order: state_write(seq_id) walks cells in cell-index order, which need not be position order
    // (find_slot's head wraps, cells are reused after seq_rm, defrag moves them); b200_kv_state_seq_positions hands out the
    // order so that the restore scatters stream row i to the cell of position pos[i]
    std::vector<bool> seen;
    for (uint32_t i = 0; i < n_cells && r.ok; ++i) {
        const int32_t  pos      = r.val<int32_t>();
        const uint32_t n_seq_id = r.val<uint32_t>();
        if (!r.ok) break;
        if (n_seq_id != 0) return B200_STATE_ERR_SEQ_AGNOSTIC;           // "invalid seq_id-agnostic kv cell" :1024-1027
        if (seen.empty()) {
            if ((int64_t) n_cells*8 > size) return B200_STATE_ERR_TRUNCATED;     // a corrupt count must not size the bitmap
            seen.assign(n_cells, false);
        }
        if (pos < 0 || (uint32_t) pos >= n_cells || seen[(size_t) pos]) return B200_STATE_ERR_POSITIONS;
        seen[(size_t) pos] = true;
    }
    const uint32_t v_trans = r.val<uint32_t>();
    const uint32_t nl      = r.val<uint32_t>();
    if (!r.ok) return B200_STATE_ERR_TRUNCATED;
    if (nl != (uint32_t) n_layer) return B200_STATE_ERR_LAYERS;          // "mismatched layer count" :1112-1115
    if (v_trans != 0) return B200_STATE_ERR_VTRANS;                      // "incompatible V transposition" :1120-1123
    for (int32_t il = 0; il < n_layer; ++il) {
        const int32_t  t   = r.val<int32_t>();
        const uint64_t row = r.val<uint64_t>();
        if (!r.ok) return B200_STATE_ERR_TRUNCATED;
        if (t != type_k) return B200_STATE_ERR_TYPE;                     // "mismatched key type" :1132-1136
        if (row != (uint64_t) k_size_row) return B200_STATE_ERR_ROW_SIZE;// "mismatched key row size" :1141-1145
        if (k_off) k_off[il] = r.n;
        if (!r.skip((int64_t) n_cells*k_size_row)) return B200_STATE_ERR_TRUNCATED;
    }
    for (int32_t il = 0; il < n_layer; ++il) {
        const int32_t  t   = r.val<int32_t>();
        const uint64_t row = r.val<uint64_t>();
        if (!r.ok) return B200_STATE_ERR_TRUNCATED;
        if (t != type_v) return B200_STATE_ERR_TYPE;
        if (row != (uint64_t) v_size_row) return B200_STATE_ERR_ROW_SIZE;
        if (v_off) v_off[il] = r.n;
        if (!r.skip((int64_t) n_cells*v_size_row)) return B200_STATE_ERR_TRUNCATED;
    }
    if (n_cells_out) *n_cells_out = (int64_t) n_cells;
    return 0;
}

B200_API int64_t b200_kv_state_seq_positions(const uint8_t *data, int64_t size, int32_t *pos_out, int64_t cap) {
    if (!data || size < 4) return -1;
    reader r{data, size};
    const uint32_t n_cells = r.val<uint32_t>();
    if ((int64_t) n_cells*8 > size - 4) return -1;
    if ((int64_t) n_cells > cap || (n_cells > 0 && !pos_out)) return -(int64_t) n_cells - 1;
    for (uint32_t i = 0; i < n_cells; ++i) {
        pos_out[i] = r.val<int32_t>();
        (void) r.val<uint32_t>();
    }
    return r.ok ? (int64_t) n_cells : -1;
}

// ---- the whole-cache stream: state_write(io, -1) / state_read(io, -1), i.e. llama_state_get_data / llama_state_set_data
// (src/llama-kv-cache.cpp:848-903; meta with sequence ids :905-923, :1038-1087).  Same framing as the single-sequence
// stream, but every used cell is written once, in cell order, with the ids of all sequences that see it.

static int32_t parse_layers(reader & r, int64_t n_cells, int32_t n_layer, int32_t type_k, int64_t k_size_row,
                            int32_t type_v, int64_t v_size_row, int64_t *k_off, int64_t *v_off) {
    const uint32_t v_trans = r.val<uint32_t>();
    const uint32_t nl      = r.val<uint32_t>();
    if (!r.ok) return B200_STATE_ERR_TRUNCATED;
    if (nl != (uint32_t) n_layer) return B200_STATE_ERR_LAYERS;
    if (v_trans != 0) return B200_STATE_ERR_VTRANS;
    for (int pass = 0; pass < 2; ++pass) {
        const int32_t t_want = pass ? type_v : type_k;
        const int64_t row_want = pass ? v_size_row : k_size_row;
        int64_t *off = pass ? v_off : k_off;
        for (int32_t il = 0; il < n_layer; ++il) {
            const int32_t  t   = r.val<int32_t>();
            const uint64_t row = r.val<uint64_t>();
            if (!r.ok) return B200_STATE_ERR_TRUNCATED;
            if (t != t_want) return B200_STATE_ERR_TYPE;
            if (row != (uint64_t) row_want) return B200_STATE_ERR_ROW_SIZE;
            if (off) off[il] = r.n;
            if (!r.skip(n_cells*row_want)) return B200_STATE_ERR_TRUNCATED;
        }
    }
    return 0;
}

B200_API int64_t b200_pc_used_cells(const b200_paged_cache *pc, int32_t *cells, int64_t cap) {
    if (!pc) return -1;
    std::vector<PagedKVCache::UsedCell> u;
    pc->impl.used_cells(u);
    if ((int64_t) u.size() > cap) return -(int64_t) u.size() - 1;
    for (size_t i = 0; i < u.size(); ++i) cells[i] = u[i].cell;
    return (int64_t) u.size();
}

B200_API int64_t b200_kv_state_all_write(uint8_t *out, int64_t cap, const b200_paged_cache *pc, int32_t n_layer,
                                         int32_t type_k, int64_t k_size_row, int32_t type_v, int64_t v_size_row,
                                         const uint8_t *k_rows, const uint8_t *v_rows) {
    if (!pc || n_layer < 0 || k_size_row < 0 || v_size_row < 0) return -1;
    std::vector<PagedKVCache::UsedCell> u;
    pc->impl.used_cells(u);
    const int64_t n_cells = (int64_t) u.size();
    int64_t need = 4 + 8 + (int64_t) n_layer*(2*(4 + 8) + n_cells*(k_size_row + v_size_row));
    for (const auto & c : u) need += 8 + 4*(int64_t) c.ids.size();
    if (n_cells > 0 && n_layer > 0 && out && (!k_rows || !v_rows)) return -1;
    if (!out || cap < need) return -need;
    writer w{out, cap};
    w.val<uint32_t>((uint32_t) n_cells);
    for (const auto & c : u) {
        w.val<int32_t>(c.pos); w.val<uint32_t>((uint32_t) c.ids.size());
        for (int32_t id : c.ids) w.val<int32_t>(id);                     // std::set order in the reference: ascending
    }
    w.val<uint32_t>(0);
    w.val<uint32_t>((uint32_t) n_layer);
    for (int32_t il = 0; il < n_layer; ++il) {
        w.val<int32_t>(type_k); w.val<uint64_t>((uint64_t) k_size_row);
        w.put(k_rows + (int64_t) il*n_cells*k_size_row, n_cells*k_size_row);
    }
    for (int32_t il = 0; il < n_layer; ++il) {
        w.val<int32_t>(type_v); w.val<uint64_t>((uint64_t) v_size_row);
        w.put(v_rows + (int64_t) il*n_cells*v_size_row, n_cells*v_size_row);
    }
    return w.n;
}

B200_API int64_t b200_kv_state_all_restore(b200_paged_cache *pc, const uint8_t *data, int64_t size, int32_t n_layer,
                                           int32_t type_k, int64_t k_size_row, int32_t type_v, int64_t v_size_row,
                                           int32_t *src_rows, int32_t *dst_cells, int64_t cap,
                                           int64_t *n_cells_out, int64_t *k_off, int64_t *v_off) {
    if (!pc || !data || size < 0 || n_layer < 0) return B200_STATE_ERR_TRUNCATED;
    reader r{data, size};
    const uint32_t n_cells = r.val<uint32_t>();
    if (!r.ok) return B200_STATE_ERR_TRUNCATED;
    std::map<int32_t, std::vector<std::pair<int32_t, int32_t>>> seqs;    // seq -> (pos, stream row)
    for (uint32_t i = 0; i < n_cells; ++i) {
        const int32_t  pos      = r.val<int32_t>();
        const uint32_t n_seq_id = r.val<uint32_t>();
        if (!r.ok || (int64_t) n_seq_id*4 > size - r.n) return B200_STATE_ERR_TRUNCATED;
        for (uint32_t j = 0; j < n_seq_id; ++j) {
            const int32_t id = r.val<int32_t>();
            if (id < 0) return B200_STATE_ERR_SEQ_ID;                    // "invalid seq_id ... out of range" :1067-1071
            seqs[id].emplace_back(pos, (int32_t) i);
        }
    }
    const int32_t rc = parse_layers(r, (int64_t) n_cells, n_layer, type_k, k_size_row, type_v, v_size_row, k_off, v_off);
    if (rc != 0) return rc;
    int64_t n_pairs = 0;
    for (auto & kv : seqs) {                                              // a paged sequence is positions 0..n-1, each once
        std::sort(kv.second.begin(), kv.second.end());
        for (size_t p = 0; p < kv.second.size(); ++p) if (kv.second[p].first != (int32_t) p) return B200_STATE_ERR_POSITIONS;
        n_pairs += (int64_t) kv.second.size();
    }
    if (n_cells_out) *n_cells_out = (int64_t) n_cells;
    if (n_pairs > cap || (n_pairs > 0 && (!src_rows || !dst_cells))) return -n_pairs - 100;
    // everything checked: from here the cache changes (the reference clears first and clears again on failure, :1047, :889-893)
    PagedKVCache & c = pc->impl;
    c.clear();
    const int64_t bs = c.block_size();
    std::vector<int32_t> done;                                            // restored sequences, candidates for sharing
    int64_t n = 0;
    for (auto & kv : seqs) {
        const auto & mine = kv.second;
        // whole blocks of the longest prefix that an already restored sequence holds in the same stream rows are shared
        // (the stream says so: one cell, several ids); the rest gets cells of its own and a copy of the rows
        int32_t best = -1; int64_t best_len = 0;
        for (int32_t other : done) {
            const auto & theirs = seqs[other];
            int64_t m = 0;
            while (m < (int64_t) mine.size() && m < (int64_t) theirs.size() && mine[(size_t) m].second == theirs[(size_t) m].second) ++m;
            m = m/bs*bs;
            if (m > best_len) { best_len = m; best = other; }
        }
        if (best >= 0) c.seq_cp(best, kv.first, 0, best_len);
        std::vector<int32_t> slots; std::vector<std::pair<int32_t, int32_t>> cow;
        if (!c.append(kv.first, (int64_t) mine.size() - best_len, slots, cow) || !cow.empty()) {
            c.clear();
            return B200_STATE_ERR_NO_ROOM;                                // "not enough cells in kv cache" :1043-1046
        }
        for (size_t p = 0; p < slots.size(); ++p) { src_rows[n] = mine[(size_t) best_len + p].second; dst_cells[n] = slots[p]; ++n; }
        done.push_back(kv.first);
    }
    return n;
}

/* restore bookkeeping = state_read_meta's single-sequence branch (:1010-1036): seq_rm(dest, -1, -1), then find_slot
 * for cell_count tokens.  Like the reference, the old contents of `seq` are gone even when no room is found
 * (state_read then throws "failed to restore kv cache" after another seq_rm, :889-896). */
B200_API int64_t b200_pc_restore_seq(b200_paged_cache *pc, int32_t seq, int64_t n_cells, int32_t *slot_mapping) {
    if (!pc || n_cells < 0 || (n_cells > 0 && !slot_mapping)) return -1;
    PagedKVCache &c = pc->impl;
    c.seq_rm(seq, -1, -1);
    std::vector<int32_t> slots;
    std::vector<std::pair<int32_t, int32_t>> cow;
    if (!c.append(seq, n_cells, slots, cow)) return -1;     // "failed to find available cells in kv cache" :1033-1036
    for (size_t i = 0; i < slots.size(); ++i) slot_mapping[i] = slots[i];
    return (int64_t) slots.size();
}

B200_API int64_t b200_pc_slots_of(const b200_paged_cache *pc, int32_t seq, int32_t *cells, int64_t cap) {
    if (!pc) return -1;
    std::vector<int32_t> v;
    if (!pc->impl.slots_of(seq, v)) return -1;
    if ((int64_t) v.size() > cap) return -(int64_t) v.size() - 1;
    for (size_t i = 0; i < v.size(); ++i) cells[i] = v[i];
    return (int64_t) v.size();
}

B200_API int64_t b200_pc_defrag(b200_paged_cache *pc, int32_t *moves, int64_t cap, int64_t max_moves) {
    if (!pc || cap < 0) return -1;
    if (max_moves < 0 || max_moves > cap) max_moves = cap;    // never plan more than the caller can receive
    std::vector<std::pair<int32_t, int32_t>> mv;
    const int64_t n = pc->impl.defrag(mv, max_moves);
    for (int64_t i = 0; i < n; ++i) { moves[2*i] = mv[(size_t) i].first; moves[2*i + 1] = mv[(size_t) i].second; }
    return n;
}

B200_API int32_t b200_pc_context_shift(b200_paged_cache *pc, int32_t seq, int64_t p0, int64_t p1,
                                       int32_t *cow_pairs, int64_t cow_cap, int32_t *row_moves, int64_t move_cap,
                                       int32_t *shift_cells, int64_t shift_cap, int64_t counts[3]) {
    if (!pc || !counts) return -1;
    PagedKVCache &c = pc->impl;
    int64_t n_cow = 0, n_moves = 0, n_shift = 0;
    if (!c.context_shift_sizes(seq, p0, p1, n_cow, n_moves, n_shift)) {
        // tell "no room for the copy-on-write copies" from a bad request
        return (c.has_seq(seq) && p0 >= 0 && p1 > p0 && p1 <= c.seq_len(seq)) ? -2 : -1;
    }
    counts[0] = n_cow; counts[1] = n_moves; counts[2] = n_shift;
    if (n_cow > cow_cap || n_moves > move_cap || n_shift > shift_cap
        || (n_cow > 0 && !cow_pairs) || (n_moves > 0 && !row_moves) || (n_shift > 0 && !shift_cells)) return -3;
    std::vector<std::pair<int32_t, int32_t>> cow, mv;
    std::vector<int32_t> cells;
    if (!c.context_shift(seq, p0, p1, cow, mv, cells)) return -1;
    for (size_t i = 0; i < cow.size(); ++i) { cow_pairs[2*i] = cow[i].first; cow_pairs[2*i + 1] = cow[i].second; }
    for (size_t i = 0; i < mv.size(); ++i)  { row_moves[2*i] = mv[i].first;  row_moves[2*i + 1] = mv[i].second; }
    for (size_t i = 0; i < cells.size(); ++i) shift_cells[i] = cells[i];
    counts[0] = (int64_t) cow.size(); counts[1] = (int64_t) mv.size(); counts[2] = (int64_t) cells.size();
    return 0;
}

} // extern "C"
