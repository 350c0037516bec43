// llama_kv_cache_paged.cpp — see src/llama_kv_cache_paged.h
#include "llama_kv_cache_paged.h"

#include <new>

#include <algorithm>

int32_t llama_kv_cache_paged::get_n_tokens() const {
    // the reference sums cells[i].seq_id.size() (src/llama-kv-cache.cpp:126-134): a shared cell counts once per sequence
    int64_t n = 0;
    for (llama_seq_id s : cache_.seq_ids()) n += cache_.seq_len(s);
    return (int32_t) n;
}

uint32_t llama_kv_cache_paged::get_used_cells() const {
    return (uint32_t) ((cache_.n_blocks() - cache_.n_free())*cache_.block_size());
}

bool llama_kv_cache_paged::defrag_if_fragmented(float defrag_thold, uint32_t padding) {
    if (!(defrag_thold > 0.0f)) return false;
    int64_t top = -1;
    for (int64_t b = cache_.n_blocks() - 1; b >= 0 && top < 0; --b) if (cache_.ref_count(b) > 0) top = b;
    if (padding == 0) padding = 1;
    uint32_t n = (uint32_t) ((top + 1)*cache_.block_size());
    n = (n + padding - 1)/padding*padding;                                // kv_self->n = GGML_PAD(cell_max(), padding), llama-context.cpp:1311
    const float frag = n >= 2048 ? 1.0f - float(get_used_cells() + padding)/float(n) : 0.0f;
    if (!((frag > 0.0f ? frag : 0.0f) > defrag_thold)) return false;
    defrag();
    return true;
}

void llama_kv_cache_paged::clear() {
    cache_.clear();
    work_ = pending_work();
    open_.clear();
    do_defrag_ = false;
    unsupported_ = false;
}

bool llama_kv_cache_paged::seq_rm(llama_seq_id seq_id, llama_pos p0, llama_pos p1) {
    if (p0 < 0) p0 = 0;
    if (seq_id < 0) {                                        // every sequence: whole ranges only (PagedKVCache::seq_rm)
        if (!cache_.seq_rm(seq_id, p0, p1)) { unsupported_ = true; return false; }
        open_.clear();
        return true;
    }
    const int64_t len = cache_.seq_len(seq_id);
    const int64_t e = (p1 < 0 || p1 > len) ? len : p1;
    if (p0 < e && e < len) {
        // tokens survive behind the removed range: remember it, the seq_add that follows closes the context shift
        if (open_.count(seq_id)) { unsupported_ = true; return false; }
        open_[seq_id] = std::make_pair(p0, (llama_pos) e);
        return true;
    }
    if (open_.count(seq_id)) { unsupported_ = true; return false; }
    return cache_.seq_rm(seq_id, p0, p1);                    // tail / whole sequence / empty range
}

void llama_kv_cache_paged::seq_cp(llama_seq_id seq_id_src, llama_seq_id seq_id_dst, llama_pos p0, llama_pos p1) {
    if (open_.count(seq_id_src) || open_.count(seq_id_dst) || !cache_.seq_cp(seq_id_src, seq_id_dst, p0, p1)) unsupported_ = true;
}

void llama_kv_cache_paged::seq_keep(llama_seq_id seq_id) {
    cache_.seq_keep(seq_id);
    for (auto it = open_.begin(); it != open_.end(); ) it = it->first == seq_id ? std::next(it) : open_.erase(it);
}

void llama_kv_cache_paged::seq_add(llama_seq_id seq_id, llama_pos p0, llama_pos p1, llama_pos delta) {
    if (delta == 0) return;                                  // src/llama-kv-cache.cpp:328-330
    if (p0 < 0) p0 = 0;
    if (p1 < 0) p1 = std::numeric_limits<llama_pos>::max();
    if (p0 == p1) return;
    auto it = open_.find(seq_id);
    const int64_t len = cache_.seq_len(seq_id);
    // the closing half of a context shift: [gap.second, end) moves down by the width of the gap
    if (it == open_.end() || p0 != it->second.second || delta != it->second.first - it->second.second || (int64_t) p1 < len) {
        unsupported_ = true;
        return;
    }
    const llama_pos g0 = it->second.first, g1 = it->second.second;
    open_.erase(it);
    if (!cache_.context_shift(seq_id, g0, g1, work_.cow_pairs, work_.row_moves, work_.shift_cells)) { unsupported_ = true; return; }
    work_.shift_delta.resize(work_.shift_cells.size(), delta);   // the reference's cells[i].delta (:360)
    // this shift's share of the lists is one group; the consumer finishes a group before it starts the next, so a later
    // shift (of this or another sequence) may reuse the blocks this one freed and read the rows this one wrote
    work_.groups.push_back({work_.cow_pairs.size(), work_.row_moves.size(), work_.shift_cells.size()});
}

void llama_kv_cache_paged::seq_div(llama_seq_id, llama_pos p0, llama_pos p1, int d) {
    if (d == 1 || p0 == p1) return;                          // src/llama-kv-cache.cpp:386-401
    unsupported_ = true;
}

llama_pos llama_kv_cache_paged::seq_pos_max(llama_seq_id seq_id) {
    // the reference starts its maximum at 0 (src/llama-kv-cache.cpp:431-441), so an absent or empty sequence answers 0, the
    // same as a sequence holding position 0 only — pinned against the reference itself in tests/test_ref_kv_cache.py
    const int64_t len = cache_.seq_len(seq_id);
    return len > 0 ? (llama_pos) (len - 1) : 0;
}

bool llama_kv_cache_paged::find_slot(llama_seq_id seq_id, int32_t n_tokens, std::vector<int32_t> & slot_mapping,
                                     std::vector<std::pair<int32_t, int32_t>> & cow_pairs) {
    // llama_context::decode runs kv_self_update() before it looks for a slot (src/llama-context.cpp:1291-1302): so must the host
    if (open_.count(seq_id) || !work_.empty()) { unsupported_ = true; return false; }
    return cache_.append(seq_id, n_tokens, slot_mapping, cow_pairs);
}

int llama_kv_cache_paged::find_slot(int32_t n_tokens, const llama_seq_id * seq_id, const llama_pos * pos,
                                    PagedKVCache::BatchInputs & out) {
    out = PagedKVCache::BatchInputs();
    if (!work_.empty()) { unsupported_ = true; return 1; }
    for (int32_t i = 0; i < n_tokens; ++i) { if (open_.count(seq_id[i])) { unsupported_ = true; return 1; } }
    return cache_.prepare_batch(n_tokens, seq_id, pos, out);
}

bool llama_kv_cache_paged::update(pending_work & out) {
    out = pending_work();
    if (!open_.empty() || unsupported_) { unsupported_ = true; return false; }
    out = std::move(work_);
    work_ = pending_work();
    if (do_defrag_) { cache_.defrag(out.defrag_moves); do_defrag_ = false; }
    return true;
}

// ---- C ABI (include/b200_host.h): the llama_kv_cache method surface for hosts that bind the library over FFI
#include "b200_host.h"

struct b200_kv_cache {
    llama_kv_cache_paged impl;
    llama_kv_cache_paged::pending_work held;                  // taken from impl.update() by the sizing call, handed out by the next
    bool holding = false;
    b200_kv_cache(int64_t n, int bs) : impl(n, bs) {}
};

extern "C" {

B200_API b200_kv_cache *b200_kvc_create(int64_t n_blocks, int32_t block_size) {
    return (n_blocks <= 0 || block_size <= 0) ? nullptr : new (std::nothrow) b200_kv_cache(n_blocks, block_size);
}
B200_API void    b200_kvc_destroy(b200_kv_cache *kv) { delete kv; }
B200_API void    b200_kvc_clear(b200_kv_cache *kv) { kv->impl.clear(); kv->held = llama_kv_cache_paged::pending_work(); kv->holding = false; }
B200_API void    b200_kvc_defrag(b200_kv_cache *kv) { kv->impl.defrag(); }
B200_API int32_t b200_kvc_seq_rm(b200_kv_cache *kv, int32_t seq, int32_t p0, int32_t p1) { return kv->impl.seq_rm(seq, p0, p1) ? 1 : 0; }
B200_API void    b200_kvc_seq_cp(b200_kv_cache *kv, int32_t src, int32_t dst, int32_t p0, int32_t p1) { kv->impl.seq_cp(src, dst, p0, p1); }
B200_API void    b200_kvc_seq_keep(b200_kv_cache *kv, int32_t seq) { kv->impl.seq_keep(seq); }
B200_API void    b200_kvc_seq_add(b200_kv_cache *kv, int32_t seq, int32_t p0, int32_t p1, int32_t delta) { kv->impl.seq_add(seq, p0, p1, delta); }
B200_API void    b200_kvc_seq_div(b200_kv_cache *kv, int32_t seq, int32_t p0, int32_t p1, int32_t d) { kv->impl.seq_div(seq, p0, p1, d); }
B200_API int32_t b200_kvc_seq_pos_max(b200_kv_cache *kv, int32_t seq) { return kv->impl.seq_pos_max(seq); }
B200_API int32_t b200_kvc_n_tokens(const b200_kv_cache *kv) { return kv->impl.get_n_tokens(); }
B200_API int32_t b200_kvc_used_cells(const b200_kv_cache *kv) { return (int32_t) kv->impl.get_used_cells(); }
B200_API int32_t b200_kvc_unsupported(const b200_kv_cache *kv) { return kv->impl.unsupported() ? 1 : 0; }
B200_API int64_t b200_kvc_slots_of(const b200_kv_cache *kv, int32_t seq, int32_t *cells, int64_t cap) {
    std::vector<int32_t> v;
    if (!kv->impl.blocks().slots_of(seq, v)) return -1;
    if ((int64_t) v.size() > cap) return -(int64_t) v.size() - 1;
    for (size_t i = 0; i < v.size(); ++i) cells[i] = v[i];
    return (int64_t) v.size();
}
B200_API int32_t b200_kvc_find_slot(b200_kv_cache *kv, int32_t n_tokens, const int32_t *seq, const int32_t *pos,
                                    int32_t *slot_mapping, int32_t *q_seq, int32_t *q_pos, int32_t *seqs, int64_t *n_seqs,
                                    int32_t *cow_pairs, int64_t cow_cap, int64_t *n_cow) {
    if (!kv || n_tokens < 0 || (n_tokens > 0 && (!seq || !slot_mapping || !q_seq || !q_pos || !seqs))) return -1;
    if (cow_cap < n_tokens) return -3;
    PagedKVCache::BatchInputs in;
    const int rc = kv->impl.find_slot(n_tokens, seq, pos, in);
    if (rc != 0) return rc;
    for (int32_t i = 0; i < n_tokens; ++i) { slot_mapping[i] = in.slot_mapping[(size_t) i]; q_seq[i] = in.q_seq[(size_t) i]; q_pos[i] = in.q_pos[(size_t) i]; }
    for (size_t r = 0; r < in.seqs.size(); ++r) seqs[r] = in.seqs[r];
    if (n_seqs) *n_seqs = (int64_t) in.seqs.size();
    for (size_t i = 0; i < in.cow_pairs.size(); ++i) { cow_pairs[2*i] = in.cow_pairs[i].first; cow_pairs[2*i + 1] = in.cow_pairs[i].second; }
    if (n_cow) *n_cow = (int64_t) in.cow_pairs.size();
    return 0;
}
// kv_self_update: counts[5] = {cow pairs, row moves, shift cells, defrag moves, groups}.  Call with NULL arrays to learn the counts
// (nothing is consumed), then with arrays of at least that size.  Returns 1, 0 when a pending request cannot be expressed
// (b200_kvc_unsupported), -3 when a capacity is too small.
B200_API int32_t b200_kvc_update(b200_kv_cache *kv, int32_t *cow_pairs, int32_t *row_moves, int32_t *shift_cells, int32_t *shift_delta,
                                 int32_t *defrag_moves, const int64_t caps[4], int64_t counts[5]) {
    if (!kv || !counts) return -1;
    llama_kv_cache_paged::pending_work & held = kv->held;
    if (!kv->holding) {
        if (!kv->impl.update(held)) { held = llama_kv_cache_paged::pending_work(); return 0; }
        kv->holding = true;
    }
    counts[0] = (int64_t) held.cow_pairs.size(); counts[1] = (int64_t) held.row_moves.size();
    counts[2] = (int64_t) held.shift_cells.size(); counts[3] = (int64_t) held.defrag_moves.size(); counts[4] = (int64_t) held.groups.size();
    const bool fits = caps && caps[0] >= counts[0] && caps[1] >= counts[1] && caps[2] >= counts[2] && caps[3] >= counts[3];
    if (!fits) { if ((counts[0] | counts[1] | counts[2] | counts[3]) == 0) { kv->holding = false; return 1; } return -3; }
    for (size_t i = 0; i < held.cow_pairs.size(); ++i) { cow_pairs[2*i] = held.cow_pairs[i].first; cow_pairs[2*i + 1] = held.cow_pairs[i].second; }
    for (size_t i = 0; i < held.row_moves.size(); ++i) { row_moves[2*i] = held.row_moves[i].first; row_moves[2*i + 1] = held.row_moves[i].second; }
    for (size_t i = 0; i < held.shift_cells.size(); ++i) { shift_cells[i] = held.shift_cells[i]; shift_delta[i] = held.shift_delta[i]; }
    for (size_t i = 0; i < held.defrag_moves.size(); ++i) { defrag_moves[2*i] = held.defrag_moves[i].first; defrag_moves[2*i + 1] = held.defrag_moves[i].second; }
    held = llama_kv_cache_paged::pending_work();
    kv->holding = false;
    return 1;
}

// the group boundaries of the work the sizing call of b200_kvc_update is holding: ends[g] = {cow_end, move_end, shift_end}
B200_API int64_t b200_kvc_update_groups(const b200_kv_cache *kv, int64_t *ends, int64_t cap) {
    if (!kv || !kv->holding) return 0;
    const auto & g = kv->held.groups;
    if ((int64_t) g.size() > cap || (!g.empty() && !ends)) return -(int64_t) g.size() - 1;
    for (size_t i = 0; i < g.size(); ++i) { ends[3*i] = (int64_t) g[i].cow_end; ends[3*i + 1] = (int64_t) g[i].move_end; ends[3*i + 2] = (int64_t) g[i].shift_end; }
    return (int64_t) g.size();
}

} // extern "C"
