// llama_kv_cache_paged.cpp — see src/llama_kv_cache_paged.h
#include "llama_kv_cache_paged.h"

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
    shifted_.clear();
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
        if (open_.count(seq_id) || shifted_.count(seq_id)) { unsupported_ = true; return false; }
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
    // one shift per sequence between two updates: the row moves of a second one would read what the first one has not
    // written yet (the lists are executed as one batch: copies, gathers, scatters, rotation)
    shifted_.insert(seq_id);
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
    shifted_.clear();
    if (do_defrag_) { cache_.defrag(out.defrag_moves); do_defrag_ = false; }
    return true;
}
