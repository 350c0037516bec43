// paged_kv_cache.cpp — see src/paged_kv_cache.h
#include "paged_kv_cache.h"

#include <algorithm>

PagedKVCache::PagedKVCache(int64_t n_blocks, int block_size)
    : block_size_(block_size > 0 ? block_size : 1), ref_((size_t) (n_blocks > 0 ? n_blocks : 0), 0), n_free_(n_blocks > 0 ? n_blocks : 0) {
    for (int32_t b = 0; b < (int32_t) ref_.size(); ++b) free_.push(b);
}

int PagedKVCache::ref_count(int64_t block) const {
    return (block < 0 || block >= (int64_t) ref_.size()) ? -1 : ref_[(size_t) block];
}

int32_t PagedKVCache::alloc_block() {
    if (free_.empty()) return -1;
    const int32_t b = free_.top(); free_.pop();
    ref_[(size_t) b] = 1; --n_free_;
    return b;
}

void PagedKVCache::unref(int32_t b) {
    if (--ref_[(size_t) b] == 0) { free_.push(b); ++n_free_; }
}

void PagedKVCache::release(Seq & s, int64_t keep_tokens) {
    const int64_t keep_blocks = (keep_tokens + block_size_ - 1)/block_size_;
    while ((int64_t) s.blocks.size() > keep_blocks) { unref(s.blocks.back()); s.blocks.pop_back(); }
    s.len = std::min(s.len, keep_tokens);
}

bool PagedKVCache::append(seq_id seq, int64_t n_new, std::vector<int32_t> & slot_mapping,
                          std::vector<std::pair<int32_t, int32_t>> & cow_pairs) {
    if (n_new < 0) return false;
    auto it = seqs_.find(seq);
    const int64_t len = it == seqs_.end() ? 0 : it->second.len;
    const int64_t have = it == seqs_.end() ? 0 : (int64_t) it->second.blocks.size();
    const int64_t off  = len % block_size_;
    const bool tail_shared = n_new > 0 && off != 0 && ref_[(size_t) it->second.blocks.back()] > 1;
    const int64_t need_blocks = (len + n_new + block_size_ - 1)/block_size_ - have;
    if (need_blocks + (tail_shared ? 1 : 0) > n_free_) return false;          // all-or-nothing
    if (n_new == 0) { if (it == seqs_.end()) seqs_[seq]; return true; }
    Seq & s = seqs_[seq];
    if (tail_shared) {                                                         // copy-on-write
        const int32_t old = s.blocks.back();
        const int32_t nb  = alloc_block();
        cow_pairs.emplace_back(old, nb);
        s.blocks.back() = nb;
        unref(old);
    }
    for (int64_t i = 0; i < need_blocks; ++i) s.blocks.push_back(alloc_block());
    for (int64_t p = len; p < len + n_new; ++p)
        slot_mapping.push_back((int32_t) ((int64_t) s.blocks[(size_t) (p/block_size_)]*block_size_ + p % block_size_));
    s.len = len + n_new;
    return true;
}

void PagedKVCache::used_cells(std::vector<UsedCell> & out) const {
    std::map<int32_t, UsedCell> by_cell;
    for (const auto & kv : seqs_) {                            // seqs_ is ordered: ids come out ascending
        const Seq & s = kv.second;
        for (int64_t p = 0; p < s.len; ++p) {
            const int32_t cell = (int32_t) ((int64_t) s.blocks[(size_t) (p/block_size_)]*block_size_ + p % block_size_);
            UsedCell & u = by_cell[cell];
            u.cell = cell; u.pos = (int32_t) p;
            u.ids.push_back(kv.first);
        }
    }
    out.clear();
    for (auto & kv : by_cell) out.push_back(std::move(kv.second));
}

int PagedKVCache::prepare_batch(int64_t n_tokens, const seq_id * seq, const int32_t * pos, BatchInputs & out) {
    out = BatchInputs();
    if (n_tokens < 0 || (n_tokens > 0 && !seq)) return -1;
    std::map<seq_id, int32_t> row;                       // sequence -> row, in order of first appearance
    std::vector<int64_t> n_new;
    for (int64_t i = 0; i < n_tokens; ++i) {
        auto it = row.find(seq[i]);
        if (it == row.end()) { it = row.emplace(seq[i], (int32_t) out.seqs.size()).first; out.seqs.push_back(seq[i]); n_new.push_back(0); }
        const int64_t want = seq_len(seq[i]) + n_new[(size_t) it->second];
        if (seq[i] < 0 || (pos && pos[i] != want)) { out = BatchInputs(); return -1; }
        out.q_seq.push_back(it->second);
        out.q_pos.push_back((int32_t) want);
        ++n_new[(size_t) it->second];
    }
    // the appends are sequential and each is all-or-nothing; undo the earlier ones when a later one fails
    const std::vector<int32_t> ref0 = ref_;
    const auto free0 = free_;
    const int64_t n_free0 = n_free_;
    const std::map<seq_id, Seq> seqs0 = seqs_;
    std::vector<std::vector<int32_t>> slots(out.seqs.size());
    for (size_t r = 0; r < out.seqs.size(); ++r) {
        if (!append(out.seqs[r], n_new[r], slots[r], out.cow_pairs)) {
            ref_ = ref0; free_ = free0; n_free_ = n_free0; seqs_ = seqs0;
            out = BatchInputs();
            return 1;
        }
    }
    std::vector<size_t> next(out.seqs.size(), 0);
    out.slot_mapping.resize((size_t) n_tokens);
    for (int64_t i = 0; i < n_tokens; ++i) {
        const size_t r = (size_t) out.q_seq[(size_t) i];
        out.slot_mapping[(size_t) i] = slots[r][next[r]++];
    }
    return 0;
}

void PagedKVCache::undo_batch(const BatchInputs & in) {
    std::vector<int64_t> first(in.seqs.size(), -1);
    for (size_t i = 0; i < in.q_seq.size(); ++i) {
        int64_t & f = first[(size_t) in.q_seq[i]];
        if (f < 0 || in.q_pos[i] < f) f = in.q_pos[i];
    }
    for (size_t r = 0; r < in.seqs.size(); ++r) {
        if (first[r] < 0) continue;
        if (first[r] == 0) seq_rm(in.seqs[r], -1, -1); else seq_rm(in.seqs[r], first[r], -1);
    }
    // A copy-on-write gave a sequence its own block `dst` for the partly filled, shared block `src` it was about to append
    // into.  Whether the device copy src -> dst ever ran is unknown here (the ubatch may have failed before it), so the table
    // entry goes back to the shared block, whose rows are certainly valid, and dst is released.  src is still allocated: it
    // was shared (ref > 1) when the pair was made and nothing but this batch has touched the cache since.
    for (const auto & pr : in.cow_pairs) {
        if (pr.first < 0 || pr.first >= (int32_t) ref_.size() || ref_[(size_t) pr.first] <= 0) continue;
        for (seq_id id : in.seqs) {
            auto it = seqs_.find(id);
            if (it == seqs_.end()) continue;
            for (int32_t & b : it->second.blocks) {
                if (b == pr.second) { b = pr.first; ++ref_[(size_t) pr.first]; unref(pr.second); }
            }
        }
    }
}

bool PagedKVCache::seq_cp(seq_id src, seq_id dst, int64_t p0, int64_t p1) {
    if (p0 > 0) return false;
    auto it = seqs_.find(src);
    if (it == seqs_.end()) return false;
    if (src == dst) return true;                                               // llama-kv-cache.cpp:245-247
    const int64_t n = (p1 < 0 || p1 > it->second.len) ? it->second.len : p1;
    const int64_t nb = (n + block_size_ - 1)/block_size_;
    std::vector<int32_t> shared(it->second.blocks.begin(), it->second.blocks.begin() + nb);
    for (int32_t b : shared) ++ref_[(size_t) b];
    auto jt = seqs_.find(dst);
    if (jt != seqs_.end()) release(jt->second, 0);
    Seq & d = seqs_[dst];
    d.blocks.swap(shared);
    d.len = n;
    return true;
}

bool PagedKVCache::seq_rm(seq_id seq, int64_t p0, int64_t p1) {
    if (seq < 0) {
        bool ok = true;
        std::vector<seq_id> ids;
        for (auto & kv : seqs_) ids.push_back(kv.first);
        for (seq_id s : ids) ok = seq_rm(s, p0, p1) && ok;
        return ok;
    }
    auto it = seqs_.find(seq);
    if (it == seqs_.end()) return true;
    if (p0 < 0) p0 = 0;
    const int64_t len = it->second.len;
    if (p1 >= 0 && p1 < len) return p0 >= p1;                                  // middle range: only the empty one
    release(it->second, std::min(p0, len));
    if (it->second.len == 0) seqs_.erase(it);
    return true;
}

bool PagedKVCache::context_shift_sizes(seq_id seq, int64_t p0, int64_t p1, int64_t & n_cow, int64_t & n_moves, int64_t & n_shift) const {
    auto it = seqs_.find(seq);
    if (it == seqs_.end()) return false;
    const Seq & s = it->second;
    if (p0 < 0 || p1 <= p0 || p1 > s.len) return false;
    const int64_t nd = p1 - p0, new_len = s.len - nd, bs = block_size_;
    const bool aligned = p0 % bs == 0 && nd % bs == 0;
    n_shift = new_len - p0;
    n_moves = aligned ? 0 : n_shift;
    n_cow = 0;
    if (n_shift > 0) {
        // blocks that will be written: aligned -> the survivors behind the gap (old indices); else -> those covering [p0, new_len)
        const int64_t b0 = aligned ? p1/bs : p0/bs, b1 = aligned ? (int64_t) s.blocks.size() : (new_len + bs - 1)/bs;
        for (int64_t b = b0; b < b1; ++b) if (ref_[(size_t) s.blocks[(size_t) b]] > 1) ++n_cow;
    }
    return n_cow <= n_free_;
}

bool PagedKVCache::context_shift(seq_id seq, int64_t p0, int64_t p1, std::vector<std::pair<int32_t, int32_t>> & cow_pairs,
                                 std::vector<std::pair<int32_t, int32_t>> & row_moves, std::vector<int32_t> & shift_cells) {
    int64_t n_cow, n_moves, n_shift;
    if (!context_shift_sizes(seq, p0, p1, n_cow, n_moves, n_shift)) return false;
    Seq & s = seqs_[seq];
    const int64_t nd = p1 - p0, new_len = s.len - nd, bs = block_size_;
    const bool aligned = p0 % bs == 0 && nd % bs == 0;
    auto cell_of = [&](const std::vector<int32_t> & blocks, int64_t p) { return (int32_t) ((int64_t) blocks[(size_t) (p/bs)]*bs + p % bs); };
    auto cow = [&](int64_t b) {                                 // private copy of table entry b (allocated before anything is freed)
        const int32_t old = s.blocks[(size_t) b];
        if (ref_[(size_t) old] <= 1) return;
        const int32_t nb = alloc_block();
        cow_pairs.emplace_back(old, nb);
        s.blocks[(size_t) b] = nb;
        unref(old);
    };
    if (aligned) {
        if (n_shift > 0) for (int64_t b = p1/bs; b < (int64_t) s.blocks.size(); ++b) cow(b);
        for (int64_t b = p0/bs; b < p1/bs; ++b) unref(s.blocks[(size_t) b]);
        s.blocks.erase(s.blocks.begin() + p0/bs, s.blocks.begin() + p1/bs);
        s.len = new_len;
    } else {
        if (n_shift > 0) for (int64_t b = p0/bs; b < (new_len + bs - 1)/bs; ++b) cow(b);
        for (int64_t q = p0; q < new_len; ++q) row_moves.emplace_back(cell_of(s.blocks, q + nd), cell_of(s.blocks, q));
        release(s, new_len);
    }
    for (int64_t q = p0; q < new_len; ++q) shift_cells.push_back(cell_of(s.blocks, q));
    if (s.len == 0) seqs_.erase(seq);
    return true;
}

void PagedKVCache::seq_keep(seq_id seq) {
    std::vector<seq_id> ids;
    for (auto & kv : seqs_) if (kv.first != seq) ids.push_back(kv.first);
    for (seq_id s : ids) seq_rm(s, 0, -1);
}

void PagedKVCache::clear() {
    seqs_.clear();
    std::fill(ref_.begin(), ref_.end(), 0);
    free_ = decltype(free_)();
    for (int32_t b = 0; b < (int32_t) ref_.size(); ++b) free_.push(b);
    n_free_ = (int64_t) ref_.size();
}

int64_t PagedKVCache::seq_len(seq_id seq) const {
    auto it = seqs_.find(seq);
    return it == seqs_.end() ? 0 : it->second.len;
}

const std::vector<int32_t> * PagedKVCache::blocks_of(seq_id seq) const {
    auto it = seqs_.find(seq);
    return it == seqs_.end() ? nullptr : &it->second.blocks;
}

bool PagedKVCache::slots_of(seq_id seq, std::vector<int32_t> & cells) const {
    auto it = seqs_.find(seq);
    if (it == seqs_.end()) return false;
    const Seq & s = it->second;
    cells.reserve(cells.size() + (size_t) s.len);
    for (int64_t p = 0; p < s.len; ++p)
        cells.push_back((int32_t) ((int64_t) s.blocks[(size_t) (p/block_size_)]*block_size_ + p % block_size_));
    return true;
}

int64_t PagedKVCache::defrag(std::vector<std::pair<int32_t, int32_t>> & moves, int64_t max_moves) {
    const int64_t n_used = n_blocks() - n_free_;
    std::vector<int32_t> remap(ref_.size(), -1);
    int64_t n_moves = 0;
    int32_t hole = 0;                                           // lowest free index below n_used
    for (int32_t src = (int32_t) ref_.size() - 1; src >= n_used; --src) {
        if (ref_[(size_t) src] == 0) continue;
        if (max_moves >= 0 && n_moves >= max_moves) break;
        while (hole < n_used && ref_[(size_t) hole] != 0) ++hole;
        if (hole >= n_used) break;
        moves.emplace_back(src, hole);
        remap[(size_t) src] = hole;
        ref_[(size_t) hole] = ref_[(size_t) src];
        ref_[(size_t) src]  = 0;
        ++n_moves;
    }
    if (n_moves == 0) return 0;
    for (auto & kv : seqs_)
        for (int32_t & b : kv.second.blocks)
            if (remap[(size_t) b] >= 0) b = remap[(size_t) b];
    free_ = decltype(free_)();
    for (int32_t b = 0; b < (int32_t) ref_.size(); ++b) if (ref_[(size_t) b] == 0) free_.push(b);
    return n_moves;
}

bool PagedKVCache::export_tables(const seq_id * seqs, int64_t n, int32_t * block_table, int64_t stride, int32_t * seq_lens) const {
    for (int64_t i = 0; i < n; ++i) {
        auto it = seqs_.find(seqs[i]);
        int32_t * row = block_table + i*stride;
        int64_t nb = 0;
        if (it != seqs_.end()) {
            nb = (int64_t) it->second.blocks.size();
            if (nb > stride) return false;
            std::copy(it->second.blocks.begin(), it->second.blocks.end(), row);
        }
        std::fill(row + nb, row + stride, 0);
        if (seq_lens) seq_lens[i] = it == seqs_.end() ? 0 : (int32_t) it->second.len;
    }
    return true;
}
