// host_capi.cpp — C ABI wrappers (include/b200_host.h) over the host classes.
#include "b200_host.h"
#include "capi_types.h"

#include <algorithm>
#include <new>
#include <vector>

extern "C" {

B200_API b200_block_manager *b200_bm_create(int64_t n) { return n < 0 ? nullptr : new (std::nothrow) b200_block_manager((size_t) n); }
B200_API void    b200_bm_destroy(b200_block_manager *bm) { delete bm; }
B200_API int32_t b200_bm_allocate(b200_block_manager *bm) { return bm->impl.allocateBlock(); }
B200_API void    b200_bm_free(b200_block_manager *bm, int32_t i) { bm->impl.freeBlock(i); }
B200_API int32_t b200_bm_evict(b200_block_manager *bm) { return bm->impl.evictBlock(); }
B200_API int64_t b200_bm_status(const b200_block_manager *bm, uint8_t *out, int64_t cap) {
    const std::vector<bool> s = bm->impl.getBlockStatus();
    for (int64_t i = 0; i < (int64_t) s.size() && i < cap; ++i) out[i] = s[(size_t) i] ? 1 : 0;
    return (int64_t) s.size();
}
B200_API void b200_fork_trace(int enabled) { g_b200_fork_trace = enabled != 0; }

B200_API b200_paged_cache *b200_pc_create(int64_t n_blocks, int32_t block_size) {
    return (n_blocks < 0 || block_size <= 0) ? nullptr : new (std::nothrow) b200_paged_cache(n_blocks, block_size);
}
B200_API void b200_pc_destroy(b200_paged_cache *pc) { delete pc; }

B200_API int64_t b200_pc_append(b200_paged_cache *pc, int32_t seq, int64_t n_new, int32_t *slot_mapping,
                                int32_t *cow_pairs, int64_t cow_cap) {
    if (cow_cap < 1 && n_new > 0) return -2;             // one append emits at most one pair
    std::vector<int32_t> slots;
    std::vector<std::pair<int32_t, int32_t>> cow;
    if (!pc->impl.append(seq, n_new, slots, cow)) return -1;
    for (size_t i = 0; i < slots.size(); ++i) slot_mapping[i] = slots[i];
    for (size_t i = 0; i < cow.size(); ++i) { cow_pairs[2*i] = cow[i].first; cow_pairs[2*i + 1] = cow[i].second; }
    return (int64_t) cow.size();
}
B200_API int32_t b200_pc_prepare_batch(b200_paged_cache *pc, int64_t n_tokens, const int32_t *seq, const int32_t *pos,
                                       int32_t *slot_mapping, int32_t *q_seq, int32_t *q_pos, int32_t *seqs, int64_t *n_seqs,
                                       int32_t *cow_pairs, int64_t cow_cap, int64_t *n_cow) {
    if (!pc || n_tokens < 0 || (n_tokens > 0 && (!seq || !slot_mapping || !q_seq || !q_pos || !seqs))) return -1;
    // at most one copy-on-write pair per sequence, at most one sequence per token: refuse before changing anything
    {
        std::vector<int32_t> distinct(seq, seq + n_tokens);
        std::sort(distinct.begin(), distinct.end());
        const int64_t n_distinct = (int64_t) (std::unique(distinct.begin(), distinct.end()) - distinct.begin());
        if (cow_cap < n_distinct) return -3;
    }
    PagedKVCache::BatchInputs in;
    const int rc = pc->impl.prepare_batch(n_tokens, seq, pos, in);
    if (rc != 0) return rc;
    for (int64_t i = 0; i < n_tokens; ++i) { slot_mapping[i] = in.slot_mapping[(size_t) i]; q_seq[i] = in.q_seq[(size_t) i]; q_pos[i] = in.q_pos[(size_t) i]; }
    for (size_t r = 0; r < in.seqs.size(); ++r) seqs[r] = in.seqs[r];
    if (n_seqs) *n_seqs = (int64_t) in.seqs.size();
    for (size_t i = 0; i < in.cow_pairs.size(); ++i) { cow_pairs[2*i] = in.cow_pairs[i].first; cow_pairs[2*i + 1] = in.cow_pairs[i].second; }
    if (n_cow) *n_cow = (int64_t) in.cow_pairs.size();
    return 0;
}
B200_API void b200_pc_undo_batch(b200_paged_cache *pc, int64_t n_tokens, const int32_t *seq, const int32_t *q_pos,
                                 const int32_t *cow_pairs, int64_t n_cow) {
    if (!pc || n_tokens <= 0 || !seq || !q_pos) return;
    PagedKVCache::BatchInputs in;
    for (int64_t i = 0; cow_pairs && i < n_cow; ++i) in.cow_pairs.emplace_back(cow_pairs[2*i], cow_pairs[2*i + 1]);
    for (int64_t i = 0; i < n_tokens; ++i) {
        size_t r = 0;
        while (r < in.seqs.size() && in.seqs[r] != seq[i]) ++r;
        if (r == in.seqs.size()) in.seqs.push_back(seq[i]);
        in.q_seq.push_back((int32_t) r);
        in.q_pos.push_back(q_pos[i]);
    }
    pc->impl.undo_batch(in);
}
B200_API int32_t b200_pc_seq_cp(b200_paged_cache *pc, int32_t s, int32_t d, int64_t p0, int64_t p1) { return pc->impl.seq_cp(s, d, p0, p1) ? 1 : 0; }
B200_API int32_t b200_pc_seq_rm(b200_paged_cache *pc, int32_t s, int64_t p0, int64_t p1) { return pc->impl.seq_rm(s, p0, p1) ? 1 : 0; }
B200_API void    b200_pc_seq_keep(b200_paged_cache *pc, int32_t s) { pc->impl.seq_keep(s); }
B200_API void    b200_pc_clear(b200_paged_cache *pc) { pc->impl.clear(); }
B200_API int64_t b200_pc_seq_len(const b200_paged_cache *pc, int32_t s) { return pc->impl.seq_len(s); }
B200_API int64_t b200_pc_n_free(const b200_paged_cache *pc) { return pc->impl.n_free(); }
B200_API int32_t b200_pc_ref_count(const b200_paged_cache *pc, int64_t b) { return pc->impl.ref_count(b); }
B200_API int32_t b200_pc_export(const b200_paged_cache *pc, const int32_t *seqs, int64_t n,
                                int32_t *bt, int64_t stride, int32_t *lens) {
    return pc->impl.export_tables(seqs, n, bt, stride, lens) ? 1 : 0;
}

} // extern "C"
