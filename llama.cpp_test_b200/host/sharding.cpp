// sharding.cpp — the two partitionings of the path over the GPUs of one box, as C entry points so that a
// C / C++ host — the only kind the reference has — does not have to re-implement the policy (include/b200_host.h).
// Pure integer bookkeeping, deterministic: every rank computes the same answer without communicating.  The reference has
// no counterpart (its multi-GPU path splits matmul rows, ggml/src/ggml-cuda/ggml-cuda.cu:598,1592).
#include "b200_host.h"

#include <algorithm>
#include <cstdint>
#include <numeric>
#include <vector>

extern "C" {

// Partition A, sequence-batch sharding: longest-processing-time greedy — sequences by context length (descending, index
// as tie-break) go to the currently lightest rank (ties: fewer sequences, then lower rank).  rank_of[i] = owner of sequence i.
B200_API int32_t b200_shard_sequences(const int32_t *ctx_lens, int64_t n_seqs, int32_t world, int32_t *rank_of) {
    if (world <= 0 || n_seqs < 0 || (n_seqs > 0 && (!ctx_lens || !rank_of))) return -1;
    std::vector<int64_t> order((size_t) n_seqs);
    std::iota(order.begin(), order.end(), 0);
    std::sort(order.begin(), order.end(), [&](int64_t a, int64_t b) { return ctx_lens[a] != ctx_lens[b] ? ctx_lens[a] > ctx_lens[b] : a < b; });
    std::vector<int64_t> load((size_t) world, 0), count((size_t) world, 0);
    for (int64_t i : order) {
        int32_t best = 0;
        for (int32_t r = 1; r < world; ++r) {
            if (load[(size_t) r] < load[(size_t) best] || (load[(size_t) r] == load[(size_t) best] && count[(size_t) r] < count[(size_t) best])) best = r;
        }
        rank_of[i] = best;
        load[(size_t) best] += ctx_lens[i];
        ++count[(size_t) best];
    }
    return 0;
}

// Partition B, KV-head sharding: GQA maps query head h to kv head h / (n_head / n_head_kv) (ggml-cpu.c:12264), so a
// contiguous range of kv heads owns a contiguous range of query heads.  out = {kv0, n_head_kv_local, q0, n_head_local}.
B200_API int32_t b200_head_shard(int32_t rank, int32_t world, int32_t n_head, int32_t n_head_kv, int32_t out[4]) {
    if (!out || world <= 0 || rank < 0 || rank >= world || n_head <= 0 || n_head_kv <= 0) return -1;
    if (n_head % n_head_kv) return -2;                       // n_head must be a multiple of n_head_kv
    if (n_head_kv % world) return -3;                        // the kv heads do not split over the ranks
    const int32_t per = n_head_kv/world, gqa = n_head/n_head_kv;
    out[0] = rank*per; out[1] = per; out[2] = rank*per*gqa; out[3] = per*gqa;
    return 0;
}

} // extern "C"
