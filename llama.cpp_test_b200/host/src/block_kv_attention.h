// block_kv_attention.h — the fork's free functions (reference: tests/test_block_kv_attention.cpp:15-57,
// tests/test_paged_attention.cpp:35-39, prebuilt ./test_apply_attention; goldens
// tests/golden/fork/test_{block_kv_attention,paged_attention,apply_attention}.stdout).
// In the fork these are host-side placeholders ("dummy attention": the query is copied to the output after
// the cache write; the "real" variant is a 4-key scalar softmax).  They are mirrored here only so callers of the
// fork API keep compiling; attention proper is b200_paged_attention (include/b200_paged_attn.h).
#pragma once
#include <cstddef>
#include <vector>
#include "block_kv_cache.h"

void blockKVAttentionFwd(const float * query, size_t query_size, BlockKVCache & cache,
                         const float * key, const float * value,
                         const std::vector<int> & slot_mapping, float * output);
void blockKVAttentionPagedFwd(const float * query, size_t query_size, BlockKVCache & cache,
                              const float * key, const float * value, float * output, size_t page_size);
void blockKVAttentionRealFwd(const float * query, size_t query_size, BlockKVCache & cache,
                             const float * key, const float * value,
                             const std::vector<int> & slot_mapping, float * output);
void applyAttention(const float * query, size_t query_size, BlockKVCache & cache,
                    const float * key, const float * value, float * output, size_t page_size, bool usePaged);
