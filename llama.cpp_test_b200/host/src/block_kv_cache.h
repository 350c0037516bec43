// block_kv_cache.h — the fork's BlockKVCache interface (reference: tests/test_block_kv_cache.cpp:17-118,
// tests/test_kv_cache.cpp:13-53, prebuilt ./test_block_kv_cache, ./test_apply_attention; SURVEY.md §2.3;
// goldens tests/golden/fork/test_block_kv_cache.stdout, test_kv_cache.stdout).
//
// Like the fork's prototype this class is a HOST-side F32 container with block granularity
// ("slot" == block index).  It exists so code written against the fork keeps compiling; the typed,
// device-resident, token-granular cache that the CUDA kernels serve is PagedKVCache (paged_kv_cache.h).
#pragma once
#include <cstddef>
#include <vector>
#include "kv_cache_manager.h"

class BlockKVCache {
public:
    BlockKVCache(size_t num_blocks, size_t block_size, size_t num_kv_heads, size_t head_size);
    ~BlockKVCache();
    BlockKVCache(const BlockKVCache &) = delete;
    BlockKVCache & operator=(const BlockKVCache &) = delete;

    void initCache();
    // i-th source block (block_size*num_kv_heads*head_size floats) -> block slot_mapping[i]; invalid slots skipped
    void writeToCache(const float * key, const float * value, const std::vector<int> & slot_mapping);
    // element-wise swap of this block i with other's block mapping[i]
    void swapBlocks(BlockKVCache & other, const std::vector<int> & mapping);
    // copy block mapping[i] of src onto the same index here
    void copyBlocks(const BlockKVCache & src, const std::vector<int> & mapping);
    void splitKVCache(const std::vector<float> & kv, std::vector<float> & k, std::vector<float> & v);
    VirtualBlockEntry insertKVBlock(const float * key, const float * value);
    void freeKVBlock(VirtualBlockEntry entry);

    const std::vector<float> & getKeyCache() const   { return key_cache_; }
    const std::vector<float> & getValueCache() const { return value_cache_; }

    size_t numBlocks() const      { return num_blocks_; }
    size_t blockDataSize() const  { return block_size_*num_kv_heads_*head_size_; }
private:
    size_t num_blocks_, block_size_, num_kv_heads_, head_size_;
    std::vector<float> key_cache_, value_cache_;
    KVCacheManager * mgr_;
};
