// kv_cache_manager.h — the fork's slot manager (reference: prebuilt ./test_apply_attention symbols
// KVCacheManager::*, PhysicalBlock::* from kv_cache_manager.cpp; exercised by tests/test_kv_cache.cpp:28-53;
// SURVEY.md §2.3; golden tests/golden/fork/test_kv_cache.stdout).
#pragma once
#include <cstddef>
#include <vector>

struct VirtualBlockEntry {
    size_t physicalBlockIndex;
    size_t slotIndex;
};

class PhysicalBlock {
public:
    explicit PhysicalBlock(size_t capacity);
    size_t allocateSlot();            // lowest free slot; throws std::runtime_error when full
    void   freeSlot(size_t slot);     // bounds-checked clear
    bool   hasFreeSlot() const;
    size_t getFreeSlots() const;
    size_t getCapacity() const;
private:
    size_t            capacity_;
    std::vector<bool> used_;
};

class KVCacheManager {
public:
    explicit KVCacheManager(size_t slotCapacity);   // starts with ONE PhysicalBlock
    ~KVCacheManager();
    KVCacheManager(const KVCacheManager &) = delete;
    KVCacheManager & operator=(const KVCacheManager &) = delete;
    void   ensureFreeSlot();                        // appends a PhysicalBlock when none has room
    VirtualBlockEntry insert();                     // first block with room, lowest slot
    void   free(VirtualBlockEntry entry);
    size_t totalPhysicalBlocks() const;
    size_t totalFreeSlots() const;
    size_t getSlotCapacity() const;
private:
    size_t                       slotCapacity_;
    std::vector<PhysicalBlock *> blocks_;
};
