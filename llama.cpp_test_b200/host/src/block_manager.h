// block_manager.h — the fork's BlockManager interface (reference: tests/test_block_manager.cpp:3-62,
// prebuilt ./test_block_manager; source absent from the reference tree, semantics recovered in
// SURVEY.md §2.3 and pinned by tests/golden/fork/test_block_manager.{stdout,stderr}, the output of the
// reference binary itself).  Same class name, method names, return conventions and log lines.
#pragma once
#include <cstddef>
#include <vector>

// set to false to silence the fork's std::cerr trace lines (default: on, like the reference)
extern bool g_b200_fork_trace;

class BlockManager {
public:
    explicit BlockManager(size_t num_blocks);
    ~BlockManager();
    // first-fit lowest free index; when full, evictBlock() and re-allocate that index; -1 if impossible
    int  allocateBlock();
    // out-of-range index: logged and ignored; freeing a free block is a silent no-op
    void freeBlock(int index);
    // frees and returns the lowest allocated index (not LRU), -1 if nothing is allocated
    int  evictBlock();
    std::vector<bool> getBlockStatus() const;
private:
    size_t            num_blocks_;
    std::vector<bool> allocated_;
};
