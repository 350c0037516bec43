// capi_types.h — the opaque handle types of include/b200_host.h (shared by the C-ABI translation units)
#pragma once
#include "block_manager.h"
#include "paged_kv_cache.h"

struct b200_block_manager { BlockManager impl; explicit b200_block_manager(size_t n) : impl(n) {} };
struct b200_paged_cache   { PagedKVCache impl; b200_paged_cache(int64_t n, int bs) : impl(n, bs) {} };
