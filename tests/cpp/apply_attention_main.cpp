// tests/cpp/apply_attention_main.cpp — the main() of the fork's prebuilt ./test_apply_attention (its
// source is absent from the reference tree), restated from what the binary prints: query = 0..15, four keys {1, 2, 3, 4} and
// values {2, 4, 6, 8}.  With usePaged = true applyAttention is the chunked dummy (query copied through); with false it is
// the 4-key scalar softmax: output[0] = mean(value) = 5, output[1] = sum_j softmax(1*key)_j value_j = 6.98531, -> 8.
// stdout must equal tests/golden/fork/test_apply_attention.stdout, the reference binary's own output.
#include "block_kv_attention.h"

#include <iostream>
#include <vector>

int main() {
    const size_t n = 16;
    std::vector<float> query(n), out(n);
    for (size_t i = 0; i < n; ++i) query[i] = (float) i;
    const std::vector<float> key = {1.0f, 2.0f, 3.0f, 4.0f}, value = {2.0f, 4.0f, 6.0f, 8.0f};
    BlockKVCache cache(/*num_blocks=*/1, /*block_size=*/4, /*num_kv_heads=*/1, /*head_size=*/1);
    for (int paged = 1; paged >= 0; --paged) {
        applyAttention(query.data(), n, cache, key.data(), value.data(), out.data(), /*page_size=*/4, paged != 0);
        std::cout << "applyAttention output (usePaged = " << (paged ? "true" : "false") << "):" << std::endl;
        for (size_t i = 0; i < n; ++i) std::cout << out[i] << " ";
        std::cout << std::endl;
    }
    return 0;
}
