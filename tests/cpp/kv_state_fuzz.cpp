// kv_state_fuzz.cpp — the session-stream parser (b200_kv_state_seq_parse: llama_kv_cache_unified::state_read_meta /
// state_read_data for one sequence, src/llama-kv-cache.cpp:1005-1240) on damaged input, built with AddressSanitizer by
// tests/test_kv_state.py: every truncation of a valid stream and seeded byte / field corruptions, each held in a heap
// buffer of exactly its own length so that any read past the end faults.  The parser must answer 0 or a B200_STATE_ERR_*
// and, when it answers 0, every offset it hands out must lie inside the buffer.
#include "b200_host.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#define CHECK(c) do { if (!(c)) { std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); std::exit(1); } } while (0)

int main() {
    const int32_t n_layer = 3, type_k = 8 /* Q8_0 */, type_v = 1 /* F16 */;
    const int64_t n_cells = 37, k_row = 34*8, v_row = 2*256;
    std::vector<uint8_t> k((size_t) (n_layer*n_cells*k_row)), v((size_t) (n_layer*n_cells*v_row));
    std::mt19937 rng(5);
    for (auto & b : k) b = (uint8_t) rng();
    for (auto & b : v) b = (uint8_t) rng();
    const int64_t size = b200_kv_state_seq_size(n_cells, n_layer, k_row, v_row);
    std::vector<uint8_t> good((size_t) size);
    CHECK(b200_kv_state_seq_write(good.data(), size, n_cells, n_layer, type_k, k_row, type_v, v_row, k.data(), v.data()) == size);
    CHECK(b200_kv_state_seq_write(nullptr, 0, n_cells, n_layer, type_k, k_row, type_v, v_row, k.data(), v.data()) == -size);

    auto parse = [&](const std::vector<uint8_t> & buf, int64_t * nc, std::vector<int64_t> & ko, std::vector<int64_t> & vo) {
        uint8_t * exact = (uint8_t *) std::malloc(buf.size() ? buf.size() : 1);       // exactly as long as the stream
        if (!buf.empty()) std::memcpy(exact, buf.data(), buf.size());
        ko.assign((size_t) n_layer, -1); vo.assign((size_t) n_layer, -1);
        const int32_t rc = b200_kv_state_seq_parse(exact, (int64_t) buf.size(), n_layer, type_k, k_row, type_v, v_row, nc, ko.data(), vo.data());
        std::free(exact);
        return rc;
    };
    int64_t nc = 0; std::vector<int64_t> ko, vo;
    CHECK(parse(good, &nc, ko, vo) == 0 && nc == n_cells);
    for (int l = 0; l < n_layer; ++l) {
        CHECK(ko[(size_t) l] >= 0 && ko[(size_t) l] + n_cells*k_row <= size && vo[(size_t) l] >= 0 && vo[(size_t) l] + n_cells*v_row <= size);
        CHECK(std::memcmp(good.data() + ko[(size_t) l], k.data() + (size_t) l*n_cells*k_row, (size_t) (n_cells*k_row)) == 0);
        CHECK(std::memcmp(good.data() + vo[(size_t) l], v.data() + (size_t) l*n_cells*v_row, (size_t) (n_cells*v_row)) == 0);
    }
    long rejected = 0, accepted = 0;
    for (int64_t len = 0; len < size; len += (len < 2048 ? 1 : 257)) {                // every short prefix, then strided
        std::vector<uint8_t> cut(good.begin(), good.begin() + len);
        const int32_t rc = parse(cut, &nc, ko, vo);
        CHECK(rc == B200_STATE_ERR_TRUNCATED);
        ++rejected;
    }
    const int64_t meta = 4 + n_cells*8 + 8;                                           // cell count, (pos, n_seq_id) per cell, v_trans + n_layer
    for (int it = 0; it < 4000; ++it) {
        std::vector<uint8_t> bad = good;
        const int n_flip = 1 + (int) (rng() % 3);
        for (int f = 0; f < n_flip; ++f) {
            // mostly the metadata and the per-layer headers, where every byte steers the parser
            const int64_t at = (rng() % 4) ? (int64_t) (rng() % (uint64_t) (meta + 64)) : (int64_t) (rng() % (uint64_t) size);
            bad[(size_t) at] = (uint8_t) rng();
        }
        if (rng() % 5 == 0) bad.resize((size_t) (rng() % (uint64_t) size));
        const int32_t rc = parse(bad, &nc, ko, vo);
        CHECK(rc <= 0 && rc >= B200_STATE_ERR_ROW_SIZE);
        if (rc == 0) {
            ++accepted;
            CHECK(nc >= 0);
            for (int l = 0; l < n_layer; ++l)
                CHECK(ko[(size_t) l] >= 0 && ko[(size_t) l] + nc*k_row <= (int64_t) bad.size() && vo[(size_t) l] >= 0 && vo[(size_t) l] + nc*v_row <= (int64_t) bad.size());
        } else ++rejected;
    }
    std::printf("rejected %ld, accepted %ld (flips inside row data)\nOK\n", rejected, accepted);
    return 0;
}
