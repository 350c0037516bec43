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
    std::printf("rejected %ld, accepted %ld (flips inside row data)\n", rejected, accepted);

    // ---- the whole-cache stream (state_write / state_read with seq_id == -1): same treatment, through the restore entry
    b200_paged_cache * pc = b200_pc_create(64, 4);
    std::vector<int32_t> slots(64); int32_t cow[2];
    CHECK(b200_pc_append(pc, 2, 21, slots.data(), cow, 1) == 0 && b200_pc_seq_cp(pc, 2, 5, 0, 13) == 1);
    CHECK(b200_pc_append(pc, 5, 6, slots.data(), cow, 1) >= 0 && b200_pc_append(pc, 0, 9, slots.data(), cow, 1) == 0);
    std::vector<int32_t> used(256);
    const int64_t n_used = b200_pc_used_cells(pc, used.data(), 256);
    CHECK(n_used == 21 + 6 + 1 + 9);                          // seq 5 shares 12 cells, copies the 13th on write, adds 6
    std::vector<uint8_t> ka((size_t) (n_layer*n_used*k_row)), va((size_t) (n_layer*n_used*v_row));
    for (auto & b : ka) b = (uint8_t) rng();
    const int64_t all_size = -b200_kv_state_all_write(nullptr, 0, pc, n_layer, type_k, k_row, type_v, v_row, nullptr, nullptr);
    std::vector<uint8_t> all((size_t) all_size);
    CHECK(b200_kv_state_all_write(all.data(), all_size, pc, n_layer, type_k, k_row, type_v, v_row, ka.data(), va.data()) == all_size);
    auto restore = [&](const std::vector<uint8_t> & buf, b200_paged_cache * into) {
        uint8_t * exact = (uint8_t *) std::malloc(buf.size() ? buf.size() : 1);
        if (!buf.empty()) std::memcpy(exact, buf.data(), buf.size());
        std::vector<int32_t> src(256), dst(256);
        ko.assign((size_t) n_layer, -1); vo.assign((size_t) n_layer, -1);
        const int64_t rc = b200_kv_state_all_restore(into, exact, (int64_t) buf.size(), n_layer, type_k, k_row, type_v, v_row,
                                                     src.data(), dst.data(), 256, &nc, ko.data(), vo.data());
        if (rc >= 0) {
            for (int64_t i = 0; i < rc; ++i) CHECK(src[(size_t) i] >= 0 && src[(size_t) i] < nc && dst[(size_t) i] >= 0 && dst[(size_t) i] < 64*4);
            for (int l = 0; l < n_layer; ++l) CHECK(ko[(size_t) l] >= 0 && ko[(size_t) l] + nc*k_row <= (int64_t) buf.size());
        }
        std::free(exact);
        return rc;
    };
    b200_paged_cache * into = b200_pc_create(64, 4);
    CHECK(restore(all, into) == 21 + (6 + 1 + 12 - 12) + 9 && b200_pc_seq_len(into, 5) == 19 && b200_pc_ref_count(into, 3) == 2);   // ids ascend: seq 0 took blocks 0..2
    long rej2 = 0, acc2 = 0;
    for (int64_t len = 0; len < all_size; len += (len < 1024 ? 1 : 389)) {
        std::vector<uint8_t> cut(all.begin(), all.begin() + len);
        CHECK(restore(cut, into) == B200_STATE_ERR_TRUNCATED);
        ++rej2;
    }
    const int64_t meta2 = 4 + n_used*12 + 64;
    for (int it = 0; it < 4000; ++it) {
        std::vector<uint8_t> bad = all;
        for (int f = 1 + (int) (rng() % 3); f > 0; --f) bad[(size_t) (rng() % (uint64_t) meta2)] = (uint8_t) rng();
        if (rng() % 6 == 0) bad.resize((size_t) (rng() % (uint64_t) all_size));
        const int64_t rc = restore(bad, into);
        CHECK(rc >= 0 || (rc <= B200_STATE_ERR_TRUNCATED && rc >= B200_STATE_ERR_SEQ_ID));
        (rc >= 0 ? acc2 : rej2) += 1;
    }
    b200_pc_destroy(pc); b200_pc_destroy(into);
    std::printf("whole-cache stream: rejected %ld, accepted %ld\nOK\n", rej2, acc2);
    return 0;
}
