// paged_cache_stress.cpp — random call streams over PagedKVCache (append / prepare_batch / seq_cp / seq_rm / seq_keep /
// context_shift / defrag), compiled with -fsanitize=address,undefined by tests/test_kv_cache_paged.py.  After every call the
// reference counts are recomputed from the block tables and compared, and the free count must close the books —
// the invariant llama_kv_cache_unified keeps per cell (`used` == number of cells with a non-empty seq set,
// src/llama-kv-cache.cpp:199-240) restated for shared blocks.
#include "paged_kv_cache.h"

#include <cstdio>
#include <cstdlib>
#include <random>

#define CHECK(c) do { if (!(c)) { std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); std::exit(1); } } while (0)

static void audit(const PagedKVCache & pc) {
    std::vector<int> refs((size_t) pc.n_blocks(), 0);
    for (PagedKVCache::seq_id s : pc.seq_ids()) {
        const std::vector<int32_t> * b = pc.blocks_of(s);
        CHECK(b != nullptr);
        const int64_t len = pc.seq_len(s), bs = pc.block_size();
        CHECK((int64_t) b->size() == (len + bs - 1)/bs);                  // no block beyond the last token
        for (int32_t x : *b) { CHECK(x >= 0 && x < pc.n_blocks()); ++refs[(size_t) x]; }
    }
    int64_t free_blocks = 0;
    for (int64_t i = 0; i < pc.n_blocks(); ++i) { CHECK(pc.ref_count(i) == refs[(size_t) i]); free_blocks += refs[(size_t) i] == 0; }
    CHECK(free_blocks == pc.n_free());
}

int main() {
    for (unsigned seed = 1; seed <= 12; ++seed) {
        std::mt19937 rng(seed);
        auto R = [&](int n) { return (int) (rng() % (unsigned) n); };
        const int bs = 1 << R(5);
        PagedKVCache pc(6 + R(40), bs);                       // small pools: the refusal paths run often
        long calls = 0, refused = 0;
        for (int step = 0; step < 1500; ++step, ++calls) {
            std::vector<int32_t> slots, shift; std::vector<std::pair<int32_t, int32_t>> cow, moves;
            const int s = R(6);
            switch (R(9)) {
            case 0: case 1: refused += !pc.append(s, R(3*bs + 2), slots, cow); break;
            case 2: case 3: {
                const int n = 1 + R(3*bs + 4);
                std::vector<int32_t> seq((size_t) n), pos;
                for (int & x : seq) x = R(6);
                PagedKVCache::BatchInputs in;
                const bool bad_pos = R(8) == 0;
                if (bad_pos) { pos.assign((size_t) n, 0); for (int i = 0; i < n; ++i) pos[(size_t) i] = (int32_t) pc.seq_len(seq[(size_t) i]) + 1 + R(3); }
                const int rc = pc.prepare_batch(n, seq.data(), bad_pos ? pos.data() : nullptr, in);
                CHECK(bad_pos ? rc == -1 : (rc == 0 || rc == 1));
                if (rc == 0 && R(5) == 0) {                               // the graph failed: llama_kv_slot_restorer::restore
                    std::vector<int64_t> before(6);
                    for (int q = 0; q < 6; ++q) before[(size_t) q] = pc.seq_len(q);
                    for (int i = 0; i < n; ++i) --before[(size_t) seq[(size_t) i]];
                    pc.undo_batch(in);
                    for (int q = 0; q < 6; ++q) CHECK(pc.seq_len(q) == before[(size_t) q]);
                } else if (rc == 0) {
                    CHECK((int) in.slot_mapping.size() == n && in.q_seq.size() == in.q_pos.size());
                    for (int i = 0; i < n; ++i) {                         // the cell of token i is the cell of its position in its row
                        const std::vector<int32_t> * b = pc.blocks_of(seq[(size_t) i]);
                        const int32_t p = in.q_pos[(size_t) i];
                        CHECK(in.seqs[(size_t) in.q_seq[(size_t) i]] == seq[(size_t) i]);
                        CHECK(in.slot_mapping[(size_t) i] == (*b)[(size_t) (p/bs)]*bs + p % bs);
                    }
                } else {
                    refused += rc == 1;
                    CHECK(in.slot_mapping.empty() && in.cow_pairs.empty());
                }
                break;
            }
            case 4: if (pc.has_seq(s)) { const int64_t n = pc.seq_len(s); pc.seq_cp(s, R(6), 0, n ? 1 + R((int) n) : 0); } break;
            case 5: if (pc.has_seq(s)) { const int64_t n = pc.seq_len(s); pc.seq_rm(s, n ? R((int) n + 1) : 0, -1); } break;
            case 6: if (R(6) == 0) pc.seq_keep(s); else pc.seq_rm(s, 0, -1); break;
            case 7: if (pc.seq_len(s) > 2) {
                const int64_t n = pc.seq_len(s), p0 = R((int) n - 1), p1 = p0 + 1 + R((int) (n - p0 - 1));
                refused += !pc.context_shift(s, p0, p1, cow, moves, shift);
            } break;
            default: pc.defrag(cow, R(2) ? -1 : 1 + R(4)); break;
            }
            for (auto & pr : cow) CHECK(pr.first != pr.second && pr.first >= 0 && pr.second >= 0);
            audit(pc);
        }
        pc.seq_rm(-1, 0, -1);
        audit(pc);
        CHECK(pc.n_free() == pc.n_blocks());
        std::printf("seed %u: block %d, %ld calls, %ld refused for lack of blocks\n", seed, bs, calls, refused);
    }
    std::printf("OK\n");
    return 0;
}
