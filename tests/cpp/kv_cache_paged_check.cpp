// tests/cpp/kv_cache_paged_check.cpp — llama_kv_cache_paged against a cell-per-token model of llama_kv_cache_unified.
//
// The model is the reference's bookkeeping restated in a few lines (src/llama-kv-cache.cpp: find_slot :622-671 hands out
// cells, seq_cp :244-293 inserts the id, seq_rm :173-242 erases it, seq_add :327-383 moves positions, seq_pos_max :437-447,
// get_n_tokens :126-134) with one token id per cell.  Both sides run the same seeded stream of llama_kv_self_* calls in the
// order the reference's callers use them (kv_self_update before every decode); after every update the adapter's work lists
// are applied to an array of token ids the way the device applies them to K/V rows, and every sequence must read back the
// model's tokens in position order.  Exit code 0 = all checks passed.
#include "llama_kv_cache_paged.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <random>
#include <vector>

#define CHECK(c) do { if (!(c)) { std::fprintf(stderr, "CHECK failed %s:%d: %s\n", __FILE__, __LINE__, #c); std::exit(1); } } while (0)

struct Tok { long id; int rot; };                            // a cached token and the K rotation its row has received so far
struct Model {                                               // seq -> tokens in position order (positions are 0..n-1)
    std::map<int, std::vector<Tok>> seqs;
    int n_tokens() const { int n = 0; for (auto & kv : seqs) n += (int) kv.second.size(); return n; }
};

static void apply(const llama_kv_cache_paged::pending_work & w, std::vector<long> & store, int bs, std::vector<int> & rotated) {
    // group by group, each to its end, as the device runs them (pending_work): copies, gather-all, scatter-all, rotation
    CHECK(w.shift_cells.size() == w.shift_delta.size());
    CHECK(w.groups.empty() ? (w.cow_pairs.empty() && w.row_moves.empty() && w.shift_cells.empty())
                           : (w.groups.back().cow_end == w.cow_pairs.size() && w.groups.back().move_end == w.row_moves.size()
                              && w.groups.back().shift_end == w.shift_cells.size()));
    size_t c0 = 0, m0 = 0, s0 = 0;
    for (auto & g : w.groups) {
        CHECK(g.cow_end >= c0 && g.move_end >= m0 && g.shift_end >= s0);
        for (size_t k = c0; k < g.cow_end; ++k) for (int i = 0; i < bs; ++i) {
            store[(size_t) w.cow_pairs[k].second*bs + i] = store[(size_t) w.cow_pairs[k].first*bs + i];
            rotated[(size_t) w.cow_pairs[k].second*bs + i] = rotated[(size_t) w.cow_pairs[k].first*bs + i];
        }
        std::vector<long> staged; std::vector<int> staged_rot;   // gather every source, then scatter
        for (size_t k = m0; k < g.move_end; ++k) { staged.push_back(store[(size_t) w.row_moves[k].first]); staged_rot.push_back(rotated[(size_t) w.row_moves[k].first]); }
        for (size_t k = m0; k < g.move_end; ++k) { store[(size_t) w.row_moves[k].second] = staged[k - m0]; rotated[(size_t) w.row_moves[k].second] = staged_rot[k - m0]; }
        for (size_t k = s0; k < g.shift_end; ++k) rotated[(size_t) w.shift_cells[k]] += w.shift_delta[k];
        c0 = g.cow_end; m0 = g.move_end; s0 = g.shift_end;
    }
    std::vector<long> s2 = store; std::vector<int> r2 = rotated;
    for (auto & d : w.defrag_moves) for (int i = 0; i < bs; ++i) {
        s2[(size_t) d.second*bs + i] = store[(size_t) d.first*bs + i];
        r2[(size_t) d.second*bs + i] = rotated[(size_t) d.first*bs + i];
    }
    store.swap(s2); rotated.swap(r2);
}

// Several context shifts closed between two update() calls (llama-server shifts several slots before one decode).  Both
// cases used to corrupt rows silently when the lists of all shifts were executed phase by phase.
static void scenario_shifts_between_updates() {
    auto read_back = [](llama_kv_cache_paged & kv, int seq, const std::vector<long> & store) {
        std::vector<int32_t> cells; std::vector<long> out;
        CHECK(kv.blocks().slots_of(seq, cells));
        for (int32_t c : cells) out.push_back(store[(size_t) c]);
        return out;
    };
    {   // 1: a block the first shift frees is the copy-on-write destination of the second while still a source of the first's moves
        llama_kv_cache_paged kv(8, 4);
        std::vector<long> store(32, -1); std::vector<int> rot(32, 0);
        std::vector<int32_t> sl; std::vector<std::pair<int32_t, int32_t>> cw;
        CHECK(kv.find_slot(0, 10, sl, cw)); for (size_t i = 0; i < sl.size(); ++i) store[(size_t) sl[i]] = (long) i;
        sl.clear(); CHECK(kv.find_slot(1, 8, sl, cw));  for (size_t i = 0; i < sl.size(); ++i) store[(size_t) sl[i]] = 100 + (long) i;
        kv.seq_cp(1, 2, 0, 8);
        CHECK(kv.seq_rm(0, 1, 4)); kv.seq_add(0, 4, -1, -3);
        CHECK(kv.seq_rm(1, 1, 2)); kv.seq_add(1, 2, -1, -1);
        CHECK(!kv.unsupported());
        llama_kv_cache_paged::pending_work w;
        CHECK(kv.update(w) && w.groups.size() == 2);
        apply(w, store, 4, rot);
        CHECK(read_back(kv, 0, store) == std::vector<long>({0, 4, 5, 6, 7, 8, 9}));
        CHECK(read_back(kv, 1, store) == std::vector<long>({100, 102, 103, 104, 105, 106, 107}));
        CHECK(read_back(kv, 2, store) == std::vector<long>({100, 101, 102, 103, 104, 105, 106, 107}));
    }
    {   // 2: a fork between two shifts shares blocks the first shift has yet to write
        llama_kv_cache_paged kv(8, 4);
        std::vector<long> store(32, -1); std::vector<int> rot(32, 0);
        std::vector<int32_t> sl; std::vector<std::pair<int32_t, int32_t>> cw;
        CHECK(kv.find_slot(0, 10, sl, cw)); for (size_t i = 0; i < sl.size(); ++i) store[(size_t) sl[i]] = (long) i;
        CHECK(kv.seq_rm(0, 1, 3)); kv.seq_add(0, 3, -1, -2);                 // 0 3 4 5 6 7 8 9
        kv.seq_cp(0, 1, 0, -1);
        CHECK(kv.seq_rm(1, 1, 2)); kv.seq_add(1, 2, -1, -1);                 // 0 4 5 6 7 8 9
        CHECK(kv.seq_rm(0, 0, 1)); kv.seq_add(0, 1, -1, -1);                 // the same sequence a second time: 3 4 5 6 7 8 9
        CHECK(!kv.unsupported());
        llama_kv_cache_paged::pending_work w;
        CHECK(kv.update(w) && w.groups.size() == 3);
        apply(w, store, 4, rot);
        CHECK(read_back(kv, 0, store) == std::vector<long>({3, 4, 5, 6, 7, 8, 9}));
        CHECK(read_back(kv, 1, store) == std::vector<long>({0, 4, 5, 6, 7, 8, 9}));
    }
    {   // 3: a sequence dropped while its shift is pending: the stale scatters must land before a later copy reuses the block
        llama_kv_cache_paged kv(6, 4);
        std::vector<long> store(24, -1); std::vector<int> rot(24, 0);
        std::vector<int32_t> sl; std::vector<std::pair<int32_t, int32_t>> cw;
        CHECK(kv.find_slot(0, 7, sl, cw));  for (size_t i = 0; i < sl.size(); ++i) store[(size_t) sl[i]] = (long) i;
        sl.clear(); CHECK(kv.find_slot(1, 8, sl, cw));  for (size_t i = 0; i < sl.size(); ++i) store[(size_t) sl[i]] = 100 + (long) i;
        kv.seq_cp(1, 2, 0, 8);
        CHECK(kv.seq_rm(0, 1, 2)); kv.seq_add(0, 2, -1, -1);
        CHECK(kv.seq_rm(0, -1, -1));                                         // sequence 0 goes away, its blocks are free again
        CHECK(kv.seq_rm(1, 2, 3)); kv.seq_add(1, 3, -1, -1);                 // copy-on-write into the freed blocks
        CHECK(!kv.unsupported());
        llama_kv_cache_paged::pending_work w;
        CHECK(kv.update(w));
        apply(w, store, 4, rot);
        CHECK(read_back(kv, 1, store) == std::vector<long>({100, 101, 103, 104, 105, 106, 107}));
        CHECK(read_back(kv, 2, store) == std::vector<long>({100, 101, 102, 103, 104, 105, 106, 107}));
    }
}

static void scenario_basic() {
    llama_kv_cache_paged kv(8, 4);
    std::vector<int32_t> slots; std::vector<std::pair<int32_t, int32_t>> cow;
    CHECK(kv.get_can_shift() && kv.seq_pos_max(0) == 0 && kv.get_n_tokens() == 0);
    CHECK(kv.find_slot(0, 10, slots, cow) && slots.size() == 10 && cow.empty());
    CHECK(kv.seq_pos_max(0) == 9 && kv.get_n_tokens() == 10 && kv.get_used_cells() == 12);
    kv.seq_cp(0, 1, 0, 6);
    CHECK(!kv.unsupported() && kv.seq_pos_max(1) == 5 && kv.get_n_tokens() == 16 && kv.get_used_cells() == 12);
    // context shift: keep 2, discard 4 — the reference's pair of calls
    CHECK(kv.seq_rm(0, 2, 6));
    CHECK(kv.seq_pos_max(0) == 9);                           // nothing happened yet: the removal is open
    kv.seq_add(0, 6, -1, -4);
    CHECK(!kv.unsupported() && kv.has_pending() && kv.seq_pos_max(0) == 5);
    slots.clear();
    CHECK(!kv.find_slot(0, 1, slots, cow));                  // decode before kv_self_update: refused
    llama_kv_cache_paged kv2(8, 4);                          // (the refusal poisons the instance; start over for the rest)
    slots.clear(); cow.clear();
    CHECK(kv2.find_slot(0, 10, slots, cow));
    kv2.seq_cp(0, 1, 0, 6);
    CHECK(kv2.seq_rm(0, 2, 6));
    kv2.seq_add(0, 6, -1, -4);
    llama_kv_cache_paged::pending_work w;
    CHECK(kv2.update(w) && !w.empty() && !kv2.has_pending());
    CHECK(w.shift_cells.size() == 4 && w.shift_delta == std::vector<int32_t>(4, -4));
    CHECK(!w.cow_pairs.empty());                             // block 0 is shared with sequence 1 and gets written
    CHECK(kv2.seq_pos_max(1) == 5);
    // requests outside the block-table model
    kv2.seq_div(0, 0, -1, 2);
    CHECK(kv2.unsupported() && !kv2.update(w) && w.empty());
    kv2.clear();
    CHECK(!kv2.unsupported() && kv2.get_n_tokens() == 0 && kv2.get_used_cells() == 0);
    CHECK(kv2.find_slot(3, 5, slots, cow));
    kv2.seq_add(3, 2, -1, 7);                                // a shift nothing asked a removal for
    CHECK(kv2.unsupported());
    kv2.clear();
    CHECK(kv2.find_slot(3, 9, slots, cow) && kv2.seq_rm(3, 1, 3));
    CHECK(!kv2.update(w));                                   // open removal never closed
    // the decode loop's fragmentation check: small contexts never defragment; a hole under a high block does above the threshold
    {
        llama_kv_cache_paged big(512, 16);                   // 8192 cells
        std::vector<int32_t> sl; std::vector<std::pair<int32_t, int32_t>> cw;
        CHECK(big.find_slot(0, 100, sl, cw) && !big.defrag_if_fragmented(0.1f));          // n = 256 < 2048
        CHECK(big.find_slot(1, 6000, sl, cw) && big.find_slot(2, 160, sl, cw));
        CHECK(!big.defrag_if_fragmented(0.1f));                                            // dense: nothing to gain
        CHECK(big.seq_rm(1, -1, -1) && !big.defrag_if_fragmented(-1.0f));                  // disabled (the default, llama-context.cpp:2250)
        CHECK(big.defrag_if_fragmented(0.1f) && big.has_pending());                        // 272 used cells under block 391
        llama_kv_cache_paged::pending_work pw;
        CHECK(big.update(pw) && pw.defrag_moves.size() == 10 && !big.defrag_if_fragmented(0.1f));
    }
    // the ubatch form: interleaved sequences, per-row inputs, refused while a context shift is pending
    llama_kv_cache_paged kv3(8, 4);
    PagedKVCache::BatchInputs in;
    const llama_seq_id sq[5] = {2, 5, 2, 5, 2}; const llama_pos ps[5] = {0, 0, 1, 1, 2};
    CHECK(kv3.find_slot(5, sq, ps, in) == 0 && in.seqs == std::vector<int32_t>({2, 5}));
    CHECK(in.q_seq == std::vector<int32_t>({0, 1, 0, 1, 0}) && in.q_pos == std::vector<int32_t>({0, 0, 1, 1, 2}));
    CHECK(in.slot_mapping == std::vector<int32_t>({0, 4, 1, 5, 2}) && kv3.seq_pos_max(2) == 2 && kv3.seq_pos_max(5) == 1);
    const llama_pos hole[1] = {7};
    CHECK(kv3.find_slot(1, sq, hole, in) == -1 && kv3.seq_pos_max(2) == 2);
    CHECK(kv3.seq_rm(2, 0, 1)); kv3.seq_add(2, 1, -1, -1);   // context shift of sequence 2, not applied yet
    CHECK(kv3.find_slot(1, sq + 1, nullptr, in) == 1 && kv3.unsupported());
}

static void scenario_random(unsigned seed) {
    std::mt19937 rng(seed);
    const int bs = 1 << (rng() % 5), n_blocks = 96;
    llama_kv_cache_paged kv(n_blocks, bs);
    Model model;
    std::vector<long> store((size_t) n_blocks*bs, -1);
    std::vector<int> rotated((size_t) n_blocks*bs, 0);       // accumulated K rotation per cell
    long next_tok = 0;
    auto R = [&](int n) { return (int) (rng() % (unsigned) n); };
    int shifts = 0, defrags = 0, multi = 0;
    for (int step = 0; step < 200; ++step) {
        // kv_self_update before every decode, as llama_context::decode does; between two decodes any number of
        // shifts, forks, removals and defrag requests may pile up (llama-server edits several slots, then decodes once)
        const int op = R(10);
        std::vector<int> live0; for (auto & s : model.seqs) live0.push_back(s.first);
        if (op < 4 || live0.empty() || R(3) == 0) {
            llama_kv_cache_paged::pending_work w;
            CHECK(kv.update(w));
            apply(w, store, bs, rotated);
            defrags += !w.defrag_moves.empty();
            multi += w.groups.size() > 1;
        }
        std::vector<int> live; for (auto & s : model.seqs) live.push_back(s.first);
        if (op < 4 || live.empty()) {                        // decode: append tokens
            const int seq = R(5), n = 1 + R(24);
            std::vector<int32_t> slots; std::vector<std::pair<int32_t, int32_t>> cow;
            if (!kv.find_slot(seq, n, slots, cow)) { CHECK(!kv.unsupported()); continue; }   // only ever "no room"
            for (auto & c : cow) for (int i = 0; i < bs; ++i) {
                store[(size_t) c.second*bs + i] = store[(size_t) c.first*bs + i];
                rotated[(size_t) c.second*bs + i] = rotated[(size_t) c.first*bs + i];
            }
            for (int32_t c : slots) { store[(size_t) c] = next_tok; rotated[(size_t) c] = 0; model.seqs[seq].push_back(Tok{next_tok++, 0}); }
        } else if (op < 6) {                                 // fork a prefix
            const int src = live[(size_t) R((int) live.size())], dst = R(5);
            if (dst == src) continue;
            const int p1 = 1 + R((int) model.seqs[src].size());
            kv.seq_cp(src, dst, 0, p1);
            CHECK(!kv.unsupported());
            model.seqs[dst].assign(model.seqs[src].begin(), model.seqs[src].begin() + p1);
        } else if (op < 8) {                                 // context shift
            const int seq = live[(size_t) R((int) live.size())], n = (int) model.seqs[seq].size();
            if (n < 3) continue;
            const int p0 = R(n - 2), p1 = p0 + 1 + R(n - p0 - 2);        // p1 < n: something survives behind the gap
            const int64_t free_before = kv.blocks().n_free();
            CHECK(kv.seq_rm(seq, p0, p1));
            kv.seq_add(seq, p1, -1, p0 - p1);
            if (kv.unsupported()) { CHECK(free_before < n_blocks); kv.clear(); model = Model(); continue; }   // no room for the copies
            auto & v = model.seqs[seq];
            for (int q = p1; q < n; ++q) v[(size_t) q].rot += p0 - p1;
            v.erase(v.begin() + p0, v.begin() + p1);
            ++shifts;
        } else if (op == 8) {                                // truncate or drop
            const int seq = live[(size_t) R((int) live.size())], n = (int) model.seqs[seq].size();
            const int p0 = R(n + 1);
            CHECK(kv.seq_rm(seq, p0, -1));
            model.seqs[seq].resize((size_t) p0);
            if (p0 == 0) model.seqs.erase(seq);
        } else {
            kv.defrag();
        }
        // bookkeeping agrees with the model at every step (seq_pos_max, token count)
        CHECK(kv.get_n_tokens() == model.n_tokens());
        for (int s = 0; s < 5; ++s) CHECK(kv.seq_pos_max(s) == (model.seqs.count(s) && !model.seqs[s].empty() ? (int) model.seqs[s].size() - 1 : 0));
    }
    llama_kv_cache_paged::pending_work w;
    CHECK(kv.update(w));
    apply(w, store, bs, rotated);
    // every sequence reads back its own tokens in order, each row rotated by exactly the deltas of ITS sequence's shifts:
    // a row shared with another sequence is copied before it is rotated (the reference rotates the shared cell for every
    // owner, src/llama-kv-cache.cpp:347-366; with blocks the other sequence keeps its rows)
    for (auto & s : model.seqs) {
        std::vector<int32_t> cells;
        CHECK(kv.blocks().slots_of(s.first, cells) && cells.size() == s.second.size());
        for (size_t i = 0; i < cells.size(); ++i) {
            CHECK(store[(size_t) cells[i]] == s.second[i].id);
            CHECK(rotated[(size_t) cells[i]] == s.second[i].rot);
        }
    }
    // ref-counts: every block is referenced by exactly the sequences whose tables name it
    std::vector<int> refs((size_t) n_blocks, 0);
    for (auto & s : model.seqs) {
        const std::vector<int32_t> * b = kv.blocks().blocks_of(s.first);
        CHECK(b != nullptr);
        for (int32_t x : *b) ++refs[(size_t) x];
    }
    int64_t free_blocks = 0;
    for (int b = 0; b < n_blocks; ++b) { CHECK(kv.blocks().ref_count(b) == refs[(size_t) b]); free_blocks += refs[(size_t) b] == 0; }
    CHECK(kv.blocks().n_free() == free_blocks);
    std::printf("seed %u: block %d, %d context shifts, %d updates with several shift groups, %d defrags, %d tokens cached\n", seed, bs, shifts, multi, defrags, model.n_tokens());
}

int main() {
    scenario_basic();
    scenario_shifts_between_updates();
    for (unsigned seed = 1; seed <= 24; ++seed) scenario_random(seed);
    std::printf("OK\n");
    return 0;
}
