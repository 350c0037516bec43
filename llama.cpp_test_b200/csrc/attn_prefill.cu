// attn_prefill.cu — chunked-prefill paged attention on the 5th-generation tensor cores (tcgen05 + TMEM).
//
// Replaces flash_attn_ext_f16 (the mma.sync tile kernel, ggml/src/ggml-cuda/fattn-mma-f16.cuh:804-936) for the
// many-query case: 512-token chunks appended into the paged cache (BASELINE config 3).  Semantics are those of
// ggml_compute_forward_flash_attn_ext_f16 (ggml/src/ggml-cpu/ggml-cpu.c:12219-12432) restricted to what a block
// table expresses: query row t of sequence s sees positions [0, min(seq_lens[s], q_pos[t]+1)); the causal limit is
// applied in-kernel, no dense mask is read (the reference fills and reads an [n_kv, n_q] F16 mask,
// src/llama-graph.cpp:403-477).
//
// Work item = (row tile, kv head).  A row tile is up to 256/gqa consecutive query rows of ONE sequence (found on
// the device by pf_plan_kernel from q_seq, so any row order is legal); with the gqa query heads of the kv head
// that is 2 x (M = 128) MMA rows: row R = token R/gqa, head R%gqa, group R/128, TMEM lane R%128.  K/V of the
// sequence stream ONCE per work item in 64-token tiles (per-row stream-K would re-read them once per query row).
//
// CTA = 10 warps, one CTA per SM; the two 128-row groups share every K/V tile, so K/V cross L2 -> shared memory
// once per 256 rows and the tensor pipe works on one group while the other's softmax runs.
//   warps 0-7  softmax (4 per group): thread r owns accumulator row r of its group (TMEM lane r): tcgen05.ld S ->
//              causal mask -> running max (lazy: the reference point only moves when the max grows by > 2^8, so O
//              is rarely rescaled) -> exp2 -> P as packed F16 over the first 32 columns of its S buffer
//              (tcgen05.st); final O * 1/l -> dst
//   warp 8     issues tcgen05.mma: S[128x64] = Q K^T (8 x K16, A and B from shared memory, K-major),
//              O[128x128] += P V (4 x K16, P from TMEM, V as the MN-major B operand straight from its row-major
//              cache layout); S is double-buffered per group; order per kv tile: PV_A(j) QK_B(j+1) PV_B(j) QK_A(j+2)
//   warp 9     gathers K/V through the block table with TMA: one box = 16 cells x 64 halves of one head
//              (128-byte swizzle), 8 boxes per 64-token stage; K ring 4 stages, V ring 3
// TMEM: all 512 columns (per group S0 64, S1 64, O 128).  Shared: Q 2 x 32 KB + K 4 x 16 KB + V 3 x 16 KB = 176 KB.
// What bounds it is measured in profiles/r1p_microbench.txt (QK^T from shared memory: 384 cycles per block).
// Work items that do not fill the persistent grid's last wave are cut into pieces of consecutive kv tiles whose
// partial (m, l, O) pf_combine_kernel merges (PF_MAX_SPLIT below); the plan / expansion / main / merge kernels are
// chained by programmatic dependent launch, the CTA set-up of the main kernel running under its predecessors.
#ifdef PF_TIMELINE
#include <cstdio>
#endif
#include "common.cuh"
#include <cuda.h>
#include <math.h>
#include <climits>

namespace b200 {

constexpr int PF_ROWS    = 128;          // MMA rows per group; a work item is two groups (256 rows)
constexpr int PF_BN      = 64;           // kv tokens per tile
constexpr int PF_D       = 128;
constexpr int PF_THREADS = 320;          // 8 softmax warps (4 per group) + MMA warp + TMA warp
constexpr int PF_MMA_WARP = 8, PF_TMA_WARP = 9;
constexpr int PF_SQ      = 0;            // [2 d-halves][128 rows][128 B]
constexpr int PF_KST     = 4, PF_VST = 3; // ring stages (a K stage is free once QK^T of its tile retired, V after PV)
constexpr int PF_SK      = 65536;        // after Q [2 groups][2 d-halves][128 rows][128 B]; K stage s at + s*16384: [2 d-halves][64 tokens][128 B]
constexpr int PF_SV      = PF_SK + PF_KST*16384;   // V stage s at + s*16384, same layout
constexpr int PF_SBAR    = PF_SV + PF_VST*16384;
constexpr int PF_SMEM    = PF_SBAR + 256;
// KV split of the tail: when the work items do not fill the last wave of the persistent grid (n_work % grid != 0, or
// fewer items than SMs: ONE sequence's 512-token chunk is 16 items), each of the leftover items is cut into up to
// PF_MAX_SPLIT pieces of consecutive kv tiles so that the pieces fill the machine; a piece leaves (m, l, unnormalised O)
// per row in the workspace and pf_combine_kernel merges them.  Pieces are never shorter than PF_MIN_PIECE tiles (a piece
// pays its own Q load, partial store and share of the merge, ~10 us: with 4-tile pieces the 512-row prompt chunks of
// llama-batched-bench at ctx <= 1k got 6 % slower); an item too short for two pieces runs whole and writes dst itself,
// the units of its other piece slots find no tiles and write nothing.
#ifndef PF_SPLIT_MAX
#define PF_SPLIT_MAX 8
#endif
#ifndef PF_PIECE_MIN
#define PF_PIECE_MIN 16
#endif
constexpr int PF_MAX_SPLIT = PF_SPLIT_MAX, PF_MIN_PIECE = PF_PIECE_MIN;

// mbarriers: SFULL[group][S buffer], PFULL[group][P buffer] (128 arrivals; the softmax may be a tile ahead of the MMA warp), ODONE[group]
enum { PF_FULLK = 0, PF_FULLV = 4, PF_EMPTYK = 7, PF_EMPTYV = 11, PF_SFULL = 14, PF_PFULL = 18, PF_ODONE = 22, PF_NBAR = 24 };

struct pf_args {
    b200_attn_desc d;
    int    gqa, tok_per_tile;            // tok_per_tile = 128/gqa
    int   *plan;                         // [0] n_tiles, [1] work counter, [2..] tile_row0[n_q], then tile_rows[n_q]
    float  qscale;                       // scale*log2(e)
    // quantised caches: the sequences' cells are expanded once per call to a dense F16 scratch (sequence s at cells
    // seq_off[s] .. + pad64(seq_len)), and the tile kernel reads that instead of the paged cache (seq_off == NULL: paged)
    int    *seq_off;                     // [n_seqs + 1]
    __half *k16, *v16;                   // [scratch_cells][n_head_kv][128]
    int64_t scratch_cells;
    int64_t cap;                         // upper bound applied to every seq_lens entry (block-table capacity, or max_seq_len)
    float  *part_o, *part_ml;            // KV-split pieces: [grid_main][256 rows][128], [grid_main][256 rows][2] {m (log2 domain), l}
    int     grid_main;                   // CTAs of pf_main_kernel (the combine kernel derives the same split from it)
    int     split_max;                   // PF_MAX_SPLIT, or 1 when the host knows no sequence is long enough for two pieces (no combine launch)
#ifdef PF_TIMELINE
    long long *dbg;                      // debug builds only: clock64 stamps of CTA 0's first work item, [3 warps][64 tiles][8]
#endif
};

#ifdef PF_TIMELINE
#define PF_T(w, j, k) do { if (blockIdx.x == 0 && first_item && (threadIdx.x & 31) == 0 && (j) < 64) a.dbg[((w)*64 + (j))*8 + (k)] = clock64(); } while (0)
#else
#define PF_T(w, j, k) do { } while (0)
#endif

__device__ __forceinline__ uint32_t pf_smem_u32(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ void pf_mbar_init(uint32_t bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void pf_mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void pf_mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void pf_mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile("{\n .reg .pred p;\nPF_WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra PF_DONE_%=;\n bra PF_WAIT_%=;\nPF_DONE_%=:\n}"
                 :: "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void pf_tma_2d(uint32_t dst, const CUtensorMap *tm, int c0, int c1, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 :: "r"(dst), "l"(tm), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
// both d-halves of one 16-cell block, issued by ONE elected lane of a converged warp: with elect.sync ptxas knows a
// single thread runs the copies and emits plain UTMALDG (an `if (lane == 0)` costs an 11-instruction vote loop each)
__device__ __forceinline__ void pf_tma_pair_elect(uint32_t dst, const CUtensorMap *tm, int c0, int cell, uint32_t bar) {
    asm volatile("{\n .reg .pred p;\n .reg .b32 d1, c1;\n elect.sync _|p, 0xffffffff;\n add.u32 d1, %0, 8192;\n add.s32 c1, %2, 64;\n"
                 " @p cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];\n"
                 " @p cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [d1], [%1, {c1, %3}], [%4];\n}"
                 :: "r"(dst), "l"(tm), "r"(c0), "r"(cell), "r"(bar) : "memory");
}
__device__ __forceinline__ void pf_fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void pf_tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void pf_tc_fence_after()  { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// shared-memory matrix descriptor, 128-byte swizzle (cute/arch/mma_sm100_desc.hpp SmemDescriptor: start >> 4 at
// [0,14), LBO >> 4 at [16,30), SBO >> 4 at [32,46), version 1 at [46,48), layout SWIZZLE_128B = 2 at [61,64))
__device__ __forceinline__ uint64_t pf_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t) ((addr & 0x3FFFFu) >> 4) | ((uint64_t) (lbo_bytes >> 4) << 16) | ((uint64_t) (sbo_bytes >> 4) << 32)
         | (1ull << 46) | (2ull << 61);
}
// instruction descriptor, kind::f16: D F32 (bit 4), A/B F16, b_major at 16, N >> 3 at [17,23), M >> 4 at [24,29)
__host__ __device__ constexpr uint32_t pf_idesc(int m, int n, int b_mn_major) {
    return (1u << 4) | ((uint32_t) b_mn_major << 16) | ((uint32_t) (n >> 3) << 17) | ((uint32_t) (m >> 4) << 24);
}
// issued by ONE elected lane of a converged warp (elect.sync lets ptxas emit the uniform-datapath instruction directly)
__device__ __forceinline__ void pf_mma(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n .reg .pred p, e;\n elect.sync _|e, 0xffffffff;\n setp.ne.b32 p, %4, 0;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
                 :: "r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
// The 8 K16 steps of one S = Q K^T tile in ONE elected block: the descriptors of step ks are the step-0 descriptors plus
// constants (32 B per step inside a 64-element swizzle atom, then the second d-half: +16 KB for Q, +8 KB for K), so the
// issuing warp spends ~25 instructions per tile instead of ~12 per MMA (it was the critical path, profiles/r1o_*).
__device__ __forceinline__ void pf_qk_block(uint32_t d_tmem, uint64_t a0, uint64_t b0, uint32_t idesc) {
    asm volatile("{\n .reg .pred e, z;\n .reg .b64 a, b;\n elect.sync _|e, 0xffffffff;\n setp.ne.b32 z, 0, 0;\n"
                 " @e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, z;\n setp.eq.b32 z, 0, 0;\n"
                 " add.u64 a, %1, 2;\n add.u64 b, %2, 2;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %3, z;\n"
                 " add.u64 a, %1, 4;\n add.u64 b, %2, 4;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %3, z;\n"
                 " add.u64 a, %1, 6;\n add.u64 b, %2, 6;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %3, z;\n"
                 " add.u64 a, %1, 1024;\n add.u64 b, %2, 512;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %3, z;\n"
                 " add.u64 a, %1, 1026;\n add.u64 b, %2, 514;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %3, z;\n"
                 " add.u64 a, %1, 1028;\n add.u64 b, %2, 516;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %3, z;\n"
                 " add.u64 a, %1, 1030;\n add.u64 b, %2, 518;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %3, z;\n}"
                 :: "r"(d_tmem), "l"(a0), "l"(b0), "r"(idesc) : "memory");
}
// The 4 K16 steps of one O += P V tile: P from TMEM (8 columns per step), V descriptor + 2 KB (16 tokens) per step
__device__ __forceinline__ void pf_pv_block(uint32_t d_tmem, uint32_t p_tmem, uint64_t b0, uint32_t idesc, uint32_t acc_first) {
    asm volatile("{\n .reg .pred e, z;\n .reg .b64 b;\n .reg .b32 t;\n elect.sync _|e, 0xffffffff;\n setp.ne.b32 z, %4, 0;\n"
                 " @e tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, z;\n setp.eq.b32 z, 0, 0;\n"
                 " add.u32 t, %1, 8;\n add.u64 b, %2, 128;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], [t], b, %3, z;\n"
                 " add.u32 t, %1, 16;\n add.u64 b, %2, 256;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], [t], b, %3, z;\n"
                 " add.u32 t, %1, 24;\n add.u64 b, %2, 384;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], [t], b, %3, z;\n}"
                 :: "r"(d_tmem), "r"(p_tmem), "l"(b0), "r"(idesc), "r"(acc_first) : "memory");
}
// A operand from TMEM (P: lane = row, 2 halves per 32-bit column, 8 columns per K16 step)
__device__ __forceinline__ void pf_mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n .reg .pred p, e;\n elect.sync _|e, 0xffffffff;\n setp.ne.b32 p, %4, 0;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n}"
                 :: "r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void pf_commit(uint32_t bar) {
    asm volatile("{\n .reg .pred e;\n elect.sync _|e, 0xffffffff;\n @e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n}" :: "r"(bar) : "memory");
}

#define PF_R32(v, o) "=r"(v[o+0]), "=r"(v[o+1]), "=r"(v[o+2]), "=r"(v[o+3]), "=r"(v[o+4]), "=r"(v[o+5]), "=r"(v[o+6]), "=r"(v[o+7]), \
                     "=r"(v[o+8]), "=r"(v[o+9]), "=r"(v[o+10]), "=r"(v[o+11]), "=r"(v[o+12]), "=r"(v[o+13]), "=r"(v[o+14]), "=r"(v[o+15]), \
                     "=r"(v[o+16]), "=r"(v[o+17]), "=r"(v[o+18]), "=r"(v[o+19]), "=r"(v[o+20]), "=r"(v[o+21]), "=r"(v[o+22]), "=r"(v[o+23]), \
                     "=r"(v[o+24]), "=r"(v[o+25]), "=r"(v[o+26]), "=r"(v[o+27]), "=r"(v[o+28]), "=r"(v[o+29]), "=r"(v[o+30]), "=r"(v[o+31])
#define PF_W32(v, o) "r"(v[o+0]), "r"(v[o+1]), "r"(v[o+2]), "r"(v[o+3]), "r"(v[o+4]), "r"(v[o+5]), "r"(v[o+6]), "r"(v[o+7]), \
                     "r"(v[o+8]), "r"(v[o+9]), "r"(v[o+10]), "r"(v[o+11]), "r"(v[o+12]), "r"(v[o+13]), "r"(v[o+14]), "r"(v[o+15]), \
                     "r"(v[o+16]), "r"(v[o+17]), "r"(v[o+18]), "r"(v[o+19]), "r"(v[o+20]), "r"(v[o+21]), "r"(v[o+22]), "r"(v[o+23]), \
                     "r"(v[o+24]), "r"(v[o+25]), "r"(v[o+26]), "r"(v[o+27]), "r"(v[o+28]), "r"(v[o+29]), "r"(v[o+30]), "r"(v[o+31])
#define PF_LIST32 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}"
#define PF_LIST32S "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32}"

// 32 consecutive FP32 columns of this thread's TMEM lane (warp w touches lanes 32*(w%4)..+31)
__device__ __forceinline__ void pf_tmem_ld32(uint32_t taddr, uint32_t (&v)[64], int o) {
    if (o == 0) asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 " PF_LIST32 ", [%32];" : PF_R32(v, 0)  : "r"(taddr) : "memory");
    else        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 " PF_LIST32 ", [%32];" : PF_R32(v, 32) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void pf_tmem_st32(uint32_t taddr, const uint32_t (&v)[64], int o) {
    if (o == 0) asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], " PF_LIST32S ";" :: "r"(taddr), PF_W32(v, 0)  : "memory");
    else        asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], " PF_LIST32S ";" :: "r"(taddr), PF_W32(v, 32) : "memory");
}
__device__ __forceinline__ void pf_tmem_st32p(uint32_t taddr, const uint32_t (&v)[32]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], " PF_LIST32S ";" :: "r"(taddr), PF_W32(v, 0) : "memory");
}
__device__ __forceinline__ void pf_tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void pf_tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ float pf_ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

// sm_100 packed-pair F32 add (FADD2): halves the subtract and row-sum instruction counts of the softmax warps.
// Measured and NOT used (profiles/r1p_microbench.txt): 3-input max (FMNMX3 issues at the XU rate, 8 cycles per warp),
// and a polynomial exp2 on the FMA pipe for a share of the columns (every variant was slower: the tile period is set by
// the tensor pipe, ~1240 of ~1350 cycles, not by the 1024 XU cycles).
__device__ __forceinline__ void pf_add2(float &d0, float &d1, float a0, float a1, float b0, float b1) {
    asm("{\n .reg .b64 a, b, d;\n mov.b64 a, {%2, %3};\n mov.b64 b, {%4, %5};\n add.rn.f32x2 d, a, b;\n mov.b64 {%0, %1}, d;\n}"
        : "=f"(d0), "=f"(d1) : "f"(a0), "f"(a1), "f"(b0), "f"(b1));
}
__device__ __forceinline__ uint32_t pf_pack_h2(float lo, float hi) {
    const __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t *>(&h);
}
// byte offset of 16-byte chunk c (0..7) of row r inside a [rows][128 B] tile with the 128-byte swizzle
__device__ __forceinline__ uint32_t pf_swz(int r, int c) { return (uint32_t) ((r >> 3)*1024 + (r & 7)*128 + ((c ^ (r & 7)) << 4)); }
__device__ __forceinline__ void pf_sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" :: "r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// ------------------------------------------------------------------------------------------------
// plan: cut the query rows into tiles of <= tok_per_tile consecutive rows of one sequence.  One CTA.
// ------------------------------------------------------------------------------------------------
template <bool MAX>
__device__ __forceinline__ int pf_block_scan_incl(int v, int *s_warp) {      // inclusive scan over the CTA (sum or max)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x = MAX ? max(x, y) : x + y; }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    if (warp == 0) {
        int w = lane < nw ? s_warp[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w = MAX ? max(w, y) : w + y; }
        s_warp[lane] = w;
    }
    __syncthreads();
    if (warp) x = MAX ? max(x, s_warp[warp - 1]) : x + s_warp[warp - 1];
    __syncthreads();
    return x;
}

__global__ void __launch_bounds__(1024)
pf_plan_kernel(const __grid_constant__ pf_args a) {
    __shared__ int s_warp[32];
    const b200_attn_desc &p = a.d;
    pdl_prologue();                                          // the plan of the previous call on this workspace may still be in use
    int *row0 = a.plan + 2, *rows = row0 + p.n_q;
    int carry_start = 0, carry_tiles = 0;                    // run start of the last row seen, tiles so far
    for (int r0 = 0; r0 < p.n_q; r0 += 1024) {
        const int r = r0 + threadIdx.x;
        int boundary = 0;
        if (r < p.n_q) {
            const int s = p.q_seq ? p.q_seq[r] : r;
            boundary = r == 0 || s != (p.q_seq ? p.q_seq[r - 1] : r - 1);
        }
        int start = pf_block_scan_incl<true>(boundary ? r : 0, s_warp);
        start = max(start, carry_start);
        const int is_tile = r < p.n_q && (boundary || (r - start) % a.tok_per_tile == 0);
        const int idx = carry_tiles + pf_block_scan_incl<false>(is_tile, s_warp) - 1;
        if (is_tile) row0[idx] = r;
        __shared__ int s_carry[2];
        if (threadIdx.x == 1023) { s_carry[0] = start; s_carry[1] = idx + 1; }
        __syncthreads();
        carry_start = s_carry[0]; carry_tiles = s_carry[1];
        __syncthreads();
    }
    const int n_tiles = carry_tiles;
    __threadfence_block();
    __syncthreads();
    for (int t = threadIdx.x; t < n_tiles; t += 1024) rows[t] = (t + 1 < n_tiles ? row0[t + 1] : p.n_q) - row0[t];
    if (threadIdx.x == 0) { a.plan[0] = n_tiles; a.plan[1] = 0; }
    if (a.seq_off) {
        const int64_t cap = a.cap;
        int carry = 0;
        for (int s0 = 0; s0 < p.n_seqs; s0 += 1024) {
            const int sq = s0 + threadIdx.x;
            int n = 0;
            if (sq < p.n_seqs) { int64_t l = p.seq_lens[sq]; l = l < 0 ? 0 : (l > cap ? cap : l); n = (int) ((l + PF_BN - 1)/PF_BN*PF_BN); }
            const int incl = pf_block_scan_incl<false>(n, s_warp);
            if (sq < p.n_seqs) a.seq_off[sq] = carry + incl - n;
            __shared__ int s_tot;
            if (threadIdx.x == 1023) s_tot = incl;
            __syncthreads();
            carry += s_tot;
            __syncthreads();
        }
        if (threadIdx.x == 0) a.seq_off[p.n_seqs] = carry;
    }
}

// ------------------------------------------------------------------------------------------------
// quantised caches: expand the sequences' cells to the dense F16 scratch.  One CTA = 16 scratch cells.
// y = q*d (dequantize_row_q8_0 / dequantize_row_q4_0, ggml/src/ggml-quants.c:349-362, :255-273) rounded to half;
// cells past seq_len (up to the 64-cell padding) are written as zeros.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void pf_dequant_block(int type, const char *row, int blk, __half *out) {
    uint4 w[4];
    __half2 *h = reinterpret_cast<__half2 *>(w);
    if (type == B200_TYPE_F16) {
        const uint4 *src = reinterpret_cast<const uint4 *>(row + 64*blk);
#pragma unroll
        for (int i = 0; i < 4; ++i) w[i] = src[i];
    } else if (type == B200_TYPE_Q8_0) {
        const char *b = row + Q8_BYTES*blk;
        const float d = h2f(ld_u16(b));
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const uint32_t x = ld_u16(b + 2 + 2*i);
            h[i] = __floats2half2_rn((float) (int8_t) (x & 0xff)*d, (float) (int8_t) (x >> 8)*d);
        }
    } else {
        const char *b = row + Q4_BYTES*blk;
        const float d = h2f(ld_u16(b));
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t x = ld_u16(b + 2 + 2*i);                        // qs[2i], qs[2i+1]
            h[i]     = __floats2half2_rn((float) ((int) (x & 0xf) - 8)*d,        (float) ((int) ((x >> 8) & 0xf) - 8)*d);
            h[8 + i] = __floats2half2_rn((float) ((int) ((x >> 4) & 0xf) - 8)*d, (float) ((int) (x >> 12) - 8)*d);
        }
    }
    uint4 *dst = reinterpret_cast<uint4 *>(out);
#pragma unroll
    for (int i = 0; i < 4; ++i) dst[i] = w[i];
}

__global__ void __launch_bounds__(256)
pf_dequant_kernel(const __grid_constant__ pf_args a) {
    const b200_attn_desc &p = a.d;
    pdl_prologue();
    const int total = a.seq_off[p.n_seqs];
    for (int64_t c0 = (int64_t) blockIdx.x*16; c0 < total; c0 += (int64_t) gridDim.x*16) {
        int lo = 0, hi = p.n_seqs;                           // sequence owning scratch cell c0: last s with seq_off[s] <= c0
        while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (a.seq_off[mid] <= c0) lo = mid; else hi = mid; }
        const int sq = lo, p0 = (int) (c0 - a.seq_off[sq]);
        int64_t seq_len = p.seq_lens[sq];
        seq_len = seq_len < 0 ? 0 : (seq_len > a.cap ? a.cap : seq_len);
        const int32_t *bt = p.block_table + (int64_t) sq*p.bt_stride;
        const int units = 16*p.n_head_kv*4;                  // (cell, head, 32-element block)
        for (int u = threadIdx.x; u < 2*units; u += blockDim.x) {
            const bool is_v = u >= units;
            const int uu = is_v ? u - units : u;
            const int blk = uu & 3, head = (uu >> 2) % p.n_head_kv, i = (uu >> 2)/p.n_head_kv;
            const int pos = p0 + i;
            __half *out = (is_v ? a.v16 : a.k16) + ((c0 + i)*p.n_head_kv + head)*PF_D + 32*blk;
            if (pos < seq_len) {
                const int64_t cell = (int64_t) bt[pos/p.block_size]*p.block_size + pos % p.block_size;
                const char *row = static_cast<const char *>(is_v ? p.v : p.k) + cell*(is_v ? p.nbv1 : p.nbk1) + (int64_t) head*(is_v ? p.nbv2 : p.nbk2);
                pf_dequant_block(is_v ? p.type_v : p.type_k, row, blk, out);
            } else {
                uint4 *dst = reinterpret_cast<uint4 *>(out);
#pragma unroll
                for (int k = 0; k < 4; ++k) dst[k] = make_uint4(0u, 0u, 0u, 0u);
            }
        }
    }
}

// which work items are split, and how: everything follows from the item count and the grid, so the main and the
// combine kernel agree without a table.  Units [0, full) are whole items; unit full + j is piece j % k of item full + j / k.
struct pf_split { int full, rem, k, n_units; };
__device__ __forceinline__ pf_split pf_split_of(int n_work, int grid, int k_max) {
    pf_split sp;
    sp.rem = n_work % grid; sp.full = n_work - sp.rem;
    sp.k = sp.rem ? grid/sp.rem : 1;
    if (sp.k > k_max) sp.k = k_max;
    sp.n_units = sp.full + sp.rem*sp.k;
    return sp;
}
// tiles [ta, tb) of piece `piece` of an item with nkt kv tiles; returns the number of pieces the item really has
__device__ __forceinline__ int pf_piece_range(int nkt, int k, int piece, int &ta, int &tb) {
    int ke = nkt/PF_MIN_PIECE; ke = ke < 1 ? 1 : (ke > k ? k : ke);
    if (piece >= ke) { ta = tb = 0; return ke; }
    ta = (int) ((int64_t) nkt*piece/ke); tb = (int) ((int64_t) nkt*(piece + 1)/ke);
    return ke;
}

// ------------------------------------------------------------------------------------------------
// main kernel: one CTA per SM, TWO 128-row groups (A = rows 0..127, B = 128..255 of a 256-row work item) share every
// K/V tile, so K/V cross L2 -> shared memory once per 256 rows, and the tensor pipe works on one group while the
// other group's softmax warps run (the S/P/O of the groups live in disjoint TMEM columns).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(PF_THREADS, 1)
pf_main_kernel(const __grid_constant__ pf_args a, const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV) {
    extern __shared__ __align__(1024) char pf_dyn[];
    const b200_attn_desc &p = a.d;
    const uint32_t sbase = pf_smem_u32(pf_dyn);
    const uint32_t sQ = sbase + PF_SQ, sK = sbase + PF_SK, sV = sbase + PF_SV, sBar = sbase + PF_SBAR;
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(pf_dyn + PF_SBAR + 208);
    volatile int      *work_slot = reinterpret_cast<volatile int *>(pf_dyn + PF_SBAR + 212);
    volatile int      *nvis_slot = reinterpret_cast<volatile int *>(pf_dyn + PF_SBAR + 216);   // [8] per-warp max n_vis
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    pdl_trigger();                                           // the CTA set-up below (barriers, TMEM) runs under the plan / expansion kernels

    if (threadIdx.x == 0) {
        for (int i = 0; i < PF_NBAR; ++i) pf_mbar_init(sBar + 8*i, (i >= PF_PFULL && i < PF_PFULL + 4) ? 128 : 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == PF_MMA_WARP) {                               // all 512 TMEM columns: per group S0 | S1 | O
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(sBar + 208) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    pf_tc_fence_before();
    __syncthreads();
    pf_tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    pdl_wait();                                              // plan, F16 scratch, the caller's tensors

    const int gqa = a.gqa, bs = p.block_size;
    const int n_blk = p.bt_stride < INT32_MAX ? (int) p.bt_stride : INT32_MAX;
    const int *tile_row0 = a.plan + 2, *tile_rows = tile_row0 + p.n_q;
    const int tokA = PF_ROWS/gqa;                            // tokens of group A (= of group B)
    uint32_t G = 0;                                          // kv tiles processed by this CTA so far (same in every role): K/V rings
    uint32_t GB = 0;                                         // ... by work items that used group B: group B's barriers and S buffers

    for (;;) {
        if (threadIdx.x == 0) *work_slot = atomicAdd(&a.plan[1], 1);
        __syncthreads();
        const int unit = *work_slot;
        const int n_tiles = a.plan[0];
        const pf_split sp = pf_split_of(n_tiles*p.n_head_kv, (int) gridDim.x, a.split_max);
        if (unit >= sp.n_units) break;
        const int pidx = unit - sp.full;                     // >= 0: a piece (slot in the partial buffers)
        const int w = pidx < 0 ? unit : sp.full + pidx/sp.k, piece = pidx < 0 ? 0 : pidx % sp.k;
        const bool split = pidx >= 0 && sp.k > 1;
        const int tile = n_tiles - 1 - w/p.n_head_kv, kvh = w % p.n_head_kv;      // later rows (longer contexts) first
        const int row0 = tile_row0[tile], nrows = tile_rows[tile];
        const bool useB = nrows > tokA;                      // group B has live rows
#ifdef PF_TIMELINE
        const bool first_item = G == 0;
#endif
        const int seq = p.q_seq ? p.q_seq[row0] : row0;
        int seq_len = p.seq_lens[seq];
        if (seq_len > a.cap) seq_len = (int) a.cap;
        if (seq_len < 0) seq_len = 0;
        const int32_t *bt = p.block_table + (int64_t) seq*p.bt_stride;

        // ---- softmax warps: this thread's row, its Q row -> F16 * scale*log2(e) -> shared (UMMA K-major, 128B swizzle)
        int n_vis = 0;
        const int R = threadIdx.x, grp = (R >> 7) & 1, r = R & 127, tl = R/gqa, hg = R - tl*gqa;
        if (warp < 8) {
            const bool live = tl < nrows;
            if (live) {
                n_vis = seq_len;
                if (p.q_pos) { const int lim = p.q_pos[row0 + tl] + 1; n_vis = lim < n_vis ? lim : n_vis; }
                if (n_vis < 0) n_vis = 0;
            }
            if (grp == 0 || useB) {
                const float *qrow = reinterpret_cast<const float *>(static_cast<const char *>(p.q) + (int64_t) (row0 + (live ? tl : 0))*p.nbq1
                                                                     + (int64_t) (kvh*gqa + hg)*p.nbq2);
#pragma unroll 8
                for (int c = 0; c < 16; ++c) {
                    float4 x = make_float4(0.f, 0.f, 0.f, 0.f), y = x;
                    if (live) { x = *reinterpret_cast<const float4 *>(qrow + 8*c); y = *reinterpret_cast<const float4 *>(qrow + 8*c + 4); }
                    pf_sts128(sQ + grp*32768 + (c >> 3)*16384 + pf_swz(r, c & 7), pf_pack_h2(x.x*a.qscale, x.y*a.qscale),
                              pf_pack_h2(x.z*a.qscale, x.w*a.qscale), pf_pack_h2(y.x*a.qscale, y.y*a.qscale), pf_pack_h2(y.z*a.qscale, y.w*a.qscale));
                }
            }
            const int wmax = __reduce_max_sync(0xffffffffu, n_vis);
            if (lane == 0) nvis_slot[warp] = wmax;
            pf_fence_async_smem();
        }
        __syncthreads();
        int nvis_max = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) nvis_max = max(nvis_max, nvis_slot[i]);
        // kv tiles [ta, ta + nkt) of this unit: all of the item's, or one piece of them
        int ta = 0, nkt = (nvis_max + PF_BN - 1)/PF_BN;
        if (split) { int tb; pf_piece_range(nkt, sp.k, piece, ta, tb); nkt = tb - ta; }

        if (warp == PF_TMA_WARP) {
            // ================= TMA producer =================
            // K (4 stages) and V (3 stages) have their own rings: a K stage is free as soon as QK^T of its tile retired,
            // and K runs one tile ahead of V — the order the MMA warp consumes them in.  Lane l holds block-table entry
            // win_base + l; the next 32 entries are fetched one window ahead, so the gather never waits on a global load
            // (the whole warp takes part, one elected lane talks to the TMA unit).
            const int dense_base = a.seq_off ? a.seq_off[seq] : -1;       // F16 scratch of a quantised cache: cells are sequence-local
            const int kc0 = a.seq_off ? kvh*PF_D : (int) ((int64_t) kvh*p.nbk2/2), vc0 = a.seq_off ? kvh*PF_D : (int) ((int64_t) kvh*p.nbv2/2);
            const int tpb = bs/16;                           // 16-cell boxes per block
            int b_cur = ta*(PF_BN/16)/tpb, o_cur = ta*(PF_BN/16) - b_cur*tpb;
            int win_base = b_cur;
            int win      = win_base + lane < n_blk ? bt[win_base + lane] : 0;
            int win_next = win_base + 32 + lane < n_blk ? bt[win_base + 32 + lane] : 0;
            int cell[4], vcell[4] = {0, 0, 0, 0};
            for (int t = 0; t <= nkt; ++t) {
                if (t < nkt) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        if (b_cur - win_base >= 32) {
                            win_base += 32; win = win_next;
                            win_next = win_base + 32 + lane < n_blk ? bt[win_base + 32 + lane] : 0;
                        }
                        cell[i] = __shfl_sync(0xffffffffu, win, b_cur - win_base)*bs + o_cur*16;
                        if (++o_cur == tpb) { o_cur = 0; ++b_cur; }
                        if (dense_base >= 0) cell[i] = dense_base + (ta + t)*PF_BN + 16*i;
                    }
                    const uint32_t g = G + t, st = g % PF_KST, sk = sK + st*16384, bk = sBar + 8*(PF_FULLK + st);
                    if (lane == 0) {
                        pf_mbar_wait(sBar + 8*(PF_EMPTYK + st), ((g / PF_KST) & 1) ^ 1);
                        pf_mbar_expect_tx(bk, 16384);
                    }
                    __syncwarp();
#pragma unroll
                    for (int i = 0; i < 4; ++i) pf_tma_pair_elect(sk + 2048*i, &tmK, kc0, cell[i], bk);
                }
                if (t > 0) {
                    const uint32_t g = G + t - 1, st = g % PF_VST, sv = sV + st*16384, bv = sBar + 8*(PF_FULLV + st);
                    if (lane == 0) {
                        pf_mbar_wait(sBar + 8*(PF_EMPTYV + st), ((g / PF_VST) & 1) ^ 1);
                        pf_mbar_expect_tx(bv, 16384);
                    }
                    __syncwarp();
#pragma unroll
                    for (int i = 0; i < 4; ++i) pf_tma_pair_elect(sv + 2048*i, &tmV, vc0, vcell[i], bv);
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) vcell[i] = cell[i];
            }
        } else if (warp == PF_MMA_WARP) {
            // ================= MMA issuer =================  (whole warp converged; one elected lane issues)
            // Issue order per kv tile j:  wait P_A(j), PV_A(j) | QK_B(j+1) | wait P_B(j), PV_B(j) | QK_A(j+2): while one
            // group's softmax works, the tensor pipe runs the other group's MMAs.
            constexpr uint32_t ID_QK = pf_idesc(PF_ROWS, PF_BN, 0), ID_PV = pf_idesc(PF_ROWS, PF_D, 1);
            const int ngrp = useB ? 2 : 1;
            auto issue_qk = [&](int gr, int j) {             // S_gr(j) = Q_gr K(j)^T
                const uint32_t g = G + j, gg = gr ? GB + j : g, st = g % PF_KST, sk = sK + st*16384;
                if (gr == 0) { pf_mbar_wait(sBar + 8*(PF_FULLK + st), (g / PF_KST) & 1); pf_tc_fence_after(); }
                pf_qk_block(tmem + gr*256 + (gg & 1)*64, pf_desc(sQ + gr*32768, 16, 1024), pf_desc(sk, 16, 1024), ID_QK);
                pf_commit(sBar + 8*(PF_SFULL + 2*gr + (gg & 1)));
                if (gr == ngrp - 1) pf_commit(sBar + 8*(PF_EMPTYK + st));       // last reader of the K stage
            };
            auto issue_pv = [&](int gr, int j) {             // O_gr += P_gr(j) V(j)
                const uint32_t g = G + j, gg = gr ? GB + j : g, st = g % PF_VST, sv = sV + st*16384;
                pf_mbar_wait(sBar + 8*(PF_PFULL + 2*gr + (gg & 1)), (gg >> 1) & 1);   // one barrier per P buffer: softmax may run a tile ahead
                if (gr == 0) pf_mbar_wait(sBar + 8*(PF_FULLV + st), (g / PF_VST) & 1);
                pf_tc_fence_after();
                // V as the MN-major B operand: LBO = next 64 d (8 KB), SBO = next 8 tokens (1 KB)
                pf_pv_block(tmem + gr*256 + 128, tmem + gr*256 + (gg & 1)*64, pf_desc(sv, 8192, 1024), ID_PV, j > 0);
                pf_commit(sBar + 8*(PF_ODONE + gr));
                if (gr == ngrp - 1) pf_commit(sBar + 8*(PF_EMPTYV + st));       // last reader of the V stage
            };
            // tcgen05.mma is only ordered against an earlier one with the same accumulator and shape, so before S_gr(t)
            // overwrites the buffer P_gr(t-2) sits in, PV_gr(t-2) must have retired (ODONE).  Group A's QK^T runs two
            // tiles ahead and group B's one: neither group's next S ever queues behind the other group's softmax, and
            // every ODONE wait has another group's MMA block in flight behind it.
            auto wait_pv = [&](int gr, int j) { pf_mbar_wait(sBar + 8*(PF_ODONE + gr), ((gr ? GB : G) + j) & 1); };
            if (nkt > 0) { issue_qk(0, 0); if (useB) issue_qk(1, 0); }
            if (nkt > 1) issue_qk(0, 1);
            for (int j = 0; j < nkt; ++j) {
                PF_T(2, j, 0);
                issue_pv(0, j);
                PF_T(2, j, 1);
                if (useB) {
                    if (j + 1 < nkt) { if (j > 0) wait_pv(1, j - 1); PF_T(2, j, 2); issue_qk(1, j + 1); }
                    PF_T(2, j, 3);
                    issue_pv(1, j);
                    PF_T(2, j, 4);
                }
                if (j + 2 < nkt) { wait_pv(0, j); PF_T(2, j, 5); issue_qk(0, j + 2); }
                PF_T(2, j, 6);
            }
        } else if (grp == 0 || useB) {
            // ================= softmax (group grp) =================
            const uint32_t lane_base = (uint32_t) (32*(warp & 3)) << 16;
            const uint32_t tS = tmem + grp*256, tO = tS + 128;
            float m_run = -INFINITY, l_run = 0.f;
            int nvis_min = __reduce_min_sync(0xffffffffu, n_vis);
            uint32_t v[64];
            const uint32_t Gg = grp ? GB : G;                // this group's tile counter (barrier parities, S buffer)
            for (int j = 0; j < nkt; ++j) {
                const uint32_t g = Gg + j, gkv = G + j;
                PF_T(grp, j, 0);
                pf_mbar_wait(sBar + 8*(PF_SFULL + 2*grp + (g & 1)), (g >> 1) & 1);
                PF_T(grp, j, 1);
                pf_tc_fence_after();
                pf_tmem_ld32(tS + (g & 1)*64 + lane_base,      v, 0);
                pf_tmem_ld32(tS + (g & 1)*64 + lane_base + 32, v, 32);
                pf_tmem_wait_ld();
                PF_T(grp, j, 2);
                const int tok0 = (ta + j)*PF_BN;
                float mx = -INFINITY, mx2 = -INFINITY;
                if (tok0 + PF_BN > nvis_min) {               // warp-uniform: only tiles that cross some row's limit mask
#pragma unroll
                    for (int c = 0; c < 64; ++c) { if (tok0 + c >= n_vis) v[c] = 0xff800000u; }
                }
#pragma unroll
                for (int c = 0; c < 64; c += 2) { mx = fmaxf(mx, __uint_as_float(v[c])); mx2 = fmaxf(mx2, __uint_as_float(v[c + 1])); }
                mx = fmaxf(mx, mx2);
                // lazy running max: move the reference point only when the max grew by more than 2^8
                const bool grow = mx > m_run + 8.0f;
                float alpha = 1.0f;
                if (grow) { alpha = pf_ex2(m_run - mx); m_run = mx; }         // m_run == -inf: alpha = 0
                const float nsub = m_run == -INFINITY ? 0.0f : -m_run;
                PF_T(grp, j, 3);
                float sum0 = 0.f, sum1 = 0.f, sum2 = 0.f, sum3 = 0.f;
                uint32_t pk[32];
#pragma unroll
                for (int c = 0; c < 32; ++c) {
                    float x0, x1;
                    pf_add2(x0, x1, __uint_as_float(v[2*c]), __uint_as_float(v[2*c + 1]), nsub, nsub);
                    const float p0 = pf_ex2(x0), p1 = pf_ex2(x1);
                    if (c & 1) pf_add2(sum2, sum3, sum2, sum3, p0, p1); else pf_add2(sum0, sum1, sum0, sum1, p0, p1);
                    pk[c] = pf_pack_h2(p0, p1);
                }
                pf_add2(sum0, sum1, sum0, sum1, sum2, sum3);
                l_run = l_run*alpha + (sum0 + sum1);
                if (j > 0 && __any_sync(0xffffffffu, grow)) {  // rare: rescale O once PV of the previous tile retired
                    pf_mbar_wait(sBar + 8*(PF_ODONE + grp), (g - 1) & 1);
                    pf_tc_fence_after();
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        pf_tmem_ld32(tO + lane_base + 64*h,      v, 0);
                        pf_tmem_ld32(tO + lane_base + 64*h + 32, v, 32);
                        pf_tmem_wait_ld();
#pragma unroll
                        for (int c = 0; c < 64; ++c) v[c] = __float_as_uint(__uint_as_float(v[c])*alpha);
                        pf_tmem_st32(tO + lane_base + 64*h,      v, 0);
                        pf_tmem_st32(tO + lane_base + 64*h + 32, v, 32);
                    }
                }
                // P (packed F16) over the first 32 columns of this tile's S buffer: S was read into registers above, and
                // the next QK^T into this buffer is issued after PV of this tile (tcgen05.mma executes in issue order)
                PF_T(grp, j, 4);
                pf_tmem_st32p(tS + (g & 1)*64 + lane_base, pk);
                pf_tmem_wait_st();
                PF_T(grp, j, 5);
                if (tok0 + PF_BN > seq_len) {                // cells past seq_len hold stale bytes: 0 * NaN must not reach O
                    const uint32_t st = gkv % PF_VST, sv = sV + st*16384;
                    pf_mbar_wait(sBar + 8*(PF_FULLV + st), (gkv / PF_VST) & 1);
                    for (int idx = r; idx < 64*16; idx += 128) {  // each group clears every stale row itself (idempotent)
                        const int tok = idx >> 4, c = idx & 15;
                        if (tok0 + tok >= seq_len) pf_sts128(sv + (c >> 3)*8192 + pf_swz(tok, c & 7), 0u, 0u, 0u, 0u);
                    }
                    pf_fence_async_smem();
                }
                pf_tc_fence_before();
                pf_mbar_arrive(sBar + 8*(PF_PFULL + 2*grp + (g & 1)));
                PF_T(grp, j, 6);
            }
            // ---- epilogue: O / l -> dst[row][head][d]; a piece: (m, l) and the unnormalised O -> its slot of the partial buffers
            const bool live = tl < nrows;
            // what kind of unit this was, worked out again from the shared slots (kept out of the tile loop's registers):
            // pslot >= 0: a piece (partial buffers); dead: a piece slot of an item that runs whole — nothing to write
            int pslot = -1;
            bool dead = false;
            {
                const pf_split sp_e = pf_split_of(a.plan[0]*p.n_head_kv, (int) gridDim.x, a.split_max);
                const int pidx_e = *work_slot - sp_e.full;
                if (pidx_e >= 0 && sp_e.k > 1) {
                    int nv = 0;
#pragma unroll
                    for (int i = 0; i < 8; ++i) nv = max(nv, nvis_slot[i]);
                    int ta_e, tb_e;
                    const int ke = pf_piece_range((nv + PF_BN - 1)/PF_BN, sp_e.k, pidx_e % sp_e.k, ta_e, tb_e);
                    if (ke > 1) pslot = pidx_e;
                    dead = pidx_e % sp_e.k >= ke;
                }
            }
            const bool is_piece = pslot >= 0;
            float *out = is_piece ? a.part_o + ((int64_t) pslot*2*PF_ROWS + R)*PF_D
                                  : static_cast<float *>(p.dst) + ((int64_t) (row0 + (live ? tl : 0))*p.n_head + kvh*gqa + hg)*PF_D;
            if (is_piece && !dead) *reinterpret_cast<float2 *>(a.part_ml + ((int64_t) pslot*2*PF_ROWS + R)*2) = make_float2(nkt > 0 ? m_run : -INFINITY, nkt > 0 ? l_run : 0.f);
            if (nkt > 0) {
                pf_mbar_wait(sBar + 8*(PF_ODONE + grp), (Gg + nkt - 1) & 1);
                pf_tc_fence_after();
                const float inv = is_piece ? 1.0f : 1.0f/l_run;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    pf_tmem_ld32(tO + lane_base + 64*h,      v, 0);
                    pf_tmem_ld32(tO + lane_base + 64*h + 32, v, 32);
                    pf_tmem_wait_ld();
                    if (live) {
#pragma unroll
                        for (int c = 0; c < 64; c += 4)
                            *reinterpret_cast<float4 *>(out + 64*h + c) = make_float4(__uint_as_float(v[c])*inv, __uint_as_float(v[c + 1])*inv,
                                                                                      __uint_as_float(v[c + 2])*inv, __uint_as_float(v[c + 3])*inv);
                    }
                }
            } else if (live && !is_piece && !dead) {
                const float nan = __int_as_float(0x7fc00000);                     // no visible cell: 0/0 like the reference (ggml-cpu.c:12418)
                for (int c = 0; c < PF_D; c += 4) *reinterpret_cast<float4 *>(out + c) = make_float4(nan, nan, nan, nan);
            }
            pf_tc_fence_before();
        }
        G += nkt;
        if (useB) GB += nkt;
        __syncthreads();                                     // O, Q and the S buffers are free for the next work item
    }

    pf_tc_fence_before();
    __syncthreads();
    if (warp == PF_MMA_WARP) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tmem) : "memory");
}

// merge the pieces of the split items: CTA (b, y) = rows 8y .. 8y+7 of split item b (the b-th of the last `rem` work
// items), one warp per row, a lane per 4 of the 128 output dims: every row's (m, l) and O loads are one round trip, all
// rows in flight at once (32 rows per warp in sequence made this kernel ~55 us of an 83 us op).  Exits at once when
// nothing was split.
__global__ void __launch_bounds__(256)
pf_combine_kernel(const __grid_constant__ pf_args a) {
    const b200_attn_desc &p = a.d;
    pdl_prologue();
    const int n_tiles = a.plan[0];
    const pf_split sp = pf_split_of(n_tiles*p.n_head_kv, a.grid_main, a.split_max);
    if (sp.k <= 1 || (int) blockIdx.x >= sp.rem) return;
    const int w = sp.full + blockIdx.x;
    const int tile = n_tiles - 1 - w/p.n_head_kv, kvh = w % p.n_head_kv;
    const int *tile_row0 = a.plan + 2, *tile_rows = tile_row0 + p.n_q;
    const int row0 = tile_row0[tile], nrows = tile_rows[tile];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // the item's kv tiles, as the main kernel counted them: from the largest visible length of its tokens
    __shared__ int s_nvis[8];
    {
        int n_vis = 0;
        if ((int) threadIdx.x < nrows) {
            const int seq = p.q_seq ? p.q_seq[row0] : row0;
            int seq_len = p.seq_lens[seq];
            if (seq_len > a.cap) seq_len = (int) a.cap;
            n_vis = seq_len < 0 ? 0 : seq_len;
            if (p.q_pos) { const int lim = p.q_pos[row0 + threadIdx.x] + 1; n_vis = lim < n_vis ? lim : n_vis; }
            if (n_vis < 0) n_vis = 0;
        }
        const int wmax = __reduce_max_sync(0xffffffffu, n_vis);
        if (lane == 0) s_nvis[warp] = wmax;
        __syncthreads();
    }
    int nvis_max = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) nvis_max = max(nvis_max, s_nvis[i]);
    int ta, tb;
    const int ke = pf_piece_range((nvis_max + PF_BN - 1)/PF_BN, sp.k, 0, ta, tb);
    if (ke <= 1) return;                                     // ran whole: the main kernel wrote dst
    const int64_t slot0 = (int64_t) blockIdx.x*sp.k;        // first piece slot of this item
    {
        const int R = blockIdx.y*8 + warp;
        const int tl = R/a.gqa, hg = R - tl*a.gqa;
        if (tl >= nrows) return;
        const float2 ml = lane < ke ? *reinterpret_cast<const float2 *>(a.part_ml + ((slot0 + lane)*2*PF_ROWS + R)*2) : make_float2(-INFINITY, 0.f);
        float4 v[PF_MAX_SPLIT];
#pragma unroll
        for (int i = 0; i < PF_MAX_SPLIT; ++i)
            v[i] = i < ke ? *reinterpret_cast<const float4 *>(a.part_o + ((slot0 + i)*2*PF_ROWS + R)*PF_D + 4*lane) : make_float4(0.f, 0.f, 0.f, 0.f);
        float M = ml.x;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, o));
        const float f = ml.x == -INFINITY ? 0.f : pf_ex2(ml.x - M);      // empty pieces (and their unwritten O) drop out
        float L = ml.y*f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) L += __shfl_xor_sync(0xffffffffu, L, o);
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int i = 0; i < PF_MAX_SPLIT; ++i) {
            const float fi = __shfl_sync(0xffffffffu, f, i);
            if (fi != 0.f) { acc.x = fmaf(v[i].x, fi, acc.x); acc.y = fmaf(v[i].y, fi, acc.y); acc.z = fmaf(v[i].z, fi, acc.z); acc.w = fmaf(v[i].w, fi, acc.w); }
        }
        const float inv = 1.0f/L;                             // no visible cell at all: 0 * inf = NaN like the reference (ggml-cpu.c:12418)
        acc.x *= inv; acc.y *= inv; acc.z *= inv; acc.w *= inv;
        *reinterpret_cast<float4 *>(static_cast<float *>(p.dst) + ((int64_t) (row0 + tl)*p.n_head + kvh*a.gqa + hg)*PF_D + 4*lane) = acc;
    }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
typedef CUresult (*pf_encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                 const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static pf_encode_fn pf_encoder() {
    static pf_encode_fn fn = nullptr;
    if (!fn) {
        void *ptr = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<pf_encode_fn>(ptr);
        else (void) cudaGetLastError();
    }
    return fn;
}

static bool pf_type_ok(int t) { return t == B200_TYPE_F16 || t == B200_TYPE_Q8_0 || t == B200_TYPE_Q4_0; }
static bool pf_needs_scratch(const b200_attn_desc &d) { return d.type_k != B200_TYPE_F16 || d.type_v != B200_TYPE_F16; }

// cells of the F16 scratch: every sequence padded to a multiple of 64 cells.  Sized from host-known bounds only
// (total_tokens is a scheduling hint and is not trusted for memory safety).
static int64_t pf_scratch_cells(const b200_attn_desc &d) {
    int64_t cap = d.bt_stride*d.block_size;
    if (d.max_seq_len > 0 && d.max_seq_len < cap) cap = d.max_seq_len;
    return (int64_t) d.n_seqs*((cap + PF_BN - 1)/PF_BN*PF_BN);
}
constexpr int64_t PF_SCRATCH_MAX = 1LL << 30;                // bytes per tensor; beyond that the per-row kernel takes the call

bool prefill_supports(const b200_attn_desc &d) {
    if (!d.block_table || !d.seq_lens || d.head_dim != PF_D || d.max_bias != 0.0f || d.logit_softcap != 0.0f) return false;
    if (!pf_type_ok(d.type_k) || !pf_type_ok(d.type_v)) return false;
    const int gqa = d.n_head/d.n_head_kv;
    if (gqa < 1 || gqa > PF_ROWS || PF_ROWS % gqa) return false;
    if (d.block_size % 16 || d.n_kv < 16) return false;                                   // one TMA box = 16 consecutive cells
    if (d.nbk1 <= 0 || d.nbv1 <= 0 || d.nbk2 <= 0 || d.nbv2 <= 0) return false;
    if (((uintptr_t) d.q | (uint64_t) d.nbq1 | (uint64_t) d.nbq2 | (uintptr_t) d.dst) & 15) return false;
    if (pf_needs_scratch(d)) {
        if (((uintptr_t) d.k | (uintptr_t) d.v | (uint64_t) d.nbk1 | (uint64_t) d.nbv1 | (uint64_t) d.nbk2 | (uint64_t) d.nbv2) & 1) return false;
        const int64_t cells = pf_scratch_cells(d);
        if (cells <= 0 || cells >= (1LL << 31) || cells*d.n_head_kv*PF_D*2 > PF_SCRATCH_MAX) return false;
        return true;
    }
    if (((uintptr_t) d.k | (uintptr_t) d.v | (uint64_t) d.nbk1 | (uint64_t) d.nbv1 | (uint64_t) d.nbk2 | (uint64_t) d.nbv2) & 15) return false;
    if ((int64_t) (d.n_head_kv - 1)*d.nbk2 + 256 > d.nbk1 || (int64_t) (d.n_head_kv - 1)*d.nbv2 + 256 > d.nbv1) return false;
    if (d.n_kv >= (1LL << 31) || d.n_kv*(d.nbk1 > d.nbv1 ? d.nbk1 : d.nbv1) >= (1LL << 40)) return false;
    return true;
}

static size_t pf_align(size_t x) { return (x + 1023) & ~(size_t) 1023; }
static size_t pf_plan_bytes(const b200_attn_desc &d) { return pf_align((2 + 2*(size_t) d.n_q + (size_t) d.n_seqs + 1)*sizeof(int)); }

static int pf_sm_count() {
    static int sms = 0;
    if (!sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); if (sms <= 0) sms = 148; }
    return sms;
}
static int64_t pf_grid(const b200_attn_desc &d) {             // persistent grid: one CTA per SM, never more than work items can exist
    const int64_t items = (int64_t) d.n_q*d.n_head_kv;        // upper bound (one row per tile)
    return items < pf_sm_count() ? items : pf_sm_count();
}
static size_t pf_part_o_bytes(const b200_attn_desc &d)  { return pf_align((size_t) pf_grid(d)*2*PF_ROWS*PF_D*sizeof(float)); }   // pieces <= grid
static size_t pf_part_ml_bytes(const b200_attn_desc &d) { return pf_align((size_t) pf_grid(d)*2*PF_ROWS*2*sizeof(float)); }

size_t prefill_workspace(const b200_attn_desc &d) {
    size_t n = pf_plan_bytes(d) + pf_part_o_bytes(d) + pf_part_ml_bytes(d);
    if (pf_needs_scratch(d)) n += 2*pf_align((size_t) pf_scratch_cells(d)*d.n_head_kv*PF_D*2);
    return n;
}

static bool pf_make_map(CUtensorMap *tm, const void *base, int64_t nb1, int64_t nb2, int n_head_kv, int64_t n_cells) {
    pf_encode_fn enc = pf_encoder();
    if (!enc) { set_error("prefill: cuTensorMapEncodeTiled not available from the driver"); return false; }
    const cuuint64_t dims[2]    = {(cuuint64_t) (((int64_t) (n_head_kv - 1)*nb2 + 256)/2), (cuuint64_t) n_cells};
    const cuuint64_t strides[1] = {(cuuint64_t) nb1};
    const cuuint32_t box[2]     = {64, 16};
    const cuuint32_t ones[2]    = {1, 1};
    const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void *>(base), dims, strides, box, ones,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("prefill: cuTensorMapEncodeTiled failed with %d", (int) r); return false; }
    return true;
}

int launch_attn_prefill(const b200_attn_desc &d, cudaStream_t st) {
    const size_t need = prefill_workspace(d);
    if (d.workspace == nullptr || d.workspace_bytes < need) {
        set_error("attention: workspace of %zu bytes needed, %zu given", need, d.workspace_bytes);
        return B200_ERR_ALLOC;
    }
    static bool configured = false;
    if (!configured) {
        const cudaError_t e = cudaFuncSetAttribute(pf_main_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PF_SMEM);
        if (e != cudaSuccess) { set_error("prefill: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return B200_ERR_CUDA; }
        configured = true;
    }
    pf_args a;
    a.d = d;
    a.gqa = d.n_head/d.n_head_kv;
    a.tok_per_tile = 2*PF_ROWS/a.gqa;
    a.plan = reinterpret_cast<int *>(d.workspace);
    a.qscale = d.scale*1.4426950408889634f;
    const bool scratch = pf_needs_scratch(d);
    a.cap = d.bt_stride*d.block_size;
    if (scratch && d.max_seq_len > 0 && d.max_seq_len < a.cap) a.cap = d.max_seq_len;
    a.seq_off = nullptr; a.k16 = a.v16 = nullptr; a.scratch_cells = 0;
    char *ws = static_cast<char *>(d.workspace) + pf_plan_bytes(d);
    a.part_o = reinterpret_cast<float *>(ws);  ws += pf_part_o_bytes(d);
    a.part_ml = reinterpret_cast<float *>(ws); ws += pf_part_ml_bytes(d);
    a.grid_main = (int) pf_grid(d);
    {                                                        // contexts the host knows to be short never split: no pieces, no combine launch
        int64_t bound = d.bt_stride*d.block_size;
        if (d.max_seq_len > 0 && d.max_seq_len < bound) bound = d.max_seq_len;
        a.split_max = bound < 2*PF_MIN_PIECE*PF_BN ? 1 : PF_MAX_SPLIT;
    }
    if (scratch) {
        a.seq_off = a.plan + 2 + 2*(size_t) d.n_q;
        a.scratch_cells = pf_scratch_cells(d);
        const size_t tb = pf_align((size_t) a.scratch_cells*d.n_head_kv*PF_D*2);
        a.k16 = reinterpret_cast<__half *>(ws);
        a.v16 = reinterpret_cast<__half *>(ws + tb);
    }
    alignas(64) CUtensorMap tmK, tmV;
    if (scratch) {
        if (!pf_make_map(&tmK, a.k16, (int64_t) d.n_head_kv*256, 256, d.n_head_kv, a.scratch_cells)) return B200_ERR_CUDA;
        if (!pf_make_map(&tmV, a.v16, (int64_t) d.n_head_kv*256, 256, d.n_head_kv, a.scratch_cells)) return B200_ERR_CUDA;
    } else {
        if (!pf_make_map(&tmK, d.k, d.nbk1, d.nbk2, d.n_head_kv, d.n_kv)) return B200_ERR_CUDA;
        if (!pf_make_map(&tmV, d.v, d.nbv1, d.nbv2, d.n_head_kv, d.n_kv)) return B200_ERR_CUDA;
    }

    launch_pdl(pf_plan_kernel, dim3(1), dim3(1024), 0, st, a);
    count_launch();
    if (!check_launch("pf_plan")) return B200_ERR_CUDA;
    const int sms = pf_sm_count();
    if (scratch) {
        int64_t g = a.scratch_cells/16;
        if (g > 8*sms) g = 8*sms;
        if (g < 1) g = 1;
        launch_pdl(pf_dequant_kernel, dim3((unsigned) g), dim3(256), 0, st, a);
        count_launch();
        if (!check_launch("pf_dequant")) return B200_ERR_CUDA;
    }
    const int64_t grid = a.grid_main;
#ifdef PF_TIMELINE
    cudaMalloc(&a.dbg, 3*64*8*sizeof(long long)); cudaMemset(a.dbg, 0, 3*64*8*sizeof(long long));
#endif
    launch_pdl(pf_main_kernel, dim3((unsigned) grid), dim3(PF_THREADS), PF_SMEM, st, a, tmK, tmV);
#ifdef PF_TIMELINE
    {
        static int dumped = 0;
        cudaStreamSynchronize(st);
        if (dumped++ == 3) {
            static long long h[3*64*8];
            cudaMemcpy(h, a.dbg, sizeof(h), cudaMemcpyDeviceToHost);
            long long t0 = h[(2*64 + 0)*8 + 0];
            for (int w = 0; w < 3; ++w)
                for (int j = 0; j < 64; ++j) {
                    fprintf(stderr, "TL w%d j%02d", w, j);
                    for (int k = 0; k < 7; ++k) fprintf(stderr, " %6lld", h[(w*64 + j)*8 + k] ? h[(w*64 + j)*8 + k] - t0 : -1);
                    fprintf(stderr, "\n");
                }
        }
        cudaFree(a.dbg);
    }
#endif
    count_launch();
    if (!check_launch("pf_main")) return B200_ERR_CUDA;
    if (a.split_max > 1) {
        launch_pdl(pf_combine_kernel, dim3((unsigned) grid, 2*PF_ROWS/8), dim3(256), 0, st, a);   // the split items' pieces -> dst (returns at once when nothing was split)
        count_launch();
        if (!check_launch("pf_combine")) return B200_ERR_CUDA;
    }
    return B200_OK;
}

} // namespace b200
