// attn_streamk.cu — GQA-grouped paged decode attention on the tensor cores with a stream-K schedule.
//
// Replaces flash_attn_ext_f16 (MMA) + flash_attn_stream_k_fixup + to_fp16_cuda + flash_attn_vec_ext_f32 +
// flash_attn_combine_results of the reference (ggml/src/ggml-cuda/fattn-mma-f16.cuh:804-936,
// fattn-common.cuh:519-651, :723-749, fattn-vec-f32.cuh:8-284); semantics are those of
// ggml_compute_forward_flash_attn_ext_f16 (ggml/src/ggml-cpu/ggml-cpu.c:12219-12432) restricted to what a
// block table can express (a row sees positions [0, n_vis) of its own sequence; no dense mask is read).
//
// Why tensor cores for decode: one cached token costs 4*gqa*128 flop; at the measured 6.5 TB/s that is
// 26 / 49 / 92 TFLOP/s for F16 / Q8_0 / Q4_0 KV with gqa 4 — at or beyond the FP32 CUDA-core peak
// (~72 TFLOP/s) before any dequantisation work (profiles/r1a_generic_ncu_full.txt: the CUDA-core kernel
// reaches 6 % of HBM with the issue slots 50 % busy).  So the contraction runs on mma.sync m16n8k16 in the
// TRANSPOSED orientation: S^T[token][head] = K[token][:] . Q[head][:]^T with M = 16 tokens, N = 8 query heads
// of one kv head (gqa <= 8), then O^T[d][head] += V^T[d][token] . P^T[token][head].  tcgen05 needs M >= 64
// rows per instruction and a TMEM round trip per softmax step; for 4..8 query rows it has no advantage
// here (its home is the chunked-prefill tile kernel).
//
// Structure:
//   * sk_plan_kernel   1 CTA: tiles per row (16 tokens each) from the DEVICE seq_lens / q_pos, prefix sums,
//                      chunk = ceil(total / n_items): item i owns global tiles [i*chunk, (i+1)*chunk).  An item
//                      crossing a row boundary yields one segment per row; segment slots are contiguous per row.
//   * sk_main_kernel   every WARP is an independent worker for (item, kv head): it streams its K/V tiles through a
//                      private ring of TMA tensor-map boxes completed on its own mbarriers (no CTA barrier in the
//                      loop; the first version used 16-byte cp.async and stalled on the LSU queues, see
//                      profiles/r1c_*), gathered through the block table: one box = the 16 cells of a tile x one
//                      head row.  F16 boxes are 3-D {128 B, 16 cells, 2 halves} with the 128-byte swizzle so that
//                      ldmatrix is bank-conflict free; Q8_0 / Q4_0 boxes are 2-D {144 | 80 B, 16 cells}, whose row
//                      pitch makes the fragment-order 32-bit loads conflict free.  Quantised tiles are
//                      dequantised in registers straight into mma A fragments (byte -> half by OR-ing the
//                      exponent of 1024; K scales applied to the FP32 scores per 32-element block, V scales
//                      fused into one HFMA2 per register).  Online softmax lives in registers (warp-shuffle
//                      max/sum).  The 8 warps of a CTA take the 8 kv heads of the same tokens, so a CTA consumes
//                      whole contiguous token rows of the flat llama.cpp cache.
//   * sk_combine_kernel merges the segments of rows that were split (rows with one segment are written
//                      directly by the main kernel).  One CTA per (row, kv head), started while the main kernel
//                      runs; it does not wait for that grid to drain but spins on the row's arrival counter,
//                      which the main kernel's warps bump after their partials are stored (sk_args::arrive).
//   Launch chain (programmatic dependent launch, common.cuh): layers 1.. of a b200_decode_step call read plan and block
//   table before they wait for their predecessor (sk_args::early); with the layer's kv_write in front
//   (b200_store_attention) the main kernel starts under the store and waits for it only in front of each row's last
//   tile, after the plan kernel has checked the slot mapping on the device (sk_args::after_store).  DESIGN.md 3.1.
#include "common.cuh"
#include <cuda.h>
#include <math.h>
#include <type_traits>
#include <climits>
#include <cstdlib>

namespace b200 {

#ifndef SK_BIAS_MMA
#define SK_BIAS_MMA 0                                         // 1: remove the V expansion bias with a constant-A MMA per m-tile
#endif
constexpr int SK_WARPS = 8;
constexpr int SK_MIN_CHUNK = 4;                              // tiles: a warp's share is never shorter (its prologue costs about two)
#ifndef SK_F16_WARPS
#define SK_F16_WARPS 8                                        // warps per CTA / ring stages of the kernels with an F16 operand
#define SK_F16_STAGES 3
#endif
#ifndef SK_Q8_WARPS
#define SK_Q8_WARPS 8                                         // warps per CTA of the Q8_0/Q8_0 kernel (8 = the three-stage form)
#endif
constexpr int SK_D     = 128;
constexpr int SK_TILE  = 16;           // tokens per tile

struct sk_args {
    b200_attn_desc d;
    int    gqa;
    int    n_items;                     // upper bound on items per kv head
    int    chunk_align;                 // chunk is rounded up to a multiple of this many tiles (1)
    int    chunk_min;                   // and is at least this many tiles
    int   *plan;                        // [4] {total_tiles, chunk, items, store slots checked} then P[n_q+1], FS[n_q+1]
    float *part_ml;                     // [max_segs][n_head][2]
    float *part_o;                      // [max_segs][n_head][D]
    float  qscale;                      // scale*log2(e), or scale/softcap when softcap != 0
    float  cap_log2e;                   // softcap*log2(e)
    uint32_t k65536;                    // 65536, read from the constant bank so ptxas keeps pack16 an IMAD (FMA pipe)
    // dense-mask mode (an unmodified ggml graph: flat cache + F16 mask): d.block_table / seq_lens / q_seq are the tile
    // lists sk_dense_scan_kernel wrote into the workspace; the mask is consulted again only on tiles it marked MIXED
    const char *mask; int64_t nb_mask1; // F16 [n_kv, >= n_q] or NULL
    float  alibi_m0, alibi_m1; int n_head_log2;   // ALiBi slopes (ggml-cpu.c:12294-12298); max_bias in d
    int   *dense_len, *dense_bt, *dense_qseq;     // scan outputs: [n_q] visible cells (16 x listed tiles), [n_q][n_kv/16] tile ids, [n_q] zeros
    int    n_tiles;                     // n_kv / 16
    // store overlap (launch_attn_streamk): the launch in front of the main kernel is this layer's kv_write of ONE new
    // token per sequence.  The plan kernel checks that slot r is the cell of sequence r's last position (plan[3]);
    // the main kernel then starts before that store has finished and waits for it only in front of a row's last tile.
    int    after_store;
    const int32_t *store_slots; int64_t store_cells;
    // early prologue (decode_step, layers 1..): plan and block table were read by the previous layer's launch already, so
    // this one reads them BEFORE it waits for its predecessor
    int    early;
    // more than 8 query heads per kv head (a multiple of 8): every 8 of them are a VIRTUAL kv head — d.n_head_kv and gqa
    // below are the virtual counts (gqa = 8), and vsplit virtual heads share one row of the cache (read once per virtual
    // head: twice the L2 traffic for gqa 16, still the tensor-core kernel instead of the CUDA-core catch-all)
    int    vsplit;
    // arrivals: [n_q][n_head_kv] segment counters.  A main-kernel warp bumps its row's counter after its stores; the
    // combine CTA of (row, kv head) spins on it instead of waiting for the whole main grid to drain, and zeroes it again
    int   *arrive;
};
constexpr int SK_MIXED = 1 << 30;       // tile-list entry flag: some cell of the tile is hidden or carries a bias

// ------------------------------------------------------------------------------------------------
// per-type tile geometry for D = 128.  WIN = bytes of one token row of the shared-memory tile.
// ------------------------------------------------------------------------------------------------
template <int T> struct sk_type;
template <> struct sk_type<B200_TYPE_F16>  { static constexpr int HEAD_BYTES = 256, WIN = 256, TILE = 4096, BLK = 0;  };
template <> struct sk_type<B200_TYPE_Q8_0> { static constexpr int HEAD_BYTES = 136, WIN = 144, TILE = 2304, BLK = 34; };
template <> struct sk_type<B200_TYPE_Q4_0> { static constexpr int HEAD_BYTES = 72,  WIN = 80,  TILE = 1280, BLK = 18; };

template <int TK, int TV>
struct sk_smem {
    static constexpr bool ANY16 = TK == B200_TYPE_F16 || TV == B200_TYPE_F16;
    static constexpr int KT = sk_type<TK>::TILE, VT = sk_type<TV>::TILE;
    static constexpr bool KFIRST = TK == B200_TYPE_F16 || TV != B200_TYPE_F16;     // F16 tiles need 1024-byte alignment
    static constexpr int KOFF = KFIRST ? 0 : VT, VOFF = KFIRST ? KT : 0;
    static constexpr int ALIGN = ANY16 ? 1024 : 128;
    static constexpr int STAGE = (KT + VT + ALIGN - 1)/ALIGN*ALIGN;                // one 16-token tile: K box + V box
    // quantised K and V: TWO tiles per loop iteration share one softmax step (the loop is instruction-issue
    // bound, profiles/r1f_*), so the ring must hold the 2 being consumed + at least 1 in flight
    static constexpr int NSUB = ANY16 ? 1 : 2;
    // With 3 stages the second tile of an iteration is requested at the top of the iteration that consumes it
    // (29 % of the warp's time in that wait, profiles/r1q_sk_q8_0_ncu_full.txt).  Tiles up to 3584 bytes take a fourth
    // stage at 8 warps.  Q8_0 K with Q8_0 V (4608) only fits four stages with 6 warps per CTA (SK_Q8_WARPS=6, register
    // cap 168): measured equal to 8 warps x 3 stages (5229 vs 5256 GB/s), so 8 stays the default
    static constexpr int WARPS = ANY16 ? SK_F16_WARPS : STAGE <= 3584 ? SK_WARPS : SK_Q8_WARPS;
    static constexpr int STAGES = ANY16 ? (STAGE <= 2560 ? 4 : SK_F16_STAGES) : (STAGE <= 2560 ? 5 : (WARPS == SK_WARPS && STAGE > 3584 ? 3 : 4));
    static constexpr int WARP = STAGE*STAGES;
    static constexpr int BAR_BYTES = ANY16 ? 1024 : 512;                          // mbarriers; F16 tiles must stay 1024-aligned
    static constexpr int META_OFF = WARPS*STAGES*8;                               // dense mode: one tile-list entry per ring slot, behind the barriers
    static_assert(META_OFF + WARPS*STAGES*4 <= BAR_BYTES, "barrier area too small for the dense-mode slot words");
    static constexpr int CTA = BAR_BYTES + WARP*WARPS + 16;                       // +16: funnel loads may touch 4 bytes past a row
    static constexpr int CTAS_PER_SM = 2*(CTA + 1024) <= 233472 ? 2 : 1;
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile("{\n .reg .pred p;\nSK_WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra SK_DONE_%=;\n bra SK_WAIT_%=;\nSK_DONE_%=:\n}"
                 :: "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *tm, int c0, int c1, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 :: "r"(dst), "l"(tm), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *tm, int c0, int c1, int c2, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 :: "r"(dst), "l"(tm), "r"(c0), "r"(c1), "r"(c2), "r"(bar) : "memory");
}

// One tile's K box + V box and their expect_tx, issued by ONE elected lane of the converged warp.  With elect.sync
// ptxas knows a single thread runs the copies and emits plain UTMALDG; `if (lane == 0)` costs an 11-instruction
// vote loop per copy (profiles/r1i_*: the issue path was 16 % of the decode loop's instructions).
template <bool K3D, bool V3D>
__device__ __forceinline__ void sk_issue_elect(uint32_t bar, uint32_t bytes, uint32_t dk, const CUtensorMap *tmK, int kc,
                                               uint32_t dv, const CUtensorMap *tmV, int vc, int cell) {
    asm volatile("{\n .reg .pred p;\n elect.sync _|p, 0xffffffff;\n"
                 " @p mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n}" :: "r"(bar), "r"(bytes) : "memory");
    if constexpr (K3D)
        asm volatile("{\n .reg .pred p;\n elect.sync _|p, 0xffffffff;\n"
                     " @p cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];\n}"
                     :: "r"(dk), "l"(tmK), "r"(0), "r"(cell), "r"(kc), "r"(bar) : "memory");
    else
        asm volatile("{\n .reg .pred p;\n elect.sync _|p, 0xffffffff;\n"
                     " @p cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];\n}"
                     :: "r"(dk), "l"(tmK), "r"(kc), "r"(cell), "r"(bar) : "memory");
    if constexpr (V3D)
        asm volatile("{\n .reg .pred p;\n elect.sync _|p, 0xffffffff;\n"
                     " @p cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];\n}"
                     :: "r"(dv), "l"(tmV), "r"(0), "r"(cell), "r"(vc), "r"(bar) : "memory");
    else
        asm volatile("{\n .reg .pred p;\n elect.sync _|p, 0xffffffff;\n"
                     " @p cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];\n}"
                     :: "r"(dv), "l"(tmV), "r"(vc), "r"(cell), "r"(bar) : "memory");
}

__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t a) {
    uint32_t d;
    asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;\n" : "=r"(d) : "r"(a));
    return d;
}
// D(16x8, f32) += A(16x16, f16, row) * B(16x8, f16, col)
__device__ __forceinline__ void mma_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_h2(float lo, float hi) {
    const __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t *>(&h);
}
__device__ __forceinline__ float2 unpack_h2(uint32_t v) { return __half22float2(*reinterpret_cast<const __half2 *>(&v)); }
__device__ __forceinline__ uint32_t hfma2_u(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d; asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d;
}
__device__ __forceinline__ uint32_t hmul2_u(uint32_t a, uint32_t b) {
    uint32_t d; asm("mul.rn.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d;
}
__device__ __forceinline__ float ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t lds32(uint32_t addr) { uint32_t v; asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr)); return v; }
__device__ __forceinline__ uint32_t lds16(uint32_t addr) { uint32_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(addr)); return v; }
__device__ __forceinline__ void sts128_zero(uint32_t addr) { asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" :: "r"(addr), "r"(0) : "memory"); }
__device__ __forceinline__ float h16f(uint32_t bits) { return __half2float(__ushort_as_half((unsigned short) bits)); }

// int8 / nibble -> half by OR-ing the exponent of 1024 (0x6400): half(0x6400 | u) == 1024 + u exactly.
// Q8_0: the word is biased to unsigned first (xor 0x80 per byte, one LOP3 per 4 elements), then one PRMT per
// register pair drops 0x64 above bytes (0, 2) / (1, 3): value = 1152 + q.  k64 is 0x64646464 held in a register.
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
    uint32_t d; asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel)); return d;
}
__device__ __forceinline__ uint32_t q8_bias(uint32_t w) { return w ^ 0x80808080u; }
__device__ __forceinline__ uint32_t q8_e02(uint32_t wb, uint32_t k64) { return prmt(wb, k64, 0x4240); }   // (b0, 0x64, b2, 0x64)
__device__ __forceinline__ uint32_t q8_e13(uint32_t wb, uint32_t k64) { return prmt(wb, k64, 0x4341); }   // (b1, 0x64, b3, 0x64)
// Q4_0: (w & mask) | 0x64006400 as ONE three-input LOP3 (mask in a register): low nibbles give 1024 + n,
// high nibbles 1024 + 16 n
__device__ __forceinline__ uint32_t lop3_and_or(uint32_t a, uint32_t m, uint32_t c) {
    uint32_t d; asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(d) : "r"(a), "r"(m), "r"(c)); return d;
}
__device__ __forceinline__ uint32_t q4_lo(uint32_t w, uint32_t mlo) { return lop3_and_or(w, mlo, 0x64006400u); }
__device__ __forceinline__ uint32_t q4_hi(uint32_t w, uint32_t mhi) { return lop3_and_or(w, mhi, 0x64006400u); }
__device__ __forceinline__ uint32_t hadd2_u(uint32_t a, uint32_t b) {
    uint32_t d; asm("add.rn.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d;
}
// a constant the optimiser cannot fold back into immediates (keeps it in one register across the loop)
__device__ __forceinline__ uint32_t sk_opaque(uint32_t v) { asm volatile("mov.b32 %0, %0;" : "+r"(v)); return v; }

// ------------------------------------------------------------------------------------------------
// plan: tiles per row, prefix sums, chunk, segment slots.  One CTA of 1024 threads.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ int sk_row_nvis(const b200_attn_desc &p, int r) {
    const int seq = p.q_seq ? p.q_seq[r] : r;
    int64_t n = p.seq_lens[seq];
    if (p.q_pos) { const int64_t lim = (int64_t) p.q_pos[r] + 1; n = lim < n ? lim : n; }
    const int64_t cap = p.bt_stride*p.block_size;
    if (n > cap) n = cap;
    return n < 0 ? 0 : (int) n;
}

// exclusive block scan of one int per thread; returns the exclusive prefix, *total gets the block sum
__device__ __forceinline__ int sk_block_scan(int v, int *s_warp, int *total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    if (warp == 0) {
        int w = lane < nw ? s_warp[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += y; }
        s_warp[lane] = w;                                   // inclusive over warps
    }
    __syncthreads();
    const int base = warp ? s_warp[warp - 1] : 0;
    *total = s_warp[nw - 1];
    __syncthreads();
    return base + x - v;
}

__global__ void __launch_bounds__(1024)
sk_plan_kernel(const __grid_constant__ sk_args a) {
    __shared__ int s_warp[32];
    pdl_prologue();
    const b200_attn_desc &p = a.d;
    int *P = a.plan + 4, *FS = P + (p.n_q + 1);
    for (int64_t i = threadIdx.x; i < (int64_t) p.n_q*p.n_head_kv; i += 1024) a.arrive[i] = 0;
    int carry = 0;
    for (int r0 = 0; r0 < p.n_q; r0 += 1024) {
        const int r = r0 + threadIdx.x;
        const int tiles = r < p.n_q ? (sk_row_nvis(p, r) + SK_TILE - 1)/SK_TILE : 0;
        int tot;
        const int ex = sk_block_scan(tiles, s_warp, &tot);
        if (r < p.n_q) P[r] = carry + ex;
        carry += tot;
    }
    const int total = carry;
    int chunk = (total + a.n_items - 1)/a.n_items;
    if (chunk < 1) chunk = 1;
    if (chunk < a.chunk_min) chunk = a.chunk_min;
    chunk = (chunk + a.chunk_align - 1)/a.chunk_align*a.chunk_align;
    // Row-aligned chunks: when some chunk up to a third longer divides every row's first tile index (equal-length
    // batches, the llama-batched-bench shape), no warp straddles two rows and pays two prologues, two partial
    // stores: 32 x 1024 Q4_0, 7 -> 8 tiles per warp, 1880 -> 2024 GB/s.  Only for short chunks, where that overhead counts.
    if (chunk <= 32 && a.chunk_align == 1) {
        __syncthreads();                                     // P[] of the whole CTA
        int best = 0;
        for (int c = chunk; c <= chunk + (chunk + 2)/3 && !best; ++c) {
            int ok = 1;
            for (int r = threadIdx.x; r < p.n_q; r += 1024) ok &= (P[r] % c == 0);
            if (__syncthreads_and(ok)) best = c;
        }
        if (best) chunk = best;
    }
    // store overlap: every stored row must land on the LAST visible cell of its own sequence (or be skipped by kv_write)
    int slots_ok = 0;
    if (a.after_store) {                                     // uniform.  Row r must be THE row of sequence r (q_seq absent or the identity):
        int ok = 1;                                          // no other row then reads the cell row r's store writes
        for (int r = threadIdx.x; r < p.n_q; r += 1024) {
            if (p.q_seq && p.q_seq[r] != r) ok = 0;
            const int64_t slot = a.store_slots ? (int64_t) a.store_slots[r] : r;
            if (slot < 0 || slot >= a.store_cells) continue;            // kv_write skips it
            const int n = sk_row_nvis(p, r);                            // with q_pos: the row's own position + 1
            ok &= n > 0 && slot == (int64_t) p.block_table[(int64_t) r*p.bt_stride + (n - 1)/p.block_size]*p.block_size + (n - 1) % p.block_size;
        }
        slots_ok = __syncthreads_and(ok);
    }
    if (threadIdx.x == 0) { P[p.n_q] = total; a.plan[0] = total; a.plan[1] = chunk; a.plan[2] = (total + chunk - 1)/chunk; a.plan[3] = slots_ok; }
    __syncthreads();
    carry = 0;
    for (int r0 = 0; r0 < p.n_q; r0 += 1024) {
        const int r = r0 + threadIdx.x;
        int nseg = 0;
        if (r < p.n_q) {
            const int b = P[r], e = (r + 1 < p.n_q) ? P[r + 1] : total;      // P[r+1] was written by this CTA
            if (e > b) nseg = (e - 1)/chunk - b/chunk + 1;
        }
        int tot;
        const int ex = sk_block_scan(nseg, s_warp, &tot);
        if (r < p.n_q) FS[r] = carry + ex;
        carry += tot;
    }
    if (threadIdx.x == 0) FS[p.n_q] = carry;
}

// ------------------------------------------------------------------------------------------------
// dense-mask mode: the mask of an unmodified ggml graph -> per-row lists of 16-cell tiles
// ------------------------------------------------------------------------------------------------
// The reference's kernels add the mask after the dot product (fattn-vec-f32.cuh:197, fattn-mma-f16.cuh), so every query
// row streams all n_kv cells of all sequences.  Here the mask (F16 [n_kv, >= n_q], filled by
// llm_graph_input_attn_kv_unified::set_input, src/llama-graph.cpp:403-477) is read ONCE: row r gets the ascending list of
// tiles holding at least one cell it may see (entry = tile id, | SK_MIXED when some cell of the tile is hidden or carries a
// non-zero bias), and the stream-K plan runs over those lists exactly as over a block table with 16-cell blocks
// (sequence = row, seq_lens[r] = 16 x listed tiles).  One CTA per row; a thread classifies 4 consecutive tiles (128
// contiguous mask bytes) per step.
__global__ void __launch_bounds__(1024)
sk_dense_scan_kernel(const __grid_constant__ sk_args a) {
    __shared__ int s_warp[32];
    pdl_prologue();
    const int row = blockIdx.x, n_tiles = a.n_tiles;
    int *out = a.dense_bt + (int64_t) row*n_tiles;
    if (a.mask == nullptr) {                                  // no mask (grid of one CTA): every row sees every cell, one identity list serves all
        for (int t = threadIdx.x; t < n_tiles; t += 1024) out[t] = t;
        for (int r = threadIdx.x; r < a.d.n_q; r += 1024) { a.dense_qseq[r] = 0; a.dense_len[r] = r == 0 ? n_tiles*SK_TILE : 0; }
        return;
    }
    if (threadIdx.x == 0) a.dense_qseq[row] = row;
    const uint4 *m = reinterpret_cast<const uint4 *>(a.mask + (int64_t) row*a.nb_mask1);
    int base = 0;
    for (int t0 = 0; t0 < n_tiles; t0 += 4096) {
        const int tile0 = t0 + 4*threadIdx.x;
        uint4 x[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) x[j] = tile0 + j/2 < n_tiles ? m[2*tile0 + j] : make_uint4(0xFC00FC00u, 0xFC00FC00u, 0xFC00FC00u, 0xFC00FC00u);
        int ent[4], cnt = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint4 u = x[2*j], v = x[2*j + 1];
            const bool hidden = u.x == 0xFC00FC00u && u.y == 0xFC00FC00u && u.z == 0xFC00FC00u && u.w == 0xFC00FC00u
                             && v.x == 0xFC00FC00u && v.y == 0xFC00FC00u && v.z == 0xFC00FC00u && v.w == 0xFC00FC00u;     // sixteen times -inf
            const bool clean = ((u.x | u.y | u.z | u.w | v.x | v.y | v.z | v.w) & 0x7FFF7FFFu) == 0;                       // sixteen times +-0
            ent[j] = hidden ? -1 : ((tile0 + j) | (clean ? 0 : SK_MIXED));
            cnt += hidden ? 0 : 1;
        }
        int tot;
        int o = base + sk_block_scan(cnt, s_warp, &tot);
#pragma unroll
        for (int j = 0; j < 4; ++j) { if (ent[j] >= 0) out[o++] = ent[j]; }
        base += tot;
    }
    if (threadIdx.x == 0) a.dense_len[row] = base*SK_TILE;
}

// ------------------------------------------------------------------------------------------------
// Q as the B operand of S^T = K.Q^T  (lane = 4n + t holds column n = query head, contraction slots of t)
// ------------------------------------------------------------------------------------------------
// F16 K: contraction index = d in natural order.  k-step s: b0 = Q[n][16s+2t, +1], b1 = Q[n][16s+2t+8, +9].
// Quantised K: 32-element block `blk` feeds k-steps 2blk (elements 0..15 of the block) and 2blk+1 (16..31); the
// byte->half expansion pairs (byte0, byte2) / (byte1, byte3) of the lane's word, so b0 = Q[e0], Q[e0+2] and
// b1 = Q[e0+1], Q[e0+3] with e0 = 32blk + 16hi + 4t.  Q4_0 high nibbles arrive as 1024 + 16n: their Q is /16.
// cb[blk][c] = the constant the biased integers add to column 2t+c of block blk (removed from the FP32 scores).
template <int TK>
struct sk_qfrag {
    static constexpr bool QUANT = TK != B200_TYPE_F16;
    uint32_t b[8][2];
    float    cb[QUANT ? 4 : 1][2];

    __device__ __forceinline__ void load(const float *qrow, bool live, float qscale, int lane) {
        const int t = lane & 3;
        if constexpr (!QUANT) {
#pragma unroll
            for (int s = 0; s < 8; ++s) {
                float2 x = make_float2(0.f, 0.f), y = make_float2(0.f, 0.f);
                if (live) {
                    x = *reinterpret_cast<const float2 *>(qrow + 16*s + 2*t);
                    y = *reinterpret_cast<const float2 *>(qrow + 16*s + 2*t + 8);
                }
                b[s][0] = pack_h2(x.x*qscale, x.y*qscale);
                b[s][1] = pack_h2(y.x*qscale, y.y*qscale);
            }
        } else {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
                const int hi = ks & 1;
                float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
                if (live) x = *reinterpret_cast<const float4 *>(qrow + 32*(ks >> 1) + 16*hi + 4*t);
                const float sc = (TK == B200_TYPE_Q4_0 && hi) ? qscale*0.0625f : qscale;
                b[ks][0] = pack_h2(x.x*sc, x.z*sc);
                b[ks][1] = pack_h2(x.y*sc, x.w*sc);
            }
#pragma unroll
            for (int blk = 0; blk < 4; ++blk) {
                const float2 a0 = unpack_h2(b[2*blk][0]), a1 = unpack_h2(b[2*blk][1]);
                const float2 c0 = unpack_h2(b[2*blk + 1][0]), c1 = unpack_h2(b[2*blk + 1][1]);
                const float slo = (a0.x + a0.y) + (a1.x + a1.y), shi = (c0.x + c0.y) + (c1.x + c1.y);
                float v = TK == B200_TYPE_Q8_0 ? 1152.0f*(slo + shi)                         // sum (1152 + q) Q
                                               : 1024.0f*(slo + shi) + 8.0f*(slo + 16.0f*shi);   // sum (1024 + s n) Q/s, and the -8
                v += __shfl_xor_sync(0xffffffffu, v, 1);
                v += __shfl_xor_sync(0xffffffffu, v, 2);
                cb[blk][0] = __shfl_sync(0xffffffffu, v, 8*t);
                cb[blk][1] = __shfl_sync(0xffffffffu, v, 8*t + 4);
            }
        }
    }
};

// ------------------------------------------------------------------------------------------------
// S^T tile (16 tokens x 8 heads), FP32, log2 domain
// ------------------------------------------------------------------------------------------------
// F16 tile: [2 halves][16 tokens][128 B], 16-byte pieces XOR-swizzled with (token & 7) (TMA SWIZZLE_128B)
__device__ __forceinline__ void sk_qk_f16(float (&S)[4], uint32_t kb, const sk_qfrag<B200_TYPE_F16> &q, int lane) {
    const int tau = (lane & 7) + 8*((lane >> 3) & 1);
    const uint32_t base = kb + tau*128;
    const uint32_t y = ((tau & 7) << 4) ^ ((lane >> 4) << 4);
    float S2[4] = {0.f, 0.f, 0.f, 0.f};                      // two independent accumulation chains
#pragma unroll
    for (int s = 0; s < 8; s += 2) {
        uint32_t a[4], b[4];
        ldmatrix_x4(a, base + (s >> 2)*2048 + ((((2*s) & 7) << 4) ^ y));
        ldmatrix_x4(b, base + ((s + 1) >> 2)*2048 + ((((2*(s + 1)) & 7) << 4) ^ y));
        mma_16816(S,  a, q.b[s][0],     q.b[s][1]);
        mma_16816(S2, b, q.b[s + 1][0], q.b[s + 1][1]);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) S[j] += S2[j];
}

// fetch 4 bytes at byte offset `off` of a row; off % 4 is 0 (ALIGNED) or 2 (funnel of two words)
template <bool ALIGNED>
__device__ __forceinline__ uint32_t sk_ld4(uint32_t row, int off) {
    if constexpr (ALIGNED) return lds32(row + off);
    else return __byte_perm(lds32(row + off - 2), lds32(row + off + 2), 0x5432);
}

// Quantised tile: [16 tokens][WIN bytes]; the head row starts `mis` (0 or 8) bytes into each token row.
// Even blocks: scale in the low half of its word, quants 2 bytes off word alignment; odd blocks: high half, aligned.
template <int TK>
__device__ __forceinline__ void sk_qk_quant(float (&S)[4], uint32_t kb, int mis, const sk_qfrag<TK> &q, uint2 kc, int lane) {
    constexpr int WIN = sk_type<TK>::WIN, BLK = sk_type<TK>::BLK;
    const int g = lane >> 2, t = lane & 3;
    const uint32_t row0 = kb + g*WIN + mis, row1 = row0 + 8*WIN;
#pragma unroll
    for (int blk = 0; blk < 4; ++blk) {
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        const int o = BLK*blk + 2 + 4*t;
        if constexpr (TK == B200_TYPE_Q8_0) {
            uint32_t A0, A1, B0, B1;
            if (blk & 1) { A0 = sk_ld4<true >(row0, o); A1 = sk_ld4<true >(row1, o); B0 = sk_ld4<true >(row0, o + 16); B1 = sk_ld4<true >(row1, o + 16); }
            else         { A0 = sk_ld4<false>(row0, o); A1 = sk_ld4<false>(row1, o); B0 = sk_ld4<false>(row0, o + 16); B1 = sk_ld4<false>(row1, o + 16); }
            A0 = q8_bias(A0); A1 = q8_bias(A1); B0 = q8_bias(B0); B1 = q8_bias(B1);
            const uint32_t a[4] = {q8_e02(A0, kc.x), q8_e02(A1, kc.x), q8_e13(A0, kc.x), q8_e13(A1, kc.x)};
            const uint32_t c[4] = {q8_e02(B0, kc.x), q8_e02(B1, kc.x), q8_e13(B0, kc.x), q8_e13(B1, kc.x)};
            mma_16816(acc, a, q.b[2*blk][0],     q.b[2*blk][1]);
            mma_16816(acc, c, q.b[2*blk + 1][0], q.b[2*blk + 1][1]);
        } else {
            uint32_t W0, W1;
            if (blk & 1) { W0 = sk_ld4<true >(row0, o); W1 = sk_ld4<true >(row1, o); }
            else         { W0 = sk_ld4<false>(row0, o); W1 = sk_ld4<false>(row1, o); }
            const uint32_t Y0 = W0 >> 8, Y1 = W1 >> 8;
            const uint32_t a[4] = {q4_lo(W0, kc.x), q4_lo(W1, kc.x), q4_lo(Y0, kc.x), q4_lo(Y1, kc.x)};
            const uint32_t c[4] = {q4_hi(W0, kc.y), q4_hi(W1, kc.y), q4_hi(Y0, kc.y), q4_hi(Y1, kc.y)};
            mma_16816(acc, a, q.b[2*blk][0],     q.b[2*blk][1]);
            mma_16816(acc, c, q.b[2*blk + 1][0], q.b[2*blk + 1][1]);
        }
        const float d0 = h16f(lds16(row0 + BLK*blk)), d1 = h16f(lds16(row1 + BLK*blk));
        S[0] = fmaf(d0, acc[0] - q.cb[blk][0], S[0]);
        S[1] = fmaf(d0, acc[1] - q.cb[blk][1], S[1]);
        S[2] = fmaf(d1, acc[2] - q.cb[blk][0], S[2]);
        S[3] = fmaf(d1, acc[3] - q.cb[blk][1], S[3]);
    }
}

// ------------------------------------------------------------------------------------------------
// O^T += V^T . P^T ; 8 m-tiles of 16 d-rows.
// Quantised V costs one HADD2 per operand register (two elements) beyond the byte/nibble -> half expansion:
//   * the block scale d[token][blk] is folded into the B operand: b = P * d (one HMUL2 per register and block, the
//     16 rows of an m-tile all belong to ONE 32-element block, so the scale depends on (token, block) only);
//   * the bias of the exponent trick (1152 + q for Q8_0; 1024 + n and 1024 + 16 n for Q4_0's low / high nibbles, whose b
//     carries an extra 1/16 for the high ones) leaves the A operand by one exact HADD2 per register: (1152 + q) - 1152,
//     (1024 + n) - 1032 = n - 8, (1024 + 16 n) - 1152 = 16 (n - 8) are all representable.  (SK_BIAS_MMA=1 is the earlier
//     form: one more MMA per m-tile with A = -bias everywhere.  ptxas rebuilt that constant 4-register operand with
//     3 MOVs per MMA under the 128-register cap, 9 % of the loop's instructions; r1v: Q4_0 3650 -> 3895 GB/s without it.)
//     Result: sum_tok (q | n - 8) * d * P, i.e. dequantize_row_*
//     (ggml/src/ggml-quants.c:255-273, :349-362) times P with P*d rounded to half once.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void sk_pv_f16(float (&O)[8][4], uint32_t vb, uint32_t p0, uint32_t p1, int lane) {
    const int tau = (lane & 7) + 8*(lane >> 4);
    const uint32_t base = vb + tau*128;
    const uint32_t y = ((tau & 7) << 4) ^ (((lane >> 3) & 1) << 4);
#pragma unroll
    for (int mu = 0; mu < 8; ++mu) {
        uint32_t a[4];
        ldmatrix_x4_trans(a, base + (mu >> 2)*2048 + ((((2*mu) & 7) << 4) ^ y));
        mma_16816(O[mu], a, p0, p1);
    }
}

// (hi << 16) | lo for lo < 65536, as an integer multiply-add: runs on the FMA pipe, which this kernel leaves
// half idle, instead of a PRMT on the ALU pipe, which bounds it (profiles/r1h_*)
// k65536 must be an opaque register: with an immediate ptxas strength-reduces the multiply back to LEA (ALU pipe)
__device__ __forceinline__ uint32_t pack16(uint32_t lo, uint32_t hi, uint32_t k65536) {
    uint32_t d; asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(hi), "r"(k65536), "r"(lo)); return d;
}

template <int TV>
__device__ __forceinline__ void sk_pv_quant(float (&O)[8][4], uint32_t vb, int mis, uint32_t p0, uint32_t p1, uint2 kc, uint32_t k64k, int lane) {
    constexpr int WIN = sk_type<TV>::WIN, BLK = sk_type<TV>::BLK;
    const int g = lane >> 2, t = lane & 3;
    const uint32_t r0 = vb + 2*t*WIN + mis, r1 = r0 + WIN, r2 = r0 + 8*WIN, r3 = r2 + WIN;   // tokens 2t, 2t+1, 2t+8, 2t+9
    const uint32_t NEG = sk_opaque(0xE480E480u);             // half2(-1152): Q8_0 bias, and Q4_0 high nibbles (72 * 16)
    const uint32_t neg[4] = {NEG, NEG, NEG, NEG};
#pragma unroll
    for (int blk = 0; blk < 4; ++blk) {
        const int so = BLK*blk;                              // block scale: 2-byte aligned
        const uint32_t b0 = hmul2_u(p0, pack16(lds16(r0 + so), lds16(r1 + so), k64k));              // P * (d[2t], d[2t+1])
        const uint32_t b1 = hmul2_u(p1, pack16(lds16(r2 + so), lds16(r3 + so), k64k));
        uint32_t a[4], c[4];
        if constexpr (TV == B200_TYPE_Q8_0) {                // lane g: bytes 4g..4g+3 of the block -> d = 32 blk + 4g + {0,1 | 2,3}
            const int o = BLK*blk + 2 + 4*g;
            uint32_t W0, W1, W2, W3;
            if (blk & 1) { W0 = sk_ld4<true >(r0, o); W1 = sk_ld4<true >(r1, o); W2 = sk_ld4<true >(r2, o); W3 = sk_ld4<true >(r3, o); }
            else         { W0 = sk_ld4<false>(r0, o); W1 = sk_ld4<false>(r1, o); W2 = sk_ld4<false>(r2, o); W3 = sk_ld4<false>(r3, o); }
            const uint32_t x01l = q8_bias(prmt(W0, W1, 0x5410)), x01h = q8_bias(prmt(W0, W1, 0x7632));
            const uint32_t x23l = q8_bias(prmt(W2, W3, 0x5410)), x23h = q8_bias(prmt(W2, W3, 0x7632));
            a[0] = q8_e02(x01l, kc.x); a[1] = q8_e13(x01l, kc.x); a[2] = q8_e02(x23l, kc.x); a[3] = q8_e13(x23l, kc.x);
            c[0] = q8_e02(x01h, kc.x); c[1] = q8_e13(x01h, kc.x); c[2] = q8_e02(x23h, kc.x); c[3] = q8_e13(x23h, kc.x);
            if constexpr (SK_BIAS_MMA) {
                mma_16816(O[2*blk],     neg, b0, b1);
                mma_16816(O[2*blk + 1], neg, b0, b1);
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) { a[j] = hadd2_u(a[j], NEG); c[j] = hadd2_u(c[j], NEG); }     // (1152 + q) - 1152, exact
            }
            mma_16816(O[2*blk],     a,   b0, b1);
            mma_16816(O[2*blk + 1], c,   b0, b1);
        } else {                                             // lane g: bytes 2g, 2g+1 -> d = 32 blk + 2g + {0,1} (low), + 16 (high nibbles)
            const int o = BLK*blk + 2 + 2*g;
            const uint32_t x01 = pack16(lds16(r0 + o), lds16(r1 + o), k64k), x23 = pack16(lds16(r2 + o), lds16(r3 + o), k64k);
            const uint32_t y01 = x01 >> 8, y23 = x23 >> 8;
            a[0] = q4_lo(x01, kc.x); a[1] = q4_lo(y01, kc.x); a[2] = q4_lo(x23, kc.x); a[3] = q4_lo(y23, kc.x);   // 1024 + n
            c[0] = q4_hi(x01, kc.y); c[1] = q4_hi(y01, kc.y); c[2] = q4_hi(x23, kc.y); c[3] = q4_hi(y23, kc.y);   // 1024 + 16 n
            const uint32_t NEGLO = 0xE408E408u;              // half2(-1032)
            const uint32_t neglo[4] = {NEGLO, NEGLO, NEGLO, NEGLO};
            const uint32_t h0 = hmul2_u(b0, 0x2C002C00u), h1 = hmul2_u(b1, 0x2C002C00u);      // * 1/16 (exact)
            if constexpr (SK_BIAS_MMA) {
                mma_16816(O[2*blk],     neglo, b0, b1);
                mma_16816(O[2*blk + 1], neg,   h0, h1);
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) { a[j] = hadd2_u(a[j], NEGLO); c[j] = hadd2_u(c[j], NEG); }   // n - 8 and 16 (n - 8), exact
            }
            mma_16816(O[2*blk],     a,     b0, b1);
            mma_16816(O[2*blk + 1], c,     h0, h1);
        }
    }
}

// d index of accumulator row (m-tile mt, half h: 0 = row g, 1 = row g+8)
template <int TV>
__device__ __forceinline__ int sk_row_d(int mt, int h, int g) {
    if constexpr (TV == B200_TYPE_F16)       return 16*mt + g + 8*h;
    else if constexpr (TV == B200_TYPE_Q8_0) return 32*(mt >> 1) + 4*g + 2*(mt & 1) + h;
    else                                     return 32*(mt >> 1) + 16*(mt & 1) + 2*g + h;
}
// ------------------------------------------------------------------------------------------------
// main kernel
// ------------------------------------------------------------------------------------------------
template <int TK, int TV, bool DENSE>
__global__ void __launch_bounds__(32*sk_smem<TK, TV>::WARPS, sk_smem<TK, TV>::CTAS_PER_SM)
sk_main_kernel(const __grid_constant__ sk_args a, const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV) {
    extern __shared__ __align__(1024) char sk_dyn[];
    using SM = sk_smem<TK, TV>;
    constexpr int STAGES = SM::STAGES;
    const b200_attn_desc &p = a.d;
    int lane = threadIdx.x & 31;
    asm volatile("mov.b32 %0, %0;" : "+r"(lane));             // pin in a register: ptxas otherwise re-reads SR_TID in the loop
    const int warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    TL_BEGIN(3, threadIdx.x == 0);        // s2 first TMA issue, s3 past the wait for the predecessor, s4 first tile consumed, s5 first row's loop end
    // Where this grid waits for its predecessor, and when it lets the combine grid in (which spins on a.arrive and may
    // only start once the PREVIOUS combine on this workspace has completed):
    //   default      wait, then trigger, at the top
    //   early        plan / block-table reads first, then wait + trigger in front of the first TMA issue
    //   after_store  trigger at the top (the kv_write in front let this grid start only after its own predecessor had
    //                completed: pdl_prologue_strict); nothing read before a row's last tile was written by that store,
    //                so the wait stands in front of that tile only
    bool waited = false;
    if (!a.after_store && !a.early) { pdl_wait(); waited = true; TL_STAMP(3); }
    if (waited || a.after_store) pdl_trigger();

    const int wid  = blockIdx.x*SM::WARPS + warp;
    const int item = wid / p.n_head_kv, kvh = wid - item*p.n_head_kv;
    // the first 128 bytes of the plan in ONE coalesced load: the header {total, chunk} and P[0..27]
    const int *P = a.plan + 4, *FS = P + (p.n_q + 1);
    const int head = lane < 4 + p.n_q + 1 ? a.plan[lane] : INT_MAX;
    const int total = __shfl_sync(0xffffffffu, head, 0), chunk = __shfl_sync(0xffffffffu, head, 1);
    if (a.after_store && !waited && __shfl_sync(0xffffffffu, head, 3) == 0) { pdl_wait(); waited = true; TL_STAMP(3); }   // the plan could not vouch for the slots
    int g0 = item*chunk;
    if (g0 >= total) { if (!waited) pdl_wait(); TL_END(); return; }
    const int g1 = min(g0 + chunk, total);

    // row containing global tile g0: last r with P[r] <= g0 (rows without tiles share P with their successor)
    // 32-ary search: each level is ONE coalesced load + ballot (a binary search is log2(n_q) DEPENDENT global loads
    // in front of the first TMA issue; on an 8 x 1024 problem the whole op lasts ~10 us).  Up to 27 rows are settled
    // by the header load itself: P is non-decreasing, so the lanes with P[r] <= g0 form a prefix.
    int lo = __popc(__ballot_sync(0xffffffffu, lane >= 4 && head <= g0)) - 1, hi = p.n_q;   // invariant: P[lo] <= g0 < P[hi]
    if (lo < 27) hi = min(hi, lo + 1);
    while (hi - lo > 1) {
        const int step = (hi - lo + 31) >> 5, idx = lo + step*(lane + 1);
        const int k = __popc(__ballot_sync(0xffffffffu, idx < hi && P[idx] <= g0));   // P is non-decreasing: a prefix of lanes
        hi = min(hi, lo + step*(k + 1));
        lo += step*k;
    }
    int row = lo;

    uint32_t smem0 = smem_u32(sk_dyn) + SM::BAR_BYTES + warp*SM::WARP;
    uint32_t bar0  = smem_u32(sk_dyn) + warp*STAGES*8;
    const uint32_t meta0 = smem_u32(sk_dyn) + SM::META_OFF + warp*STAGES*4;      // DENSE: the tile-list entry of each ring slot
    asm volatile("mov.b32 %0, %0;" : "+r"(smem0));
    asm volatile("mov.b32 %0, %0;" : "+r"(bar0));
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) mbar_init(bar0 + 8*s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();

    // head-row placement: quantised rows are only 8-byte aligned, the TMA box starts `mis` bytes earlier
    const int64_t koff = (int64_t) (kvh/a.vsplit)*p.nbk2, voff = (int64_t) (kvh/a.vsplit)*p.nbv2;
    const int kmis = TK == B200_TYPE_F16 ? 0 : (int) (koff & 15), vmis = TV == B200_TYPE_F16 ? 0 : (int) (voff & 15);
    const int kc = TK == B200_TYPE_F16 ? (int) (koff >> 7) : (int) (koff - kmis);     // TMA coordinate of the head row
    const int vc = TV == B200_TYPE_F16 ? (int) (voff >> 7) : (int) (voff - vmis);
    const int bs = p.block_size, tpb = bs/SK_TILE;                                    // tiles per block (bs % 16 == 0)
    const int n_blk = p.bt_stride < INT32_MAX ? (int) p.bt_stride : INT32_MAX;
    constexpr int NSUB = SM::NSUB;
    int cs = 0, is = 0;                 // ring slots of the next tile to consume / to issue (equal at row boundaries)
    uint32_t cph = 0;                   // phase parity of the consume slot's current use
    // expansion constants, one register each: Q8_0 {0x64646464}, Q4_0 {low-nibble mask, high-nibble mask}
    const uint2 kcon = make_uint2(sk_opaque(TK == B200_TYPE_Q8_0 ? 0x64646464u : 0x000f000fu), sk_opaque(0x00f000f0u));
    const uint2 vcon = TK == TV ? kcon : make_uint2(sk_opaque(TV == B200_TYPE_Q8_0 ? 0x64646464u : 0x000f000fu), sk_opaque(0x00f000f0u));

    const uint32_t k64k = a.k65536;

    while (g0 < g1) {
        // P[row], P[row + 1]: the header load holds P[0..27]; a shuffle keeps one global round trip out of the chain
        // plan -> block table -> first TMA box that every row-segment starts with
        int pb, pe;
        if (row < 27) { pb = __shfl_sync(0xffffffffu, head, 4 + row); pe = __shfl_sync(0xffffffffu, head, 5 + row); }
        else          { pb = P[row]; pe = P[row + 1]; }
        if (pe <= g0) { ++row; continue; }                  // row without tiles
        const int t0 = g0 - pb, t1 = min(g1, pe) - pb;       // tiles [t0, t1) of this row
        const int seq = p.q_seq ? p.q_seq[row] : row;
        const int n_vis = sk_row_nvis(p, row);
        const int32_t *bt = p.block_table + (int64_t) seq*p.bt_stride;
        const int seg = FS[row] + (item - pb/chunk);
        const bool single = (FS[row + 1] - FS[row]) == 1;
        // DENSE: this row's mask line and the ALiBi slopes of the lane's two query heads, in the log2 domain
        const __half *mrow = nullptr;
        float sl0 = 1.4426950408889634f, sl1 = 1.4426950408889634f;
        if constexpr (DENSE) {
            mrow = reinterpret_cast<const __half *>(a.mask + (int64_t) row*a.nb_mask1);
            if (p.max_bias > 0.0f) {
                const int hh = kvh*a.gqa + 2*t;
                sl0 *= hh     < a.n_head_log2 ? powf(a.alibi_m0, (float) (hh + 1)) : powf(a.alibi_m1, (float) (2*(hh     - a.n_head_log2) + 1));
                sl1 *= hh + 1 < a.n_head_log2 ? powf(a.alibi_m0, (float) (hh + 2)) : powf(a.alibi_m1, (float) (2*(hh + 1 - a.n_head_log2) + 1));
            }
        }

        float O[8][4];
#pragma unroll
        for (int mu = 0; mu < 8; ++mu) { O[mu][0] = O[mu][1] = O[mu][2] = O[mu][3] = 0.f; }
        float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;   // columns 2t, 2t+1

        // block-table window: lane l holds entry win_base + l; the next window is fetched 32 blocks ahead, so the
        // gather never waits on a global load inside the tile loop
        int b_cur = t0/tpb, o_cur = t0 - b_cur*tpb, win_base = b_cur;
        int win      = win_base + lane      < n_blk ? bt[win_base + lane]      : 0;
        int win_next = win_base + 32 + lane < n_blk ? bt[win_base + 32 + lane] : 0;

        const int nt = t1 - t0;
        int issued = 0;
        const int t_new = (n_vis - 1)/SK_TILE - t0;         // the tile holding the row's newest cell, relative to t0
        auto issue = [&]() {                                // all lanes (shuffle); lane 0 talks to the TMA unit
            if (!waited && issued >= t_new) { pdl_wait(); waited = true; TL_STAMP_ONCE(3); }   // the one tile the store in front of this launch wrote into
            if (b_cur - win_base >= 32) {
                win_base += 32; win = win_next;
                win_next = win_base + 32 + lane < n_blk ? bt[win_base + 32 + lane] : 0;
            }
            const int blk = __shfl_sync(0xffffffffu, win, b_cur - win_base);
            const int cell = (DENSE ? (blk & (SK_MIXED - 1)) : blk)*bs + o_cur*SK_TILE;
            if (++o_cur == tpb) { o_cur = 0; ++b_cur; }
            if constexpr (DENSE) { if (lane == 0) asm volatile("st.shared.b32 [%0], %1;" :: "r"(meta0 + 4*is), "r"(blk) : "memory"); }
            {
                const uint32_t bar = bar0 + 8*is, sb = smem0 + is*SM::STAGE;
                sk_issue_elect<TK == B200_TYPE_F16, TV == B200_TYPE_F16>(bar, SM::KT + SM::VT, sb + SM::KOFF, &tmK, kc,
                                                                        sb + SM::VOFF, &tmV, vc, cell);
            }
            if (++is == STAGES) is = 0;
        };

        if (!waited && !a.after_store) { pdl_wait(); pdl_trigger(); waited = true; TL_STAMP(3); }      // early: the tables are here already
        TL_STAMP_ONCE(2);
#pragma unroll
        for (int s = 0; s < STAGES - NSUB; ++s) { if (issued < nt) { issue(); ++issued; } }

        // ---- Q fragments (this kv head's query heads in the 8 mma columns); after the first boxes are on their way,
        // so the Q round trip runs under theirs instead of in front of it
        sk_qfrag<TK> qf;
        {
            const bool live = g < a.gqa;
            const float *qrow = reinterpret_cast<const float *>(static_cast<const char *>(p.q) + (int64_t) row*p.nbq1
                                                                 + (int64_t) (kvh*a.gqa + (live ? g : 0))*p.nbq2);
            qf.load(qrow, live, a.qscale, lane);
        }

        // One loop iteration = NSUB tiles sharing one softmax step.  FULL iterations (all tiles complete and inside the
        // row's visible range, no softcap) run branch-free between the barrier waits and the rescale vote, so the
        // independent dequantise -> MMA chains of the tiles interleave; the row's last iteration takes the general path.
        auto body = [&](auto full_c, int i) {
            constexpr bool FULL = decltype(full_c)::value;
            const int nsub = FULL ? NSUB : min(NSUB, nt - i);        // tiles of this iteration (warp-uniform)
            // FULL iterations wait for their first tile only; the later ones were requested last and are waited for
            // right before their own Q.K^T, behind the first tile's
            uint32_t kbs[NSUB], vbs[NSUB], tbar[NSUB], tph[NSUB], ment[NSUB];
#pragma unroll
            for (int u = 0; u < NSUB; ++u) {
                kbs[u] = vbs[u] = tbar[u] = tph[u] = ment[u] = 0;
                if (FULL || u < nsub) {
                    if constexpr (DENSE && !FULL) ment[u] = lds32(meta0 + 4*cs);
                    if (!FULL || u == 0) mbar_wait(bar0 + 8*cs, cph);
                    tbar[u] = bar0 + 8*cs; tph[u] = cph;
                    const uint32_t sb = smem0 + cs*SM::STAGE;
                    kbs[u] = sb + SM::KOFF; vbs[u] = sb + SM::VOFF;
                    if (++cs == STAGES) { cs = 0; cph ^= 1; }
                }
            }
            if constexpr (!FULL) {
#pragma unroll
                for (int u = 0; u < NSUB; ++u) {
                    const int tok0 = (t0 + i + u)*SK_TILE;
                    if (u < nsub && tok0 + SK_TILE > n_vis) {     // tail tile: cells past n_vis hold stale bytes; 0 * NaN must not reach O
                        constexpr int VW = sk_type<TV>::WIN;
                        for (int tau = n_vis - tok0; tau < SK_TILE; ++tau) {
                            if constexpr (TV == B200_TYPE_F16) { if (lane < 16) sts128_zero(vbs[u] + ((lane >> 3)*16 + tau)*128 + (lane & 7)*16); }
                            else                               { if (lane < VW/16) sts128_zero(vbs[u] + tau*VW + lane*16); }
                        }
                    }
                }
                __syncwarp();
            }

            float S[NSUB][4];
#pragma unroll
            for (int u = 0; u < NSUB; ++u) {
                S[u][0] = S[u][1] = S[u][2] = S[u][3] = FULL ? 0.f : -INFINITY;
                if (FULL || u < nsub) {
                    float A[4] = {0.f, 0.f, 0.f, 0.f};
                    if (FULL && u > 0) mbar_wait(tbar[u], tph[u]);
                    if constexpr (TK == B200_TYPE_F16) sk_qk_f16(A, kbs[u], qf, lane);
                    else                               sk_qk_quant<TK>(A, kbs[u], kmis, qf, kcon, lane);
                    if constexpr (!FULL) {
                        if (p.logit_softcap != 0.0f) {
#pragma unroll
                            for (int j = 0; j < 4; ++j) A[j] = a.cap_log2e*tanhf(A[j]);
                        }
                    }
                    if constexpr (DENSE && !FULL) {               // mask values of the tile's cells: -inf hides, anything else is a bias
                        const int c0 = (int) (ment[u] & (SK_MIXED - 1))*SK_TILE + g;
                        const float mv0 = __half2float(mrow[c0]), mv1 = __half2float(mrow[c0 + 8]);
                        A[0] = fmaf(sl0, mv0, A[0]); A[1] = fmaf(sl1, mv0, A[1]); A[2] = fmaf(sl0, mv1, A[2]); A[3] = fmaf(sl1, mv1, A[3]);
                    }
                    S[u][0] = A[0]; S[u][1] = A[1]; S[u][2] = A[2]; S[u][3] = A[3];
                    if constexpr (!FULL) {
                        const int tok = (t0 + i + u)*SK_TILE + g;
                        if (tok >= n_vis)     { S[u][0] = -INFINITY; S[u][1] = -INFINITY; }
                        if (tok + 8 >= n_vis) { S[u][2] = -INFINITY; S[u][3] = -INFINITY; }
                    }
                }
            }

            float x0 = fmaxf(S[0][0], S[0][2]), x1 = fmaxf(S[0][1], S[0][3]);
#pragma unroll
            for (int u = 1; u < NSUB; ++u) { x0 = fmaxf(x0, fmaxf(S[u][0], S[u][2])); x1 = fmaxf(x1, fmaxf(S[u][1], S[u][3])); }
#pragma unroll
            for (int o = 4; o < 32; o <<= 1) {
                x0 = fmaxf(x0, __shfl_xor_sync(0xffffffffu, x0, o));
                x1 = fmaxf(x1, __shfl_xor_sync(0xffffffffu, x1, o));
            }
            const float n0 = fmaxf(m0, x0), n1 = fmaxf(m1, x1);      // finite: the first token of the first tile is always visible
            const float al0 = ex2(m0 - n0), al1 = ex2(m1 - n1);
            m0 = n0; m1 = n1;
            l0 *= al0; l1 *= al1;
            if (__any_sync(0xffffffffu, al0 != 1.0f || al1 != 1.0f)) {
#pragma unroll
                for (int mu = 0; mu < 8; ++mu) { O[mu][0] *= al0; O[mu][1] *= al1; O[mu][2] *= al0; O[mu][3] *= al1; }
            }
#pragma unroll
            for (int u = 0; u < NSUB; ++u) {
                if (FULL || u < nsub) {
                    const float p00 = ex2(S[u][0] - n0), p01 = ex2(S[u][1] - n1), p10 = ex2(S[u][2] - n0), p11 = ex2(S[u][3] - n1);
                    l0 += p00 + p10;
                    l1 += p01 + p11;
                    const uint32_t pt0 = movmatrix_trans(pack_h2(p00, p01));
                    const uint32_t pt1 = movmatrix_trans(pack_h2(p10, p11));
                    if constexpr (TV == B200_TYPE_F16) sk_pv_f16(O, vbs[u], pt0, pt1, lane);
                    else                               sk_pv_quant<TV>(O, vbs[u], vmis, pt0, pt1, vcon, k64k, lane);
                }
            }
        };

        const bool no_cap = p.logit_softcap == 0.0f;
        for (int i = 0; i < nt; i += NSUB) {
            __syncwarp();                                    // every lane is done with the slots about to be refilled
#pragma unroll
            for (int u = 0; u < NSUB; ++u) { if (issued < nt) { issue(); ++issued; } }
            bool fast = no_cap && i + NSUB <= nt && (t0 + i + NSUB)*SK_TILE <= n_vis;
            if constexpr (DENSE) {                           // and no tile of the iteration is marked MIXED
                if (fast) {
                    uint32_t any = 0;
#pragma unroll
                    for (int u = 0; u < NSUB; ++u) any |= lds32(meta0 + 4*((cs + u) % STAGES));
                    fast = !(any & SK_MIXED);
                }
            }
            if (fast) body(std::true_type{}, i);
            else      body(std::false_type{}, i);
            TL_STAMP_ONCE(4);
        }
        TL_STAMP_ONCE(5);
        TL_EXTRA(nt);

        // ---- segment epilogue
#pragma unroll
        for (int o = 4; o < 32; o <<= 1) {                   // row sums over the 8 token groups
            l0 += __shfl_xor_sync(0xffffffffu, l0, o);
            l1 += __shfl_xor_sync(0xffffffffu, l1, o);
        }
        const int h0 = kvh*a.gqa + 2*t;                          // query heads of columns 2t, 2t+1
        const bool w0 = 2*t < a.gqa, w1 = 2*t + 1 < a.gqa;
        float *o0 = single ? static_cast<float *>(p.dst) + ((int64_t) row*p.n_head + h0)*SK_D
                           : a.part_o + ((int64_t) seg*p.n_head + h0)*SK_D;
        const float i0 = single ? 1.0f/l0 : 1.0f, i1 = single ? 1.0f/l1 : 1.0f;
#pragma unroll
        for (int mu = 0; mu < 8; ++mu) {
            const int da = sk_row_d<TV>(mu, 0, g), db = sk_row_d<TV>(mu, 1, g);
            if (w0) { o0[da] = O[mu][0]*i0; o0[db] = O[mu][2]*i0; }
            if (w1) { o0[SK_D + da] = O[mu][1]*i1; o0[SK_D + db] = O[mu][3]*i1; }
        }
        if (!single && g == 0) {
            float *ml = a.part_ml + ((int64_t) seg*p.n_head + h0)*2;
            if (w0) { ml[0] = m0; ml[1] = l0; }
            if (w1) { ml[2] = m1; ml[3] = l1; }
        }
        __threadfence();                                     // every lane's stores, device-wide, before the arrival
        __syncwarp();
        if (lane == 0) atomicAdd(a.arrive + (int64_t) row*p.n_head_kv + kvh, 1);
        g0 = pb + t1;
        ++row;
    }
    if (!waited) pdl_wait();             // this grid's completion must imply the store's: the successor only waits for this grid
    TL_END();
}

// merge the segments of split rows; one CTA per (row, kv head), one warp per query head of it, log2-domain maxima.
// The grid is let in while the main kernel runs (its trigger) and does not wait for that grid to drain: warp by warp it
// spins on the row's arrival counter, which the main kernel's warps bump after their partials are stored and fenced
// (the boundary "main grid complete -> griddepcontrol.wait returns" is ~1.4 us on a 10 us layer-op,
// profiles/r2q_timeline_*).  No deadlock: the main grid triggers only from running CTAs, so all of them are resident
// or done before this grid's first CTA starts.  Rows with one segment were written by the main kernel; their CTA
// still sees the arrival, so this grid's completion implies every store of the main grid.
__device__ __forceinline__ int sk_ld_acquire(const int *p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// one warp: the segments [s0, s0 + ns) of (row, head h) -> dst
__device__ __forceinline__ void sk_merge_segments(const sk_args &a, int row, int h, int s0, int ns, int lane) {
    const b200_attn_desc &p = a.d;
    // Latency matters here, not bandwidth (a row has up to ~40 segments and the whole kernel moves a few hundred KB):
    // the segments' (m, l) are read lane-parallel (once for rows of up to 32 segments), the rescale factors travel by
    // shuffle, and the O partials of sixteen segments are in flight at once — the first sixteen requested together with
    // the (m, l), in front of the first use of either: ONE L2 round trip for a row of up to 16 segments.  (One dependent
    // global load per segment per pass cost 9 us of a 20 us layer-op.)
    const float2 *ML = reinterpret_cast<const float2 *>(a.part_ml) + (int64_t) s0*p.n_head + h;      // segment s at ML[s*n_head]
    const float  *PO = a.part_o + ((int64_t) s0*p.n_head + h)*SK_D + 4*lane;                         // segment s at PO + s*n_head*SK_D
    const float2 ml0 = lane < ns ? __ldcg(ML + (int64_t) lane*p.n_head) : make_float2(-INFINITY, 0.f);   // kept for the first batch; .cg: written by other SMs during this grid
    float4 v[16];
#pragma unroll
    for (int u = 0; u < 16; ++u) v[u] = ns > 0 ? __ldcg(reinterpret_cast<const float4 *>(PO + (int64_t) min(u, ns - 1)*p.n_head*SK_D)) : make_float4(0.f, 0.f, 0.f, 0.f);
    float M = ml0.x;
    for (int s = lane + 32; s < ns; s += 32) M = fmaxf(M, __ldcg(ML + (int64_t) s*p.n_head).x);
#pragma unroll
    for (int o_ = 16; o_ > 0; o_ >>= 1) M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, o_));
    float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
    float L = 0.f;
    for (int c = 0; c < ns; c += 32) {
        float f = 0.f;
        if (c + lane < ns) { const float2 ml = c ? __ldcg(ML + (int64_t) (c + lane)*p.n_head) : ml0; f = ex2(ml.x - M); L = fmaf(ml.y, f, L); }
        const int cnt = ns - c < 32 ? ns - c : 32;
        // branch-free batches of 16 partials in flight (one L2 round trip for a row of up to 16 segments): segments
        // past the end are clamped to a valid address and enter with the zero factor of their lane
        for (int j0 = 0; j0 < cnt; j0 += 16) {
            if (c + j0 > 0) {
#pragma unroll
                for (int u = 0; u < 16; ++u) v[u] = __ldcg(reinterpret_cast<const float4 *>(PO + (int64_t) min(c + j0 + u, ns - 1)*p.n_head*SK_D));
            }
#pragma unroll
            for (int u = 0; u < 16; ++u) {
                const float fj = __shfl_sync(0xffffffffu, f, j0 + u);       // lanes >= cnt hold f == 0
                o.x = fmaf(v[u].x, fj, o.x); o.y = fmaf(v[u].y, fj, o.y); o.z = fmaf(v[u].z, fj, o.z); o.w = fmaf(v[u].w, fj, o.w);
            }
        }
    }
#pragma unroll
    for (int o_ = 16; o_ > 0; o_ >>= 1) L += __shfl_xor_sync(0xffffffffu, L, o_);
    const float inv = 1.0f/L;                                          // ns == 0: 0 * inf = NaN like the reference (ggml-cpu.c:12418)
    o.x *= inv; o.y *= inv; o.z *= inv; o.w *= inv;
    *reinterpret_cast<float4 *>(static_cast<float *>(p.dst) + ((int64_t) row*p.n_head + h)*SK_D + 4*lane) = o;
}

__global__ void __launch_bounds__(256)
sk_combine_kernel(const __grid_constant__ sk_args a) {
    const b200_attn_desc &p = a.d;
    TL_BEGIN(4, threadIdx.x == 0);
    pdl_trigger();
    const int lane = threadIdx.x & 31;
    const int row = (int) (blockIdx.x / p.n_head_kv), h = (int) (blockIdx.x - (unsigned) row*p.n_head_kv)*a.gqa + (threadIdx.x >> 5);
    const int *FS = a.plan + 4 + (p.n_q + 1);
    const int s0 = FS[row], ns = FS[row + 1] - s0;
    int *arrived = a.arrive + blockIdx.x;
    if (lane == 0) {
        unsigned long long t0 = 0;
        while (sk_ld_acquire(arrived) < ns) {
            __nanosleep(32);
            unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t0 == 0) t0 = t;
            else if (t - t0 > 4000000000ull) __trap();                // 4 s: a lost arrival must fail loudly, not hang the device
        }
    }
    __syncwarp();
    TL_STAMP(2);
    if (ns != 1) sk_merge_segments(a, row, h, s0, ns, lane);             // one segment: written by the main kernel
    __syncthreads();                                                   // every warp of the CTA has seen the count
    if (threadIdx.x == 0) *arrived = 0;                                  // ready for the next launch on this plan
    TL_END();
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static int sk_sm_count() {
    static int n = 0;
    if (!n) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev); if (n <= 0) n = 148; }
    return n;
}

static bool sk_type_ok(int t) { return t == B200_TYPE_F16 || t == B200_TYPE_Q8_0 || t == B200_TYPE_Q4_0; }

// dense-mask mode (block_table == NULL): the flat cache of an unmodified ggml graph, scanned into tile lists first
static bool sk_dense(const b200_attn_desc &d) { return d.block_table == nullptr; }

bool streamk_supports(const b200_attn_desc &d) {
    if (d.head_dim != SK_D) return false;
    if (!sk_type_ok(d.type_k) || !sk_type_ok(d.type_v)) return false;
    const int gqa = d.n_head/d.n_head_kv;
    if (gqa < 1 || gqa*d.n_head_kv != d.n_head || (gqa > 8 && (gqa % 8 || gqa > 64))) return false;   // > 8: virtual kv heads of 8 (sk_args::vsplit)
    if (sk_dense(d)) {
        if (d.n_kv % SK_TILE || d.n_kv < SK_TILE || d.n_q < 1) return false;                // whole 16-cell tiles
        if (d.mask) {
            if (d.type_k != d.type_v) return false;                                         // the DENSE kernels exist for equal K/V types
            if (((uintptr_t) d.mask | (uint64_t) d.nb_mask1) & 15) return false;            // 16-byte loads of the mask rows
        }
        if ((int64_t) d.n_q*(d.n_kv/SK_TILE) > (1LL << 28)) return false;                    // tile lists: n_q x n_kv/16 ints of workspace
    } else {
        if (d.max_bias != 0.0f) return false;                                               // ALiBi needs the mask's positions
        if (d.block_size % SK_TILE || d.n_kv < SK_TILE) return false;                       // one TMA box = 16 consecutive cells
    }
    if (((uintptr_t) d.k | (uintptr_t) d.v | (uint64_t) d.nbk1 | (uint64_t) d.nbv1) & 15) return false;
    if ((d.nbk2 & (d.type_k == B200_TYPE_F16 ? 127 : 7)) || (d.nbv2 & (d.type_v == B200_TYPE_F16 ? 127 : 7))) return false;
    if (d.nbk1 <= 0 || d.nbv1 <= 0 || d.nbk2 <= 0 || d.nbv2 <= 0) return false;
    // heads inside the token row (the llama.cpp cache, cell-major).  An F16 tensor may also be head-major ([d, n_kv, n_head]
    // contiguous, as tests/test-backend-ops.cpp builds K and V): its 3-D map addresses 128-byte pieces by (cell, piece) with
    // strides (nb1, 128), which is indifferent to the order of the two.  Quantised head-major rows are 136 / 72 bytes apart
    // (not a legal TMA stride) and stay with the CUDA-core kernel.
    if ((int64_t) d.n_head_kv*d.nbk2 > d.nbk1 && d.type_k != B200_TYPE_F16) return false;
    if ((int64_t) d.n_head_kv*d.nbv2 > d.nbv1 && d.type_v != B200_TYPE_F16) return false;
    const int64_t span_k = d.n_kv*d.nbk1 + (int64_t) d.n_head_kv*d.nbk2, span_v = d.n_kv*d.nbv1 + (int64_t) d.n_head_kv*d.nbv2;
    if ((span_k > span_v ? span_k : span_v) >= (1LL << 40) || d.n_kv >= (1LL << 31)) return false;
    if (((uintptr_t) d.q | (uint64_t) d.nbq1 | (uint64_t) d.nbq2) & 15) return false;      // float4 loads of Q
    if ((uintptr_t) d.dst & 15) return false;
    return true;
}

template <typename F> static auto sk_dispatch(int tk, int tv, F &&f) {
#define SK_CASE(K, V) if (tk == K && tv == V) return f(std::integral_constant<int, K>{}, std::integral_constant<int, V>{});
    SK_CASE(B200_TYPE_F16,  B200_TYPE_F16)  SK_CASE(B200_TYPE_F16,  B200_TYPE_Q8_0) SK_CASE(B200_TYPE_F16,  B200_TYPE_Q4_0)
    SK_CASE(B200_TYPE_Q8_0, B200_TYPE_F16)  SK_CASE(B200_TYPE_Q8_0, B200_TYPE_Q8_0) SK_CASE(B200_TYPE_Q8_0, B200_TYPE_Q4_0)
    SK_CASE(B200_TYPE_Q4_0, B200_TYPE_F16)  SK_CASE(B200_TYPE_Q4_0, B200_TYPE_Q8_0)
#undef SK_CASE
    return f(std::integral_constant<int, B200_TYPE_Q4_0>{}, std::integral_constant<int, B200_TYPE_Q4_0>{});
}

static int sk_vsplit(const b200_attn_desc &d) { const int gqa = d.n_head/d.n_head_kv; return gqa > 8 ? gqa/8 : 1; }
static int sk_nkv(const b200_attn_desc &d)    { return d.n_head_kv*sk_vsplit(d); }                   // virtual kv heads

static int sk_items(const b200_attn_desc &d) {
    const int wps = sk_dispatch(d.type_k, d.type_v, [](auto k, auto v) {         // resident warps per SM
        using SM = sk_smem<decltype(k)::value, decltype(v)::value>; return SM::CTAS_PER_SM*SM::WARPS; });
    const int64_t warps = (int64_t) sk_sm_count()*wps;
    int64_t items = warps/sk_nkv(d);
    if (items < 1) items = 1;
    // always the whole machine: small problems are kept in few warps by the plan's minimum chunk (SK_MIN_CHUNK tiles),
    // the surplus CTAs leave after the header load.  (A grid cut to the work, 129 CTAs for 8 x 1024, measured 4 % slower
    // than 148 with the same 128 active: profiles/r1v_plan_ab.txt.)
    return (int) items;
}

static size_t sk_align(size_t x) { return (x + 255) & ~(size_t) 255; }

// dense mode: [n_q] visible cells per row, [n_q] q_seq, [n_q][n_kv/16] tile lists, in front of the usual layout
static size_t sk_dense_bytes(const b200_attn_desc &d) {
    return sk_dense(d) ? sk_align((2*(size_t) d.n_q + (size_t) d.n_q*(size_t) (d.n_kv/SK_TILE))*sizeof(int)) : 0;
}

size_t streamk_workspace(const b200_attn_desc &d) {
    const size_t max_segs = (size_t) sk_items(d) + d.n_q;
    return sk_dense_bytes(d) + sk_align((4 + 2*((size_t) d.n_q + 1))*sizeof(int)) + sk_align((size_t) d.n_q*sk_nkv(d)*sizeof(int))
         + sk_align(max_segs*d.n_head*2*sizeof(float)) + max_segs*d.n_head*SK_D*sizeof(float);
}

typedef CUresult (*sk_encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                 const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static sk_encode_fn sk_encoder() {
    static sk_encode_fn fn = nullptr;
    if (!fn) {
        void *ptr = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<sk_encode_fn>(ptr);
        else (void) cudaGetLastError();
    }
    return fn;
}

// One box = the 16 cells of a tile x one head row (see the file header).  The flat cache is cell-major:
// cell c at base + c*nb1, head h at + h*nb2.
static bool sk_make_map(CUtensorMap *tm, int type, const void *base, int64_t nb1, int64_t nb2, int n_head_kv, int64_t n_cells) {
    sk_encode_fn enc = sk_encoder();
    if (!enc) { set_error("stream-K: cuTensorMapEncodeTiled not available from the driver"); return false; }
    const cuuint32_t ones[3] = {1, 1, 1};
    CUresult r;
    if (type == B200_TYPE_F16) {
        const cuuint64_t dims[3]    = {64, (cuuint64_t) n_cells, (cuuint64_t) (((int64_t) (n_head_kv - 1)*nb2 + 256)/128)};
        const cuuint64_t strides[2] = {(cuuint64_t) nb1, 128};
        const cuuint32_t box[3]     = {64, SK_TILE, 2};
        r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void *>(base), dims, strides, box, ones,
                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else {
        const int hb = type == B200_TYPE_Q8_0 ? 136 : 72, win = type == B200_TYPE_Q8_0 ? 144 : 80;
        const cuuint64_t dims[2]    = {(cuuint64_t) ((int64_t) (n_head_kv - 1)*nb2 + hb), (cuuint64_t) n_cells};
        const cuuint64_t strides[1] = {(cuuint64_t) nb1};
        const cuuint32_t box[2]     = {(cuuint32_t) win, SK_TILE};
        r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(base), dims, strides, box, ones,
                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    }
    if (r != CUDA_SUCCESS) { set_error("stream-K: cuTensorMapEncodeTiled failed with %d", (int) r); return false; }
    return true;
}

template <int TK, int TV, bool DENSE = false>
static int sk_launch(const sk_args &a, const CUtensorMap &tmK, const CUtensorMap &tmV, cudaStream_t st) {
    using SM = sk_smem<TK, TV>;
    static bool configured = false;
    if (!configured) {
        const cudaError_t e = cudaFuncSetAttribute(sk_main_kernel<TK, TV, DENSE>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM::CTA);
        if (e != cudaSuccess) { set_error("stream-K: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return B200_ERR_CUDA; }
        configured = true;
    }
    const int64_t warps = (int64_t) a.n_items*a.d.n_head_kv;
    const unsigned grid = (unsigned) ((warps + SM::WARPS - 1)/SM::WARPS);
    launch_pdl(sk_main_kernel<TK, TV, DENSE>, dim3(grid), dim3(32*SM::WARPS), SM::CTA, st, a, tmK, tmV);
    return B200_OK;
}

int launch_kv_write(const b200_kv_write_desc &d, cudaStream_t stream);          // kv_write.cu

// store != NULL: this layer's kv_write, launched between the plan and the main kernel (the schedule does not depend on
// it).  store_checked: with B200_ATTN_REUSE_PLAN, the plan in the workspace was computed with these very slots.
// chained: b200_decode_step set B200_ATTN_REUSE_PLAN itself — the previous launch on this stream was the previous layer of
// the same call, with this plan and these tables (sk_args::early).
int launch_attn_streamk(const b200_attn_desc &d, cudaStream_t st, const b200_kv_write_desc *store, bool store_checked, bool chained) {
    const size_t need = streamk_workspace(d);
    if (d.workspace == nullptr || d.workspace_bytes < need) {
        set_error("attention: workspace of %zu bytes needed, %zu given", need, d.workspace_bytes);
        return B200_ERR_ALLOC;
    }
    TL_ARM();
    sk_args a;
    a.d = d;
    a.vsplit = sk_vsplit(d);
    a.d.n_head_kv = sk_nkv(d);                               // virtual kv heads from here on (the tensor maps keep the real count)
    a.gqa = d.n_head/a.d.n_head_kv;
    a.k65536 = 65536u;
    a.n_items = sk_items(d);
    a.chunk_align = 1;
    a.chunk_min = SK_MIN_CHUNK;
    const size_t max_segs = (size_t) sk_items(d) + d.n_q;    // the layout streamk_workspace() promised
    char *ws = static_cast<char *>(d.workspace);
    const bool dense = sk_dense(d);
    a.mask = nullptr; a.nb_mask1 = 0; a.alibi_m0 = a.alibi_m1 = 1.0f; a.n_head_log2 = 1;
    a.dense_len = a.dense_bt = a.dense_qseq = nullptr; a.n_tiles = 0;
    const bool reuse = !dense && (d.flags & B200_ATTN_REUSE_PLAN);
    // store overlap: ONE new token per sequence, row r == sequence r == stored row r (a decode ubatch; with q_seq / q_pos
    // given the plan kernel checks that they say so), launches chained by PDL
    a.after_store = store && store->n_tokens > 0 && store->n_tokens == d.n_q && d.n_q <= d.n_seqs && !dense
                 && pdl_enabled() && (!reuse || store_checked) ? 1 : 0;
    a.store_slots = a.after_store ? store->slot_mapping : nullptr;
    a.store_cells = a.after_store ? store->n_cells : 0;
    a.early = reuse && chained && pdl_enabled() ? 1 : 0;
    if (dense) {                                             // the tile lists stand in for block table / seq_lens / q_seq
        a.n_tiles    = (int) (d.n_kv/SK_TILE);
        a.dense_len  = reinterpret_cast<int *>(ws);
        a.dense_qseq = a.dense_len + d.n_q;
        a.dense_bt   = a.dense_qseq + d.n_q;
        ws += sk_dense_bytes(d);
        a.mask = static_cast<const char *>(d.mask); a.nb_mask1 = d.nb_mask1;
        a.n_head_log2 = 1 << (int) floor(log2((double) d.n_head));                       // ggml-cpu.c:12294-12298
        a.alibi_m0 = powf(2.0f, -(d.max_bias       )/a.n_head_log2);
        a.alibi_m1 = powf(2.0f, -(d.max_bias / 2.0f)/a.n_head_log2);
        a.d.block_table = a.dense_bt; a.d.bt_stride = a.n_tiles; a.d.block_size = SK_TILE;
        a.d.seq_lens = a.dense_len; a.d.n_seqs = d.n_q;
        a.d.q_seq = a.dense_qseq; a.d.q_pos = nullptr;
    }
    a.plan = reinterpret_cast<int *>(ws);
    ws += sk_align((4 + 2*((size_t) d.n_q + 1))*sizeof(int));
    a.arrive = reinterpret_cast<int *>(ws);
    ws += sk_align((size_t) d.n_q*sk_nkv(d)*sizeof(int));
    a.part_ml = reinterpret_cast<float *>(ws);
    ws += sk_align(max_segs*d.n_head*2*sizeof(float));
    a.part_o = reinterpret_cast<float *>(ws);
    const float log2e = 1.4426950408889634f;
    if (d.logit_softcap != 0.0f) { a.qscale = d.scale/d.logit_softcap; a.cap_log2e = d.logit_softcap*log2e; }
    else                         { a.qscale = d.scale*log2e;           a.cap_log2e = 0.0f; }

    alignas(64) CUtensorMap tmK, tmV;
    if (!sk_make_map(&tmK, d.type_k, d.k, d.nbk1, d.nbk2, d.n_head_kv, d.n_kv)) return B200_ERR_CUDA;
    if (!sk_make_map(&tmV, d.type_v, d.v, d.nbv1, d.nbv2, d.n_head_kv, d.n_kv)) return B200_ERR_CUDA;

    if (dense) {
        launch_pdl(sk_dense_scan_kernel, dim3(d.mask ? (unsigned) d.n_q : 1u), dim3(1024), 0, st, a);
        count_launch();
        if (!check_launch("sk_dense_scan")) return B200_ERR_CUDA;
    }
    if (!reuse) {                                            // layers 1.. of a step run on layer 0's schedule
        launch_pdl(sk_plan_kernel, dim3(1), dim3(1024), 0, st, a);
        count_launch();
        if (!check_launch("sk_plan")) return B200_ERR_CUDA;
    }
    if (store) { const int rc = launch_kv_write(*store, st); if (rc != B200_OK) return rc; }
    int rc;
    if (dense && d.mask) {                              // equal K/V types (streamk_supports)
        rc = d.type_k == B200_TYPE_F16  ? sk_launch<B200_TYPE_F16,  B200_TYPE_F16,  true>(a, tmK, tmV, st)
           : d.type_k == B200_TYPE_Q8_0 ? sk_launch<B200_TYPE_Q8_0, B200_TYPE_Q8_0, true>(a, tmK, tmV, st)
                                        : sk_launch<B200_TYPE_Q4_0, B200_TYPE_Q4_0, true>(a, tmK, tmV, st);
    } else {
        rc = sk_dispatch(d.type_k, d.type_v, [&](auto k, auto v) { return sk_launch<decltype(k)::value, decltype(v)::value>(a, tmK, tmV, st); });
    }
    if (rc != B200_OK) return rc;
    count_launch();
    if (!check_launch("sk_main")) return B200_ERR_CUDA;
    launch_pdl(sk_combine_kernel, dim3((unsigned) ((int64_t) d.n_q*a.d.n_head_kv)), dim3(32*a.gqa), 0, st, a);
    count_launch();
    if (!check_launch("sk_combine")) return B200_ERR_CUDA;
    return B200_OK;
}

} // namespace b200
