// attn_decode_dq.cuh — stream-K paged decode for a Q8_0 KV cache on tcgen05 (included by attn_streamk.cu: same plan,
// same combine, same workspace; only the main kernel differs).
//
// Why a second formulation.  The mma.sync kernel builds every A-fragment register with ALU instructions and is
// issue-bound on quantised caches (0.74-0.77 of the HBM peak for Q8_0).  Here the cache bytes are
// expanded ONCE into the UMMA operand layouts by the 256 threads of a CTA (two threads per token row; 16-byte loads
// and stores, ~7 instructions per 4 elements) and the two contractions run as tcgen05.mma with tokens on the M axis:
//     S^T[128 tokens][16] = K16[128 x 128] . Q16^T        (A = the expanded rows in TMEM, lane = token; B from shared memory)
//     O^T[128 d][16]     += V16^T[128 d x 128 tokens] . P   (A = the row-major V tile read MN-major, B = P^T K-major)
// so the softmax has one TOKEN per thread pair (TMEM lane; 4 head columns each), and the output one d per thread pair.
// Work item = (stream-K item, kv head) per CTA; a super-tile is 8 consecutive 16-token tiles of one row (8 TMA boxes
// through the block table for K, 8 for V), staged two deep.  2 CTAs per SM (112 KB of shared memory, 128 TMEM columns each).
//
// Element order: a thread expands its token row word by word, PRMT pairs bytes (0, 2) and (1, 3) of a word, so
// position p of the expanded row holds d = (p & ~3) | ((p & 1) << 1) | ((p >> 1) & 1).  The contraction over d does not
// care as long as Q16 uses the same order; for V the order permutes the rows of O^T and is undone at the final store.
//
// Arithmetic: q*d is formed exactly as dequantize_row_q8_0 does (ggml/src/ggml-quants.c:349-362) and rounded to half
// once (what the reference's CUDA path does before its F16 MMA kernel, ggml/src/ggml-cuda/fattn-common.cuh:723-749);
// scores, softmax and the accumulators are FP32.

constexpr int DQ_WORKERS = 256;                      // 8 worker warps: warp w owns tokens 32*(w&3).. and half (w>>2) of the row / of the heads
constexpr int DQ_THREADS = DQ_WORKERS + 32;          // + the control warp: TMA gathers and tcgen05.mma issue
constexpr int DQ_TOK     = 128;                       // tokens per super-tile
constexpr int DQ_WIN     = 144;                       // staged bytes per token row (136-byte head row + alignment slack)
constexpr int DQ_SLOT    = DQ_TOK*DQ_WIN;             // 18432
constexpr int DQ_NSLOT   = 2;                         // staged super-tiles per operand: one being expanded, one in flight
constexpr int DQ_OFF_VF  = 0;                         // V16 [2 d-halves][128 tokens][128 B], 128-byte swizzle
constexpr int DQ_OFF_Q   = 32768;                     // Q16 [2 d-halves][16 rows][128 B]
constexpr int DQ_OFF_P   = 36864;                     // P^T [2 token-halves][16 rows][128 B]
constexpr int DQ_OFF_KS  = 40960;                     // staged K rows, DQ_NSLOT slots
constexpr int DQ_OFF_VS  = DQ_OFF_KS + DQ_NSLOT*DQ_SLOT;
constexpr int DQ_OFF_BAR = DQ_OFF_VS + DQ_NSLOT*DQ_SLOT;   // kfull[2], vfull[2], S, PV, K-ready, P-ready, tmem slot, reduction scratch
constexpr int DQ_SMEM    = DQ_OFF_BAR + 256;          // 114944: two CTAs per SM
// TMEM columns (128 allocated): S^T 0..15, O^T 16..31, K16 32..95 (the A operand of S^T = K16 . Q16^T: lane = token, two
// halves per column; it never touches shared memory)
constexpr int DQ_TM_S = 0, DQ_TM_O = 16, DQ_TM_K = 32, DQ_TM_COLS = 128;

__device__ __forceinline__ void dq_fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void dq_tc_fence_before()  { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void dq_tc_fence_after()   { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// shared-memory matrix descriptor, 128-byte swizzle (start >> 4 at [0,14), LBO >> 4 at [16,30), SBO >> 4 at [32,46),
// version 1 at [46,48), SWIZZLE_128B = 2 at [61,64))
__device__ __forceinline__ uint64_t dq_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t) ((addr & 0x3FFFFu) >> 4) | ((uint64_t) (lbo_bytes >> 4) << 16) | ((uint64_t) (sbo_bytes >> 4) << 32)
         | (1ull << 46) | (2ull << 61);
}
// instruction descriptor, kind::f16: D F32 (bit 4), A/B F16, a_major at 15, b_major at 16, N >> 3 at [17,23), M >> 4 at [24,29)
__host__ __device__ constexpr uint32_t dq_idesc(int m, int n, int a_mn_major, int b_mn_major) {
    return (1u << 4) | ((uint32_t) a_mn_major << 15) | ((uint32_t) b_mn_major << 16) | ((uint32_t) (n >> 3) << 17) | ((uint32_t) (m >> 4) << 24);
}
__device__ __forceinline__ void dq_mma(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
                 :: "r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void dq_mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n}"
                 :: "r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
#define DQ_W32(v, o) "r"(v[o+0]), "r"(v[o+1]), "r"(v[o+2]), "r"(v[o+3]), "r"(v[o+4]), "r"(v[o+5]), "r"(v[o+6]), "r"(v[o+7]), \
                     "r"(v[o+8]), "r"(v[o+9]), "r"(v[o+10]), "r"(v[o+11]), "r"(v[o+12]), "r"(v[o+13]), "r"(v[o+14]), "r"(v[o+15]), \
                     "r"(v[o+16]), "r"(v[o+17]), "r"(v[o+18]), "r"(v[o+19]), "r"(v[o+20]), "r"(v[o+21]), "r"(v[o+22]), "r"(v[o+23]), \
                     "r"(v[o+24]), "r"(v[o+25]), "r"(v[o+26]), "r"(v[o+27]), "r"(v[o+28]), "r"(v[o+29]), "r"(v[o+30]), "r"(v[o+31])
#define DQ_LIST32S "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32}"
// this thread's expanded half row (32 registers, two halves each) into 32 consecutive columns of its TMEM lane
__device__ __forceinline__ void dq_tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], " DQ_LIST32S ";" :: "r"(taddr), DQ_W32(v, 0) : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void dq_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar) : "memory");
}
// The 8 K16 steps of S^T = K16 . Q16^T and the commit in ONE elected block of a converged warp: with elect.sync ptxas
// emits the uniform-datapath instructions directly (an `if (lane == 0)` costs a vote loop per MMA), and the step
// descriptors are the step-0 ones plus constants: K16 8 TMEM columns per step; Q16 32 B per step, +2 KB for the second d-half.
__device__ __forceinline__ void dq_qk_block(uint32_t d_tmem, uint32_t a_tmem, uint64_t b0, uint32_t idesc, uint32_t bar) {
    asm volatile("{\n .reg .pred e, z;\n .reg .b64 b;\n .reg .b32 t;\n elect.sync _|e, 0xffffffff;\n setp.ne.b32 z, 0, 0;\n"
                 " @e tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, z;\n setp.eq.b32 z, 0, 0;\n"
                 " add.u32 t, %1, 8;\n  add.u64 b, %2, 2;\n   @e tcgen05.mma.cta_group::1.kind::f16 [%0], [t], b, %3, z;\n"
                 " add.u32 t, %1, 16;\n add.u64 b, %2, 4;\n   @e tcgen05.mma.cta_group::1.kind::f16 [%0], [t], b, %3, z;\n"
                 " add.u32 t, %1, 24;\n add.u64 b, %2, 6;\n   @e tcgen05.mma.cta_group::1.kind::f16 [%0], [t], b, %3, z;\n"
                 " add.u32 t, %1, 32;\n add.u64 b, %2, 128;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], [t], b, %3, z;\n"
                 " add.u32 t, %1, 40;\n add.u64 b, %2, 130;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], [t], b, %3, z;\n"
                 " add.u32 t, %1, 48;\n add.u64 b, %2, 132;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], [t], b, %3, z;\n"
                 " add.u32 t, %1, 56;\n add.u64 b, %2, 134;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], [t], b, %3, z;\n"
                 " @e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%4];\n}"
                 :: "r"(d_tmem), "r"(a_tmem), "l"(b0), "r"(idesc), "r"(bar) : "memory");
}
// The 8 K16 steps (16 tokens each) of O^T += V16^T . P and the commit: V16 +2 KB per step; P^T 32 B per step, +2 KB for the
// second token half.  acc_first = 0 overwrites O^T (first super-tile of a segment).
__device__ __forceinline__ void dq_pv_block(uint32_t d_tmem, uint64_t a0, uint64_t b0, uint32_t idesc, uint32_t acc_first, uint32_t bar) {
    asm volatile("{\n .reg .pred e, z;\n .reg .b64 a, b;\n elect.sync _|e, 0xffffffff;\n setp.ne.b32 z, %4, 0;\n"
                 " @e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, z;\n setp.eq.b32 z, 0, 0;\n"
                 " add.u64 a, %1, 128;\n add.u64 b, %2, 2;\n   @e tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %3, z;\n"
                 " add.u64 a, %1, 256;\n add.u64 b, %2, 4;\n   @e tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %3, z;\n"
                 " add.u64 a, %1, 384;\n add.u64 b, %2, 6;\n   @e tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %3, z;\n"
                 " add.u64 a, %1, 512;\n add.u64 b, %2, 128;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %3, z;\n"
                 " add.u64 a, %1, 640;\n add.u64 b, %2, 130;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %3, z;\n"
                 " add.u64 a, %1, 768;\n add.u64 b, %2, 132;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %3, z;\n"
                 " add.u64 a, %1, 896;\n add.u64 b, %2, 134;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %3, z;\n"
                 " @e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%5];\n}"
                 :: "r"(d_tmem), "l"(a0), "l"(b0), "r"(idesc), "r"(acc_first), "r"(bar) : "memory");
}
__device__ __forceinline__ void dq_tmem_ld8(uint32_t taddr, float (&v)[8]) {
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void dq_tmem_st8(uint32_t taddr, const float (&v)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 :: "r"(taddr), "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
                    "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])) : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void dq_tmem_ld4(uint32_t taddr, float (&v)[4]) {
    uint32_t r[4];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void dq_tmem_st4(uint32_t taddr, const float (&v)[4]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%2,%3,%4};"
                 :: "r"(taddr), "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])) : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void dq_lds128(uint32_t addr, uint32_t (&w)[4]) {
    asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "r"(addr));
}
__device__ __forceinline__ void dq_sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" :: "r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void dq_sts16(uint32_t addr, uint32_t v) {
    asm volatile("st.shared.u16 [%0], %1;" :: "r"(addr), "h"((unsigned short) v) : "memory");
}
__device__ __forceinline__ uint32_t dq_hsub2(uint32_t a, uint32_t b) {
    uint32_t d; asm("sub.rn.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d;
}
// position p of an expanded row <-> element d of the head row (see the file header)
__device__ __forceinline__ int dq_pos_to_d(int p) { return (p & ~3) | ((p & 1) << 1) | ((p >> 1) & 1); }

// Expand half HF (blocks 2HF, 2HF+1 = positions 64HF .. 64HF+63) of this thread's staged token row (136 bytes starting
// MIS bytes into its 144-byte window) into 32 registers: r[i] = positions (64HF + 2i, +1) as half2.
// zero: the token is not visible (stale bytes must not reach the MMA).
template <int MIS, int HF>
__device__ __forceinline__ void dq_expand_half(uint32_t src_row, uint32_t (&r)[32], bool zero) {
    constexpr int BASE = HF ? 64 : 0;                                        // the 80 bytes [BASE, BASE + 80) hold both blocks
    uint32_t w[20];
#pragma unroll
    for (int i = 0; i < 5; ++i) { uint32_t q[4]; dq_lds128(src_row + BASE + 16*i, q); w[4*i] = q[0]; w[4*i + 1] = q[1]; w[4*i + 2] = q[2]; w[4*i + 3] = q[3]; }
    const uint32_t k64 = 0x64646464u, bias = 0x64806480u;                    // half2(1152, 1152)
#pragma unroll
    for (int bb = 0; bb < 2; ++bb) {
        const int o = MIS + 34*(2*HF + bb) - BASE;                           // byte offset of the block: scale, then 32 quants
        const uint32_t sw = w[o >> 2];
        const uint32_t d2 = zero ? 0u : ((o & 2) ? __byte_perm(sw, sw, 0x3232) : __byte_perm(sw, sw, 0x1010));   // (d, d) as half2
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int qo = o + 2 + 4*i;                                      // byte offset of this quant word (multiple of 2)
            const uint32_t qw = (qo & 2) ? __byte_perm(w[qo >> 2], w[(qo >> 2) + 1], 0x5432) : w[qo >> 2];
            const uint32_t ub = qw ^ 0x80808080u;                            // q + 128 per byte
            const uint32_t e02 = __byte_perm(ub, k64, 0x4240), e13 = __byte_perm(ub, k64, 0x4341);       // half2(1152 + q)
            r[16*bb + 2*i]     = hmul2_u(dq_hsub2(e02, bias), d2);           // (q0, q2) * d: the integer is exact first
            r[16*bb + 2*i + 1] = hmul2_u(dq_hsub2(e13, bias), d2);           // (q1, q3) * d
        }
    }
}
template <int HF>
__device__ __forceinline__ void dq_expand(uint32_t src_row, int mis, uint32_t (&r)[32], bool zero) {
    if (mis) dq_expand_half<8, HF>(src_row, r, zero);
    else     dq_expand_half<0, HF>(src_row, r, zero);
}
// d-half HF of row tk of a [2][128][128 B] swizzled tile <- the 8 chunks of an expanded half row
__device__ __forceinline__ void dq_store_half(uint32_t dst_tile, int tk, int hf, const uint32_t (&r)[32]) {
    const uint32_t rowbase = dst_tile + (uint32_t) (hf*16384 + (tk >> 3)*1024 + (tk & 7)*128), x7 = (uint32_t) (tk & 7) << 4;
#pragma unroll
    for (int c = 0; c < 8; ++c) dq_sts128(rowbase + (((uint32_t) c << 4) ^ x7), r[4*c], r[4*c + 1], r[4*c + 2], r[4*c + 3]);
}

__device__ __forceinline__ void dq_mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar) : "memory"); }
__device__ __forceinline__ void dq_workers_sync() { asm volatile("bar.sync 1, %0;" :: "n"(DQ_WORKERS) : "memory"); }

__global__ void __launch_bounds__(DQ_THREADS, 2)
dq_main_kernel(const __grid_constant__ sk_args a, const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV) {
    extern __shared__ __align__(1024) char dq_dyn[];
    const b200_attn_desc &p = a.d;
    const uint32_t sbase = smem_u32(dq_dyn);
    const uint32_t sVf = sbase + DQ_OFF_VF, sQ = sbase + DQ_OFF_Q, sP = sbase + DQ_OFF_P;
    const uint32_t sKs = sbase + DQ_OFF_KS, sVs = sbase + DQ_OFF_VS, sBar = sbase + DQ_OFF_BAR;
    // kfull / vfull: TMA -> workers, one per slot.  S / PV: tcgen05.commit -> workers.  KR / PR: the 256 workers -> the control
    // warp ("K16 (and Q16) are in place", "V16 and P are in place"): an arrive, so no worker ever waits for the control warp.
    const uint32_t bKfull = sBar, bVfull = sBar + 16, bS = sBar + 32, bPV = sBar + 40, bKR = sBar + 48, bPR = sBar + 56;
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(dq_dyn + DQ_OFF_BAR + 64);
    float *red = reinterpret_cast<float *>(dq_dyn + DQ_OFF_BAR + 96);           // [2 halves][4 quadrants][4 heads]
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const bool control = warp == DQ_WORKERS/32;
    const int hf = (warp >> 2) & 1, tk = 32*(warp & 3) + lane;                  // worker: token tk of the super-tile, half hf
    pdl_prologue();

    if (t == 0) {
        mbar_init(bKfull, 1); mbar_init(bKfull + 8, 1); mbar_init(bVfull, 1); mbar_init(bVfull + 8, 1); mbar_init(bS, 1); mbar_init(bPV, 1);
        mbar_init(bKR, DQ_WORKERS); mbar_init(bPR, DQ_WORKERS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (control) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 128;" :: "r"(sBar + 64) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // Q16 rows >= gqa and P rows 8..15 stay zero for the whole kernel
    for (int i = t; i < (4096 + 4096)/16; i += DQ_THREADS) dq_sts128(sQ + 16*i, 0u, 0u, 0u, 0u);
    dq_tc_fence_before();
    __syncthreads();
    dq_tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t tlane = tmem + ((uint32_t) (32*(warp & 3)) << 16);
    const uint32_t tS = tlane + DQ_TM_S + 4*hf, tO = tlane + DQ_TM_O + 4*hf, tK = tlane + DQ_TM_K + 32*hf;

    const int item = blockIdx.x / p.n_head_kv, kvh = blockIdx.x - item*p.n_head_kv;
    const int *P = a.plan + 4, *FS = P + (p.n_q + 1);
    const int total = a.plan[0], chunk = a.plan[1];
    int g0 = item*chunk;
    const int g1 = min(g0 + chunk, total);
    int row = 0;
    if (g0 < g1) {
        int lo = 0, hi = p.n_q;                               // last r with P[r] <= g0
        while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (P[mid] <= g0) lo = mid; else hi = mid; }
        row = lo;
    }
    const int64_t koff = (int64_t) kvh*p.nbk2, voff = (int64_t) kvh*p.nbv2;
    const int kmis = (int) (koff & 15), vmis = (int) (voff & 15);
    const int kc = (int) (koff - kmis), vc = (int) (voff - vmis);
    const int bs = p.block_size, tpb = bs/SK_TILE;
    const int n_blk = p.bt_stride < INT32_MAX ? (int) p.bt_stride : INT32_MAX;
    constexpr uint32_t ID_QK = dq_idesc(128, 16, 0, 0), ID_PV = dq_idesc(128, 16, 1, 0);
    uint32_t G = 0;                                           // super-tiles this CTA has processed: slots and barrier phases

    while (g0 < g1) {
        const int pb = P[row], pe = P[row + 1];
        if (pe <= g0) { ++row; continue; }
        const int t0 = g0 - pb, t1 = min(g1, pe) - pb;        // 16-token tiles [t0, t1) of this row
        const int nst = (t1 - t0 + 7) >> 3;                   // super-tiles of this segment

        if (control) {
            // ================= control warp: gathers through the block table, issues the MMAs =================
            const int seq = p.q_seq ? p.q_seq[row] : row;
            const int32_t *bt = p.block_table + (int64_t) seq*p.bt_stride;
            // block-table entries come from a register window (lane l: entry win_base + l, the next 32 one window ahead)
            int win_base = t0/tpb;
            int win      = win_base + lane      < n_blk ? bt[win_base + lane]      : 0;
            int win_next = win_base + 32 + lane < n_blk ? bt[win_base + 32 + lane] : 0;
            // the 8 boxes of super-tile j (lanes 0..7: one box each), K or V; g: its CTA-wide index (slot = g & 1)
            auto issue = [&](int j, uint32_t g, bool is_v) {
                const int tile0 = t0 + 8*j, nb = min(8, t1 - tile0);
                const uint32_t bar = (is_v ? bVfull : bKfull) + 8*(g & 1), dst = (is_v ? sVs : sKs) + (g & 1)*DQ_SLOT;
                if (tile0/tpb - win_base >= 32) {             // K and V of a super-tile, and super-tiles, come in order
                    win_base += 32; win = win_next;
                    win_next = win_base + 32 + lane < n_blk ? bt[win_base + 32 + lane] : 0;
                }
                if (lane == 0) mbar_expect_tx(bar, (uint32_t) (nb*SK_TILE*DQ_WIN));
                const int ti = tile0 + (lane & 7), b = ti/tpb, idx = b - win_base;
                const int e0 = __shfl_sync(0xffffffffu, win, idx & 31), e1 = __shfl_sync(0xffffffffu, win_next, idx & 31);
                const int blk = idx < 32 ? e0 : e1;
                if (lane < nb) {
                    const int cell = blk*bs + (ti - b*tpb)*SK_TILE;
                    tma_load_2d(dst + (uint32_t) (lane*SK_TILE*DQ_WIN), is_v ? &tmV : &tmK, is_v ? vc : kc, cell, bar);
                }
            };
            issue(0, G, false); issue(0, G, true);
            if (nst > 1) { issue(1, G + 1, false); issue(1, G + 1, true); }
            const uint64_t bQ = dq_desc(sQ, 16, 1024), aV = dq_desc(sVf, 16384, 1024), bP = dq_desc(sP, 16, 1024);
            for (int j = 0; j < nst; ++j, ++G) {
                mbar_wait(bKR, G & 1);                        // K16 in TMEM, Q16 in shared memory, staged K rows free
                dq_tc_fence_after();
                dq_qk_block(tmem + DQ_TM_S, tmem + DQ_TM_K, bQ, ID_QK, bS);
                if (j + 2 < nst) issue(j + 2, G + 2, false);
                mbar_wait(bPR, G & 1);                        // V16 and P in shared memory, O rescaled, staged V rows free
                dq_tc_fence_after();
                dq_pv_block(tmem + DQ_TM_O, aV, bP, ID_PV, j > 0 ? 1u : 0u, bPV);
                if (j + 2 < nst) issue(j + 2, G + 2, true);
            }
        } else {
            // ================= worker warps =================
            const int n_vis = sk_row_nvis(p, row);
            const int seg = FS[row] + (item - pb/chunk);
            const bool single = (FS[row + 1] - FS[row]) == 1;
            // ---- Q16: thread t fills positions [8*(t&15), +8) of row t>>4 (rows >= gqa stay zero).  Its previous readers
            //      (the QK MMAs of the last segment) retired before the epilogue's PV wait.
            {
                const int h = t >> 4, p0 = 8*(t & 15);
                if (h < a.gqa) {
                    const float *qrow = reinterpret_cast<const float *>(static_cast<const char *>(p.q) + (int64_t) row*p.nbq1
                                                                         + (int64_t) (kvh*a.gqa + h)*p.nbq2);
                    uint32_t r[4];
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        r[k] = pack_h2(qrow[dq_pos_to_d(p0 + 2*k)]*a.qscale, qrow[dq_pos_to_d(p0 + 2*k + 1)]*a.qscale);
                    const int c = (p0 & 63) >> 3;
                    dq_sts128(sQ + (uint32_t) ((p0 >> 6)*2048 + (h >> 3)*1024 + (h & 7)*128 + ((c ^ (h & 7)) << 4)), r[0], r[1], r[2], r[3]);
                }
            }
            float ref[4], lsum[4];                            // heads 4hf .. 4hf+3
#pragma unroll
            for (int h = 0; h < 4; ++h) { ref[h] = -INFINITY; lsum[h] = 0.f; }

            for (int j = 0; j < nst; ++j, ++G) {
                const int tok0 = (t0 + 8*j)*SK_TILE;          // first token of the super-tile within the row
                const bool dead = tok0 + tk >= n_vis || tk >= min(8, t1 - (t0 + 8*j))*SK_TILE;
                const uint32_t slot = G & 1, par = (G >> 1) & 1;
                // ---- K: expand into TMEM (QK of the previous super-tile retired: bS was waited on)
                mbar_wait(bKfull + 8*slot, par);
                {
                    uint32_t r[32];
                    const uint32_t src = sKs + slot*DQ_SLOT + (uint32_t) (tk*DQ_WIN);
                    if (hf) dq_expand<1>(src, kmis, r, false); else dq_expand<0>(src, kmis, r, false);
                    dq_tmem_st32(tK, r);
                }
                dq_fence_async_smem();                        // Q16 (first super-tile of a segment)
                dq_tc_fence_before();
                dq_mbar_arrive(bKR);
                // ---- V: expand into shared memory (the previous P.V must have retired: it reads V16 and P)
                if (G > 0) mbar_wait(bPV, (G - 1) & 1);
                mbar_wait(bVfull + 8*slot, par);
                {
                    uint32_t r[32];
                    const uint32_t src = sVs + slot*DQ_SLOT + (uint32_t) (tk*DQ_WIN);
                    if (hf) dq_expand<1>(src, vmis, r, dead); else dq_expand<0>(src, vmis, r, dead);
                    dq_store_half(sVf, tk, hf, r);
                }
                // ---- softmax: this thread's token, its 4 head columns
                mbar_wait(bS, G & 1);
                dq_tc_fence_after();
                float s[4];
                dq_tmem_ld4(tS, s);
                float mx[4];
#pragma unroll
                for (int h = 0; h < 4; ++h) { if (dead) s[h] = -INFINITY; mx[h] = s[h]; }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
                    for (int h = 0; h < 4; ++h) mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], o));
                }
                // `red` is free: its readers of the previous super-tile arrived on PR before PV was issued, and bS of this
                // one completed after every worker arrived on KR
                if (lane == 0) *reinterpret_cast<float4 *>(red + 4*warp) = make_float4(mx[0], mx[1], mx[2], mx[3]);
                dq_workers_sync();
                bool move = false;                            // uniform over the 4 warps of this half
                float alpha[4];
                {
                    const float4 m0 = *reinterpret_cast<const float4 *>(red + 16*hf),     m1 = *reinterpret_cast<const float4 *>(red + 16*hf + 4);
                    const float4 m2 = *reinterpret_cast<const float4 *>(red + 16*hf + 8), m3 = *reinterpret_cast<const float4 *>(red + 16*hf + 12);
                    const float m[4] = {fmaxf(fmaxf(m0.x, m1.x), fmaxf(m2.x, m3.x)), fmaxf(fmaxf(m0.y, m1.y), fmaxf(m2.y, m3.y)),
                                        fmaxf(fmaxf(m0.z, m1.z), fmaxf(m2.z, m3.z)), fmaxf(fmaxf(m0.w, m1.w), fmaxf(m2.w, m3.w))};
#pragma unroll
                    for (int h = 0; h < 4; ++h) {
                        alpha[h] = 1.0f;
                        // lazy reference: it only moves when the maximum outgrows it by 2^8 (P stays below 256: exact enough in F16)
                        if (m[h] > ref[h] + 8.0f) { alpha[h] = ex2(ref[h] - m[h]); ref[h] = m[h]; move = true; }   // ref = -inf: alpha = 0
                    }
                }
                if (move) {
#pragma unroll
                    for (int h = 0; h < 4; ++h) lsum[h] *= alpha[h];
                    if (j > 0) {                              // O holds earlier tiles of this segment: rescale this lane's 4 columns (PV retired above)
                        float o4[4];
                        dq_tmem_ld4(tO, o4);
#pragma unroll
                        for (int h = 0; h < 4; ++h) o4[h] *= alpha[h];
                        dq_tmem_st4(tO, o4);
                    }
                }
                {
                    const uint32_t pbase = sP + (uint32_t) ((tk >> 6)*2048 + (tk & 7)*2), c = (uint32_t) ((tk & 63) >> 3);
#pragma unroll
                    for (int h = 0; h < 4; ++h) {
                        const float pv = dead ? 0.0f : ex2(s[h] - ref[h]);
                        lsum[h] += pv;
                        const uint32_t hh = (uint32_t) (4*hf + h);
                        dq_sts16(pbase + hh*128 + ((c ^ hh) << 4), (uint32_t) __half_as_ushort(__float2half_rn(pv)));
                    }
                }
                dq_fence_async_smem();
                dq_tc_fence_before();
                dq_mbar_arrive(bPR);
            }

            // ---- segment epilogue: O^T row tk = position tk of this half's heads; row sums over the 128 token lanes
            mbar_wait(bPV, (G - 1) & 1);
            dq_tc_fence_after();
            float o4[4];
            dq_tmem_ld4(tO, o4);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
                for (int h = 0; h < 4; ++h) lsum[h] += __shfl_xor_sync(0xffffffffu, lsum[h], o);
            }
            dq_workers_sync();                                // the last readers of `red` (this segment's last maxima) are done
            if (lane == 0) *reinterpret_cast<float4 *>(red + 4*warp) = make_float4(lsum[0], lsum[1], lsum[2], lsum[3]);
            dq_workers_sync();
            const int dd = dq_pos_to_d(tk);
#pragma unroll
            for (int h = 0; h < 4; ++h) {
                if (4*hf + h < a.gqa) {
                    const float l = (red[16*hf + h] + red[16*hf + 4 + h]) + (red[16*hf + 8 + h] + red[16*hf + 12 + h]);
                    const int head = kvh*a.gqa + 4*hf + h;
                    if (single) static_cast<float *>(p.dst)[((int64_t) row*p.n_head + head)*SK_D + dd] = o4[h]*(1.0f/l);
                    else {
                        a.part_o[((int64_t) seg*p.n_head + head)*SK_D + dd] = o4[h];
                        if (tk == 0) { float *ml = a.part_ml + ((int64_t) seg*p.n_head + head)*2; ml[0] = ref[h]; ml[1] = l; }
                    }
                }
            }
            dq_tc_fence_before();
            dq_workers_sync();                                // `red` and the TMEM columns are free for the next segment
        }
        g0 = pb + t1;
        ++row;
    }

    dq_tc_fence_before();
    __syncthreads();
    if (control) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 128;" :: "r"(tmem) : "memory");
}
