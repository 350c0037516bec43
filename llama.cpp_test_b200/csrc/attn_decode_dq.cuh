// attn_decode_dq.cuh — stream-K paged decode for a Q8_0 KV cache on tcgen05 (included by attn_streamk.cu: same plan,
// same combine, same workspace; only the main kernel differs).
//
// Why a second formulation.  The mma.sync kernel builds every A-fragment register with ALU instructions and is
// issue-bound on quantised caches (0.74-0.77 of the HBM peak for Q8_0).  Here the cache bytes are
// expanded ONCE into the UMMA shared-memory layouts by all 128 threads of a CTA (16-byte loads and stores, ~7
// instructions per 4 elements) and the two contractions run as tcgen05.mma with tokens on the M axis:
//     S^T[128 tokens][16] = K16[128 x 128] . Q16^T        (A, B from shared memory, K-major, 128-byte swizzle)
//     O^T[128 d][16]     += V16^T[128 d x 128 tokens] . P   (A = the row-major V tile read MN-major, B = P^T K-major)
// so the softmax has one TOKEN per thread (TMEM lane) and 8 head columns, and the output one d per thread.
// Work item = (stream-K item, kv head) per CTA; a super-tile is 8 consecutive 16-token tiles of one row (8 TMA boxes
// through the block table for K, 8 for V).  2 CTAs per SM (108 KB of shared memory, 32 TMEM columns each).
//
// Element order: a thread expands its token row word by word, PRMT pairs bytes (0, 2) and (1, 3) of a word, so
// position p of the expanded row holds d = (p & ~3) | ((p & 1) << 1) | ((p >> 1) & 1).  The contraction over d does not
// care as long as Q16 uses the same order; for V the order permutes the rows of O^T and is undone at the final store.
//
// Arithmetic: q*d is formed exactly as dequantize_row_q8_0 does (ggml/src/ggml-quants.c:349-362) and rounded to half
// once (what the reference's CUDA path does before its F16 MMA kernel, ggml/src/ggml-cuda/fattn-common.cuh:723-749);
// scores, softmax and the accumulators are FP32.

constexpr int DQ_THREADS = 128;
constexpr int DQ_TOK     = 128;                       // tokens per super-tile
constexpr int DQ_WIN     = 144;                       // staged bytes per token row (136-byte head row + alignment slack)
constexpr int DQ_SLOT    = DQ_TOK*DQ_WIN;             // 18432
constexpr int DQ_OFF_KF  = 0;                         // [2 d-halves][128 tokens][128 B], 128-byte swizzle
constexpr int DQ_OFF_VF  = 32768;
constexpr int DQ_OFF_Q   = 65536;                     // [2 d-halves][16 rows][128 B]
constexpr int DQ_OFF_P   = 69632;                     // [2 token-halves][16 rows][128 B]
constexpr int DQ_OFF_KS  = 73728;                     // staged K rows
constexpr int DQ_OFF_VS  = DQ_OFF_KS + DQ_SLOT;
constexpr int DQ_OFF_BAR = DQ_OFF_VS + DQ_SLOT;       // kfull, vfull, sbar, pvbar, tmem slot, reduction scratch
constexpr int DQ_SMEM    = DQ_OFF_BAR + 256;

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
__device__ __forceinline__ void dq_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar) : "memory");
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

// Expand this thread's staged token row (136 bytes starting MIS bytes into its 144-byte window) into the 16 chunks of
// row `t` of a [2][128][128 B] swizzled tile.  zero: the token is not visible (stale bytes must not reach the MMA).
template <int MIS>
__device__ __forceinline__ void dq_expand_row(uint32_t src_row, uint32_t dst_tile, int t, bool zero) {
    uint32_t w[36];
#pragma unroll
    for (int i = 0; i < 9; ++i) { uint32_t q[4]; dq_lds128(src_row + 16*i, q); w[4*i] = q[0]; w[4*i + 1] = q[1]; w[4*i + 2] = q[2]; w[4*i + 3] = q[3]; }
    const uint32_t k64 = 0x64646464u, bias = 0x64806480u;                    // half2(1152, 1152)
    const uint32_t rowbase = dst_tile + (uint32_t) ((t >> 3)*1024 + (t & 7)*128), x7 = (uint32_t) (t & 7) << 4;
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        const int o = MIS + 34*b;                                            // byte offset of the block: scale, then 32 quants
        const uint32_t sw = w[o >> 2];
        const uint32_t d2 = (o & 2) ? __byte_perm(sw, sw, 0x3232) : __byte_perm(sw, sw, 0x1010);   // (d, d) as half2
#pragma unroll
        for (int ii = 0; ii < 4; ++ii) {
            uint32_t r[4];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int qo = o + 2 + 8*ii + 4*u;                           // byte offset of this quant word (multiple of 2)
                const uint32_t qw = (qo & 2) ? __byte_perm(w[qo >> 2], w[(qo >> 2) + 1], 0x5432) : w[qo >> 2];
                const uint32_t ub = qw ^ 0x80808080u;                        // q + 128 per byte
                const uint32_t e02 = __byte_perm(ub, k64, 0x4240), e13 = __byte_perm(ub, k64, 0x4341);   // half2(1152 + q)
                r[2*u]     = hmul2_u(dq_hsub2(e02, bias), d2);               // (q0, q2) * d, exact integer first
                r[2*u + 1] = hmul2_u(dq_hsub2(e13, bias), d2);               // (q1, q3) * d
            }
            if (zero) { r[0] = r[1] = r[2] = r[3] = 0u; }
            const int c = ((b & 1) << 2) + ii;                               // 16-byte chunk inside the 128-byte row
            dq_sts128(rowbase + (uint32_t) ((b >> 1)*16384) + (((uint32_t) c << 4) ^ x7), r[0], r[1], r[2], r[3]);
        }
    }
}

__global__ void __launch_bounds__(DQ_THREADS, 2)
dq_main_kernel(const __grid_constant__ sk_args a, const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV) {
    extern __shared__ __align__(1024) char dq_dyn[];
    const b200_attn_desc &p = a.d;
    const uint32_t sbase = smem_u32(dq_dyn);
    const uint32_t sKf = sbase + DQ_OFF_KF, sVf = sbase + DQ_OFF_VF, sQ = sbase + DQ_OFF_Q, sP = sbase + DQ_OFF_P;
    const uint32_t sKs = sbase + DQ_OFF_KS, sVs = sbase + DQ_OFF_VS, sBar = sbase + DQ_OFF_BAR;
    const uint32_t bKfull = sBar, bVfull = sBar + 8, bS = sBar + 16, bPV = sBar + 24;
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(dq_dyn + DQ_OFF_BAR + 32);
    float *red = reinterpret_cast<float *>(dq_dyn + DQ_OFF_BAR + 64);           // [4 warps][8]
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    pdl_prologue();

    if (t == 0) {
        mbar_init(bKfull, 1); mbar_init(bVfull, 1); mbar_init(bS, 1); mbar_init(bPV, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 32;" :: "r"(sBar + 32) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // Q16 rows >= gqa and P rows 8..15 stay zero for the whole kernel
    for (int i = t; i < (4096 + 4096)/16; i += DQ_THREADS) dq_sts128(sQ + 16*i, 0u, 0u, 0u, 0u);
    dq_tc_fence_before();
    __syncthreads();
    dq_tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t tS = tmem + ((uint32_t) (32*warp) << 16), tO = tS + 16;

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
    uint32_t G = 0;                                           // super-tiles this CTA has processed: barrier phases

    while (g0 < g1) {
        const int pb = P[row], pe = P[row + 1];
        if (pe <= g0) { ++row; continue; }
        const int t0 = g0 - pb, t1 = min(g1, pe) - pb;        // 16-token tiles [t0, t1) of this row
        const int seq = p.q_seq ? p.q_seq[row] : row;
        const int n_vis = sk_row_nvis(p, row);
        const int32_t *bt = p.block_table + (int64_t) seq*p.bt_stride;
        const int seg = FS[row] + (item - pb/chunk);
        const bool single = (FS[row + 1] - FS[row]) == 1;
        const int nst = (t1 - t0 + 7) >> 3;                   // super-tiles of this segment

        // the 8 boxes of super-tile j (warp 0, lanes 0..7: one box each), K or V
        auto issue = [&](int j, bool is_v) {
            const int tile0 = t0 + 8*j, nb = min(8, t1 - tile0);
            const uint32_t bar = is_v ? bVfull : bKfull, dst = is_v ? sVs : sKs;
            if (lane == 0) mbar_expect_tx(bar, (uint32_t) (nb*SK_TILE*DQ_WIN));
            __syncwarp();
            if (lane < nb) {
                const int ti = tile0 + lane, b = ti/tpb;
                const int blk = b < n_blk ? bt[b] : 0;
                const int cell = blk*bs + (ti - b*tpb)*SK_TILE;
                tma_load_2d(dst + (uint32_t) (lane*SK_TILE*DQ_WIN), is_v ? &tmV : &tmK, is_v ? vc : kc, cell, bar);
            }
        };
        if (warp == 0) { issue(0, false); issue(0, true); }

        // ---- Q16: thread t fills positions [16*(t&7), +16) of row t>>3 (rows >= gqa stay zero)
        {
            const int h = t >> 3, part = t & 7;
            if (h < a.gqa) {
                const float *qrow = reinterpret_cast<const float *>(static_cast<const char *>(p.q) + (int64_t) row*p.nbq1
                                                                     + (int64_t) (kvh*a.gqa + h)*p.nbq2);
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    const int p0 = 16*part + 8*cc;            // 8 positions = one 16-byte chunk
                    uint32_t r[4];
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        r[k] = pack_h2(qrow[dq_pos_to_d(p0 + 2*k)]*a.qscale, qrow[dq_pos_to_d(p0 + 2*k + 1)]*a.qscale);
                    const int c = (p0 & 63) >> 3;
                    dq_sts128(sQ + (uint32_t) ((p0 >> 6)*2048 + (h >> 3)*1024 + (h & 7)*128 + ((c ^ (h & 7)) << 4)), r[0], r[1], r[2], r[3]);
                }
            }
        }

        float ref[8], lsum[8];
#pragma unroll
        for (int h = 0; h < 8; ++h) { ref[h] = -INFINITY; lsum[h] = 0.f; }

        for (int j = 0; j < nst; ++j, ++G) {
            const int tok0 = (t0 + 8*j)*SK_TILE;              // first token of the super-tile within the row
            const bool dead = tok0 + t >= n_vis || t >= min(8, t1 - (t0 + 8*j))*SK_TILE;
            // ---- K: expand, then S^T = K16 . Q16^T
            mbar_wait(bKfull, G & 1);
            if (kmis) dq_expand_row<8>(sKs + (uint32_t) (t*DQ_WIN), sKf, t, false);
            else      dq_expand_row<0>(sKs + (uint32_t) (t*DQ_WIN), sKf, t, false);
            dq_fence_async_smem();
            __syncthreads();                                  // K16 (and Q16) complete, staged K rows free
            if (warp == 0) {
                if (j + 1 < nst) issue(j + 1, false);
                if (lane == 0) {
                    dq_tc_fence_after();
                    const uint64_t a0 = dq_desc(sKf, 16, 1024), b0 = dq_desc(sQ, 16, 1024);
#pragma unroll
                    for (int ks = 0; ks < 8; ++ks)
                        dq_mma(tmem, a0 + (uint64_t) ((ks >> 2)*1024 + (ks & 3)*2), b0 + (uint64_t) ((ks >> 2)*128 + (ks & 3)*2), ID_QK, ks > 0);
                    dq_commit(bS);
                }
                __syncwarp();
            }
            // ---- V: expand (the previous P.V must have retired: it reads V16 and P)
            if (G > 0) mbar_wait(bPV, (G - 1) & 1);
            mbar_wait(bVfull, G & 1);
            if (vmis) dq_expand_row<8>(sVs + (uint32_t) (t*DQ_WIN), sVf, t, dead);
            else      dq_expand_row<0>(sVs + (uint32_t) (t*DQ_WIN), sVf, t, dead);
            // ---- softmax: this thread's token, 8 head columns
            mbar_wait(bS, G & 1);
            dq_tc_fence_after();
            float s[8];
            dq_tmem_ld8(tS, s);
            float mx[8];
#pragma unroll
            for (int h = 0; h < 8; ++h) { if (dead) s[h] = -INFINITY; mx[h] = s[h]; }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
                for (int h = 0; h < 8; ++h) mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], o));
            }
            if (lane == 0) {                                  // `red` is free: its last readers passed a barrier since
#pragma unroll
                for (int h = 0; h < 8; ++h) red[8*warp + h] = mx[h];
            }
            __syncthreads();
            bool move = false;
            float alpha[8];
#pragma unroll
            for (int h = 0; h < 8; ++h) {
                const float m = fmaxf(fmaxf(red[h], red[8 + h]), fmaxf(red[16 + h], red[24 + h]));
                alpha[h] = 1.0f;
                // lazy reference: it only moves when the maximum outgrows it by 2^8 (P stays below 256: exact enough in F16)
                if (m > ref[h] + 8.0f) { alpha[h] = ex2(ref[h] - m); ref[h] = m; move = true; }   // ref = -inf: alpha = 0
            }
            if (move) {                                       // CTA-uniform
#pragma unroll
                for (int h = 0; h < 8; ++h) lsum[h] *= alpha[h];
                if (j > 0) {                                  // O holds earlier tiles of this segment: rescale row t (PV retired above)
                    float o8[8];
                    dq_tmem_ld8(tO, o8);
#pragma unroll
                    for (int h = 0; h < 8; ++h) o8[h] *= alpha[h];
                    dq_tmem_st8(tO, o8);
                }
            }
            {
                const uint32_t pbase = sP + (uint32_t) ((t >> 6)*2048 + (t & 7)*2), c = (uint32_t) ((t & 63) >> 3);
#pragma unroll
                for (int h = 0; h < 8; ++h) {
                    const float pv = dead ? 0.0f : ex2(s[h] - ref[h]);
                    lsum[h] += pv;
                    dq_sts16(pbase + (uint32_t) (h*128) + ((c ^ (uint32_t) h) << 4), (uint32_t) __half_as_ushort(__float2half_rn(pv)));
                }
            }
            dq_fence_async_smem();
            dq_tc_fence_before();
            __syncthreads();                                  // V16 and P complete, staged V rows free, O rescaled
            if (warp == 0) {
                if (j + 1 < nst) issue(j + 1, true);
                if (lane == 0) {
                    dq_tc_fence_after();
                    const uint64_t a0 = dq_desc(sVf, 16384, 1024), b0 = dq_desc(sP, 16, 1024);
#pragma unroll
                    for (int ks = 0; ks < 8; ++ks)
                        dq_mma(tmem + 16, a0 + (uint64_t) (ks*128), b0 + (uint64_t) ((ks >> 2)*128 + (ks & 3)*2), ID_PV, (j > 0 || ks > 0) ? 1u : 0u);
                    dq_commit(bPV);
                }
                __syncwarp();
            }
        }

        // ---- segment epilogue: O^T row t = position t of every head; row sums over the 128 token lanes
        mbar_wait(bPV, (G - 1) & 1);
        dq_tc_fence_after();
        float o8[8];
        dq_tmem_ld8(tO, o8);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
            for (int h = 0; h < 8; ++h) lsum[h] += __shfl_xor_sync(0xffffffffu, lsum[h], o);
        }
        __syncthreads();
        if (lane == 0) {
#pragma unroll
            for (int h = 0; h < 8; ++h) red[8*warp + h] = lsum[h];
        }
        __syncthreads();
        const int dd = dq_pos_to_d(t);
#pragma unroll
        for (int h = 0; h < 8; ++h) {
            if (h < a.gqa) {
                const float l = (red[h] + red[8 + h]) + (red[16 + h] + red[24 + h]);
                const int head = kvh*a.gqa + h;
                if (single) static_cast<float *>(p.dst)[((int64_t) row*p.n_head + head)*SK_D + dd] = o8[h]*(1.0f/l);
                else {
                    a.part_o[((int64_t) seg*p.n_head + head)*SK_D + dd] = o8[h];
                    if (t == 0) { float *ml = a.part_ml + ((int64_t) seg*p.n_head + head)*2; ml[0] = ref[h]; ml[1] = l; }
                }
            }
        }
        dq_tc_fence_before();
        __syncthreads();                                      // red, Q16 and the TMEM columns are free for the next segment
        g0 = pb + t1;
        ++row;
    }

    dq_tc_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;" :: "r"(tmem) : "memory");
}
