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
//   * sk_main_kernel   every WARP is an independent worker for (item, kv head): it streams its K/V tiles with a
//                      private 16-byte cp.async ring (no CTA barrier in the loop), gathers through the block table,
//                      places each tile in shared memory as [16-byte chunk][token][16 B] so that ldmatrix (F16)
//                      and the fragment loads of the quantised types are bank-conflict free, dequantises in
//                      registers straight into mma A fragments, and keeps the online softmax in registers
//                      (warp-shuffle max/sum).  The 8 warps of a CTA take the 8 kv heads of the same tokens, so a
//                      CTA consumes whole contiguous token rows of the flat llama.cpp cache.
//   * sk_combine_kernel merges the segments of rows that were split (rows with one segment are written
//                      directly by the main kernel).
#include "common.cuh"
#include <math.h>

namespace b200 {

constexpr int SK_WARPS  = 8;
constexpr int SK_D      = 128;
constexpr int SK_TILE   = 16;          // tokens per tile
constexpr int SK_STAGES = 3;

struct sk_args {
    b200_attn_desc d;
    int    gqa;
    int    n_items;                     // upper bound on items per kv head
    int   *plan;                        // [4] {total_tiles, chunk, -, -} then P[n_q+1], FS[n_q+1]
    float *part_ml;                     // [max_segs][n_head][2]
    float *part_o;                      // [max_segs][n_head][D]
    float  qscale;                      // scale*log2(e), or scale/softcap when softcap != 0
    float  cap_log2e;                   // softcap*log2(e)
};

// ------------------------------------------------------------------------------------------------
// type traits for D = 128: bytes of one head row and the most 16-byte chunks a (mis-aligned) row can span
// ------------------------------------------------------------------------------------------------
template <int T> struct sk_type;
template <> struct sk_type<B200_TYPE_F16>  { static constexpr int HEAD_BYTES = 256; static constexpr int NCH = 16; };
template <> struct sk_type<B200_TYPE_Q8_0> { static constexpr int HEAD_BYTES = 136; static constexpr int NCH = 9;  };
template <> struct sk_type<B200_TYPE_Q4_0> { static constexpr int HEAD_BYTES = 72;  static constexpr int NCH = 5;  };
template <int T> constexpr int sk_tile_bytes() { return sk_type<T>::NCH*SK_TILE*16; }

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src, int src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" :: "r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" :: "n"(N) : "memory"); }

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
__device__ __forceinline__ float ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t lds32(uint32_t addr) { uint32_t v; asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr)); return v; }

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
    const b200_attn_desc &p = a.d;
    int *P = a.plan + 4, *FS = P + (p.n_q + 1);
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
    if (threadIdx.x == 0) { P[p.n_q] = total; a.plan[0] = total; a.plan[1] = chunk; a.plan[2] = (total + chunk - 1)/chunk; a.plan[3] = 0; }
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
// tile copy: 16 tokens x NCH 16-byte chunks of one head row -> smem [chunk][token][16 B]
// lane handles tokens (lane&7) and (lane&7)+8; quarter-warps write 128 contiguous bytes (no bank conflict)
// and read 64 contiguous bytes per token from HBM/L2.
// ------------------------------------------------------------------------------------------------
template <int NCH>
__device__ __forceinline__ void sk_copy_tile(uint32_t dst, const char *row0, const char *row1, bool ok0, bool ok1,
                                             int nch, int last_bytes, int lane) {
    constexpr int ITERS = 2*((NCH + 3)/4);
    const int tq = lane & 7, cq = lane >> 3;
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
        const int half = it & 1, chunk = cq + 4*(it >> 1);
        if (chunk < nch) {
            const char *row = half ? row1 : row0;
            const bool ok = half ? ok1 : ok0;
            const int bytes = ok ? (chunk == nch - 1 ? last_bytes : 16) : 0;
            cp_async16(dst + chunk*(SK_TILE*16) + (tq + 8*half)*16, row + chunk*16, bytes);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// per-type fragment producers
// ------------------------------------------------------------------------------------------------
// Q as the B operand of S^T = K.Q^T.  For F16 K the contraction index is d in natural order:
// k-step s, lane (n = lane>>2, t = lane&3): b0 = Q[n][16s+2t, +1], b1 = Q[n][16s+2t+8, +9].
struct sk_qfrag_f16 {
    uint32_t b[8][2];
    __device__ __forceinline__ void load(const float *qrow, bool live, float qscale, int lane) {
        const int t = lane & 3;
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
    }
};

// S^T tile (16 tokens x 8 heads) for an F16 K tile at smem address kb
__device__ __forceinline__ void sk_qk_f16(float (&S)[4], uint32_t kb, const sk_qfrag_f16 &q, int lane) {
    const uint32_t base = kb + (lane >> 4)*(SK_TILE*16) + ((lane & 7) + 8*((lane >> 3) & 1))*16;
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        uint32_t a[4];
        ldmatrix_x4(a, base + 2*s*(SK_TILE*16));
        mma_16816(S, a, q.b[s][0], q.b[s][1]);
    }
}

// O^T += V^T . P^T for an F16 V tile; m-tile mu covers d = 16mu .. 16mu+15 in natural order
__device__ __forceinline__ void sk_pv_f16(float (&O)[8][4], uint32_t vb, uint32_t p0, uint32_t p1, int lane) {
    const uint32_t base = vb + ((lane >> 3) & 1)*(SK_TILE*16) + ((lane & 7) + 8*(lane >> 4))*16;
#pragma unroll
    for (int mu = 0; mu < 8; ++mu) {
        uint32_t a[4];
        ldmatrix_x4_trans(a, base + 2*mu*(SK_TILE*16));
        mma_16816(O[mu], a, p0, p1);
    }
}

// ------------------------------------------------------------------------------------------------
// main kernel
// ------------------------------------------------------------------------------------------------
template <int TK, int TV>
struct sk_smem {
    static constexpr int KB = sk_tile_bytes<TK>();
    static constexpr int VB = sk_tile_bytes<TV>();
    static constexpr int STAGE = KB + VB;
    static constexpr int WARP = STAGE*SK_STAGES;
    static constexpr int CTA = WARP*SK_WARPS;
};

template <int TK, int TV>
__global__ void __launch_bounds__(32*SK_WARPS)
sk_main_kernel(const __grid_constant__ sk_args a) {
    extern __shared__ __align__(128) char sk_dyn[];
    using SM = sk_smem<TK, TV>;
    const b200_attn_desc &p = a.d;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;

    const int wid  = blockIdx.x*SK_WARPS + warp;
    const int item = wid / p.n_head_kv, kvh = wid - item*p.n_head_kv;
    const int total = a.plan[0], chunk = a.plan[1];
    const int *P = a.plan + 4, *FS = P + (p.n_q + 1);
    int g0 = item*chunk;
    if (g0 >= total) return;
    const int g1 = min(g0 + chunk, total);

    // row containing global tile g0: last r with P[r] <= g0 (rows without tiles share P with their successor)
    int lo = 0, hi = p.n_q;                                  // invariant: P[lo] <= g0 < P[hi]
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (P[mid] <= g0) lo = mid; else hi = mid; }
    int row = lo;

    const uint32_t smem0 = smem_u32(sk_dyn) + warp*SM::WARP;

    // head-row placement inside the 16-byte chunk grid (quantised rows are only 8-byte aligned)
    const int64_t koff = (int64_t) kvh*p.nbk2, voff = (int64_t) kvh*p.nbv2;
    const int kmis = (int) (koff & 15), vmis = (int) (voff & 15);
    const int knch = (kmis + sk_type<TK>::HEAD_BYTES + 15) >> 4, vnch = (vmis + sk_type<TV>::HEAD_BYTES + 15) >> 4;
    const int klast = kmis + sk_type<TK>::HEAD_BYTES - 16*(knch - 1), vlast = vmis + sk_type<TV>::HEAD_BYTES - 16*(vnch - 1);
    const char *kbase = static_cast<const char *>(p.k) + koff - kmis;
    const char *vbase = static_cast<const char *>(p.v) + voff - vmis;
    const int bs = p.block_size;
    const bool tile_in_block = (bs % SK_TILE) == 0;

    while (g0 < g1) {
        const int pb = P[row], pe = P[row + 1];
        if (pe <= g0) { ++row; continue; }                  // row without tiles
        const int t0 = g0 - pb, t1 = min(g1, pe) - pb;       // tiles [t0, t1) of this row
        const int seq = p.q_seq ? p.q_seq[row] : row;
        const int n_vis = sk_row_nvis(p, row);
        const int32_t *bt = p.block_table + (int64_t) seq*p.bt_stride;
        const int seg = FS[row] + (item - pb/chunk);
        const bool single = (FS[row + 1] - FS[row]) == 1;

        // ---- Q fragments (this kv head's query heads in the 8 mma columns)
        sk_qfrag_f16 qf;
        {
            const bool live = g < a.gqa;
            const float *qrow = reinterpret_cast<const float *>(static_cast<const char *>(p.q) + (int64_t) row*p.nbq1
                                                                 + (int64_t) (kvh*a.gqa + (live ? g : 0))*p.nbq2);
            qf.load(qrow, live, a.qscale, lane);
        }

        float O[8][4];
#pragma unroll
        for (int mu = 0; mu < 8; ++mu) { O[mu][0] = O[mu][1] = O[mu][2] = O[mu][3] = 0.f; }
        float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;   // columns 2t, 2t+1

        auto issue = [&](int tile, int slot) {
            const int pos0 = tile*SK_TILE + (lane & 7), pos1 = pos0 + 8;
            const bool ok0 = pos0 < n_vis, ok1 = pos1 < n_vis;
            int64_t c0 = 0, c1 = 0;
            if (tile_in_block) {
                const int tp = tile*SK_TILE;
                const int64_t cb = (int64_t) bt[tp/bs]*bs + tp % bs;     // same address for the whole warp
                c0 = cb + (lane & 7); c1 = c0 + 8;
            } else {
                if (ok0) c0 = (int64_t) bt[pos0/bs]*bs + pos0 % bs;
                if (ok1) c1 = (int64_t) bt[pos1/bs]*bs + pos1 % bs;
            }
            const uint32_t sb = smem0 + slot*SM::STAGE;
            sk_copy_tile<sk_type<TK>::NCH>(sb,          kbase + c0*p.nbk1, kbase + c1*p.nbk1, ok0, ok1, knch, klast, lane);
            sk_copy_tile<sk_type<TV>::NCH>(sb + SM::KB, vbase + c0*p.nbv1, vbase + c1*p.nbv1, ok0, ok1, vnch, vlast, lane);
        };

        const int nt = t1 - t0;
#pragma unroll
        for (int s = 0; s < SK_STAGES - 1; ++s) { if (s < nt) issue(t0 + s, s); cp_async_commit(); }

        for (int i = 0; i < nt; ++i) {
            cp_async_wait<SK_STAGES - 2>();
            __syncwarp();
            if (i + SK_STAGES - 1 < nt) issue(t0 + i + SK_STAGES - 1, (i + SK_STAGES - 1) % SK_STAGES);
            cp_async_commit();

            const uint32_t sb = smem0 + (i % SK_STAGES)*SM::STAGE;
            float S[4] = {0.f, 0.f, 0.f, 0.f};
            sk_qk_f16(S, sb, qf, lane);

            if (p.logit_softcap != 0.0f) {
#pragma unroll
                for (int j = 0; j < 4; ++j) S[j] = a.cap_log2e*tanhf(S[j]);
            }
            const int tok = (t0 + i)*SK_TILE + g;
            if (tok >= n_vis)     { S[0] = -INFINITY; S[1] = -INFINITY; }
            if (tok + 8 >= n_vis) { S[2] = -INFINITY; S[3] = -INFINITY; }

            float x0 = fmaxf(S[0], S[2]), x1 = fmaxf(S[1], S[3]);
#pragma unroll
            for (int o = 4; o < 32; o <<= 1) {
                x0 = fmaxf(x0, __shfl_xor_sync(0xffffffffu, x0, o));
                x1 = fmaxf(x1, __shfl_xor_sync(0xffffffffu, x1, o));
            }
            const float n0 = fmaxf(m0, x0), n1 = fmaxf(m1, x1);      // finite: token t0*16 of a tile is always visible
            const float al0 = ex2(m0 - n0), al1 = ex2(m1 - n1);
            m0 = n0; m1 = n1;
            const float p00 = ex2(S[0] - n0), p01 = ex2(S[1] - n1), p10 = ex2(S[2] - n0), p11 = ex2(S[3] - n1);
            l0 = l0*al0 + p00 + p10;
            l1 = l1*al1 + p01 + p11;
            if (__any_sync(0xffffffffu, al0 != 1.0f || al1 != 1.0f)) {
#pragma unroll
                for (int mu = 0; mu < 8; ++mu) { O[mu][0] *= al0; O[mu][1] *= al1; O[mu][2] *= al0; O[mu][3] *= al1; }
            }
            const uint32_t pt0 = movmatrix_trans(pack_h2(p00, p01));
            const uint32_t pt1 = movmatrix_trans(pack_h2(p10, p11));
            sk_pv_f16(O, sb + SM::KB, pt0, pt1, lane);
        }
        cp_async_wait<0>();
        __syncwarp();

        // ---- segment epilogue: row sums over the 8 token groups, then output or partial
#pragma unroll
        for (int o = 4; o < 32; o <<= 1) {
            l0 += __shfl_xor_sync(0xffffffffu, l0, o);
            l1 += __shfl_xor_sync(0xffffffffu, l1, o);
        }
        const int h0 = kvh*a.gqa + 2*t;                          // query heads of columns 2t, 2t+1
        const bool w0 = 2*t < a.gqa, w1 = 2*t + 1 < a.gqa;
        if (single) {
            float *o0 = static_cast<float *>(p.dst) + ((int64_t) row*p.n_head + h0)*SK_D;
            const float i0 = 1.0f/l0, i1 = 1.0f/l1;
#pragma unroll
            for (int mu = 0; mu < 8; ++mu) {
                if (w0) { o0[16*mu + g] = O[mu][0]*i0; o0[16*mu + g + 8] = O[mu][2]*i0; }
                if (w1) { o0[SK_D + 16*mu + g] = O[mu][1]*i1; o0[SK_D + 16*mu + g + 8] = O[mu][3]*i1; }
            }
        } else {
            float *o0 = a.part_o + ((int64_t) seg*p.n_head + h0)*SK_D;
#pragma unroll
            for (int mu = 0; mu < 8; ++mu) {
                if (w0) { o0[16*mu + g] = O[mu][0]; o0[16*mu + g + 8] = O[mu][2]; }
                if (w1) { o0[SK_D + 16*mu + g] = O[mu][1]; o0[SK_D + 16*mu + g + 8] = O[mu][3]; }
            }
            if (g == 0) {
                float *ml = a.part_ml + ((int64_t) seg*p.n_head + h0)*2;
                if (w0) { ml[0] = m0; ml[1] = l0; }
                if (w1) { ml[2] = m1; ml[3] = l1; }
            }
        }
        g0 = pb + t1;
        ++row;
    }
}

// merge the segments of split rows; one warp per (row, query head), log2-domain maxima
__global__ void __launch_bounds__(128)
sk_combine_kernel(const __grid_constant__ sk_args a) {
    const b200_attn_desc &p = a.d;
    const int lane = threadIdx.x & 31;
    const int64_t idx = (int64_t) blockIdx.x*4 + (threadIdx.x >> 5);
    if (idx >= (int64_t) p.n_q*p.n_head) return;
    const int row = (int) (idx / p.n_head), h = (int) (idx - (int64_t) row*p.n_head);
    const int *FS = a.plan + 4 + (p.n_q + 1);
    const int s0 = FS[row], ns = FS[row + 1] - s0;
    if (ns == 1) return;                                              // written by the main kernel
    float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
    float M = -INFINITY, L = 0.f;
    for (int s = 0; s < ns; ++s) M = fmaxf(M, a.part_ml[((int64_t) (s0 + s)*p.n_head + h)*2]);
    for (int s = 0; s < ns; ++s) {
        const float *ml = a.part_ml + ((int64_t) (s0 + s)*p.n_head + h)*2;
        const float f = ex2(ml[0] - M);
        const float4 v = *reinterpret_cast<const float4 *>(a.part_o + ((int64_t) (s0 + s)*p.n_head + h)*SK_D + 4*lane);
        o.x = fmaf(v.x, f, o.x); o.y = fmaf(v.y, f, o.y); o.z = fmaf(v.z, f, o.z); o.w = fmaf(v.w, f, o.w);
        L = fmaf(ml[1], f, L);
    }
    const float inv = 1.0f/L;                                          // ns == 0: 0 * inf = NaN like the reference (ggml-cpu.c:12418)
    o.x *= inv; o.y *= inv; o.z *= inv; o.w *= inv;
    *reinterpret_cast<float4 *>(static_cast<float *>(p.dst) + ((int64_t) row*p.n_head + h)*SK_D + 4*lane) = o;
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static int sk_sm_count() {
    static int n = 0;
    if (!n) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev); if (n <= 0) n = 148; }
    return n;
}

static bool sk_type_ok(int t) { return t == B200_TYPE_F16; }

bool streamk_supports(const b200_attn_desc &d) {
    if (!d.block_table || d.head_dim != SK_D || d.max_bias != 0.0f) return false;
    if (!sk_type_ok(d.type_k) || !sk_type_ok(d.type_v)) return false;
    const int gqa = d.n_head/d.n_head_kv;
    if (gqa < 1 || gqa > 8) return false;
    if (((uintptr_t) d.k | (uintptr_t) d.v | (uint64_t) d.nbk1 | (uint64_t) d.nbv1) & 15) return false;
    const int ka = d.type_k == B200_TYPE_F16 ? 15 : 7, va = d.type_v == B200_TYPE_F16 ? 15 : 7;
    if ((d.nbk2 & ka) || (d.nbv2 & va)) return false;
    if (((uintptr_t) d.q | (uint64_t) d.nbq1 | (uint64_t) d.nbq2) & 7) return false;      // float2 loads of Q
    if ((uintptr_t) d.dst & 15) return false;
    return true;
}

template <int TK, int TV> static int sk_ctas_per_sm() {
    const int by_smem = (227*1024)/(sk_smem<TK, TV>::CTA + 1024);
    return by_smem < 1 ? 1 : (by_smem > 2 ? 2 : by_smem);
}

static int sk_ctas_per_sm_rt(int tk, int tv) {
    (void) tk; (void) tv;
    return sk_ctas_per_sm<B200_TYPE_F16, B200_TYPE_F16>();
}

static int sk_items(const b200_attn_desc &d) {
    const int64_t warps = (int64_t) sk_sm_count()*sk_ctas_per_sm_rt(d.type_k, d.type_v)*SK_WARPS;
    int64_t items = warps/d.n_head_kv;
    if (items < 1) items = 1;
    // no more items than 4-tile pieces of the (host-estimated) work: tiny problems stay in few warps
    int64_t tok = d.total_tokens > 0 ? d.total_tokens
                : (int64_t) d.n_q*(d.max_seq_len > 0 ? d.max_seq_len : d.bt_stride*d.block_size);
    int64_t cap = tok/(4*SK_TILE) + 1;
    if (cap < d.n_q) cap = d.n_q;
    if (items > cap) items = cap;
    return (int) items;
}

static size_t sk_align(size_t x) { return (x + 255) & ~(size_t) 255; }

size_t streamk_workspace(const b200_attn_desc &d) {
    const size_t max_segs = (size_t) sk_items(d) + d.n_q;
    return sk_align((4 + 2*((size_t) d.n_q + 1))*sizeof(int)) + sk_align(max_segs*d.n_head*2*sizeof(float))
         + max_segs*d.n_head*SK_D*sizeof(float);
}

template <int TK, int TV>
static int sk_launch(const sk_args &a, cudaStream_t st) {
    using SM = sk_smem<TK, TV>;
    static bool configured = false;
    if (!configured) {
        const cudaError_t e = cudaFuncSetAttribute(sk_main_kernel<TK, TV>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM::CTA);
        if (e != cudaSuccess) { set_error("stream-K: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return B200_ERR_CUDA; }
        configured = true;
    }
    const int64_t warps = (int64_t) a.n_items*a.d.n_head_kv;
    const unsigned grid = (unsigned) ((warps + SK_WARPS - 1)/SK_WARPS);
    sk_main_kernel<TK, TV><<<grid, 32*SK_WARPS, SM::CTA, st>>>(a);
    return B200_OK;
}

int launch_attn_streamk(const b200_attn_desc &d, cudaStream_t st) {
    const size_t need = streamk_workspace(d);
    if (d.workspace == nullptr || d.workspace_bytes < need) {
        set_error("attention: workspace of %zu bytes needed, %zu given", need, d.workspace_bytes);
        return B200_ERR_ALLOC;
    }
    sk_args a;
    a.d = d;
    a.gqa = d.n_head/d.n_head_kv;
    a.n_items = sk_items(d);
    const size_t max_segs = (size_t) a.n_items + d.n_q;
    char *ws = static_cast<char *>(d.workspace);
    a.plan = reinterpret_cast<int *>(ws);
    ws += sk_align((4 + 2*((size_t) d.n_q + 1))*sizeof(int));
    a.part_ml = reinterpret_cast<float *>(ws);
    ws += sk_align(max_segs*d.n_head*2*sizeof(float));
    a.part_o = reinterpret_cast<float *>(ws);
    const float log2e = 1.4426950408889634f;
    if (d.logit_softcap != 0.0f) { a.qscale = d.scale/d.logit_softcap; a.cap_log2e = d.logit_softcap*log2e; }
    else                         { a.qscale = d.scale*log2e;           a.cap_log2e = 0.0f; }

    sk_plan_kernel<<<1, 1024, 0, st>>>(a);
    count_launch();
    if (!check_launch("sk_plan")) return B200_ERR_CUDA;
    const int rc = sk_launch<B200_TYPE_F16, B200_TYPE_F16>(a, st);
    if (rc != B200_OK) return rc;
    count_launch();
    if (!check_launch("sk_main")) return B200_ERR_CUDA;
    const int64_t pairs = (int64_t) d.n_q*d.n_head;
    sk_combine_kernel<<<(unsigned) ((pairs + 3)/4), 128, 0, st>>>(a);
    count_launch();
    if (!check_launch("sk_combine")) return B200_ERR_CUDA;
    return B200_OK;
}

} // namespace b200
