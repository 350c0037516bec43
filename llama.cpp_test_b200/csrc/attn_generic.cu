// attn_generic.cu — CUDA-core attention over the flat / paged KV cache: any strides, any GQA ratio,
// dense-mask mode (unmodified ggml graphs) and paged mode, decode or multi-token rows.
//
// This is the path north_star assigns to "single-query decode on CUDA cores" and the catch-all for
// shapes the stream-K tensor-core kernel (attn_streamk.cu) does not take.  It replaces
// flash_attn_vec_ext_f32 + flash_attn_combine_results (ggml/src/ggml-cuda/fattn-vec-f32.cuh:8-284,
// fattn-common.cuh:609-651); semantics are those of ggml_compute_forward_flash_attn_ext_f16
// (ggml/src/ggml-cpu/ggml-cpu.c:12219-12432): mask -inf cells skipped, scale, softcap, ALiBi slope on the
// mask, online softmax, GQA broadcast ik2 = iq2/rk2, dst [d, n_head, n_q].
//
// Unlike the reference kernels masked / foreign cells are never read in paged mode:
// a row walks only its own sequence's blocks through the block table.
#include "common.cuh"
#include <math.h>

namespace b200 {

constexpr int AG_WARPS = 4;

struct attn_generic_args {
    b200_attn_desc d;
    int   rk2;            // n_head / n_head_kv
    int   n_splits;
    float *partial;       // [n_q][n_head][n_splits][D+2] when n_splits > 1
    float m0, m1; int n_head_log2;   // ALiBi, ggml-cpu.c:12294-12298
};

__device__ __forceinline__ float alibi_slope(const attn_generic_args &a, int h) {
    if (!(a.d.max_bias > 0.0f)) return 1.0f;
    return h < a.n_head_log2 ? powf(a.m0, (float) (h + 1)) : powf(a.m1, (float) (2*(h - a.n_head_log2) + 1));
}

// grid (n_q, n_head/G, n_splits), block 32*AG_WARPS.  One CTA: one query row, G query heads that share
// one kv head, one slice of the row's visible positions; warps stride over positions, lanes split D.
template <int D, int G, int TK, int TV>
__global__ void __launch_bounds__(32*AG_WARPS)
attn_generic_kernel(const __grid_constant__ attn_generic_args a) {
    // elements per lane; head sizes 80 / 96 / 112 put 4 on each of the first D/4 lanes (never across a quant block), the
    // other lanes carry zeros through the dot products.  80 and 112 are no multiple of the quant block: F16 / F32 rows only
    constexpr int EPL = (D == 64 || D == 128 || D == 256) ? D/32 : 4;
    constexpr int LANES = D/EPL;
    const b200_attn_desc &p = a.d;
    const int iq = blockIdx.x, h0 = blockIdx.y*G, split = blockIdx.z;
    const int kvh = h0/a.rk2;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool paged = p.block_table != nullptr;
    const bool active = lane < LANES;

    // visible range of this row
    int64_t n_vis; int seq = 0;
    if (paged) {
        seq = p.q_seq ? p.q_seq[iq] : iq;
        n_vis = p.seq_lens[seq];
        if (p.q_pos) { const int64_t lim = (int64_t) p.q_pos[iq] + 1; n_vis = lim < n_vis ? lim : n_vis; }
        if (n_vis < 0) n_vis = 0;
    } else {
        n_vis = p.n_kv;
    }
    int64_t chunk = (n_vis + a.n_splits - 1)/a.n_splits;
    chunk = (chunk + 31)/32*32;
    const int64_t i_begin = (int64_t) split*chunk;
    const int64_t i_end = i_begin + chunk < n_vis ? i_begin + chunk : n_vis;

    float q[G][EPL], acc[G][EPL], m[G], l[G], slope[G];
    float scale = p.scale;
    if (p.logit_softcap != 0.0f) scale /= p.logit_softcap;          // ggml-cpu.c:12290-12292
#pragma unroll
    for (int g = 0; g < G; ++g) {
        const float *qp = reinterpret_cast<const float *>(static_cast<const char *>(p.q) + (int64_t) iq*p.nbq1 + (int64_t) (h0 + g)*p.nbq2);
#pragma unroll
        for (int e = 0; e < EPL; ++e) { q[g][e] = active ? qp[lane*EPL + e]*scale : 0.0f; acc[g][e] = 0.0f; }
        m[g] = -INFINITY; l[g] = 0.0f;
        slope[g] = alibi_slope(a, h0 + g);
    }
    // ALiBi through a block table: the bias the reference's mask would carry for this row, -|pos(cell) - pos(query)| rounded to
    // half (llm_graph_input_attn_kv_unified::set_input, src/llama-graph.cpp:403-477, fills an F16 mask); cell i of a sequence
    // holds position i, the query sits at q_pos or at the sequence's newest position
    const bool alibi_paged = paged && p.max_bias > 0.0f;
    const int64_t qpos = p.q_pos ? (int64_t) p.q_pos[iq] : n_vis - 1;
    const __half *mrow = (!paged && p.mask) ? reinterpret_cast<const __half *>(static_cast<const char *>(p.mask) + (int64_t) iq*p.nb_mask1) : nullptr;
    const int32_t *bt = paged ? p.block_table + (int64_t) seq*p.bt_stride : nullptr;
    const char *kbase = static_cast<const char *>(p.k) + (int64_t) kvh*p.nbk2;
    const char *vbase = static_cast<const char *>(p.v) + (int64_t) kvh*p.nbv2;

    for (int64_t i = i_begin + warp; i < i_end; i += AG_WARPS) {
        int64_t cell; float mval = 0.0f;
        if (paged) {
            cell = (int64_t) bt[i/p.block_size]*p.block_size + i % p.block_size;
            if (alibi_paged) mval = __half2float(__float2half_rn(-fabsf((float) (i - qpos))));
        } else {
            cell = i;
            if (mrow) {
                mval = __half2float(mrow[i]);
                if (mval == -INFINITY) continue;                      // ggml-cpu.c:12350 (warp-uniform)
            }
        }
        float kf[EPL], vf[EPL];
        if (LANES == 32 || active) {
            load_row_elems<TK, EPL>(kbase + cell*p.nbk1, lane*EPL, kf);
            load_row_elems<TV, EPL>(vbase + cell*p.nbv1, lane*EPL, vf);
        } else {
#pragma unroll
            for (int e = 0; e < EPL; ++e) kf[e] = vf[e] = 0.0f;
        }
#pragma unroll
        for (int g = 0; g < G; ++g) {
            float s = 0.0f;
#pragma unroll
            for (int e = 0; e < EPL; ++e) s = fmaf(q[g][e], kf[e], s);
            s = warp_sum(s);
            if (p.logit_softcap != 0.0f) s = p.logit_softcap*tanhf(s);     // :12361-12363
            s += slope[g]*mval;                                            // :12349, :12365
            const float m_new = fmaxf(m[g], s);
            const float ms = __expf(m[g] - m_new);                          // 0 when m was -inf
            const float vs = __expf(s - m_new);
            l[g] = l[g]*ms + vs;
#pragma unroll
            for (int e = 0; e < EPL; ++e) acc[g][e] = fmaf(acc[g][e], ms, vs*vf[e]);
            m[g] = m_new;
        }
    }

    // merge the warps
    __shared__ float s_acc[AG_WARPS][G][D];
    __shared__ float s_m[AG_WARPS][G], s_l[AG_WARPS][G];
#pragma unroll
    for (int g = 0; g < G; ++g) {
#pragma unroll
        for (int e = 0; e < EPL; ++e) if (LANES == 32 || active) s_acc[warp][g][lane*EPL + e] = acc[g][e];
        if (lane == 0) { s_m[warp][g] = m[g]; s_l[warp][g] = l[g]; }
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < G*D; idx += 32*AG_WARPS) {
        const int g = idx / D, e = idx - g*D;
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < AG_WARPS; ++w) M = fmaxf(M, s_m[w][g]);
        float o = 0.0f, L = 0.0f;
#pragma unroll
        for (int w = 0; w < AG_WARPS; ++w) {
            const float f = s_m[w][g] == -INFINITY ? 0.0f : __expf(s_m[w][g] - M);
            o = fmaf(s_acc[w][g][e], f, o);
            L = fmaf(s_l[w][g], f, L);
        }
        const int h = h0 + g;
        if (a.n_splits == 1) {
            static_cast<float *>(p.dst)[((int64_t) iq*p.n_head + h)*D + e] = o/L;   // 0/0 = NaN like 1.0f/S, :12418
        } else {
            float *pp = a.partial + (((int64_t) iq*p.n_head + h)*a.n_splits + split)*(D + 2);
            pp[e] = o;
            if (e == 0) { pp[D] = M; pp[D + 1] = L; }
        }
    }
}

// grid (n_q*n_head), block D: merge split partials (replaces flash_attn_combine_results)
template <int D>
__global__ void __launch_bounds__(D)
attn_combine_kernel(const float *partial, float *dst, int n_splits) {
    const int64_t row = blockIdx.x;
    const float *pp = partial + row*n_splits*(D + 2);
    float M = -INFINITY;
    for (int s = 0; s < n_splits; ++s) M = fmaxf(M, pp[s*(D + 2) + D]);
    float o = 0.0f, L = 0.0f;
    for (int s = 0; s < n_splits; ++s) {
        const float ms = pp[s*(D + 2) + D];
        const float f = ms == -INFINITY ? 0.0f : __expf(ms - M);
        o = fmaf(pp[s*(D + 2) + threadIdx.x], f, o);
        L = fmaf(pp[s*(D + 2) + D + 1], f, L);
    }
    dst[row*D + threadIdx.x] = o/L;
}

template <int D, int G, int TK>
static void launch_tv(const attn_generic_args &a, dim3 grid, cudaStream_t st) {
    if constexpr (D % 32 != 0) {                     // no quantised rows of these lengths exist (generic_supports)
        if (a.d.type_v == B200_TYPE_F32) attn_generic_kernel<D, G, TK, B200_TYPE_F32><<<grid, 32*AG_WARPS, 0, st>>>(a);
        else                             attn_generic_kernel<D, G, TK, B200_TYPE_F16><<<grid, 32*AG_WARPS, 0, st>>>(a);
    } else {
        switch (a.d.type_v) {
            case B200_TYPE_F32:  attn_generic_kernel<D, G, TK, B200_TYPE_F32 ><<<grid, 32*AG_WARPS, 0, st>>>(a); break;
            case B200_TYPE_F16:  attn_generic_kernel<D, G, TK, B200_TYPE_F16 ><<<grid, 32*AG_WARPS, 0, st>>>(a); break;
            case B200_TYPE_Q8_0: attn_generic_kernel<D, G, TK, B200_TYPE_Q8_0><<<grid, 32*AG_WARPS, 0, st>>>(a); break;
            default:             attn_generic_kernel<D, G, TK, B200_TYPE_Q4_0><<<grid, 32*AG_WARPS, 0, st>>>(a); break;
        }
    }
}
template <int D, int G>
static void launch_tk(const attn_generic_args &a, dim3 grid, cudaStream_t st) {
    if constexpr (D % 32 != 0) {
        if (a.d.type_k == B200_TYPE_F32) launch_tv<D, G, B200_TYPE_F32>(a, grid, st);
        else                             launch_tv<D, G, B200_TYPE_F16>(a, grid, st);
    } else {
        switch (a.d.type_k) {
            case B200_TYPE_F32:  launch_tv<D, G, B200_TYPE_F32 >(a, grid, st); break;
            case B200_TYPE_F16:  launch_tv<D, G, B200_TYPE_F16 >(a, grid, st); break;
            case B200_TYPE_Q8_0: launch_tv<D, G, B200_TYPE_Q8_0>(a, grid, st); break;
            default:             launch_tv<D, G, B200_TYPE_Q4_0>(a, grid, st); break;
        }
    }
}
template <int D>
static void launch_g(const attn_generic_args &a, int G, dim3 grid, cudaStream_t st) {
    if (G == 8)      launch_tk<D, 8>(a, grid, st);
    else if (G == 4) launch_tk<D, 4>(a, grid, st);
    else             launch_tk<D, 1>(a, grid, st);
}

bool generic_supports(const b200_attn_desc &d) {
    const bool plain = (d.type_k == B200_TYPE_F16 || d.type_k == B200_TYPE_F32) && (d.type_v == B200_TYPE_F16 || d.type_v == B200_TYPE_F32);
    const bool hs_ok = d.head_dim == 64 || d.head_dim == 96 || d.head_dim == 128 || d.head_dim == 256
                    || ((d.head_dim == 80 || d.head_dim == 112) && plain);
    return hs_ok && type_ok(d.type_k) && type_ok(d.type_v) && d.n_head_kv > 0 && d.n_head % d.n_head_kv == 0;
}

static int generic_splits(const b200_attn_desc &d, int G) {
    int64_t max_len = d.block_table ? (d.max_seq_len > 0 ? d.max_seq_len : d.bt_stride*d.block_size) : d.n_kv;
    const int64_t ctas = (int64_t) d.n_q*(d.n_head/G);
    const int64_t target = 4LL*148;                                  // ~4 CTAs of 4 warps per SM
    int64_t s = (target + ctas - 1)/ctas;
    const int64_t max_s = (max_len + 255)/256;                       // at least 256 positions per split
    if (s > max_s) s = max_s;
    if (s < 1) s = 1;
    if (s > 64) s = 64;
    return (int) s;
}

static int generic_group(const b200_attn_desc &d) {
    const int rk2 = d.n_head/d.n_head_kv;
    return rk2 % 8 == 0 ? 8 : rk2 % 4 == 0 ? 4 : 1;
}

size_t generic_workspace(const b200_attn_desc &d) {
    const int s = generic_splits(d, generic_group(d));
    return s > 1 ? (size_t) d.n_q*d.n_head*s*(d.head_dim + 2)*sizeof(float) : 0;
}

int launch_attn_generic(const b200_attn_desc &d, cudaStream_t st) {
    attn_generic_args a;
    a.d = d;
    a.rk2 = d.n_head/d.n_head_kv;
    const int G = generic_group(d);
    a.n_splits = generic_splits(d, G);
    a.partial = static_cast<float *>(d.workspace);
    if (a.n_splits > 1 && (d.workspace == nullptr || d.workspace_bytes < generic_workspace(d))) {
        set_error("attention: workspace of %zu bytes needed, %zu given", generic_workspace(d), d.workspace_bytes);
        return B200_ERR_ALLOC;
    }
    a.n_head_log2 = 1 << (int) floor(log2((double) d.n_head));
    a.m0 = powf(2.0f, -(d.max_bias       )/a.n_head_log2);
    a.m1 = powf(2.0f, -(d.max_bias / 2.0f)/a.n_head_log2);
    const dim3 grid((unsigned) d.n_q, (unsigned) (d.n_head/G), (unsigned) a.n_splits);
    switch (d.head_dim) {
        case 128: launch_g<128>(a, G, grid, st); break;
        case 256: launch_g<256>(a, G, grid, st); break;
        case 80:  launch_g<80 >(a, G, grid, st); break;
        case 96:  launch_g<96 >(a, G, grid, st); break;
        case 112: launch_g<112>(a, G, grid, st); break;
        default:  launch_g<64 >(a, G, grid, st); break;
    }
    count_launch();
    if (!check_launch("attn_generic")) return B200_ERR_CUDA;
    if (a.n_splits > 1) {
        const unsigned rows = (unsigned) d.n_q*d.n_head;
        float *dst = static_cast<float *>(d.dst);
        switch (d.head_dim) {
            case 128: attn_combine_kernel<128><<<rows, 128, 0, st>>>(a.partial, dst, a.n_splits); break;
            case 256: attn_combine_kernel<256><<<rows, 256, 0, st>>>(a.partial, dst, a.n_splits); break;
            case 80:  attn_combine_kernel<80 ><<<rows, 80,  0, st>>>(a.partial, dst, a.n_splits); break;
            case 96:  attn_combine_kernel<96 ><<<rows, 96,  0, st>>>(a.partial, dst, a.n_splits); break;
            case 112: attn_combine_kernel<112><<<rows, 112, 0, st>>>(a.partial, dst, a.n_splits); break;
            default:  attn_combine_kernel<64 ><<<rows, 64,  0, st>>>(a.partial, dst, a.n_splits); break;
        }
        count_launch();
        if (!check_launch("attn_combine")) return B200_ERR_CUDA;
    }
    return B200_OK;
}

} // namespace b200
