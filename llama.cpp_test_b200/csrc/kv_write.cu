// kv_write.cu — slot-mapped KV-cache write with fused F32 -> {F32,F16,Q8_0,Q4_0} conversion, and the
// batched block copy used for copy-on-write / defrag moves.
//
// Replaces (reference paths): ggml_cuda_cpy + cpy_f32_f16 / cpy_f32_q<cpy_blck_f32_q8_0|q4_0>
// (ggml/src/ggml-cuda/cpy.cu:34-127, :290-402, :510-584) as scheduled by build_attn
// (src/llama-graph.cpp:1375-1397), and the per-move cudaMemcpyAsync of defrag (cpy.cu:544-547,
// src/llama-context.cpp:653-711).  The reference launches ONE single-thread CTA per 32-element quant
// block (cpy.cu:372-381); here one CTA owns one token row, 8 lanes cooperate on each quant block with
// shuffles, the converted row is staged in shared memory and leaves the SM as coalesced 16-byte stores
// (prefill-sized stores: four token rows per CTA, all their source loads in flight before the first use).
//
// Bit-exactness: arithmetic is the IEEE sequence of quantize_row_q8_0_ref / quantize_row_q4_0_ref
// (ggml/src/ggml-quants.c:194-216, :31-66) with explicit _rn intrinsics so nvcc cannot contract
// x*id + 8.5f into an FMA; half conversion is RNE like ggml_fp32_to_fp16_row (ggml/src/ggml.c:387).
#include "common.cuh"
#include <cstdlib>

namespace b200 {

constexpr int KVW_THREADS = 256;

// 8-lane group reductions (lanes of a group are contiguous and 8-aligned)
__device__ __forceinline__ float group8_max(float v) {
    v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
    v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
    v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 4));
    return v;
}

// first element (in index order) with the largest |v|: quantize_row_q4_0_ref keeps `max` of the
// first strictly-greater amax (ggml-quants.c:42-48), so ties resolve to the lower index.
__device__ __forceinline__ void group8_absmax_first(float &amax, float &vmax, int lane) {
#pragma unroll
    for (int o = 1; o <= 4; o <<= 1) {
        const float oa = __shfl_xor_sync(0xffffffffu, amax, o);
        const float ov = __shfl_xor_sync(0xffffffffu, vmax, o);
        const bool partner_is_lower = ((lane ^ o) < lane);
        // lo = lower lane's candidate, hi = higher lane's; result = (lo.amax < hi.amax) ? hi : lo
        const float lo_a = partner_is_lower ? oa : amax, hi_a = partner_is_lower ? amax : oa;
        const float lo_v = partner_is_lower ? ov : vmax, hi_v = partner_is_lower ? vmax : ov;
        const bool take_hi = lo_a < hi_a;
        amax = take_hi ? hi_a : lo_a;
        vmax = take_hi ? hi_v : lo_v;
    }
}

// Convert 4 consecutive floats (element index e0 = 4*i of the token row) into the staged row.
template <int TYPE>
__device__ __forceinline__ void convert4(const float4 x, int e0, char *stage, int lane, bool active) {
    if constexpr (TYPE == B200_TYPE_F32) {
        if (!active) return;
        *reinterpret_cast<float4 *>(stage + 4*e0) = x;
    } else if constexpr (TYPE == B200_TYPE_F16) {
        if (!active) return;
        const __half2 a = __floats2half2_rn(x.x, x.y), b = __floats2half2_rn(x.z, x.w);
        uint2 w;
        w.x = *reinterpret_cast<const uint32_t *>(&a);
        w.y = *reinterpret_cast<const uint32_t *>(&b);
        *reinterpret_cast<uint2 *>(stage + 2*e0) = w;
    } else if constexpr (TYPE == B200_TYPE_Q8_0) {
        // ggml-quants.c:199-214
        float amax = fmaxf(fmaxf(fabsf(x.x), fabsf(x.y)), fmaxf(fabsf(x.z), fabsf(x.w)));
        amax = group8_max(amax);
        const float d  = __fdiv_rn(amax, 127.0f);
        const float id = d ? __fdiv_rn(1.0f, d) : 0.0f;
        const int q0 = (int) roundf(__fmul_rn(x.x, id)), q1 = (int) roundf(__fmul_rn(x.y, id));
        const int q2 = (int) roundf(__fmul_rn(x.z, id)), q3 = (int) roundf(__fmul_rn(x.w, id));
        if (!active) return;
        char *blk = stage + (e0/QK)*Q8_BYTES;
        uint16_t *qs = reinterpret_cast<uint16_t *>(blk + 2 + (e0 % QK));      // 2-byte aligned
        qs[0] = (uint16_t) ((q0 & 0xff) | ((q1 & 0xff) << 8));
        qs[1] = (uint16_t) ((q2 & 0xff) | ((q3 & 0xff) << 8));
        if ((lane & 7) == 0) *reinterpret_cast<__half *>(blk) = __float2half_rn(d);
    } else {                                                                   // Q4_0, ggml-quants.c:38-64
        float amax = 0.0f, vmax = 0.0f;
        if (amax < fabsf(x.x)) { amax = fabsf(x.x); vmax = x.x; }
        if (amax < fabsf(x.y)) { amax = fabsf(x.y); vmax = x.y; }
        if (amax < fabsf(x.z)) { amax = fabsf(x.z); vmax = x.z; }
        if (amax < fabsf(x.w)) { amax = fabsf(x.w); vmax = x.w; }
        group8_absmax_first(amax, vmax, lane);
        const float d  = __fdiv_rn(vmax, -8.0f);
        const float id = d ? __fdiv_rn(1.0f, d) : 0.0f;
        auto q4 = [&](float v) -> uint32_t {
            const int t = (int) (int8_t) (int) __fadd_rn(__fmul_rn(v, id), 8.5f);   // (int8_t)(x0 + 8.5f)
            return (uint32_t) (t < 15 ? t : 15) & 0xffu;                            // MIN(15, .)
        };
        // 4 codes, one per byte lane of a 32-bit word
        uint32_t mine = q4(x.x) | (q4(x.y) << 8) | (q4(x.z) << 16) | (q4(x.w) << 24);
        const uint32_t theirs = __shfl_xor_sync(0xffffffffu, mine, 4);
        if (!active) return;
        char *blk = stage + (e0/QK)*Q4_BYTES;
        const int t = lane & 7;
        if (t < 4) {            // elements 4t..4t+3 are the low nibbles of qs[4t..4t+3]; partner holds +16
            const uint32_t w = (mine & 0x0f0f0f0fu) | ((theirs & 0x0f0f0f0fu) << 4);
            uint16_t *qs = reinterpret_cast<uint16_t *>(blk + 2 + 4*t);
            qs[0] = (uint16_t) (w & 0xffffu);
            qs[1] = (uint16_t) (w >> 16);
        }
        if (t == 0) *reinterpret_cast<__half *>(blk) = __float2half_rn(d);
    }
}

// the staged (converted, heads packed) row -> cache + cell*nb_token; the whole CTA takes part, after a barrier behind the staging
template <int TYPE>
__device__ __forceinline__ void store_staged_row(const char *stage, char *cache, int64_t cell, int n_head_kv, int head_dim,
                                                 int64_t nb_token, int64_t nb_head) {
    const int64_t head_bytes = row_bytes(TYPE, head_dim);
    char *dst = cache + cell*nb_token;
    if (nb_head == head_bytes && (((uintptr_t) dst | (uintptr_t) (head_bytes*n_head_kv)) & 15) == 0) {
        const int n16 = (int) (head_bytes*n_head_kv/16);
        for (int i = threadIdx.x; i < n16; i += KVW_THREADS)
            reinterpret_cast<uint4 *>(dst)[i] = reinterpret_cast<const uint4 *>(stage)[i];
    } else {                                   // arbitrary head stride / alignment: 2-byte units
        const int per_head = (int) (head_bytes/2);
        for (int i = threadIdx.x; i < per_head*n_head_kv; i += KVW_THREADS) {
            const int h = i / per_head, o = i - h*per_head;
            reinterpret_cast<uint16_t *>(dst + h*nb_head)[o] = reinterpret_cast<const uint16_t *>(stage)[i];
        }
    }
}

template <int TYPE>
__device__ void write_row(const float *src, char *cache, int64_t cell, int n_head_kv, int head_dim,
                          int64_t nb_token, int64_t nb_head, char *stage) {
    const int row_elems = n_head_kv*head_dim;
    const int lane = threadIdx.x & 31;
    // uniform trip count for the whole CTA: lanes past the row still take part in the shuffles
    for (int base = 0; base < row_elems; base += 4*KVW_THREADS) {
        const int e0 = base + 4*threadIdx.x;
        const bool active = e0 < row_elems;
        const float4 x = active ? *reinterpret_cast<const float4 *>(src + e0) : make_float4(0.f, 0.f, 0.f, 0.f);
        convert4<TYPE>(x, e0, stage, lane, active);
    }
    __syncthreads();
    store_staged_row<TYPE>(stage, cache, cell, n_head_kv, head_dim, nb_token, nb_head);
}

// KVW_ROWS token rows of one CTA, every row's source loads (and slots) requested before the first is used: a CTA that
// takes one row has one 4 KB load in flight, 8 CTAs per SM cover a third of the bytes HBM latency asks for
// (config 3's 4096-token store ran at 0.59 of HBM that way).  Rows of up to 4*KVW_THREADS elements only.
constexpr int KVW_ROWS = 4;

template <int TYPE>
__device__ __forceinline__ void write_rows(const b200_kv_write_desc &p, const float *src0, char *cache, int64_t t0, int64_t nbt, int64_t nbh, char *stage) {
    const int row_elems = p.n_head_kv*p.head_dim;
    const int lane = threadIdx.x & 31, e0 = 4*threadIdx.x;
    const bool active = e0 < row_elems;
    const int stage_row = (int) row_bytes(TYPE, row_elems);
    float4 x[KVW_ROWS];
    int64_t cell[KVW_ROWS];
#pragma unroll
    for (int r = 0; r < KVW_ROWS; ++r) {
        const int64_t t = t0 + r;
        const bool live = t < p.n_tokens;
        x[r] = live && active ? *reinterpret_cast<const float4 *>(reinterpret_cast<const char *>(src0) + t*p.src_nb1 + 4*e0) : make_float4(0.f, 0.f, 0.f, 0.f);
        cell[r] = !live ? -1 : p.slot_mapping ? (int64_t) p.slot_mapping[t] : t;
    }
#pragma unroll
    for (int r = 0; r < KVW_ROWS; ++r) convert4<TYPE>(x[r], e0, stage + r*stage_row, lane, active);
    __syncthreads();
#pragma unroll
    for (int r = 0; r < KVW_ROWS; ++r)
        if (cell[r] >= 0 && cell[r] < p.n_cells) store_staged_row<TYPE>(stage + r*stage_row, cache, cell[r], p.n_head_kv, p.head_dim, nbt, nbh);
}

// grid (ceil(n_tokens / rows), 2): y == 0 writes K, y == 1 writes V; ROWS == 1: one token per CTA, rows of any length
template <int ROWS>
__global__ void __launch_bounds__(KVW_THREADS)
kv_write_kernel(const b200_kv_write_desc p) {
    extern __shared__ __align__(16) char stage[];
    TL_BEGIN(1, threadIdx.x == 0);
    pdl_prologue_strict();                                            // the attention behind this store may start early (attn_streamk.cu)
    TL_STAMP(2);
    const bool is_v = blockIdx.y == 1;
    const float *src = is_v ? p.v_src : p.k_src;
    if (src == nullptr) { TL_END(); return; }
    char *cache = static_cast<char *>(is_v ? p.v_cache : p.k_cache);
    const int type = is_v ? p.type_v : p.type_k;
    const int64_t nbt = is_v ? p.nbv_token : p.nbk_token, nbh = is_v ? p.nbv_head : p.nbk_head;
    if constexpr (ROWS > 1) {
        const int64_t t0 = (int64_t) blockIdx.x*ROWS;
        switch (type) {
            case B200_TYPE_F32:  write_rows<B200_TYPE_F32 >(p, src, cache, t0, nbt, nbh, stage); break;
            case B200_TYPE_F16:  write_rows<B200_TYPE_F16 >(p, src, cache, t0, nbt, nbh, stage); break;
            case B200_TYPE_Q8_0: write_rows<B200_TYPE_Q8_0>(p, src, cache, t0, nbt, nbh, stage); break;
            default:             write_rows<B200_TYPE_Q4_0>(p, src, cache, t0, nbt, nbh, stage); break;
        }
    } else {
        const int64_t t = blockIdx.x;
        const int64_t cell = p.slot_mapping ? (int64_t) p.slot_mapping[t] : t;
        if (cell < 0 || cell >= p.n_cells) { TL_END(); return; }     // writeToCache skips invalid slots
        src = reinterpret_cast<const float *>(reinterpret_cast<const char *>(src) + t*p.src_nb1);
        switch (type) {
            case B200_TYPE_F32:  write_row<B200_TYPE_F32 >(src, cache, cell, p.n_head_kv, p.head_dim, nbt, nbh, stage); break;
            case B200_TYPE_F16:  write_row<B200_TYPE_F16 >(src, cache, cell, p.n_head_kv, p.head_dim, nbt, nbh, stage); break;
            case B200_TYPE_Q8_0: write_row<B200_TYPE_Q8_0>(src, cache, cell, p.n_head_kv, p.head_dim, nbt, nbh, stage); break;
            default:             write_row<B200_TYPE_Q4_0>(src, cache, cell, p.n_head_kv, p.head_dim, nbt, nbh, stage); break;
        }
    }
    TL_END();
}

int launch_kv_write(const b200_kv_write_desc &d, cudaStream_t stream) {
    if (d.n_tokens == 0) return B200_OK;
    const int64_t row_elems = (int64_t) d.n_head_kv*d.head_dim;
    const int64_t sk = row_bytes(d.type_k, row_elems), sv = row_bytes(d.type_v, row_elems);
    const size_t smem = (size_t) (sk > sv ? sk : sv);
    if (smem > 48*1024) { set_error("kv_write: token row of %lld bytes exceeds 48 KiB staging", (long long) smem); return B200_ERR_UNSUPPORTED; }
    TL_ARM();
    // several rows per CTA once one-row CTAs would no longer be resident all at once (8 per SM); decode-sized stores keep one
    static int sms = 0;
    if (!sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); if (sms <= 0) sms = 148; }
    if (row_elems <= 4*KVW_THREADS && 2*d.n_tokens > 8*(int64_t) sms && smem*KVW_ROWS <= 48*1024) {
        const dim3 grid((unsigned) ((d.n_tokens + KVW_ROWS - 1)/KVW_ROWS), 2);
        launch_pdl(kv_write_kernel<KVW_ROWS>, grid, dim3(KVW_THREADS), smem*KVW_ROWS, stream, d);
    } else {
        const dim3 grid((unsigned) d.n_tokens, 2);
        launch_pdl(kv_write_kernel<1>, grid, dim3(KVW_THREADS), smem, stream, d);
    }
    count_launch();
    return check_launch("kv_write") ? B200_OK : B200_ERR_CUDA;
}

// ---------------------------------------------------------------------------------------------
// batched block copy: grid (n_pairs, n_caches); each CTA streams one block with 16-byte accesses
// ---------------------------------------------------------------------------------------------
constexpr int BC_MAX_CACHES = 128;
struct block_copy_args {
    char   *caches[BC_MAX_CACHES];
    int64_t block_bytes, n_blocks;
    const int32_t *pairs;
};

__global__ void __launch_bounds__(256)
block_copy_kernel(const __grid_constant__ block_copy_args a) {
    const int32_t s = a.pairs[2*blockIdx.x], d = a.pairs[2*blockIdx.x + 1];
    if (s == d || s < 0 || d < 0 || s >= a.n_blocks || d >= a.n_blocks) return;
    char *base = a.caches[blockIdx.y];
    const uint4 *src = reinterpret_cast<const uint4 *>(base + (int64_t) s*a.block_bytes);
    uint4       *dst = reinterpret_cast<uint4 *>(base + (int64_t) d*a.block_bytes);
    const int n16 = (int) (a.block_bytes/16);
    int i = threadIdx.x;
    for (; i + 3*256 < n16; i += 4*256) {                            // 4 independent 16-byte loads in flight
        const uint4 v0 = src[i], v1 = src[i + 256], v2 = src[i + 512], v3 = src[i + 768];
        dst[i] = v0; dst[i + 256] = v1; dst[i + 512] = v2; dst[i + 768] = v3;
    }
    for (; i < n16; i += 256) dst[i] = src[i];
}

int launch_block_copy(void *const *caches, int n_caches, int64_t block_bytes, int64_t n_blocks,
                      const int32_t *pairs, int64_t n_pairs, cudaStream_t stream) {
    if (n_pairs == 0 || n_caches == 0) return B200_OK;
    if (block_bytes <= 0 || (block_bytes & 15)) { set_error("block_copy: block_bytes %lld not a multiple of 16", (long long) block_bytes); return B200_ERR_INVALID; }
    for (int c0 = 0; c0 < n_caches; c0 += BC_MAX_CACHES) {
        block_copy_args a;
        const int n = n_caches - c0 < BC_MAX_CACHES ? n_caches - c0 : BC_MAX_CACHES;
        for (int i = 0; i < n; ++i) {
            a.caches[i] = static_cast<char *>(caches[c0 + i]);
            if (((uintptr_t) a.caches[i]) & 15) { set_error("block_copy: cache %d not 16-byte aligned", c0 + i); return B200_ERR_INVALID; }
        }
        a.block_bytes = block_bytes; a.n_blocks = n_blocks; a.pairs = pairs;
        block_copy_kernel<<<dim3((unsigned) n_pairs, (unsigned) n), 256, 0, stream>>>(a);
        count_launch();
        if (!check_launch("block_copy")) return B200_ERR_CUDA;
    }
    return B200_OK;
}

// ------------------------------------------------------------------------------------------------
// token-ordered rows <-> scattered cells (session state save / load on the paged layout)
// ------------------------------------------------------------------------------------------------
struct kv_rows_args {
    char   *caches[BC_MAX_CACHES];
    char   *staging;                 // [n_caches of this launch][n_rows][row_bytes]
    int64_t nb_token, row_bytes, n_cells, n_rows;
    const int32_t *slots;            // [n_rows] cell of row i; out of range: row skipped
};

template <bool SCATTER>
__global__ void __launch_bounds__(128)
kv_rows_kernel(const __grid_constant__ kv_rows_args a) {
    const int64_t n16 = a.row_bytes/16;
    char *base = a.caches[blockIdx.y], *stage = a.staging + (int64_t) blockIdx.y*a.n_rows*a.row_bytes;
    for (int64_t r = blockIdx.x; r < a.n_rows; r += gridDim.x) {
        const int32_t c = a.slots[r];
        if (c < 0 || c >= a.n_cells) continue;
        uint4 *cell = reinterpret_cast<uint4 *>(base + (int64_t) c*a.nb_token), *row = reinterpret_cast<uint4 *>(stage + r*a.row_bytes);
        for (int64_t i = threadIdx.x; i < n16; i += 128) { if (SCATTER) cell[i] = row[i]; else row[i] = cell[i]; }
    }
}

int launch_kv_rows(bool scatter, void *const *caches, int n_caches, int64_t nb_token, int64_t row_bytes, int64_t n_cells,
                   const int32_t *slots, int64_t n_rows, void *staging, cudaStream_t stream) {
    if (n_rows == 0 || n_caches == 0) return B200_OK;
    if (row_bytes <= 0 || (row_bytes & 15) || (nb_token & 15) || nb_token < row_bytes || (((uintptr_t) staging) & 15)) {
        set_error("kv_rows: row_bytes %lld / nb_token %lld / staging must be multiples of 16 with row_bytes <= nb_token", (long long) row_bytes, (long long) nb_token);
        return B200_ERR_INVALID;
    }
    const unsigned gx = (unsigned) (n_rows < 148*32 ? n_rows : 148*32);
    for (int c0 = 0; c0 < n_caches; c0 += BC_MAX_CACHES) {
        kv_rows_args a;
        const int n = n_caches - c0 < BC_MAX_CACHES ? n_caches - c0 : BC_MAX_CACHES;
        for (int i = 0; i < n; ++i) {
            a.caches[i] = static_cast<char *>(caches[c0 + i]);
            if (((uintptr_t) a.caches[i]) & 15) { set_error("kv_rows: cache %d not 16-byte aligned", c0 + i); return B200_ERR_INVALID; }
        }
        a.staging = static_cast<char *>(staging) + (int64_t) c0*n_rows*row_bytes;
        a.nb_token = nb_token; a.row_bytes = row_bytes; a.n_cells = n_cells; a.n_rows = n_rows; a.slots = slots;
        if (scatter) kv_rows_kernel<true><<<dim3(gx, (unsigned) n), 128, 0, stream>>>(a);
        else         kv_rows_kernel<false><<<dim3(gx, (unsigned) n), 128, 0, stream>>>(a);
        count_launch();
        if (!check_launch("kv_rows")) return B200_ERR_CUDA;
    }
    return B200_OK;
}

// ------------------------------------------------------------------------------------------------
// K-shift: cached K rows re-rotated in place after a position change (context shift / seq_add / seq_div).
// One CTA per cell and group of layers (the cos/sin table of a cell is the same for every layer and costs more than a
// row: up to 63 dependent multiplications plus a large-argument sincos; built once, it serves the whole group); each row is
// expanded to F32 in shared memory (the reference's ggml_cast of a
// quantised cache, src/llama-context.cpp:469-471; FP16_TO_FP32 per element for F16, ggml-cpu.c:10955+), rotated
// with the cos/sin table of its position (ggml_rope_cache_init, ggml-cpu.c:10595-10611: theta advances by repeated
// multiplication, YaRN mix per rope_yarn :10578-10593), then written back through the KV-write path above (the
// reference's ggml_cpy back into the cache, llama-context.cpp:487).  _rn intrinsics keep the reference's unfused
// float sequence; cosf / sinf are CUDA's (<= 2 ulp from libm: results may differ in the last half ulp / one code).
// ------------------------------------------------------------------------------------------------
struct k_shift_args {
    char *caches[BC_MAX_CACHES];
    b200_k_shift_desc d;                     // .caches is a host pointer: not used on the device
    float theta_scale, corr0, corr1, mscale; // powf(freq_base, -2/n_rot), ggml_rope_yarn_corr_dims, attn_factor [* YaRN magnitude]
    int   n;                                 // caches in this launch; CTA (x, y) serves caches y, y + gridDim.y, ... of row x
};

template <int TYPE>
__device__ __forceinline__ float k_shift_load(const char *head, int i) {
    if constexpr (TYPE == B200_TYPE_F32)       return reinterpret_cast<const float *>(head)[i];
    else if constexpr (TYPE == B200_TYPE_F16)  return __half2float(reinterpret_cast<const __half *>(head)[i]);
    else if constexpr (TYPE == B200_TYPE_Q8_0) {       // dequantize_row_q8_0, ggml-quants.c:349-362
        const char *blk = head + (i/QK)*Q8_BYTES;
        return __fmul_rn((float) reinterpret_cast<const int8_t *>(blk + 2)[i % QK], __half2float(*reinterpret_cast<const __half *>(blk)));
    } else {                                           // dequantize_row_q4_0, ggml-quants.c:255-273
        const char *blk = head + (i/QK)*Q4_BYTES;
        const int j = i % QK;
        const int b = reinterpret_cast<const uint8_t *>(blk + 2)[j & 15];
        const int q = (j < 16 ? (b & 0x0f) : (b >> 4)) - 8;
        return __fmul_rn((float) q, __half2float(*reinterpret_cast<const __half *>(blk)));
    }
}

template <int TYPE>
__global__ void __launch_bounds__(KVW_THREADS)
k_shift_kernel(const __grid_constant__ k_shift_args a) {
    extern __shared__ __align__(16) char ks_smem[];
    const b200_k_shift_desc &p = a.d;
    const int hd = p.head_dim, row_elems = p.n_head_kv*hd, half_rot = p.n_rot/2;
    float *x  = reinterpret_cast<float *>(ks_smem);
    float *cs = x + row_elems;                                         // (cos, sin) per rotated pair
    char  *stage = ks_smem + (((size_t) (row_elems + p.n_rot)*4 + 15) & ~(size_t) 15);
    const bool neox = (p.mode & 2) != 0;
    const int64_t n_rows = p.cells ? p.n_rows : p.n_cells;
    for (int64_t r = blockIdx.x; r < n_rows; r += gridDim.x) {
        const int64_t cell = p.cells ? (int64_t) p.cells[r] : r;
        if (cell < 0 || cell >= p.n_cells) continue;                 // CTA-uniform
        const float pos = (float) (int64_t) p.shift[cell];
        for (int ic = threadIdx.x; ic < half_rot; ic += KVW_THREADS) {
            float theta = pos;
            for (int k = 0; k < ic; ++k) theta = __fmul_rn(theta, a.theta_scale);
            const float ff = p.freq_factors ? p.freq_factors[ic] : 1.0f;
            const float extrap = __fdiv_rn(theta, ff);
            const float interp = __fmul_rn(p.freq_scale, extrap);
            float th = interp;
            if (p.ext_factor != 0.0f) {
                const float y = __fdiv_rn(__fsub_rn((float) ic, a.corr0), fmaxf(0.001f, __fsub_rn(a.corr1, a.corr0)));
                const float ramp_mix = __fmul_rn(__fsub_rn(1.0f, fminf(1.0f, fmaxf(0.0f, y))), p.ext_factor);
                th = __fadd_rn(__fmul_rn(interp, __fsub_rn(1.0f, ramp_mix)), __fmul_rn(extrap, ramp_mix));
            }
            cs[2*ic]     = __fmul_rn(cosf(th), a.mscale);
            cs[2*ic + 1] = __fmul_rn(sinf(th), a.mscale);
        }
        // (round 2: fetching the NEXT layer's row into registers while this one is rotated and written back was measured and
        // made the launch slower, 0.42 -> 0.54 ms for F16, profiles/r2l_kshift.jsonl; the plain form stays)
        for (int ci = blockIdx.y; ci < a.n; ci += gridDim.y) {
            char *cache = a.caches[ci];
            const char *row = cache + cell*p.nb_token;
            for (int e = threadIdx.x; e < row_elems; e += KVW_THREADS) {
                const int h = e/hd;
                x[e] = k_shift_load<TYPE>(row + h*p.nb_head, e - h*hd);
            }
            __syncthreads();                                             // also: the cos/sin table is complete
            for (int pr = threadIdx.x; pr < p.n_head_kv*half_rot; pr += KVW_THREADS) {
                const int h = pr/half_rot, ic = pr - h*half_rot;
                float *xh = x + h*hd;
                const int i0 = neox ? ic : 2*ic, i1 = neox ? ic + half_rot : 2*ic + 1;
                const float c = cs[2*ic], sn = cs[2*ic + 1], x0 = xh[i0], x1 = xh[i1];
                xh[i0] = __fsub_rn(__fmul_rn(x0, c),  __fmul_rn(x1, sn));
                xh[i1] = __fadd_rn(__fmul_rn(x0, sn), __fmul_rn(x1, c));
            }
            __syncthreads();
            write_row<TYPE>(x, cache, cell, p.n_head_kv, hd, p.nb_token, p.nb_head, stage);
            __syncthreads();
        }
    }
}

// Fast form for head_dim 128 (one warp = one head row, a lane = 4 consecutive elements): the general kernel above is
// instruction-bound, not latency- or byte-bound — per element it pays an integer division for the head, a 2-byte load, a
// shared-memory round trip and two more divisions in the rotation loop, ~300 instructions per thread and layer, which at 8
// CTAs per SM is the whole 0.42 ms of a 32-layer shift (round 1: 0.19 of HBM for F16, the same time for every type).  Here
// a lane loads its 4 elements with one or two vector loads, rotates them in registers (NORM pairs live inside the lane; NEOX
// partners are 64 elements = 16 lanes away: one shuffle each) with its cos / sin kept in registers across the layers, and
// hands them straight to convert4: no F32 row in shared memory, two barriers per layer instead of three, ~60 instructions.
// Same arithmetic sequence as the general kernel (and ggml_compute_forward_rope_f32 / _f16): results are bit-identical.
template <int TYPE>
__device__ __forceinline__ float4 k_shift_load4(const char *head, int i) {          // elements i..i+3 of one head row, i % 4 == 0
    if constexpr (TYPE == B200_TYPE_F32) return *reinterpret_cast<const float4 *>(head + 4*i);
    else if constexpr (TYPE == B200_TYPE_F16) {
        const uint2 w = *reinterpret_cast<const uint2 *>(head + 2*i);
        const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&w.x)), b = __half22float2(*reinterpret_cast<const __half2 *>(&w.y));
        return make_float4(a.x, a.y, b.x, b.y);
    } else if constexpr (TYPE == B200_TYPE_Q8_0) {
        const char *blk = head + (i/QK)*Q8_BYTES;
        const float d = __half2float(*reinterpret_cast<const __half *>(blk));
        const uint32_t lo = ld_u16(blk + 2 + (i % QK)), hi = ld_u16(blk + 4 + (i % QK));
        return make_float4(__fmul_rn((float) (int8_t) (lo & 0xff), d), __fmul_rn((float) (int8_t) (lo >> 8), d),
                           __fmul_rn((float) (int8_t) (hi & 0xff), d), __fmul_rn((float) (int8_t) (hi >> 8), d));
    } else {
        const char *blk = head + (i/QK)*Q4_BYTES;
        const float d = __half2float(*reinterpret_cast<const __half *>(blk));
        const int j = i % QK, sh = j < 16 ? 0 : 4;
        const uint32_t lo = ld_u16(blk + 2 + (j & 15)), hi = ld_u16(blk + 4 + (j & 15));
        auto deq = [&](uint32_t b) { return __fmul_rn((float) ((int) ((b >> sh) & 0xf) - 8), d); };
        return make_float4(deq(lo & 0xff), deq(lo >> 8), deq(hi & 0xff), deq(hi >> 8));
    }
}

template <int TYPE, bool NEOX>
__global__ void __launch_bounds__(KVW_THREADS)
k_shift_fast_kernel(const __grid_constant__ k_shift_args a) {
    extern __shared__ __align__(16) char ks_smem[];
    const b200_k_shift_desc &p = a.d;
    constexpr int HD = 128;
    const int half_rot = p.n_rot/2, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float *cs = reinterpret_cast<float *>(ks_smem);                   // (cos, sin) per rotated pair
    char  *stage = ks_smem + (((size_t) p.n_rot*4 + 15) & ~(size_t) 15);
    const int64_t n_rows = p.cells ? p.n_rows : p.n_cells;
    for (int64_t r = blockIdx.x; r < n_rows; r += gridDim.x) {
        const int64_t cell = p.cells ? (int64_t) p.cells[r] : r;
        if (cell < 0 || cell >= p.n_cells) continue;                 // CTA-uniform
        const float pos = (float) (int64_t) p.shift[cell];
        __syncthreads();                                             // the previous row's readers of cs are done
        for (int ic = threadIdx.x; ic < half_rot; ic += KVW_THREADS) {
            float theta = pos;
            for (int k = 0; k < ic; ++k) theta = __fmul_rn(theta, a.theta_scale);
            const float ff = p.freq_factors ? p.freq_factors[ic] : 1.0f;
            const float extrap = __fdiv_rn(theta, ff);
            const float interp = __fmul_rn(p.freq_scale, extrap);
            float th = interp;
            if (p.ext_factor != 0.0f) {
                const float y = __fdiv_rn(__fsub_rn((float) ic, a.corr0), fmaxf(0.001f, __fsub_rn(a.corr1, a.corr0)));
                const float ramp_mix = __fmul_rn(__fsub_rn(1.0f, fminf(1.0f, fmaxf(0.0f, y))), p.ext_factor);
                th = __fadd_rn(__fmul_rn(interp, __fsub_rn(1.0f, ramp_mix)), __fmul_rn(extrap, ramp_mix));
            }
            cs[2*ic]     = __fmul_rn(cosf(th), a.mscale);
            cs[2*ic + 1] = __fmul_rn(sinf(th), a.mscale);
        }
        __syncthreads();
        // this lane's cos / sin, kept across the layers.  NORM: pairs (4l, 4l+1), (4l+2, 4l+3) -> ic = 2l, 2l+1.
        // NEOX (n_rot == 128): element 4(l%16)+k pairs with the one 64 further on -> ic = 4(l%16)+k for both partners.
        float c[4], sn[4];
        bool rot[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int ic = NEOX ? 4*(lane & 15) + k : 2*lane + (k >> 1);
            rot[k] = ic < half_rot;
            c[k]  = rot[k] ? cs[2*ic] : 1.0f;
            sn[k] = rot[k] ? cs[2*ic + 1] : 0.0f;
        }
        for (int ci = blockIdx.y; ci < a.n; ci += gridDim.y) {
            char *cache = a.caches[ci];
            const char *row = cache + cell*p.nb_token;
            for (int h = warp; h < p.n_head_kv; h += KVW_THREADS/32) {
                float4 x = k_shift_load4<TYPE>(row + h*p.nb_head, 4*lane);
                if constexpr (NEOX) {
                    const float4 y = make_float4(__shfl_xor_sync(0xffffffffu, x.x, 16), __shfl_xor_sync(0xffffffffu, x.y, 16),
                                                 __shfl_xor_sync(0xffffffffu, x.z, 16), __shfl_xor_sync(0xffffffffu, x.w, 16));
                    float xv[4] = {x.x, x.y, x.z, x.w};
                    const float yv[4] = {y.x, y.y, y.z, y.w};
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        if (!rot[k]) continue;
                        xv[k] = lane < 16 ? __fsub_rn(__fmul_rn(xv[k], c[k]), __fmul_rn(yv[k], sn[k]))      // x0' = x0 c - x1 s
                                          : __fadd_rn(__fmul_rn(yv[k], sn[k]), __fmul_rn(xv[k], c[k]));     // x1' = x0 s + x1 c
                    }
                    x = make_float4(xv[0], xv[1], xv[2], xv[3]);
                } else {
                    if (rot[0]) { const float x0 = x.x, x1 = x.y; x.x = __fsub_rn(__fmul_rn(x0, c[0]), __fmul_rn(x1, sn[0])); x.y = __fadd_rn(__fmul_rn(x0, sn[0]), __fmul_rn(x1, c[0])); }
                    if (rot[2]) { const float x0 = x.z, x1 = x.w; x.z = __fsub_rn(__fmul_rn(x0, c[2]), __fmul_rn(x1, sn[2])); x.w = __fadd_rn(__fmul_rn(x0, sn[2]), __fmul_rn(x1, c[2])); }
                }
                convert4<TYPE>(x, HD*h + 4*lane, stage, lane, true);
            }
            __syncthreads();
            store_staged_row<TYPE>(stage, cache, cell, p.n_head_kv, HD, p.nb_token, p.nb_head);
            __syncthreads();                                             // the stage is free again
        }
    }
}

// ggml_rope_yarn_corr_dims, ggml/src/ggml.c:3735-3750 (host side, the same libm calls as the reference)
static void k_shift_corr_dims(int n_dims, int n_ctx_orig, float freq_base, float beta_fast, float beta_slow, float dims[2]) {
    const float pi = 3.14159265358979323846f;
    auto corr_dim = [&](float n_rot) { return n_dims*logf(n_ctx_orig/(n_rot*2*pi))/(2*logf(freq_base)); };
    const float start = floorf(corr_dim(beta_fast)), end = ceilf(corr_dim(beta_slow));
    dims[0] = start > 0 ? start : 0;
    dims[1] = end < n_dims - 1 ? end : (float) (n_dims - 1);
}

int launch_k_shift(const b200_k_shift_desc &d, cudaStream_t stream) {
    const int64_t n_rows = d.cells ? d.n_rows : d.n_cells;
    if (n_rows <= 0 || d.n_caches <= 0) return B200_OK;
    const bool quant = d.type_k == B200_TYPE_Q8_0 || d.type_k == B200_TYPE_Q4_0;
    if (!(quant || d.type_k == B200_TYPE_F16 || d.type_k == B200_TYPE_F32)) { set_error("k_shift: unsupported cache type %d", d.type_k); return B200_ERR_UNSUPPORTED; }
    if (d.mode & ~2) { set_error("k_shift: rope mode %d (only NORM = 0 and NEOX = 2)", d.mode); return B200_ERR_UNSUPPORTED; }
    if (!d.caches || !d.shift || d.n_head_kv <= 0 || d.head_dim <= 0 || (d.head_dim & 1) || d.n_rot <= 0 || (d.n_rot & 1) || d.n_rot > d.head_dim
        || (quant && d.head_dim % QK)) {
        set_error("k_shift: bad arguments (head_dim %d, n_rot %d)", d.head_dim, d.n_rot);
        return B200_ERR_INVALID;
    }
    const int64_t row_elems = (int64_t) d.n_head_kv*d.head_dim;
    const size_t smem = (((size_t) (row_elems + d.n_rot)*4 + 15) & ~(size_t) 15) + (size_t) row_bytes(d.type_k, row_elems);
    if (smem > 48*1024) { set_error("k_shift: token row needs %zu bytes of shared memory (limit 48 KiB)", smem); return B200_ERR_UNSUPPORTED; }
    k_shift_args a;
    a.d = d;
    a.theta_scale = powf(d.freq_base, -2.0f/d.n_rot);
    float corr[2];
    k_shift_corr_dims(d.n_rot, d.n_ctx_orig, d.freq_base, d.beta_fast, d.beta_slow, corr);
    a.corr0 = corr[0]; a.corr1 = corr[1];
    a.mscale = d.attn_factor;
    if (d.ext_factor != 0.0f) a.mscale *= 1.0f + 0.1f*logf(1.0f/d.freq_scale);
    const unsigned gx = (unsigned) (n_rows < 148*64 ? n_rows : 148*64);
    for (int c0 = 0; c0 < d.n_caches; c0 += BC_MAX_CACHES) {
        const int n = d.n_caches - c0 < BC_MAX_CACHES ? d.n_caches - c0 : BC_MAX_CACHES;
        for (int i = 0; i < n; ++i) a.caches[i] = static_cast<char *>(d.caches[c0 + i]);
        a.n = n;
        // enough CTAs to fill the GPU (~16 per SM), as few layer groups as that allows: the table is rebuilt per group
        int gy = (int) ((148*16 + (int64_t) gx - 1)/gx);
        if (gy > n) gy = n;
        if (gy < 1) gy = 1;
        const dim3 grid(gx, (unsigned) gy);
        // head_dim 128 with aligned heads: the register form (NEOX needs its partners 16 lanes away: n_rot == 128)
        const bool neox = (d.mode & 2) != 0;
        const int64_t hb = row_bytes(d.type_k, 128);
        const bool aligned = d.type_k == B200_TYPE_F32 ? ((d.nb_head | d.nb_token) % 16 == 0 && ((uintptr_t) a.caches[0] & 15) == 0)
                           : d.type_k == B200_TYPE_F16 ? ((d.nb_head | d.nb_token) % 8 == 0 && ((uintptr_t) a.caches[0] & 7) == 0)
                                                       : ((d.nb_head | d.nb_token) % 2 == 0 && ((uintptr_t) a.caches[0] & 1) == 0);
        bool all_aligned = aligned;
        for (int i = 0; i < n && all_aligned; ++i) all_aligned = ((uintptr_t) a.caches[i] & 15) == 0;
        if (d.head_dim == 128 && d.nb_head >= hb && all_aligned && d.n_rot % 4 == 0 && (!neox || d.n_rot == 128) && !getenv("B200_KSHIFT_GENERAL")) {
            const size_t smem_fast = (((size_t) d.n_rot*4 + 15) & ~(size_t) 15) + (size_t) row_bytes(d.type_k, row_elems);
#define KS_FAST(T) do { if (neox) k_shift_fast_kernel<T, true><<<grid, KVW_THREADS, smem_fast, stream>>>(a); \
                        else      k_shift_fast_kernel<T, false><<<grid, KVW_THREADS, smem_fast, stream>>>(a); } while (0)
            switch (d.type_k) {
                case B200_TYPE_F32:  KS_FAST(B200_TYPE_F32);  break;
                case B200_TYPE_F16:  KS_FAST(B200_TYPE_F16);  break;
                case B200_TYPE_Q8_0: KS_FAST(B200_TYPE_Q8_0); break;
                default:             KS_FAST(B200_TYPE_Q4_0); break;
            }
#undef KS_FAST
            count_launch();
            if (!check_launch("k_shift")) return B200_ERR_CUDA;
            continue;
        }
        switch (d.type_k) {
            case B200_TYPE_F32:  k_shift_kernel<B200_TYPE_F32 ><<<grid, KVW_THREADS, smem, stream>>>(a); break;
            case B200_TYPE_F16:  k_shift_kernel<B200_TYPE_F16 ><<<grid, KVW_THREADS, smem, stream>>>(a); break;
            case B200_TYPE_Q8_0: k_shift_kernel<B200_TYPE_Q8_0><<<grid, KVW_THREADS, smem, stream>>>(a); break;
            default:             k_shift_kernel<B200_TYPE_Q4_0><<<grid, KVW_THREADS, smem, stream>>>(a); break;
        }
        count_launch();
        if (!check_launch("k_shift")) return B200_ERR_CUDA;
    }
    return B200_OK;
}

// ---------------------------------------------------------------------------------------------
// GGML_OP_CPY in general (ggml_cuda_cpy, ggml/src/ggml-cuda/cpy.cu:510-584; cases of tests/test-backend-ops.cpp:3992-4009):
// strided / permuted sources, F16 sources, quantised sources to F32.  One CTA per row of ne0 elements: the row is brought
// to F32 in shared memory (element-wise through the source strides, so a source that is not contiguous along dim 0 works
// too), then leaves through the KV-write path (same converters, staged, 16-byte stores).  Row r of the source, in its
// logical order (i1, i2, i3), goes to row r of the destination: contiguous, or with the source's shape and its own strides.
// ---------------------------------------------------------------------------------------------
struct cpy_rows_args {
    const char *src; char *dst;
    int type_src, type_dst;
    int64_t ne0, ne1, ne2, ne3;
    int64_t snb0, snb1, snb2, snb3;          // source strides in bytes (snb0 == element size: the row is contiguous)
    int64_t dnb1, dnb2, dnb3;                // destination row strides in bytes (rows themselves are contiguous)
};

__global__ void __launch_bounds__(KVW_THREADS)
cpy_rows_kernel(const __grid_constant__ cpy_rows_args a) {
    extern __shared__ __align__(16) char cr_smem[];
    float *x = reinterpret_cast<float *>(cr_smem);
    char *stage = cr_smem + a.ne0*4;
    const int64_t r = blockIdx.x;
    const int64_t i3 = r/(a.ne1*a.ne2), i2 = (r - i3*a.ne1*a.ne2)/a.ne1, i1 = r - i3*a.ne1*a.ne2 - i2*a.ne1;
    const char *srow = a.src + i1*a.snb1 + i2*a.snb2 + i3*a.snb3;
    for (int e = threadIdx.x; e < a.ne0; e += KVW_THREADS) {
        float v;
        switch (a.type_src) {
            case B200_TYPE_F32:  v = *reinterpret_cast<const float *>(srow + e*a.snb0); break;
            case B200_TYPE_F16:  v = __half2float(*reinterpret_cast<const __half *>(srow + e*a.snb0)); break;
            case B200_TYPE_Q8_0: v = k_shift_load<B200_TYPE_Q8_0>(srow, e); break;
            default:             v = k_shift_load<B200_TYPE_Q4_0>(srow, e); break;
        }
        x[e] = v;
    }
    __syncthreads();
    char *drow = a.dst + i1*a.dnb1 + i2*a.dnb2 + i3*a.dnb3;
    const int hd = (int) a.ne0;
    switch (a.type_dst) {
        case B200_TYPE_F32:  write_row<B200_TYPE_F32 >(x, drow, 0, 1, hd, 0, row_bytes(B200_TYPE_F32,  hd), stage); break;
        case B200_TYPE_F16:  write_row<B200_TYPE_F16 >(x, drow, 0, 1, hd, 0, row_bytes(B200_TYPE_F16,  hd), stage); break;
        case B200_TYPE_Q8_0: write_row<B200_TYPE_Q8_0>(x, drow, 0, 1, hd, 0, row_bytes(B200_TYPE_Q8_0, hd), stage); break;
        default:             write_row<B200_TYPE_Q4_0>(x, drow, 0, 1, hd, 0, row_bytes(B200_TYPE_Q4_0, hd), stage); break;
    }
}

int launch_cpy_rows(const void *src, void *dst, int type_src, int type_dst, const int64_t ne[4], const int64_t snb[4],
                    const int64_t dnb[4], cudaStream_t stream) {
    cpy_rows_args a;
    a.src = static_cast<const char *>(src); a.dst = static_cast<char *>(dst); a.type_src = type_src; a.type_dst = type_dst;
    a.ne0 = ne[0]; a.ne1 = ne[1]; a.ne2 = ne[2]; a.ne3 = ne[3];
    a.snb0 = snb[0]; a.snb1 = snb[1]; a.snb2 = snb[2]; a.snb3 = snb[3];
    a.dnb1 = dnb[1]; a.dnb2 = dnb[2]; a.dnb3 = dnb[3];
    const int64_t rows = ne[1]*ne[2]*ne[3];
    if (rows <= 0 || ne[0] <= 0) return B200_OK;
    const size_t smem = (size_t) ne[0]*4 + (size_t) row_bytes(type_dst, ne[0]);
    if (smem > 48*1024 || rows >= (1LL << 31)) { set_error("cpy: row of %lld elements / %lld rows not supported", (long long) ne[0], (long long) rows); return B200_ERR_UNSUPPORTED; }
    cpy_rows_kernel<<<(unsigned) rows, KVW_THREADS, smem, stream>>>(a);
    count_launch();
    return check_launch("cpy_rows") ? B200_OK : B200_ERR_CUDA;
}

} // namespace b200
