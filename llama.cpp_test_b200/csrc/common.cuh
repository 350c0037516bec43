// common.cuh — shared device helpers for the sm_100a paged-attention kernels.
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <utility>

#include "b200_paged_attn.h"

namespace b200 {

constexpr int QK       = 32;   // elements per quant block (QK4_0 == QK8_0, ggml/src/ggml-common.h:161,203)
constexpr int Q8_BYTES = 34;   // block_q8_0: half d + 32 int8
constexpr int Q4_BYTES = 18;   // block_q4_0: half d + 16 nibble bytes

__host__ __device__ constexpr int64_t row_bytes(int type, int64_t n) {
    return type == B200_TYPE_F32  ? n*4
         : type == B200_TYPE_F16  ? n*2
         : type == B200_TYPE_Q8_0 ? (n/QK)*Q8_BYTES
         : type == B200_TYPE_Q4_0 ? (n/QK)*Q4_BYTES : 0;
}

__host__ __device__ constexpr bool type_ok(int type) {
    return type == B200_TYPE_F32 || type == B200_TYPE_F16 || type == B200_TYPE_Q8_0 || type == B200_TYPE_Q4_0;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__device__ __forceinline__ uint32_t ld_u16(const void *p) { return *reinterpret_cast<const uint16_t *>(p); }
__device__ __forceinline__ float    h2f(uint32_t bits)    { return __half2float(__ushort_as_half((unsigned short) bits)); }

// Load N (2 or 4 or 8) consecutive elements e0..e0+N-1 of one head row of `TYPE` as floats.
// e0 is a multiple of N, so the group never straddles a quant block (N <= 8 divides 16 and 32).
// Dequantisation follows dequantize_row_q8_0 / dequantize_row_q4_0 (ggml/src/ggml-quants.c:349-362,
// :255-273): y = q*d with d the block's half scale; Q4_0 element j<16 is the low nibble of qs[j],
// element j>=16 the high nibble of qs[j-16], both minus 8.
template <int TYPE, int N>
__device__ __forceinline__ void load_row_elems(const char *row, int e0, float (&out)[N]) {
    if constexpr (TYPE == B200_TYPE_F32) {
        const float *p = reinterpret_cast<const float *>(row) + e0;
#pragma unroll
        for (int i = 0; i < N; ++i) out[i] = p[i];
    } else if constexpr (TYPE == B200_TYPE_F16) {
        const __half2 *p = reinterpret_cast<const __half2 *>(row + 2*e0);
#pragma unroll
        for (int i = 0; i < N/2; ++i) { const float2 f = __half22float2(p[i]); out[2*i] = f.x; out[2*i+1] = f.y; }
    } else if constexpr (TYPE == B200_TYPE_Q8_0) {
        const char *blk = row + (e0/QK)*Q8_BYTES;
        const float d = h2f(ld_u16(blk));
        const char *qs = blk + 2 + (e0 % QK);            // 2-byte aligned
#pragma unroll
        for (int i = 0; i < N/2; ++i) {
            const uint32_t w = ld_u16(qs + 2*i);
            out[2*i]   = (float) (int8_t) (w & 0xff) * d;
            out[2*i+1] = (float) (int8_t) (w >> 8)   * d;
        }
    } else {                                             // Q4_0
        const char *blk = row + (e0/QK)*Q4_BYTES;
        const float d = h2f(ld_u16(blk));
        const int j = e0 % QK;
        const char *qs = blk + 2 + (j & 15);
        const int sh = (j & 16) ? 4 : 0;
#pragma unroll
        for (int i = 0; i < N/2; ++i) {
            const uint32_t w = ld_u16(qs + 2*i);
            out[2*i]   = (float) ((int) ((w >> sh)       & 0xf) - 8) * d;
            out[2*i+1] = (float) ((int) ((w >> (8 + sh)) & 0xf) - 8) * d;
        }
    }
}

// Programmatic dependent launch: every kernel of the decode chain (kv_write -> plan -> main -> combine -> next layer's
// kv_write ...) lets its successor be scheduled at once (launch_dependents) and waits for its predecessor's memory
// before touching global memory (wait), so launch latency and CTA ramp-up hide under the predecessor's tail instead of
// adding ~1-2 us per kernel boundary to a 20 us layer-op (config 1).  Without the launch attribute both are no-ops.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait()    { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_prologue() {
    pdl_trigger();
    pdl_wait();
}
// The other order, for a kernel whose SUCCESSOR starts work before its own wait (kv_write in front of the stream-K
// main kernel, attn_streamk.cu "store overlap"): the successor is scheduled only once this kernel's predecessor has
// completed, so everything two or more launches back is finished and visible by the time the successor's first
// instruction runs.
__device__ __forceinline__ void pdl_prologue_strict() {
    pdl_wait();
    pdl_trigger();
}

bool pdl_enabled();                       // shim.cu: false when B200_NO_PDL=1

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args &&...args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(std::forward<Args>(args))...);
}

// -DB200_TIMELINE (diagnosis builds only, profiles/run_timeline.sh; never the product library): thread 0 of the first
// and last CTA of every kernel of the decode chain appends one record of %globaltimer stamps to a device buffer that
// shim.cu writes out at exit.  Record: {kernel id | block << 8, entry, s2, s3, s4, s5, end, extra}.
#ifdef B200_TIMELINE
unsigned long long *timeline_buffer();                        // shim.cu: [0] = records so far, records from [8]
constexpr unsigned long long TL_MAX = 1u << 16;
static __device__ unsigned long long *tl_buf;                 // one per translation unit, armed by tl_arm()
static inline void tl_arm() {
    static bool done = false;
    if (!done) { unsigned long long *p = timeline_buffer(); cudaMemcpyToSymbol(tl_buf, &p, sizeof p); done = true; }
}
struct tl_rec {
    unsigned long long v[8]; bool on;
    __device__ tl_rec(int id, bool thread_on) {
        on = thread_on && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1) && blockIdx.y == 0;
        for (int i = 0; i < 8; ++i) v[i] = 0;
        v[0] = (unsigned long long) id | ((unsigned long long) blockIdx.x << 8);
        stamp(1);
    }
    __device__ void stamp(int i) { if (on) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); v[i] = t; } }
    __device__ void stamp_once(int i) { if (on && v[i] == 0) stamp(i); }
    __device__ void flush() {
        if (!on) return;
        stamp(6);
        const unsigned long long k = atomicAdd(tl_buf, 1ULL);
        if (k < TL_MAX) for (int i = 0; i < 8; ++i) tl_buf[8 + 8*k + i] = v[i];
        on = false;
    }
};
#define TL_BEGIN(id, thread_on) tl_rec tl_((id), (thread_on))
#define TL_STAMP(i)      tl_.stamp(i)
#define TL_STAMP_ONCE(i) tl_.stamp_once(i)
#define TL_EXTRA(x)      (tl_.v[7] = (unsigned long long) (x))
#define TL_END()         tl_.flush()
#define TL_ARM()         tl_arm()
#else
#define TL_BEGIN(id, thread_on) do { } while (0)
#define TL_STAMP(i)      do { } while (0)
#define TL_STAMP_ONCE(i) do { } while (0)
#define TL_EXTRA(x)      do { } while (0)
#define TL_END()         do { } while (0)
#define TL_ARM()         do { } while (0)
#endif

// process-wide bookkeeping (defined in shim.cu)
void set_error(const char *fmt, ...);
void count_launch(int n = 1);
bool check_launch(const char *what);

} // namespace b200
