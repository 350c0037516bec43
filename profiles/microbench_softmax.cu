// microbench_softmax.cu — per-warp issue cost of the instructions the prefill softmax warps are made of
// (one warp per SM sub-partition unless stated).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mb microbench_softmax.cu
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>

__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ void add2(float &d0, float &d1, float a0, float a1, float b0, float b1) {
    asm volatile("{\n .reg .b64 a, b, d;\n mov.b64 a, {%2, %3};\n mov.b64 b, {%4, %5};\n add.rn.f32x2 d, a, b;\n mov.b64 {%0, %1}, d;\n}"
        : "=f"(d0), "=f"(d1) : "f"(a0), "f"(a1), "f"(b0), "f"(b1));
}
__device__ __forceinline__ uint32_t pack(float lo, float hi) { uint32_t r; asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo)); return r; }
__device__ __forceinline__ float max3(float a, float b, float c) { float d; asm volatile("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }

template <int MODE>
__global__ void k(float *out, long long *cyc, float seed) {
    float v[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) v[i] = seed*(i + threadIdx.x)*1e-3f - 3.0f;
    uint32_t pk[32];
    float s0 = 0, s1 = 0, s2 = 0, s3 = 0, mx = -1e30f;
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < 64; ++it) {
        if (MODE == 0) {            // 64 independent ex2
#pragma unroll
            for (int i = 0; i < 64; ++i) v[i] = ex2(v[i]);
        } else if (MODE == 1) {     // 32 packs
#pragma unroll
            for (int i = 0; i < 32; ++i) { pk[i] = pack(v[2*i], v[2*i + 1]); v[2*i] = __uint_as_float(pk[i]); }
        } else if (MODE == 2) {     // 32 add2 (independent)
#pragma unroll
            for (int i = 0; i < 32; ++i) add2(v[2*i], v[2*i + 1], v[2*i], v[2*i + 1], seed, seed);
        } else if (MODE == 3) {     // 32 max3, 4 chains
            float m0 = mx, m1 = mx, m2 = mx, m3 = mx;
#pragma unroll
            for (int i = 0; i < 32; i += 4) { m0 = max3(m0, v[2*i], v[2*i+1]); m1 = max3(m1, v[2*i+2], v[2*i+3]); m2 = max3(m2, v[2*i+4], v[2*i+5]); m3 = max3(m3, v[2*i+6], v[2*i+7]); }
            mx = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3)); v[0] += mx*1e-9f;
        } else if (MODE == 4) {     // the softmax tile body: sub, ex2, sum, pack
#pragma unroll
            for (int c = 0; c < 32; ++c) {
                float x0, x1;
                add2(x0, x1, v[2*c], v[2*c + 1], seed, seed);
                const float p0 = ex2(x0), p1 = ex2(x1);
                if (c & 1) add2(s2, s3, s2, s3, p0, p1); else add2(s0, s1, s0, s1, p0, p1);
                pk[c] = pack(p0, p1);
                v[2*c] = __uint_as_float(pk[c]);
            }
        } else if (MODE == 6) {     // 64 fmaxf, 2 chains
            float m0 = mx, m1 = mx;
#pragma unroll
            for (int i = 0; i < 32; ++i) { asm volatile("max.f32 %0, %0, %1;" : "+f"(m0) : "f"(v[2*i])); asm volatile("max.f32 %0, %0, %1;" : "+f"(m1) : "f"(v[2*i+1])); }
            mx = fmaxf(m0, m1); v[0] += mx*1e-9f;
        } else if (MODE == 7) {     // 32 fma.f32x2 (independent)
#pragma unroll
            for (int i = 0; i < 32; ++i)
                asm volatile("{\n .reg .b64 a, b;\n mov.b64 a, {%0, %1};\n mov.b64 b, {%2, %2};\n fma.rn.f32x2 a, a, b, b;\n mov.b64 {%0, %1}, a;\n}" : "+f"(v[2*i]), "+f"(v[2*i+1]) : "f"(seed));
        } else if (MODE == 8) {     // 64 scalar fma (independent)
#pragma unroll
            for (int i = 0; i < 64; ++i) asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(v[i]) : "f"(seed));
        } else if (MODE == 9) {     // softmax body, compiler-scheduled (non-volatile asm)
#pragma unroll
            for (int c = 0; c < 32; ++c) {
                float x0 = v[2*c] + seed, x1 = v[2*c+1] + seed, p0, p1;
                asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p0) : "f"(x0)); asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p1) : "f"(x1));
                if (c & 1) { s2 += p0; s3 += p1; } else { s0 += p0; s1 += p1; }
                asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(pk[c]) : "f"(p1), "f"(p0));
                v[2*c] = __uint_as_float(pk[c]) ; v[2*c+1] = x1*0.5f;
            }
        } else if (MODE == 5) {     // 32 serial max3
#pragma unroll
            for (int i = 0; i < 32; ++i) mx = max3(mx, v[2*i], v[2*i+1]);
            v[0] += mx*1e-9f;
        }
    }
    const long long t1 = clock64();
    float acc = s0 + s1 + s2 + s3 + mx;
#pragma unroll
    for (int i = 0; i < 64; ++i) acc += v[i];
    out[blockIdx.x*blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = (t1 - t0)/64;
}

int main() {
    float *out; long long *cyc, h;
    cudaMalloc(&out, 148*1024*4); cudaMalloc(&cyc, 8);
    const char *names[] = {"64 ex2", "32 cvt.f16x2 packs", "32 add.f32x2", "32 max3 (4 chains)", "softmax body (32 pairs)", "32 max3 (serial)", "64 max.f32 (2 chains)", "32 fma.f32x2", "64 fma.f32", "softmax body, scheduled"};
    for (int threads = 128; threads <= 512; threads *= 2) {
        printf("threads/SM = %d (%d warps per sub-partition)\n", threads, threads/128);
#define RUN(M) k<M><<<148, threads>>>(out, cyc, 1.0f); cudaDeviceSynchronize(); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); printf("  %-26s %5lld cycles per iteration\n", names[M], h);
        RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5) RUN(6) RUN(7) RUN(8) RUN(9)
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
