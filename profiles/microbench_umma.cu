// microbench_umma.cu — tensor-pipe time of the prefill kernel's MMA blocks on one SM (one CTA per SM, one issuing warp):
// is S = Q K^T with both operands in shared memory (SS) bound by the operand reads?  Operand contents are irrelevant
// (zeros); descriptors are the prefill kernel's.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mbu microbench_umma.cu
#include <cstdio>
#include <cstdint>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t) ((addr & 0x3FFFFu) >> 4) | ((uint64_t) (lbo >> 4) << 16) | ((uint64_t) (sbo >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
__host__ __device__ constexpr uint32_t idesc(int m, int n, int b_mn) {
    return (1u << 4) | ((uint32_t) b_mn << 16) | ((uint32_t) (n >> 3) << 17) | ((uint32_t) (m >> 4) << 24);
}
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t id, uint32_t acc) {
    asm volatile("{\n .reg .pred p, e;\n elect.sync _|e, 0xffffffff;\n setp.ne.b32 p, %4, 0;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
                 :: "r"(d), "l"(a), "l"(b), "r"(id), "r"(acc) : "memory");
}
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t id, uint32_t acc) {
    asm volatile("{\n .reg .pred p, e;\n elect.sync _|e, 0xffffffff;\n setp.ne.b32 p, %4, 0;\n @e tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n}"
                 :: "r"(d), "r"(a), "l"(b), "r"(id), "r"(acc) : "memory");
}
__device__ __forceinline__ void commit(uint32_t bar) {
    asm volatile("{\n .reg .pred e;\n elect.sync _|e, 0xffffffff;\n @e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n}" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile("{\n .reg .pred p;\nW_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra D_%=;\n bra W_%=;\nD_%=:\n}" :: "r"(bar), "r"(parity) : "memory");
}

// MODE 0: QK SS N=64 (8 K16 steps)   1: QK SS N=128   2: QK TS N=64 (A = Q in TMEM)   3: PV TS N=128 (4 steps, V MN-major)
// 4: the kernel's mix per kv tile: 2x(QK SS N=64) + 2x(PV TS)    5: 2x(QK TS N=64) + 2x(PV TS)   6: QK SS N=256
template <int MODE>
__global__ void __launch_bounds__(32, 1) k(long long *cyc, int iters) {
    extern __shared__ __align__(1024) unsigned char dyn[];
    __shared__ uint32_t tmem_slot;
    __shared__ __align__(8) unsigned long long bar;
    const uint32_t base = (smem_u32(dyn) + 1023) & ~1023u;
    const uint32_t sQ = base, sK = base + 65536, sV = sK + 65536, sBar = smem_u32(&bar);
    for (int i = threadIdx.x; i < (65536*3)/4; i += 32) reinterpret_cast<uint32_t *>(dyn + (base - smem_u32(dyn)))[i] = 0;
    if (threadIdx.x == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(sBar) : "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(&tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_slot;
    auto qk_ss = [&](uint32_t d, uint32_t q, uint32_t kk, int n) {
        const uint32_t id = idesc(128, n, 0);
        const uint64_t a0 = desc(q, 16, 1024), b0 = desc(kk, 16, 1024);
        for (int s = 0; s < 8; ++s) mma_ss(d, a0 + (s & 3)*2 + (s >> 2)*1024, b0 + (s & 3)*2 + (s >> 2)*(n*128/16), id, s > 0);
    };
    auto qk_ts = [&](uint32_t d, uint32_t qt, uint32_t kk, int n) {
        const uint32_t id = idesc(128, n, 0);
        const uint64_t b0 = desc(kk, 16, 1024);
        for (int s = 0; s < 8; ++s) mma_ts(d, qt + 8*s, b0 + (s & 3)*2 + (s >> 2)*(n*128/16), id, s > 0);
    };
    auto pv_ts = [&](uint32_t d, uint32_t pt, uint32_t vv) {
        const uint32_t id = idesc(128, 128, 1);
        const uint64_t b0 = desc(vv, 8192, 1024);
        for (int s = 0; s < 4; ++s) mma_ts(d, pt + 8*s, b0 + 128*s, id, 1);
    };
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) qk_ss(tmem + (it & 1)*64, sQ, sK + (it & 3)*16384, 64);
        if (MODE == 1) qk_ss(tmem + (it & 1)*128, sQ, sK + (it & 1)*32768, 128);
        if (MODE == 2) qk_ts(tmem + (it & 1)*64, tmem + 384, sK + (it & 3)*16384, 64);
        if (MODE == 3) pv_ts(tmem + 128, tmem + (it & 1)*64, sV + (it & 3)*16384);
        if (MODE == 4) { qk_ss(tmem + (it & 1)*64, sQ, sK + (it & 3)*16384, 64); pv_ts(tmem + 128, tmem + ((it + 1) & 1)*64, sV + (it & 3)*16384);
                         qk_ss(tmem + 256 + (it & 1)*64, sQ + 32768, sK + (it & 3)*16384, 64); pv_ts(tmem + 384, tmem + 256 + ((it + 1) & 1)*64, sV + (it & 3)*16384); }
        if (MODE == 5) { qk_ts(tmem + (it & 1)*64, tmem + 448, sK + (it & 3)*16384, 64); pv_ts(tmem + 128, tmem + ((it + 1) & 1)*64, sV + (it & 3)*16384);
                         qk_ts(tmem + 256 + (it & 1)*64, tmem + 448, sK + (it & 3)*16384, 64); pv_ts(tmem + 320, tmem + 256 + ((it + 1) & 1)*64, sV + (it & 3)*16384); }
        if (MODE == 6) qk_ss(tmem, sQ, sK, 256);
    }
    commit(sBar);
    mbar_wait(sBar, 0);
    const long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tmem) : "memory");
}

int main() {
    long long *cyc, h;
    cudaMalloc(&cyc, 8);
    const int iters = 2000, smem = 3*65536 + 1024;
    const char *names[] = {"QK SS  M128 N64  K128 (8 MMAs)", "QK SS  M128 N128 K128 (8 MMAs)", "QK TS  M128 N64  K128 (Q in TMEM)", "PV TS  M128 N128 K64  (4 MMAs)",
                           "kv tile: 2x QK SS N64 + 2x PV TS", "kv tile: 2x QK TS N64 + 2x PV TS", "QK SS  M128 N256 K128 (8 MMAs)"};
#define RUN(M) cudaFuncSetAttribute(k<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem); k<M><<<148, 32, smem>>>(cyc, iters); \
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("mode %d: %s\n", M, cudaGetErrorString(cudaGetLastError())); return 1; } \
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); printf("  %-36s %7.1f cycles per block\n", names[M], (double) h/iters);
    RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5) RUN(6)
    return 0;
}
