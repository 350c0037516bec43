// peer_gather.cu — the one exchange step of the KV-head-sharded path (partition B) over NVLink peer
// memory instead of a library collective.
//
// Every rank g of G computes the query heads of its own kv heads, out_g [n_q][n_head/G][d] F32, and every rank needs
// all of them (the wo matmul that follows consumes [n_q][n_head][d]).  The payload is tiny (config 4: 128 KiB per
// rank per layer, 1 MiB in total), so the cost of an all-gather is latency: a library call costs ~25 us per layer
// against ~82 us of attention.  Here each rank PUSHES its slice straight into every peer's gathered buffer with plain
// 16-byte stores through peer-mapped pointers (all peers are one NVSwitch hop away), publishes an epoch flag per
// peer with a system-scope release, and a one-warp kernel waits for the G flags with acquire loads:
//
//     b200_peer_publish   gathered_p[slot][rank] = local   for every peer p (its own copy included);   flag_p[rank] = epoch
//     b200_peer_wait      spin until flag_own[r] >= epoch for every r                          (bounded: ~2 s, then error)
//
// Both kernels take the epoch from a device counter the publish kernel advances itself, so a whole multi-layer step
// is CUDA-graph capturable and replays stay in step across ranks.  The caller names the destination slot (e.g. the
// layer index) and must not use the same slot for two consecutive exchanges: a rank can only start exchange e+2 after
// it saw every peer's flag e+1, which each peer raised after (in stream order) consuming the slot of exchange e.
//
// The reference has no counterpart: its multi-GPU path splits matmul rows and moves them with cudaMemcpyPeerAsync
// (ggml/src/ggml-cuda/ggml-cuda.cu:598,1592); attention is never sharded there.
//
// Memory comes from cudaMalloc + CUDA IPC handles (the processes are one per GPU); handles are exchanged by the host
// side over torch.distributed (plumbing), see llama.cpp_test_b200/sharding.py PeerGather.
#include "common.cuh"

namespace b200 {

constexpr int PG_MAX_WORLD = 16;

struct pg_args {
    const uint4 *local;                  // this rank's slice
    int64_t n16;                         // slice size in 16-byte units
    char     *peer_buf[PG_MAX_WORLD];    // peer p's gathered buffer: [n_slots][world][slice]
    uint32_t *peer_flag[PG_MAX_WORLD];   // peer p's flags: [world] (slot r written by rank r)
    uint32_t *state;                     // local: [0] epoch of the last publish, [1] CTAs done, [2] wait error flag
    int rank, world, slot;
};

// WAIT: the CTA that raises the flags also waits for the peers' flags, so publish + wait is ONE launch (a separate wait
// kernel costs a full launch boundary, ~4 us of a ~100 us layer at G = 8)
template <bool WAIT>
__global__ void __launch_bounds__(256)
pg_publish_kernel(const __grid_constant__ pg_args a) {
    // blockIdx.y = destination peer; blockIdx.x strides over the slice
    pdl_prologue();
    const uint32_t epoch = a.state[0] + 1;                               // every CTA reads the pre-launch value
    const int p = blockIdx.y;
    uint4 *dst = reinterpret_cast<uint4 *>(a.peer_buf[p]) + ((int64_t) a.slot*a.world + a.rank)*a.n16;
    for (int64_t i = (int64_t) blockIdx.x*256 + threadIdx.x; i < a.n16; i += (int64_t) gridDim.x*256) dst[i] = a.local[i];
    __shared__ int s_last;
    __threadfence_system();                                              // this thread's peer stores are visible system-wide ...
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&a.state[1], 1u) == gridDim.x*gridDim.y - 1;
    __syncthreads();
    if (!s_last) return;
    // ... before the last CTA of the grid raises the flags: lane q tells peer q (fence + relaxed store = release; the
    // eight remote stores go out in parallel instead of eight serial release stores)
    if ((int) threadIdx.x < a.world) {
        __threadfence_system();
        asm volatile("st.relaxed.sys.global.u32 [%0], %1;" :: "l"(a.peer_flag[threadIdx.x] + a.rank), "r"(epoch) : "memory");
    }
    if (threadIdx.x == 0) { a.state[1] = 0; a.state[0] = epoch; }
    if (WAIT && (int) threadIdx.x < a.world) {                           // and waits for theirs
        const uint32_t *flags = a.peer_flag[a.rank];
        const long long t0 = clock64();
        uint32_t v;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + threadIdx.x) : "memory");
            if ((int32_t) (v - epoch) >= 0) break;
            if (clock64() - t0 > 4000000000LL) { a.state[2] = 1 + threadIdx.x; break; }
        } while (true);
    }
}

__global__ void __launch_bounds__(32)
pg_wait_kernel(uint32_t *flags, uint32_t *state, int world) {
    const uint32_t epoch = state[0];                                     // the publish before us (stream order) set it
    if ((int) threadIdx.x < world) {
        const long long t0 = clock64();
        uint32_t v;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + threadIdx.x) : "memory");
            if ((int32_t) (v - epoch) >= 0) break;
            if (clock64() - t0 > 4000000000LL) { state[2] = 1 + threadIdx.x; break; }      // ~2 s: a peer died; do not hang the GPU
        } while (true);
    }
}

int launch_peer_publish(const void *local, int64_t bytes, void *const *peer_bufs, void *const *peer_flags, void *state,
                        int slot, int rank, int world, bool wait, cudaStream_t stream) {
    if (slot < 0 || world < 1 || world > PG_MAX_WORLD || rank < 0 || rank >= world || bytes <= 0 || (bytes & 15) || (((uintptr_t) local) & 15)) {
        set_error("peer_publish: bad arguments (world %d, rank %d, bytes %lld)", world, rank, (long long) bytes);
        return B200_ERR_INVALID;
    }
    pg_args a;
    a.local = static_cast<const uint4 *>(local); a.n16 = bytes/16; a.state = static_cast<uint32_t *>(state); a.rank = rank; a.world = world; a.slot = slot;
    for (int p = 0; p < world; ++p) {
        a.peer_buf[p] = static_cast<char *>(peer_bufs[p]); a.peer_flag[p] = static_cast<uint32_t *>(peer_flags[p]);
        if (!a.peer_buf[p] || !a.peer_flag[p] || (((uintptr_t) a.peer_buf[p]) & 15)) { set_error("peer_publish: peer %d not mapped", p); return B200_ERR_INVALID; }
    }
    int64_t gx = (a.n16 + 256*4 - 1)/(256*4);                            // ~4 stores per thread
    if (gx < 1) gx = 1;
    if (gx > 16) gx = 16;                                                // 16 x world CTAs: latency-bound, not bandwidth-bound
    if (wait) launch_pdl(pg_publish_kernel<true>,  dim3((unsigned) gx, (unsigned) world), dim3(256), 0, stream, a);
    else      launch_pdl(pg_publish_kernel<false>, dim3((unsigned) gx, (unsigned) world), dim3(256), 0, stream, a);
    count_launch();
    return check_launch("peer_publish") ? B200_OK : B200_ERR_CUDA;
}

// ------------------------------------------------------------------------------------------------
// low-latency form: every 8-byte store carries 4 bytes of payload and the epoch, so the receiver needs no separate
// flag and the sender no fence — the latency is one NVLink flight plus a poll.  (8-byte aligned vector stores are
// single transactions; the same property NCCL's LL protocol rests on.)  Twice the bytes on the wire, which for a
// 128 KiB slice is still ~2 us.  One launch does both directions: push my slice to every peer's LL buffer, then
// poll the peers' words in my LL buffer and unpack them into the gathered buffer (my own slice is copied directly).
// ------------------------------------------------------------------------------------------------
struct pg_ll_args {
    const uint32_t *local;               // this rank's slice
    int64_t n32;                         // slice size in 4-byte words
    uint2    *peer_ll[PG_MAX_WORLD];     // peer p's LL buffer: [n_slots][world][n32] of (payload, epoch)
    uint32_t *gathered;                  // local: [n_slots][world][n32]
    uint32_t *state;                     // local: [0] epoch, [1] CTAs done, [2] wait error flag
    int rank, world, slot;
};

__global__ void __launch_bounds__(256)
pg_ll_kernel(const __grid_constant__ pg_ll_args a) {
    pdl_prologue();
    const uint32_t epoch = a.state[0] + 1;
    const int64_t stride = (int64_t) gridDim.x*256, t = (int64_t) blockIdx.x*256 + threadIdx.x;
    const int64_t slot_off = (int64_t) a.slot*a.world;
    // push: word w of my slice -> LL_p[slot][rank][w] for every other rank p; my own copy goes straight to gathered
    for (int64_t w = t; w < a.n32; w += stride) {
        const uint32_t x = a.local[w];
        a.gathered[(slot_off + a.rank)*a.n32 + w] = x;
        for (int d = 1; d < a.world; ++d) {
            const int p = (a.rank + d) % a.world;                        // stagger the destinations
            asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" :: "l"(a.peer_ll[p] + (slot_off + a.rank)*a.n32 + w), "r"(x), "r"(epoch) : "memory");
        }
    }
    // pull: poll the peers' words in my own LL buffer.  The (peer, word) pairs are one flat index space strided over the grid
    // and taken EIGHT AT A TIME with independent loads: a system-scope load is a ~0.7 us round trip, and polling the pairs one
    // after another (the first version: 7 peers x 4 words per thread, each load behind the previous one's epoch check) cost
    // 17 us per exchange at 8 GPUs against 6 us at 2 (profiles/r2i_bench_n8.jsonl); in batches the wait is one round trip
    // plus however late the slowest peer is.
    const uint2 *mine = a.peer_ll[a.rank] + slot_off*a.n32;
    const int64_t total = (int64_t) (a.world - 1)*a.n32;
    const long long t0 = clock64();
    constexpr int U = 8;
    for (int64_t i0 = t; i0 < total; i0 += stride*U) {
        uint32_t x[U], e[U];
        int64_t src[U], dst[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t i = i0 + (int64_t) u*stride;
            const int r = (a.rank + 1 + (int) (i / a.n32)) % a.world;    // peers in staggered order, as they were pushed
            const int64_t w = i % a.n32;
            src[u] = i < total ? (int64_t) r*a.n32 + w : -1;
            dst[u] = (slot_off + r)*a.n32 + w;
            e[u] = epoch; x[u] = 0;
            if (src[u] >= 0) asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(x[u]), "=r"(e[u]) : "l"(mine + src[u]) : "memory");
        }
        bool dead = false;
        for (;;) {                                                       // re-poll, again all at once, whatever has not arrived
            bool pending = false;
#pragma unroll
            for (int u = 0; u < U; ++u) pending |= e[u] != epoch;
            if (!pending) break;
            if (clock64() - t0 > 4000000000LL) {                         // ~2 s: a peer died; do not hang the GPU
#pragma unroll
                for (int u = 0; u < U; ++u) if (e[u] != epoch) a.state[2] = 1 + (int) (src[u] / a.n32);
                dead = true;
                break;
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (e[u] != epoch) asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(x[u]), "=r"(e[u]) : "l"(mine + src[u]) : "memory");
            }
        }
        if (dead) break;
#pragma unroll
        for (int u = 0; u < U; ++u) { if (src[u] >= 0) a.gathered[dst[u]] = x[u]; }
    }
    __shared__ int s_last;
    __syncthreads();
    if (threadIdx.x == 0) {
        s_last = atomicAdd(&a.state[1], 1u) == gridDim.x - 1;
        if (s_last) { a.state[1] = 0; a.state[0] = epoch; }              // every CTA has read the old epoch by now
    }
}

int launch_peer_exchange_ll(const void *local, int64_t bytes, void *const *peer_ll, void *gathered, void *state,
                            int slot, int rank, int world, cudaStream_t stream) {
    if (slot < 0 || world < 1 || world > PG_MAX_WORLD || rank < 0 || rank >= world || bytes <= 0 || (bytes & 15) || (((uintptr_t) local) & 3) || !gathered || !state) {
        set_error("peer_exchange_ll: bad arguments (world %d, rank %d, bytes %lld)", world, rank, (long long) bytes);
        return B200_ERR_INVALID;
    }
    pg_ll_args a;
    a.local = static_cast<const uint32_t *>(local); a.n32 = bytes/4; a.gathered = static_cast<uint32_t *>(gathered);
    a.state = static_cast<uint32_t *>(state); a.rank = rank; a.world = world; a.slot = slot;
    for (int p = 0; p < world; ++p) {
        a.peer_ll[p] = static_cast<uint2 *>(peer_ll[p]);
        if (!a.peer_ll[p] || (((uintptr_t) a.peer_ll[p]) & 7)) { set_error("peer_exchange_ll: peer %d not mapped", p); return B200_ERR_INVALID; }
    }
    int64_t gx = (a.n32 + 256 - 1)/256;                                  // one word per thread to push, world - 1 to pull
    if (gx < 1) gx = 1;
    if (gx > 148) gx = 148;                                              // nobody waits on a local CTA, so co-residency is not required
    launch_pdl(pg_ll_kernel, dim3((unsigned) gx), dim3(256), 0, stream, a);
    count_launch();
    return check_launch("peer_exchange_ll") ? B200_OK : B200_ERR_CUDA;
}

int launch_peer_wait(void *flags, void *state, int world, cudaStream_t stream) {
    if (world < 1 || world > PG_MAX_WORLD || !flags || !state) { set_error("peer_wait: bad arguments"); return B200_ERR_INVALID; }
    pg_wait_kernel<<<1, 32, 0, stream>>>(static_cast<uint32_t *>(flags), static_cast<uint32_t *>(state), world);
    count_launch();
    return check_launch("peer_wait") ? B200_OK : B200_ERR_CUDA;
}

} // namespace b200
