// shim.cu — the C ABI of libb200attn.so (include/b200_paged_attn.h): validation, dispatch, bookkeeping.
// No CPU fallback anywhere: every compute entry launches CUDA kernels or returns an error.
#include "common.cuh"

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <vector>

namespace b200 {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char *fmt, ...) {
    va_list ap; va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
bool pdl_enabled() {
    static int on = -1;
    if (on < 0) { const char *e = getenv("B200_NO_PDL"); on = (e && e[0] == '1') ? 0 : 1; }
    return on == 1;
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
bool check_launch(const char *what) {
    const cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) { set_error("%s: %s", what, cudaGetErrorString(e)); (void) cudaGetLastError(); return false; }
    return true;
}

// kernels (kv_write.cu, attn_generic.cu, attn_streamk.cu)
int    launch_kv_write(const b200_kv_write_desc &d, cudaStream_t stream);
int    launch_block_copy(void *const *caches, int n_caches, int64_t block_bytes, int64_t n_blocks,
                         const int32_t *pairs, int64_t n_pairs, cudaStream_t stream);
int    launch_kv_rows(bool scatter, void *const *caches, int n_caches, int64_t nb_token, int64_t row_bytes, int64_t n_cells,
                      const int32_t *slots, int64_t n_rows, void *staging, cudaStream_t stream);
int    launch_k_shift(const b200_k_shift_desc &d, cudaStream_t stream);
int    launch_peer_publish(const void *local, int64_t bytes, void *const *peer_bufs, void *const *peer_flags, void *state,
                           int slot, int rank, int world, bool wait, cudaStream_t stream);
int    launch_peer_wait(void *flags, void *state, int world, cudaStream_t stream);
int    launch_peer_exchange_ll(const void *local, int64_t bytes, void *const *peer_ll, void *gathered, void *state,
                               int slot, int rank, int world, cudaStream_t stream);
bool   generic_supports(const b200_attn_desc &d);
size_t generic_workspace(const b200_attn_desc &d);
int    launch_attn_generic(const b200_attn_desc &d, cudaStream_t st);
bool   streamk_supports(const b200_attn_desc &d);
size_t streamk_workspace(const b200_attn_desc &d);
int    launch_attn_streamk(const b200_attn_desc &d, cudaStream_t st, const b200_kv_write_desc *store, bool store_checked, bool chained);
bool   prefill_supports(const b200_attn_desc &d);
size_t prefill_workspace(const b200_attn_desc &d);
int    launch_attn_prefill(const b200_attn_desc &d, cudaStream_t st);

static bool have_device() {
    static int state = -1;                       // -1 unknown, 0 none, 1 ok
    if (state < 0) {
        int n = 0;
        const cudaError_t e = cudaGetDeviceCount(&n);
        state = (e == cudaSuccess && n > 0) ? 1 : 0;
        if (!state) { (void) cudaGetLastError(); }
    }
    if (!state) set_error("no CUDA device: libb200attn has no CPU fallback");
    return state == 1;
}

static int validate_attn(const b200_attn_desc *d) {
    if (!d) { set_error("attention: NULL descriptor"); return B200_ERR_INVALID; }
    if (!d->q || !d->k || !d->v || !d->dst) { set_error("attention: NULL tensor pointer"); return B200_ERR_INVALID; }
    if (d->n_q < 0 || d->n_head <= 0 || d->n_head_kv <= 0 || d->head_dim <= 0) { set_error("attention: bad shape"); return B200_ERR_INVALID; }
    if (d->n_head % d->n_head_kv) { set_error("attention: n_head %d not a multiple of n_head_kv %d", d->n_head, d->n_head_kv); return B200_ERR_INVALID; }
    if (d->block_table) {
        if (!d->seq_lens || d->n_seqs <= 0 || d->block_size <= 0 || d->bt_stride <= 0) { set_error("attention: paged mode needs seq_lens, n_seqs, block_size, bt_stride"); return B200_ERR_INVALID; }
        if (!d->q_seq && d->n_q > d->n_seqs) { set_error("attention: n_q %d > n_seqs %d without q_seq", d->n_q, d->n_seqs); return B200_ERR_INVALID; }
    } else if (d->n_kv < 0) { set_error("attention: negative n_kv"); return B200_ERR_INVALID; }
    return B200_OK;
}

static int pick_kernel(const b200_attn_desc &d) {
    if (d.kernel == B200_KERNEL_GENERIC) return B200_KERNEL_GENERIC;
    if (d.kernel == B200_KERNEL_STREAMK) return B200_KERNEL_STREAMK;
    if (d.kernel == B200_KERNEL_PREFILL) return B200_KERNEL_PREFILL;
    // many causal query rows per sequence: the tcgen05 tile kernel reads K/V once per 128/gqa rows instead of once per row
    // ... when a sequence brings enough rows to fill a tile: a ubatch interleaved position by position over many sequences
    // (llama-batched-bench's prompt order: 512 rows = 2 positions x 256 sequences) has a few rows per sequence, which the
    // per-row kernel serves better (measured through libllama, profiles/r2g_*: 464 vs 555 ms of attention per prompt)
    if (d.q_pos && d.n_q >= 128 && (int64_t) d.n_q >= 8*(int64_t) d.n_seqs && prefill_supports(d)) return B200_KERNEL_PREFILL;
    return streamk_supports(d) ? B200_KERNEL_STREAMK : B200_KERNEL_GENERIC;
}

#ifdef B200_TIMELINE
unsigned long long *timeline_buffer() {
    static unsigned long long *buf = nullptr;
    if (!buf) {
        cudaMalloc(&buf, (8 + 8*TL_MAX)*sizeof(unsigned long long));
        cudaMemset(buf, 0, (8 + 8*TL_MAX)*sizeof(unsigned long long));
        atexit([] {
            cudaDeviceSynchronize();
            std::vector<unsigned long long> h(8 + 8*TL_MAX);
            if (cudaMemcpy(h.data(), buf, h.size()*sizeof(unsigned long long), cudaMemcpyDeviceToHost) != cudaSuccess) return;
            const char *path = getenv("B200_TIMELINE_OUT");
            FILE *f = path ? fopen(path, "w") : stderr;
            if (!f) return;
            const unsigned long long n = h[0] < TL_MAX ? h[0] : TL_MAX;
            fprintf(f, "# %llu records: kernel(1 kv_write 2 plan 3 main 4 combine) block entry s2 s3 s4 s5 end extra  (ns, absolute)\n", n);
            for (unsigned long long k = 0; k < n; ++k) {
                const unsigned long long *v = &h[8 + 8*k];
                fprintf(f, "%llu %llu %llu %llu %llu %llu %llu %llu %llu\n", v[0] & 255, v[0] >> 8, v[1], v[2], v[3], v[4], v[5], v[6], v[7]);
            }
            if (path) fclose(f);
        });
    }
    return buf;
}
#endif

} // namespace b200

using namespace b200;

extern "C" {

B200_API const char *b200_last_error(void) { return g_err; }
B200_API const char *b200_version(void)    { return "b200attn 0.1 (sm_100a)"; }
B200_API int64_t     b200_launch_count(void) { return g_launches.load(); }

B200_API int b200_device_sm_count(void) {
    if (!have_device()) return 0;
    int dev = 0, n = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    return n;
}

B200_API int b200_supports_attention(const b200_attn_desc *d) {
    if (validate_attn(d) != B200_OK) return 0;
    if (d->kernel == B200_KERNEL_STREAMK) return streamk_supports(*d) ? 1 : 0;
    if (d->kernel == B200_KERNEL_STREAMK_DQ) return 0;                                   // removed in round 2 (include/b200_paged_attn.h)
    if (d->kernel == B200_KERNEL_PREFILL) return prefill_supports(*d) ? 1 : 0;
    if (d->kernel == B200_KERNEL_GENERIC) return generic_supports(*d) ? 1 : 0;
    return (generic_supports(*d) || streamk_supports(*d) || prefill_supports(*d)) ? 1 : 0;
}

B200_API size_t b200_paged_attention_workspace(const b200_attn_desc *d) {
    if (validate_attn(d) != B200_OK) return 0;
    size_t a = generic_supports(*d) ? generic_workspace(*d) : 0;
    size_t b = streamk_supports(*d) ? streamk_workspace(*d) : 0;
    size_t c = prefill_supports(*d) ? prefill_workspace(*d) : 0;
    if (d->kernel == B200_KERNEL_GENERIC) b = c = 0;
    if (d->kernel == B200_KERNEL_STREAMK) a = c = 0;
    if (d->kernel == B200_KERNEL_PREFILL) a = b = 0;
    if (b > a) a = b;
    return a > c ? a : c;
}

static int validate_kv_write(const b200_kv_write_desc *d);

// The attention of one layer, with that layer's KV store (or NULL) in front of it.  The stream-K path places the store
// between its plan and main kernels and lets the main kernel start under it (attn_streamk.cu, "store overlap"); every
// other kernel runs store, then attention.  store_checked, chained: see launch_attn_streamk.
static int attention_with_store(const b200_attn_desc *d, const b200_kv_write_desc *store, bool store_checked, bool chained, void *stream) {
    const int rc = validate_attn(d);
    if (rc != B200_OK) return rc;
    if (store) { const int rs = validate_kv_write(store); if (rs != B200_OK) return rs; }
    if (!have_device()) return B200_ERR_CUDA;
    if (d->n_q == 0 || pick_kernel(*d) != B200_KERNEL_STREAMK || !streamk_supports(*d) || d->kernel == B200_KERNEL_STREAMK_DQ) {
        if (store) { const int rs = launch_kv_write(*store, static_cast<cudaStream_t>(stream)); if (rs != B200_OK) return rs; }
        store = nullptr;
    }
    if (d->n_q == 0) return B200_OK;
    if (d->kernel == B200_KERNEL_STREAMK_DQ) { set_error("attention: kernel id 4 (tcgen05 decode formulation) was removed; use AUTO or STREAMK"); return B200_ERR_UNSUPPORTED; }
    const int k = pick_kernel(*d);
    if (k == B200_KERNEL_PREFILL) {
        if (!prefill_supports(*d)) { set_error("attention: prefill tile kernel does not take this shape"); return B200_ERR_UNSUPPORTED; }
        return launch_attn_prefill(*d, static_cast<cudaStream_t>(stream));
    }
    if (k == B200_KERNEL_STREAMK) {
        if (!streamk_supports(*d)) {
            set_error("attention: stream-K kernel does not take this shape"); return B200_ERR_UNSUPPORTED;
        }
        return launch_attn_streamk(*d, static_cast<cudaStream_t>(stream), store, store_checked, chained);
    }
    if (!generic_supports(*d)) {
        set_error("attention: unsupported shape/type (head_dim %d, type_k %d, type_v %d)", d->head_dim, d->type_k, d->type_v);
        return B200_ERR_UNSUPPORTED;
    }
    return launch_attn_generic(*d, static_cast<cudaStream_t>(stream));
}

B200_API int b200_paged_attention(const b200_attn_desc *d, void *stream) { return attention_with_store(d, nullptr, false, false, stream); }

B200_API int b200_store_attention(const b200_kv_write_desc *store, const b200_attn_desc *d, void *stream) {
    return attention_with_store(d, store, false, false, stream);
}

static int validate_kv_write(const b200_kv_write_desc *d) {
    if (!d) { set_error("kv_write: NULL descriptor"); return B200_ERR_INVALID; }
    if (d->n_tokens < 0 || d->n_head_kv <= 0 || d->head_dim <= 0) { set_error("kv_write: bad shape"); return B200_ERR_INVALID; }
    if ((d->k_src && !d->k_cache) || (d->v_src && !d->v_cache)) { set_error("kv_write: source without cache"); return B200_ERR_INVALID; }
    if (!type_ok(d->type_k) || !type_ok(d->type_v)) { set_error("kv_write: unsupported cache type %d/%d", d->type_k, d->type_v); return B200_ERR_UNSUPPORTED; }
    if (d->head_dim % QK) { set_error("kv_write: head_dim %d not a multiple of 32", d->head_dim); return B200_ERR_UNSUPPORTED; }
    if ((d->src_nb1 & 15) || ((uintptr_t) d->k_src & 15) || ((uintptr_t) d->v_src & 15)) { set_error("kv_write: source rows must be 16-byte aligned"); return B200_ERR_INVALID; }
    if ((d->nbk_token | d->nbk_head | d->nbv_token | d->nbv_head | (int64_t) (uintptr_t) d->k_cache | (int64_t) (uintptr_t) d->v_cache) & 1) { set_error("kv_write: cache strides must be even"); return B200_ERR_INVALID; }
    return B200_OK;
}

B200_API int b200_kv_write(const b200_kv_write_desc *d, void *stream) {
    const int rc = validate_kv_write(d);
    if (rc != B200_OK) return rc;
    if (!have_device()) return B200_ERR_CUDA;
    return launch_kv_write(*d, static_cast<cudaStream_t>(stream));
}

// two layers of one step: everything the schedule is computed from is the same, only the tensors differ
static bool same_plan_inputs(const b200_attn_desc &x, const b200_attn_desc &y) {
    return x.n_q == y.n_q && x.n_head == y.n_head && x.n_head_kv == y.n_head_kv && x.head_dim == y.head_dim
        && x.type_k == y.type_k && x.type_v == y.type_v && x.n_kv == y.n_kv
        && x.block_table != nullptr && x.block_table == y.block_table && x.bt_stride == y.bt_stride
        && x.seq_lens == y.seq_lens && x.n_seqs == y.n_seqs && x.block_size == y.block_size && x.q_seq == y.q_seq && x.q_pos == y.q_pos
        && x.max_seq_len == y.max_seq_len && x.total_tokens == y.total_tokens
        && x.workspace == y.workspace && x.workspace_bytes == y.workspace_bytes && x.kernel == y.kernel;
}

B200_API int b200_decode_step(const b200_kv_write_desc *writes, const b200_attn_desc *attns, int32_t n_layers, void *stream) {
    if (n_layers < 0 || (n_layers > 0 && !attns)) { set_error("decode_step: bad arguments"); return B200_ERR_INVALID; }
    const int32_t *planned_slots = nullptr; int64_t planned_tokens = -1, planned_cells = -1;   // the store the standing plan was checked against
    for (int32_t l = 0; l < n_layers; ++l) {
        const b200_kv_write_desc *w = writes ? &writes[l] : nullptr;
        int rc;
        // the schedule of layer l-1 stands only if layer l-1 really ran the stream-K plan kernel and layer l runs stream-K too:
        // pick_kernel also looks at pointers and strides (alignment), which same_plan_inputs does not compare
        if (l > 0 && !(attns[l].flags & B200_ATTN_REUSE_PLAN) && same_plan_inputs(attns[l - 1], attns[l])
            && attns[l - 1].logit_softcap == attns[l].logit_softcap
            && validate_attn(&attns[l - 1]) == B200_OK && pick_kernel(attns[l - 1]) == B200_KERNEL_STREAMK
            && validate_attn(&attns[l]) == B200_OK && pick_kernel(attns[l]) == B200_KERNEL_STREAMK) {
            b200_attn_desc d = attns[l];                  // the schedule of layer l-1 stands: skip the plan launch
            d.flags |= B200_ATTN_REUSE_PLAN;
            rc = attention_with_store(&d, w, w && planned_tokens >= 0 && w->slot_mapping == planned_slots && w->n_tokens == planned_tokens && w->n_cells == planned_cells, true, stream);
        } else {
            rc = attention_with_store(&attns[l], w, false, false, stream);
            const bool planned = w && !(attns[l].flags & B200_ATTN_REUSE_PLAN);       // this call ran the plan kernel, with w's slots
            planned_slots = planned ? w->slot_mapping : nullptr; planned_tokens = planned ? w->n_tokens : -1; planned_cells = planned ? w->n_cells : -1;
        }
        if (rc != B200_OK) return rc;
    }
    return B200_OK;
}

B200_API int b200_block_copy(void *const *caches, int32_t n_caches, int64_t block_bytes, int64_t n_blocks,
                             const int32_t *pairs, int64_t n_pairs, void *stream) {
    if (n_caches < 0 || n_pairs < 0 || (n_caches > 0 && !caches) || (n_pairs > 0 && !pairs)) { set_error("block_copy: bad arguments"); return B200_ERR_INVALID; }
    if (!have_device()) return B200_ERR_CUDA;
    return launch_block_copy(caches, n_caches, block_bytes, n_blocks, pairs, n_pairs, static_cast<cudaStream_t>(stream));
}

B200_API int b200_kv_rows_gather(void *const *caches, int32_t n_caches, int64_t nb_token, int64_t row_bytes, int64_t n_cells,
                                 const int32_t *slots, int64_t n_rows, void *staging, void *stream) {
    if (n_caches < 0 || n_rows < 0 || (n_caches > 0 && !caches) || (n_rows > 0 && (!slots || !staging))) { set_error("kv_rows_gather: bad arguments"); return B200_ERR_INVALID; }
    if (!have_device()) return B200_ERR_CUDA;
    return launch_kv_rows(false, caches, n_caches, nb_token, row_bytes, n_cells, slots, n_rows, staging, static_cast<cudaStream_t>(stream));
}

B200_API int b200_kv_rows_scatter(void *const *caches, int32_t n_caches, int64_t nb_token, int64_t row_bytes, int64_t n_cells,
                                  const int32_t *slots, int64_t n_rows, const void *staging, void *stream) {
    if (n_caches < 0 || n_rows < 0 || (n_caches > 0 && !caches) || (n_rows > 0 && (!slots || !staging))) { set_error("kv_rows_scatter: bad arguments"); return B200_ERR_INVALID; }
    if (!have_device()) return B200_ERR_CUDA;
    return launch_kv_rows(true, caches, n_caches, nb_token, row_bytes, n_cells, slots, n_rows, const_cast<void *>(staging), static_cast<cudaStream_t>(stream));
}

B200_API int b200_k_shift(const b200_k_shift_desc *d, void *stream) {
    if (!d) { set_error("k_shift: null descriptor"); return B200_ERR_INVALID; }
    if (d->n_caches < 0 || d->n_cells < 0 || (d->cells && d->n_rows < 0)) { set_error("k_shift: bad arguments"); return B200_ERR_INVALID; }
    if (!have_device()) return B200_ERR_CUDA;
    return launch_k_shift(*d, static_cast<cudaStream_t>(stream));
}

B200_API void *b200_peer_alloc(size_t bytes) {
    if (!have_device() || bytes == 0) return nullptr;
    void *p = nullptr;
    if (cudaMalloc(&p, bytes) != cudaSuccess || cudaMemset(p, 0, bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
        set_error("peer_alloc: %s", cudaGetErrorString(cudaGetLastError()));
        return nullptr;
    }
    return p;
}
B200_API int b200_peer_free(void *p) { return (!p || cudaFree(p) == cudaSuccess) ? B200_OK : B200_ERR_CUDA; }
B200_API int b200_peer_export(void *p, unsigned char handle[64]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaIpcMemHandle_t h;
    if (!p || !handle || cudaIpcGetMemHandle(&h, p) != cudaSuccess) { set_error("peer_export: %s", cudaGetErrorString(cudaGetLastError())); return B200_ERR_CUDA; }
    memcpy(handle, &h, 64);
    return B200_OK;
}
B200_API void *b200_peer_open(const unsigned char handle[64]) {
    if (!have_device() || !handle) return nullptr;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    void *p = nullptr;
    if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { set_error("peer_open: %s", cudaGetErrorString(cudaGetLastError())); return nullptr; }
    return p;
}
B200_API int b200_peer_close(void *p) { return (!p || cudaIpcCloseMemHandle(p) == cudaSuccess) ? B200_OK : B200_ERR_CUDA; }
B200_API int b200_peer_publish(const void *local, int64_t bytes, void *const *peer_bufs, void *const *peer_flags,
                               void *state, int32_t slot, int32_t rank, int32_t world, void *stream) {
    if (!local || !peer_bufs || !peer_flags || !state) { set_error("peer_publish: bad arguments"); return B200_ERR_INVALID; }
    if (!have_device()) return B200_ERR_CUDA;
    return launch_peer_publish(local, bytes, peer_bufs, peer_flags, state, slot, rank, world, false, static_cast<cudaStream_t>(stream));
}
B200_API int b200_peer_exchange(const void *local, int64_t bytes, void *const *peer_bufs, void *const *peer_flags,
                                void *state, int32_t slot, int32_t rank, int32_t world, void *stream) {
    if (!local || !peer_bufs || !peer_flags || !state) { set_error("peer_exchange: bad arguments"); return B200_ERR_INVALID; }
    if (!have_device()) return B200_ERR_CUDA;
    return launch_peer_publish(local, bytes, peer_bufs, peer_flags, state, slot, rank, world, true, static_cast<cudaStream_t>(stream));
}
B200_API int b200_peer_exchange_ll(const void *local, int64_t bytes, void *const *peer_ll, void *gathered, void *state,
                                   int32_t slot, int32_t rank, int32_t world, void *stream) {
    if (!local || !peer_ll) { set_error("peer_exchange_ll: bad arguments"); return B200_ERR_INVALID; }
    if (!have_device()) return B200_ERR_CUDA;
    return launch_peer_exchange_ll(local, bytes, peer_ll, gathered, state, slot, rank, world, static_cast<cudaStream_t>(stream));
}
B200_API int b200_peer_wait(void *flags, void *state, int32_t world, void *stream) {
    if (!have_device()) return B200_ERR_CUDA;
    return launch_peer_wait(flags, state, world, static_cast<cudaStream_t>(stream));
}
B200_API int b200_peer_read_state(const void *state, uint32_t out[4]) {
    if (!state || !out || !have_device()) return B200_ERR_INVALID;
    return cudaMemcpy(out, state, 16, cudaMemcpyDeviceToHost) == cudaSuccess ? B200_OK : B200_ERR_CUDA;
}

} // extern "C"
