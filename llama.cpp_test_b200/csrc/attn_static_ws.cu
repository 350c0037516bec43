// attn_static_ws.cu — b200_paged_attention_static_ws: the stream-K decode path for callers that cannot provide
// (or allocate) a workspace: code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc.
// No counterpart in the reference.
//
// The scratch the stream-K kernels need (the device-side plan and the (m, l, O) partials of split rows, see
// attn_streamk.cu) lives in ONE file-scope __device__ array of B200_STATIC_WS_BYTES.  Every GPU that loads the library
// holds that array until the process exits.  In host code the array's name is only a host-side stand-in for the
// symbol, never a device address: the address is asked for once per device, with that device current, through
// cudaGetSymbolAddress, and that pointer is what the kernels receive.
//
// Nothing in the array has to start at zero: sk_plan_kernel rewrites the whole plan (header, prefix sums, segment
// slots) before sk_main_kernel reads it, and a partial is always written by sk_main_kernel before sk_combine_kernel
// reads it — so the previous call's contents are simply overwritten.  Because EVERY call on a device shares the one
// array, two calls that overlap on the GPU (different streams, or two host threads on one device) race on it; calls
// must be ordered on one stream (or by events).  B200_ATTN_REUSE_PLAN works as with a caller-owned workspace: the
// previous call through this entry point on the same stream left its plan here.
#include "common.cuh"

#include <atomic>
#include <mutex>

namespace b200 {

bool   streamk_supports(const b200_attn_desc &d);
size_t streamk_workspace(const b200_attn_desc &d);
int    launch_attn_streamk(const b200_attn_desc &d, cudaStream_t st, const b200_kv_write_desc *store, bool store_checked, bool chained);

static __device__ __align__(256) unsigned char g_static_ws[B200_STATIC_WS_BYTES];

constexpr int STATIC_WS_MAX_DEVICES = 64;
static std::mutex         g_static_ws_mutex;
static std::atomic<void *> g_static_ws_addr[STATIC_WS_MAX_DEVICES];

// device address of g_static_ws on the CURRENT device (set up once per device); nullptr + error text on failure
static void *static_ws_address() {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= STATIC_WS_MAX_DEVICES) {
        (void) cudaGetLastError();
        set_error("static workspace: no current CUDA device (or device index above %d)", STATIC_WS_MAX_DEVICES - 1);
        return nullptr;
    }
    void *p = g_static_ws_addr[dev].load(std::memory_order_acquire);
    if (p) return p;
    std::lock_guard<std::mutex> lock(g_static_ws_mutex);
    p = g_static_ws_addr[dev].load(std::memory_order_relaxed);
    if (p) return p;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_static_ws);       // resolves the symbol in the current device's context
    if (e != cudaSuccess || !p) {
        (void) cudaGetLastError();
        set_error("static workspace: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
        return nullptr;
    }
    g_static_ws_addr[dev].store(p, std::memory_order_release);
    return p;
}

} // namespace b200

using namespace b200;

extern "C" {

B200_API size_t b200_static_workspace_bytes(void) { return (size_t) B200_STATIC_WS_BYTES; }

B200_API int b200_paged_attention_static_ws(const b200_attn_desc *d, void *stream) {
    if (!d) { set_error("attention (static workspace): NULL descriptor"); return B200_ERR_INVALID; }
    if (!d->q || !d->k || !d->v || !d->dst) { set_error("attention (static workspace): NULL tensor pointer"); return B200_ERR_INVALID; }
    if (d->n_q < 0 || d->n_head <= 0 || d->n_head_kv <= 0 || d->head_dim <= 0 || d->n_head % d->n_head_kv) {
        set_error("attention (static workspace): bad shape"); return B200_ERR_INVALID;
    }
    if (!d->block_table || !d->seq_lens || d->n_seqs <= 0 || d->block_size <= 0 || d->bt_stride <= 0) {
        set_error("attention (static workspace): paged mode only (block_table, seq_lens, n_seqs, block_size, bt_stride)");
        return B200_ERR_INVALID;
    }
    if (!d->q_seq && d->n_q > d->n_seqs) { set_error("attention (static workspace): n_q %d > n_seqs %d without q_seq", d->n_q, d->n_seqs); return B200_ERR_INVALID; }
    if (d->kernel != B200_KERNEL_AUTO && d->kernel != B200_KERNEL_STREAMK) {
        set_error("attention (static workspace): only the stream-K decode kernel runs on the static workspace"); return B200_ERR_UNSUPPORTED;
    }
    if (!streamk_supports(*d)) { set_error("attention (static workspace): the stream-K kernel does not take this shape"); return B200_ERR_UNSUPPORTED; }
    const size_t need = streamk_workspace(*d);
    if (need > (size_t) B200_STATIC_WS_BYTES) {
        set_error("attention (static workspace): this call needs %zu bytes of scratch, the static workspace holds %zu "
                  "(B200_STATIC_WS_BYTES); use b200_paged_attention with a caller-provided workspace", need, (size_t) B200_STATIC_WS_BYTES);
        return B200_ERR_ALLOC;
    }
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) {
        (void) cudaGetLastError();
        set_error("no CUDA device: libb200attn has no CPU fallback");
        return B200_ERR_CUDA;
    }
    if (d->n_q == 0) return B200_OK;
    void *ws = static_ws_address();
    if (!ws) return B200_ERR_CUDA;
    b200_attn_desc x = *d;
    x.kernel = B200_KERNEL_STREAMK;
    x.workspace = ws;
    x.workspace_bytes = (size_t) B200_STATIC_WS_BYTES;
    return launch_attn_streamk(x, static_cast<cudaStream_t>(stream), nullptr, false, false);
}

} // extern "C"
