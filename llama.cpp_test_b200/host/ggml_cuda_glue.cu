// ggml_cuda_glue.cu — INTEGRATION.md section 1 as a translation unit: the reference CUDA backend's two hot-path
// calls routed to libb200attn.so.
//
// This file is compiled ONLY together with the reference tree (oracle/Makefile, target `refcuda`; it includes the
// reference's ggml-cuda/common.cuh for ggml_backend_cuda_context and its pool) and is not part of libb200attn.so.
// The reference's ggml-cuda.cu is compiled where it lies with
//     -Dggml_cuda_flash_attn_ext=b200_glue_flash_attn_ext -Dggml_cuda_cpy=b200_glue_cpy
//     -Dggml_cuda_dup=b200_glue_dup -Dggml_cuda_cpy_fn=b200_glue_cpy_fn
// so that the calls in ggml_cuda_compute_forward (ggml/src/ggml-cuda/ggml-cuda.cu:2123-2130 DUP / CPY / CONT, :2298
// FLASH_ATTN_EXT) and in the CUDA-graph bookkeeping (:2470) land here, while fattn.cu / cpy.cu keep defining the
// reference's own functions under their own names — which are the fall-back for every node b200_supports_op declines.
// Nothing of the reference is modified or copied; the same effect is the ~40-line patch of INTEGRATION.md.
//
// B200_GLUE=0 in the environment sends everything to the reference's kernels (A/B runs with one binary);
// B200_GLUE_STATS=1 prints, at exit, how many nodes each side computed; B200_GLUE_TIMING=1 brackets every FLASH_ATTN_EXT
// node (either side) with CUDA events and prints, at exit, the GPU time they took per ubatch size — the attention share
// of a llama-batched-bench run (SURVEY 8d config 5 asks for it "from cb_eval / nsys"; neither is needed this way).
#include "common.cuh"                  // the reference's ggml/src/ggml-cuda/common.cuh
#include "b200_ggml_abi.h"

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <vector>

// the reference's own entry points (ggml/src/ggml-cuda/fattn.cu:244, cpy.cu:510, :586, :591)
void   ggml_cuda_flash_attn_ext(ggml_backend_cuda_context & ctx, ggml_tensor * dst);
void   ggml_cuda_cpy(ggml_backend_cuda_context & ctx, const ggml_tensor * src0, ggml_tensor * src1);
void   ggml_cuda_dup(ggml_backend_cuda_context & ctx, ggml_tensor * dst);
void * ggml_cuda_cpy_fn(const ggml_tensor * src0, ggml_tensor * src1);

static_assert(sizeof(ggml_tensor) == sizeof(b200_ggml_tensor), "b200_ggml_tensor is not the reference's ggml_tensor");

namespace {

std::atomic<long long> n_fa_b200{0}, n_fa_ref{0}, n_cpy_b200{0}, n_cpy_ref{0};

struct fa_span { cudaEvent_t e0, e1; int64_t n_tokens; };
std::vector<fa_span> g_spans;

bool timing_on() {
    static const bool on = [] {
        const char * e = getenv("B200_GLUE_TIMING");
        const bool v = e && e[0] == '1';
        if (v) {
            atexit([] {
                cudaDeviceSynchronize();
                std::map<int64_t, std::pair<long long, double>> by;
                for (auto & sp : g_spans) { float ms = 0; if (cudaEventElapsedTime(&ms, sp.e0, sp.e1) == cudaSuccess) { by[sp.n_tokens].first++; by[sp.n_tokens].second += ms; } }
                for (auto & kv : by) fprintf(stderr, "b200 glue timing: FLASH_ATTN_EXT nodes with %lld query rows: %lld nodes, %.3f ms of GPU time\n",
                                             (long long) kv.first, kv.second.first, kv.second.second);
            });
        }
        return v;
    }();
    return on;
}

bool glue_on() {
    static const bool on = [] {
        const char * e = getenv("B200_GLUE");
        const bool v = !(e && e[0] == '0');
        const char * s = getenv("B200_GLUE_STATS");
        if (s && s[0] == '1') {
            atexit([] {
                fprintf(stderr, "b200 glue: FLASH_ATTN_EXT nodes b200=%lld reference=%lld; CPY/DUP/CONT nodes b200=%lld reference=%lld; "
                                "b200 kernel launches=%lld\n", n_fa_b200.load(), n_fa_ref.load(), n_cpy_b200.load(), n_cpy_ref.load(),
                        (long long) b200_launch_count());
            });
        }
        return v;
    }();
    return on;
}

// what b200_supports_op looks at for a copy: an op node whose src[0] / src[1] are the two tensors
bool b200_takes_cpy(const ggml_tensor * src0, const ggml_tensor * src1) {
    if (!glue_on()) return false;
    ggml_tensor node = {};
    node.op = GGML_OP_CPY;
    node.src[0] = const_cast<ggml_tensor *>(src0);
    node.src[1] = const_cast<ggml_tensor *>(src1);
    return b200_supports_op(reinterpret_cast<const b200_ggml_tensor *>(&node)) != 0;
}

} // namespace

static void fa_dispatch(ggml_backend_cuda_context & ctx, ggml_tensor * dst);

void b200_glue_flash_attn_ext(ggml_backend_cuda_context & ctx, ggml_tensor * dst) {
    if (!timing_on()) return fa_dispatch(ctx, dst);
    fa_span sp; sp.n_tokens = dst->src[0]->ne[1];
    cudaEventCreate(&sp.e0); cudaEventCreate(&sp.e1);
    cudaEventRecord(sp.e0, ctx.stream());
    fa_dispatch(ctx, dst);
    cudaEventRecord(sp.e1, ctx.stream());
    g_spans.push_back(sp);
}

static void fa_dispatch(ggml_backend_cuda_context & ctx, ggml_tensor * dst) {
    const b200_ggml_tensor * node = reinterpret_cast<const b200_ggml_tensor *>(dst);
    if (dst->src[4] != nullptr && !(glue_on() && b200_supports_op(node))) {
        GGML_ABORT("a paged FLASH_ATTN_EXT node (src[4] block table) needs libb200attn: %s", glue_on() ? b200_last_error() : "B200_GLUE=0");
    }
    if (glue_on() && b200_supports_op(node)) {
        const size_t ws = b200_flash_attn_ext_workspace(node);
        ggml_cuda_pool_alloc<char> scratch(ctx.pool(), ws > 0 ? ws : 1);       // the backend's pool, as launch_fattn does (fattn-common.cuh:708-711)
        const int rc = b200_flash_attn_ext(node, scratch.get(), ws, ctx.stream());
        if (rc != B200_OK) GGML_ABORT("b200_flash_attn_ext: %s", b200_last_error());
        n_fa_b200++;
        return;
    }
    n_fa_ref++;
    ggml_cuda_flash_attn_ext(ctx, dst);
}

void b200_glue_cpy(ggml_backend_cuda_context & ctx, const ggml_tensor * src0, ggml_tensor * src1) {
    if (b200_takes_cpy(src0, src1)) {
        const int rc = b200_cpy(reinterpret_cast<const b200_ggml_tensor *>(src0), reinterpret_cast<b200_ggml_tensor *>(src1), ctx.stream());
        if (rc != B200_OK) GGML_ABORT("b200_cpy: %s", b200_last_error());
        n_cpy_b200++;
        return;
    }
    n_cpy_ref++;
    ggml_cuda_cpy(ctx, src0, src1);
}

void b200_glue_dup(ggml_backend_cuda_context & ctx, ggml_tensor * dst) {       // cpy.cu:586-589
    b200_glue_cpy(ctx, dst->src[0], dst);
}

// The backend's CUDA-graph replay patches the destination pointer of the reference's copy KERNELS by function address
// (ggml-cuda.cu:2466-2480, :2574-2582).  A copy this library computes has no such kernel node, so the backend is told
// "unsupported copy op" and leaves CUDA graphs off for that graph (they are on only for single-sequence decode anyway,
// :2456-2463); the slot-mapped b200_kv_write form needs no patching at all — its destinations are read from device memory.
void * b200_glue_cpy_fn(const ggml_tensor * src0, ggml_tensor * src1) {
    if (b200_takes_cpy(src0, src1)) return nullptr;
    return ggml_cuda_cpy_fn(src0, src1);
}
