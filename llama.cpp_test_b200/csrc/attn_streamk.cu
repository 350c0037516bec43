// attn_streamk.cu — placeholder until the stream-K tensor-core decode kernel lands (next commit).
#include "common.cuh"
namespace b200 {
bool   streamk_supports(const b200_attn_desc &) { return false; }
size_t streamk_workspace(const b200_attn_desc &) { return 0; }
int    launch_attn_streamk(const b200_attn_desc &, cudaStream_t) { set_error("stream-K kernel not built"); return B200_ERR_UNSUPPORTED; }
}
