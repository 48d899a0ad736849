// One translation unit per (q dtype, fp8 KV) pair so the instantiations build
// in parallel: compile with -DB200ATTN_INST_QT=<type> -DB200ATTN_INST_FP8=<0|1>.
#include "decode_kernel.cuh"

namespace b200attn {
B200ATTN_DEFINE_LAUNCH(B200ATTN_INST_QT, (B200ATTN_INST_FP8 != 0))
}  // namespace b200attn
