// One translation unit per (q dtype, fp8 KV, head dim) triple so the instantiations build in parallel:
// compile with -DB200ATTN_INST_QT=<type> -DB200ATTN_INST_FP8=<0|1> -DB200ATTN_INST_D=<head dim>.
#include "decode_kernel.cuh"

namespace b200attn {
B200ATTN_DEFINE_LAUNCH_D(B200ATTN_INST_QT, (B200ATTN_INST_FP8 != 0), B200ATTN_INST_D)
}  // namespace b200attn
