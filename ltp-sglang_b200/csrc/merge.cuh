#pragma once
#include "common.cuh"

namespace b200attn {

struct MergeParams {
  const void* plan;
  const float* ws_o;
  const float* ws_lse;
  void* o;
  int64_t o_sb, o_sh;
  float* lse;
  int32_t batch, hq, ng, ts, dv;
};

cudaError_t launch_merge(const MergeParams& p, bool half, cudaStream_t stream);

}  // namespace b200attn
