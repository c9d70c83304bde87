// Scalar activation math shared by all kernels (reference: models/model_res_sigmoid.py:75-126,146).
#pragma once
#include <cuda_runtime.h>

namespace pntf {

// softplus with beta = 10 and torch's threshold rule (returns y when 10*y > 20), plus its first
// two derivatives s = sigmoid(10 y), s' = 10 s (1-s).
struct Act { float a, s, sp; };

__device__ __forceinline__ float sigmoidf_acc(float x) { return 1.0f / (1.0f + expf(-x)); }

__device__ __forceinline__ Act act_eval(float y) {
  Act r;
  const float z = 10.0f * y;
  r.a = (z > 20.0f) ? y : 0.1f * log1pf(expf(z));
  r.s = sigmoidf_acc(z);
  r.sp = 10.0f * r.s * (1.0f - r.s);
  return r;
}
// third derivative factor s'' = 10 s' (1 - 2 s)
__device__ __forceinline__ float act_spp(const Act& r) { return 10.0f * r.sp * (1.0f - 2.0f * r.s); }

// output activation o = sigmoid(0.1 y); o' = 0.1 o (1-o); o'' = 0.1 o' (1-2o);
// o''' = 0.1 (0.1 o' (1-2o)^2 - 2 o'^2)
struct ActOut { float o, o1, o2, o3; };
__device__ __forceinline__ ActOut actout_eval(float y) {
  ActOut r;
  r.o = sigmoidf_acc(0.1f * y);
  r.o1 = 0.1f * r.o * (1.0f - r.o);
  r.o2 = 0.1f * r.o1 * (1.0f - 2.0f * r.o);
  const float u = 1.0f - 2.0f * r.o;
  r.o3 = 0.1f * (0.1f * r.o1 * u * u - 2.0f * r.o1 * r.o1);
  return r;
}

}  // namespace pntf
