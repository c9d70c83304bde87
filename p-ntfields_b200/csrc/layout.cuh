// Parameter layout of the travel-time network (reference: models/model_res_sigmoid.py:159-181,
// models/model_res_sigmoid_multi.py:153-175).  The flat fp32 buffer holds the 30 tensors in
// state_dict order; the 14 layers a forward pass actually runs are listed in execution order.
#pragma once
#include <cstdint>

namespace pntf {

constexpr int kParamCount = 576129;
constexpr int kNumTensors = 30;
constexpr int kNumLayers = 14;   // executed layers (encoder1.0 is allocated but never executed)
constexpr int kNumFreq = 128;

struct TensorDesc { int offset, rows, cols, used; };

// state_dict order
__host__ __device__ constexpr TensorDesc tensor_desc(int i) {
  constexpr TensorDesc t[kNumTensors] = {
      {0, 128, 256, 1},      {32768, 128, 1, 1},     // encoder.0
      {32896, 128, 128, 1},  {49280, 128, 1, 1},     // encoder.1
      {49408, 128, 128, 1},  {65792, 128, 1, 1},     // encoder.2
      {65920, 128, 128, 1},  {82304, 128, 1, 1},     // encoder.3
      {82432, 128, 256, 0},  {115200, 128, 1, 0},    // encoder1.0 (unused)
      {115328, 128, 128, 1}, {131712, 128, 1, 1},    // encoder1.1
      {131840, 128, 128, 1}, {148224, 128, 1, 1},    // encoder1.2
      {148352, 256, 256, 1}, {213888, 256, 1, 1},    // generator.0
      {214144, 256, 256, 1}, {279680, 256, 1, 1},    // generator.1
      {279936, 256, 256, 1}, {345472, 256, 1, 1},    // generator.2
      {345728, 128, 256, 1}, {378496, 128, 1, 1},    // generator.3
      {378624, 1, 128, 1},   {378752, 1, 1, 1},      // generator.4
      {378753, 256, 256, 1}, {444289, 256, 1, 1},    // generator1.0
      {444545, 256, 256, 1}, {510081, 256, 1, 1},    // generator1.1
      {510337, 256, 256, 1}, {575873, 256, 1, 1},    // generator1.2
  };
  return t[i];
}

// Executed layers: y = x W^T + b with W[M][K].  Layers 0..5 run per endpoint (encoder),
// 6..13 per pair (generator).  `packed` = offset of the layer inside a packed weight copy
// (16-byte aligned, sizes M*K).
struct LayerDesc { int w, b, M, K, packed; };

__host__ __device__ constexpr LayerDesc layer_desc(int l) {
  constexpr LayerDesc t[kNumLayers] = {
      {0, 32768, 128, 256, 0},              // 0  encoder.0
      {32896, 49280, 128, 128, 32768},      // 1  encoder.1
      {115328, 131712, 128, 128, 49152},    // 2  encoder1.1 (+skip)
      {49408, 65792, 128, 128, 65536},      // 3  encoder.2
      {131840, 148224, 128, 128, 81920},    // 4  encoder1.2 (+skip)
      {65920, 82304, 128, 128, 98304},      // 5  encoder.3 (no activation)
      {148352, 213888, 256, 256, 114688},   // 6  generator.0
      {378753, 444289, 256, 256, 180224},   // 7  generator1.0 (+skip)
      {214144, 279680, 256, 256, 245760},   // 8  generator.1
      {444545, 510081, 256, 256, 311296},   // 9  generator1.1 (+skip)
      {279936, 345472, 256, 256, 376832},   // 10 generator.2
      {510337, 575873, 256, 256, 442368},   // 11 generator1.2 (+skip)
      {345728, 378496, 128, 256, 507904},   // 12 generator.3
      {378624, 378752, 1, 128, 540672},     // 13 generator.4 (output activation)
  };
  return t[l];
}
constexpr int kPackedCount = 540672 + 128;   // floats in one packed copy

}  // namespace pntf
