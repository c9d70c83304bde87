// Ground-truth speed sampling: unsigned distance from query points to a triangle mesh and the
// clipped, normalised speed the training data stores.
// Reference: dataprocessing/speed_sampling_gpu.py:325-336 (point_obstacle_distance: sqrt of the squared distance
// returned by the third-party bvh_distance_queries.BVH, absent from the reference tree) and :409-419
// (speed = clip(distance, offset, margin) / margin).  The distance of a point to a triangle soup does not depend on
// the search structure, so this is an exact restatement: closest point on each candidate triangle by Voronoi-region
// classification (Ericson, Real-Time Collision Detection 5.1.5), minimum over the mesh.
//
// Layout: a prep kernel turns the mesh into tiles of 128 triangles (a, b-a, c-a and their three dot products: three
// float4 per triangle) with one axis-aligned box per tile.  The query kernel gives every thread one point: a first
// sweep over the tile boxes yields an upper bound (the farthest corner of the nearest box), the second sweep stages
// in shared memory only the tiles whose box some point of the CTA can still improve on.  Any triangle / point order
// is correct; spatially sorted inputs (the host side sorts both along a Morton curve) make whole CTAs skip most tiles.
#include <cfloat>
#include <cstdint>

#include <cuda_runtime.h>

#include "../../include/pntf_b200.h"
#include "tc_api.cuh"

namespace pntf {

constexpr int kMeshTile = 128;       // triangles per tile
constexpr int kMeshThreads = 256;    // points per CTA

struct MeshTiles {
  float4* tri;     // [T * 128][3]: (a.xyz, ab.ab), (ab.xyz, ab.ac), (ac.xyz, ac.ac)
  float* box;      // [T][6]: lo.xyz, hi.xyz
};

__global__ void __launch_bounds__(kMeshTile) mesh_prep_kernel(const float* __restrict__ tri, int64_t M, MeshTiles out) {
  __shared__ float red[6][kMeshTile];
  const int64_t t = int64_t(blockIdx.x) * kMeshTile + threadIdx.x;
  // the last tile is padded with copies of the mesh's last triangle
  const float* p = tri + (t < M ? t : M - 1) * 9;
  float a[3], ab[3], ac[3], lo[3], hi[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    a[i] = p[i];
    const float b = p[3 + i], c = p[6 + i];
    ab[i] = b - a[i];
    ac[i] = c - a[i];
    lo[i] = fminf(a[i], fminf(b, c));
    hi[i] = fmaxf(a[i], fmaxf(b, c));
  }
  const float d00 = ab[0] * ab[0] + ab[1] * ab[1] + ab[2] * ab[2];
  const float d01 = ab[0] * ac[0] + ab[1] * ac[1] + ab[2] * ac[2];
  const float d11 = ac[0] * ac[0] + ac[1] * ac[1] + ac[2] * ac[2];
  float4* o = out.tri + t * 3;
  o[0] = make_float4(a[0], a[1], a[2], d00);
  o[1] = make_float4(ab[0], ab[1], ab[2], d01);
  o[2] = make_float4(ac[0], ac[1], ac[2], d11);
#pragma unroll
  for (int i = 0; i < 3; ++i) { red[i][threadIdx.x] = lo[i]; red[3 + i][threadIdx.x] = hi[i]; }
  __syncthreads();
  for (int s = kMeshTile / 2; s > 0; s >>= 1) {
    if (threadIdx.x < s) {
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        red[i][threadIdx.x] = fminf(red[i][threadIdx.x], red[i][threadIdx.x + s]);
        red[3 + i][threadIdx.x] = fmaxf(red[3 + i][threadIdx.x], red[3 + i][threadIdx.x + s]);
      }
    }
    __syncthreads();
  }
  if (threadIdx.x < 6) out.box[int64_t(blockIdx.x) * 6 + threadIdx.x] = red[threadIdx.x][0];
}

// squared distance from p to the triangle (a, a+ab, a+ac)
__device__ __forceinline__ float tri_dist2(const float px, const float py, const float pz, const float4 A, const float4 B,
                                           const float4 C) {
  const float apx = px - A.x, apy = py - A.y, apz = pz - A.z;
  const float d00 = A.w, d01 = B.w, d11 = C.w;
  const float d1 = B.x * apx + B.y * apy + B.z * apz;   // ab . ap
  const float d2 = C.x * apx + C.y * apy + C.z * apz;   // ac . ap
  const float d3 = d1 - d00, d4 = d2 - d01;             // ab . bp, ac . bp
  const float d5 = d1 - d01, d6 = d2 - d11;             // ab . cp, ac . cp
  const float vc = d1 * d4 - d3 * d2;
  const float vb = d5 * d2 - d1 * d6;
  const float va = d3 * d6 - d5 * d4;
  float v, w;
  if (d1 <= 0.f && d2 <= 0.f) { v = 0.f; w = 0.f; }                              // vertex a
  else if (d3 >= 0.f && d4 <= d3) { v = 1.f; w = 0.f; }                          // vertex b
  else if (vc <= 0.f && d1 >= 0.f && d3 <= 0.f) { v = d1 / (d1 - d3); w = 0.f; } // edge ab
  else if (d6 >= 0.f && d5 <= d6) { v = 0.f; w = 1.f; }                          // vertex c
  else if (vb <= 0.f && d2 >= 0.f && d6 <= 0.f) { v = 0.f; w = d2 / (d2 - d6); } // edge ac
  else if (va <= 0.f && (d4 - d3) >= 0.f && (d5 - d6) >= 0.f) {                  // edge bc
    w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
    v = 1.f - w;
  } else {                                                                       // interior
    const float den = 1.f / (va + vb + vc);
    v = vb * den;
    w = vc * den;
  }
  const float rx = apx - v * B.x - w * C.x, ry = apy - v * B.y - w * C.y, rz = apz - v * B.z - w * C.z;
  return rx * rx + ry * ry + rz * rz;
}

__global__ void __launch_bounds__(kMeshThreads) mesh_distance_kernel(MeshTiles m, int T, const float* __restrict__ pts,
                                                                     int64_t N, float* __restrict__ dist, float offset,
                                                                     float margin, float* __restrict__ speed) {
  __shared__ float4 tile[kMeshTile * 3];
  __shared__ float boxes[kMeshThreads][6];
  const int64_t i = int64_t(blockIdx.x) * kMeshThreads + threadIdx.x;
  const bool live = i < N;
  const float* p = pts + (live ? i : N - 1) * 3;
  const float px = p[0], py = p[1], pz = p[2];
  // sweep 1: upper bound = squared distance to the farthest corner of the best tile box
  float best = FLT_MAX;
  for (int t0 = 0; t0 < T; t0 += kMeshThreads) {
    __syncthreads();
    if (t0 + threadIdx.x < T) {
#pragma unroll
      for (int k = 0; k < 6; ++k) boxes[threadIdx.x][k] = m.box[int64_t(t0 + threadIdx.x) * 6 + k];
    }
    __syncthreads();
    const int n = min(kMeshThreads, T - t0);
    for (int t = 0; t < n; ++t) {
      const float fx = fmaxf(fabsf(px - boxes[t][0]), fabsf(px - boxes[t][3]));
      const float fy = fmaxf(fabsf(py - boxes[t][1]), fabsf(py - boxes[t][4]));
      const float fz = fmaxf(fabsf(pz - boxes[t][2]), fabsf(pz - boxes[t][5]));
      best = fminf(best, fx * fx + fy * fy + fz * fz);
    }
  }
  best = best * 1.0001f + 1e-12f;   // the bound only has to stay above the true minimum
  // sweep 2: exact distances over the tiles some point of this CTA can still improve on
  for (int t0 = 0; t0 < T; t0 += kMeshThreads) {
    __syncthreads();
    if (t0 + threadIdx.x < T) {
#pragma unroll
      for (int k = 0; k < 6; ++k) boxes[threadIdx.x][k] = m.box[int64_t(t0 + threadIdx.x) * 6 + k];
    }
    __syncthreads();
    const int n = min(kMeshThreads, T - t0);
    for (int t = 0; t < n; ++t) {
      const float gx = fmaxf(fmaxf(boxes[t][0] - px, px - boxes[t][3]), 0.f);
      const float gy = fmaxf(fmaxf(boxes[t][1] - py, py - boxes[t][4]), 0.f);
      const float gz = fmaxf(fmaxf(boxes[t][2] - pz, pz - boxes[t][5]), 0.f);
      // slack above the rounding error of a computed triangle distance (about 1e-7 absolute on the distance): a tile
      // is skipped only if none of its triangles could come out below `best`, so the minimum -- bit for bit -- does
      // not depend on the order of tiles or points
      const bool need = live && (gx * gx + gy * gy + gz * gz <= fmaf(best, 1.001f, 1e-9f));
      if (!__syncthreads_or(need)) continue;
      const float4* src = m.tri + int64_t(t0 + t) * kMeshTile * 3;
      for (int e = threadIdx.x; e < kMeshTile * 3; e += kMeshThreads) tile[e] = __ldg(src + e);
      __syncthreads();
      if (need) {
#pragma unroll 4
        for (int j = 0; j < kMeshTile; ++j)
          best = fminf(best, tri_dist2(px, py, pz, tile[3 * j], tile[3 * j + 1], tile[3 * j + 2]));   // fminf drops NaN
      }
      __syncthreads();
    }
  }
  if (live) {
    const float dd = sqrtf(best);
    dist[i] = dd;
    if (speed) speed[i] = fminf(fmaxf(dd, offset), margin) / margin;
  }
}

static size_t mesh_tiles(int64_t M) { return size_t((M + kMeshTile - 1) / kMeshTile); }
static size_t mesh_tri_bytes(int64_t M) { return (mesh_tiles(M) * kMeshTile * 3 * sizeof(float4) + 255) / 256 * 256; }

}  // namespace pntf

using namespace pntf;

extern "C" size_t pntf_mesh_workspace_bytes(int64_t M) {
  if (M <= 0) return 0;
  return mesh_tri_bytes(M) + mesh_tiles(M) * 6 * sizeof(float);
}

extern "C" int pntf_mesh_distance(const float* triangles, int64_t M, const float* points, int64_t N, float* distance,
                                  float offset, float margin, float* speed, void* workspace, size_t workspace_bytes,
                                  void* stream) {
  if (M <= 0 || N < 0) return fail(PNTF_ERR_ARG, "pntf_mesh_distance: need M > 0 triangles and N >= 0 points");
  if (N == 0) return PNTF_OK;
  if (!triangles || !points || !distance) return fail(PNTF_ERR_ARG, "pntf_mesh_distance: null pointer");
  if (speed && !(margin > 0.f)) return fail(PNTF_ERR_ARG, "pntf_mesh_distance: margin must be positive");
  if (mesh_tiles(M) > size_t(INT32_MAX)) return fail(PNTF_ERR_ARG, "pntf_mesh_distance: mesh too large");
  const size_t need = pntf_mesh_workspace_bytes(M);
  if (!workspace || workspace_bytes < need)
    return fail(PNTF_ERR_WORKSPACE, "workspace too small (%zu < %zu)", workspace_bytes, need);
  if ((reinterpret_cast<uintptr_t>(workspace) & 255) != 0) return fail(PNTF_ERR_ARG, "workspace must be 256-byte aligned");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  MeshTiles m;
  m.tri = static_cast<float4*>(workspace);
  m.box = reinterpret_cast<float*>(static_cast<uint8_t*>(workspace) + mesh_tri_bytes(M));
  const int T = int(mesh_tiles(M));
  mesh_prep_kernel<<<T, kMeshTile, 0, s>>>(triangles, M, m);
  if (int rc = launch_check("mesh_prep")) return rc;
  const int64_t blocks = (N + kMeshThreads - 1) / kMeshThreads;
  if (blocks > INT32_MAX) return fail(PNTF_ERR_ARG, "pntf_mesh_distance: too many points");
  mesh_distance_kernel<<<int(blocks), kMeshThreads, 0, s>>>(m, T, points, N, distance, offset, margin, speed);
  return launch_check("mesh_distance");
}
