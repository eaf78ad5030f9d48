// tile_pass.cuh -- the one kernel skeleton every element-parallel operation uses.
//
// k_tile<D, Op>:   one CTA per tile
//   1. stage the tile's vertices (Op::load_vertex) in shared memory,
//   2. every thread computes Op::element for its elements and stages the
//      per-element contribution matrix in shared memory (plane-major, so the
//      stores are conflict-free),
//   3. every thread gathers the contributions of one local vertex through the
//      tile's jagged-diagonal incidence list in a FIXED order (deterministic --
//      no float atomics), then either finishes the vertex (Op::vertex, interior
//      vertices: touched by this tile only) or writes one partial-sum slot
//      (boundary vertices).
// k_boundary<D, Op>: one thread per boundary vertex sums its partial slots in
//   slot order and finishes the vertex with the same Op::vertex.
#pragma once
#include "dev.cuh"
#include "launch.h"

#include <stdexcept>
#include <string>

namespace femb {

enum GatherMode {
  G_STAR = 0,     // corner j>=1 gets column j-1 of the staged D x D matrix, corner 0 gets -(sum of columns)
  G_UNIFORM = 1   // every corner gets the staged D-vector, corner 0 weighted by Op::w0()
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Deterministic CTA reduction: lanes -> warp (shuffle tree) -> warp 0 sums the
// warp partials in warp order.  Result valid in thread 0.
__device__ __forceinline__ float block_sum(float v, float* s_warp) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  if (lane == 0) s_warp[warp] = v;
  __syncthreads();
  float r = 0.f;
  if (warp == 0) {
    const int nw = (blockDim.x + 31) >> 5;
    r = lane < nw ? s_warp[lane] : 0.f;
    r = warp_sum(r);
  }
  return r;
}

template <int D, int MODE, int NV>
__device__ __forceinline__ void gather_local_vertex(const DevTiling& tl, const TileDesc& td, const float* stage,
                                                    int l, float w0, float (&acc)[D]) {
  const int S = tl.stage_stride;
#pragma unroll
  for (int i = 0; i < D; ++i) acc[i] = 0.f;
  const uint16_t* jds = tl.jds + td.jds_off;
  const uint32_t* rounds = tl.rounds + td.rounds_off;
  int off = 0;
  for (int k = 0; k < td.n_rounds; ++k) {
    const int cnt = (int)rounds[k];
    if (l >= cnt) break;  // valences are sorted descending: no later round includes l
    const uint32_t ent = jds[off + l];
    const int e = ent & 0xfffu;
    const int c = ent >> 12;
    if (MODE == G_STAR) {
      if (c) {
        const float* s = stage + (c - 1) * D * S + e;
#pragma unroll
        for (int i = 0; i < D; ++i) acc[i] += s[i * S];
      } else {
#pragma unroll
        for (int i = 0; i < D; ++i) {
          float t = stage[i * S + e];
#pragma unroll
          for (int j = 1; j < D; ++j) t += stage[(j * D + i) * S + e];
          acc[i] -= t;
        }
      }
    } else {
      const float w = c ? 1.f : w0;
#pragma unroll
      for (int i = 0; i < NV; ++i) acc[i] += w * stage[i * S + e];
    }
    off += cnt;
  }
}

template <int D, class Op>
__global__ void __launch_bounds__(Op::BLOCK, Op::MIN_CTAS) k_tile(DevTiling tl, typename Op::Args a) {
  extern __shared__ float4 smem4[];
  __shared__ float s_warp[32];
  if (Op::early_exit(a)) return;
  float4* sv = smem4;
  float* stage = reinterpret_cast<float*>(smem4 + tl.max_lv);
  const TileDesc td = tl.tiles[blockIdx.x];
  const int S = tl.stage_stride;
  const LocalVert* lvt = tl.lv + td.lv_off;

  for (int l = threadIdx.x; l < td.n_lv; l += blockDim.x) sv[l] = Op::load_vertex(a, lvt[l].gid);
  __syncthreads();

  float red = 0.f;
  for (int e = threadIdx.x; e < td.n_elem; e += blockDim.x) {
    float out[Op::NSTAGE];
    Op::element(a, td.slot_off + e, sv, out, red);
#pragma unroll
    for (int k = 0; k < Op::NSTAGE; ++k) stage[k * S + e] = out[k];
  }
  __syncthreads();

  for (int l = threadIdx.x; l < td.n_lv; l += blockDim.x) {
    float acc[D];
    gather_local_vertex<D, Op::GATHER, (Op::NSTAGE < D ? Op::NSTAGE : D)>(tl, td, stage, l, Op::w0(), acc);
    const LocalVert v = lvt[l];
    if (v.part < 0) {
      Op::vertex(a, v.gid, sv[l], acc, red);
    } else {
      float4 p;
      p.x = acc[0]; p.y = acc[1]; p.z = D == 3 ? acc[D - 1] : 0.f; p.w = 0.f;
      Op::partial_buf(a)[v.part] = p;
    }
  }
  if (Op::REDUCE) {
    const float r = block_sum(red, s_warp);
    if (threadIdx.x == 0) Op::red_buf(a)[blockIdx.x] = (double)r;
  }
}

constexpr int kBoundaryBlock = 256;

template <int D, class Op>
__global__ void __launch_bounds__(kBoundaryBlock) k_boundary(DevTiling tl, typename Op::Args a, int red_base) {
  __shared__ float s_warp[32];
  if (Op::early_exit(a)) return;
  float red = 0.f;
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < tl.n_boundary) {
    const float4* part = Op::partial_buf(a);
    const int s0 = tl.bptr[b], s1 = tl.bptr[b + 1];
    float acc[D];
#pragma unroll
    for (int i = 0; i < D; ++i) acc[i] = 0.f;
    for (int s = s0; s < s1; ++s) {
      const float4 p = part[s];
      acc[0] += p.x;
      acc[1] += p.y;
      if (D == 3) acc[D - 1] += p.z;
    }
    const int gid = tl.n_interior + b;
    Op::vertex(a, gid, Op::load_vertex(a, gid), acc, red);
  }
  if (Op::REDUCE) {
    const float r = block_sum(red, s_warp);
    if (threadIdx.x == 0) Op::red_buf(a)[red_base + blockIdx.x] = (double)r;
  }
}

template <int D, class Op>
inline void run_tile_pass(const char* name, const LaunchCfg& c, const DevTiling& tl, const typename Op::Args& a) {
  static int attr_set[64] = {};  // per device (the attribute lives in the device's context): bytes allowed so far
  if (c.smem > attr_set[c.device & 63]) {
    const cudaError_t e = cudaFuncSetAttribute(k_tile<D, Op>, cudaFuncAttributeMaxDynamicSharedMemorySize, c.smem);
    if (e != cudaSuccess) {
      cudaGetLastError();
      throw std::runtime_error(std::string(name) + ": cannot reserve shared memory: " + cudaGetErrorString(e));
    }
    attr_set[c.device & 63] = c.smem;
  }
  k_tile<D, Op><<<tl.n_tiles, Op::BLOCK, c.smem, c.stream>>>(tl, a);
  note_launch(name, c.stream);
  if (c.ev_mid) cudaEventRecord(c.ev_mid, c.stream);
  if (tl.n_boundary > 0) {
    const int nb = (tl.n_boundary + kBoundaryBlock - 1) / kBoundaryBlock;
    k_boundary<D, Op><<<nb, kBoundaryBlock, 0, c.stream>>>(tl, a, tl.n_tiles);
    note_launch(name, c.stream);
  }
}

}  // namespace femb
