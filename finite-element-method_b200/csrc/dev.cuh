// dev.cuh -- device-side views and small dense algebra shared by the kernels.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

#include "tiling.h"

namespace femb {

// Persistent tile kernels (per operation family): 1 = one CTA per SM slot loops over tiles t, t+G, t+2G, ...
// and overlaps the list / vertex fetches of its next tile with the arithmetic of the current one;
// 0 = one CTA per tile.  Measured on C4 (B200): the light operator passes gain ~8 %, the register-heavy
// explicit / assemble passes lose (spills under the 64-register cap of 4 CTAs/SM).
#ifndef FEMB_PERSIST_APPLY
#define FEMB_PERSIST_APPLY 1
#endif
#ifndef FEMB_PERSIST_EXPLICIT
#define FEMB_PERSIST_EXPLICIT 0
#endif
#ifndef FEMB_PERSIST_ASSEMBLE
#define FEMB_PERSIST_ASSEMBLE 0
#endif
// The diagnostics build (make diag -> femb200/libfemb200_diag.so, FEMB200_LIB selects it) adds the kernel timeline
// (femb_trace_enable, tools/trace_steps.py) and the "one tile kernel" exchange form (FEMB_PEER_MODE=one).  Their
// checks cost the explicit tile kernel ~35 instructions per CTA-thread: 1.4 % in a same-box A/B (0.2579 vs 0.2540 ms,
// profiles/r02_ab_old.log), so the default build carries neither.
#ifndef FEMB_DIAG_BUILD
#define FEMB_DIAG_BUILD 0
#endif
constexpr bool kDiagBuild = FEMB_DIAG_BUILD != 0;

// In-tile completion of boundary vertices (tile_pass.cuh: itc_arrive) is compiled in only on request: measured slower
// when used (profiles/r02_ab_itc.log), and its mere presence cost the persistent operator kernel 7 % (0.2352 ->
// 0.2517 ms per pass on C4: register pressure, a 32-byte spill) and the explicit kernel 1.6 %, found by a same-box
// A/B against the build before it (profiles/r02_ab_old.log).  make EXTRA=-DFEMB_ITC_BUILD=1, then FEMB_ITC=1.
#ifndef FEMB_ITC_BUILD
#define FEMB_ITC_BUILD 0
#endif
constexpr bool kItcBuild = FEMB_ITC_BUILD != 0;

// Two elements per thread in packed fp32 (FFMA2) for the explicit element kernels, 128-thread CTAs, element planes
// read from a pair-interleaved copy.  Measured on C4 (B200, same box): 25 % fewer warp instructions (129.5 M vs
// 173.1 M), issue utilisation 45 % vs 61 %, but half the warps per SM -- 0.2807 ms per step against 0.2752 for the
// one-element-per-thread kernel; with the persistent tile loop on top 0.2760 (tile kernel alone 0.2548 vs 0.2597).
// The three forms land within 2 % of each other, so the default stays the scalar kernel; FEMB_PAIR_EXPLICIT=1
// (optionally with FEMB_PERSIST_EXPLICIT=1) builds the packed one, which passes the same GPU parity suite.
#ifndef FEMB_PAIR_EXPLICIT
#define FEMB_PAIR_EXPLICIT 0
#endif
// ... with the element planes of the thread's next pair requested before the current pair is worked on
#ifndef FEMB_PAIR_PREFETCH
#define FEMB_PAIR_PREFETCH 1
#endif

// Device tile descriptor: the host TileDesc plus what the owning CTA needs to know about ITS NEXT tile
// (tile index + grid size) before that tile's own descriptor has been fetched.
struct DevTile {           // 16 x int32 = 64 bytes, 16-byte aligned (bulk-copied)
  int32_t slot_off, n_elem, lv_off, n_lv, jds_off, rounds_off, n_rounds, n_int, int_base;
  int32_t nx_slot_off, nx_n_elem, nx_lv_off, nx_n_lv, nx_jds_off, nx_int_base, nx_n_int;
};

struct HaloDev;   // launch.h: device-resident description of the peer-memory halo exchange

// Device view of the tiling + element planes (all pointers are device memory).
struct DevTiling {
  const DevTile* tiles;
  const LocalVert* lv;
  const uint16_t* jds;
  const uint32_t* rounds;
  const int32_t* bptr;   // [n_boundary+1]
  int n_tiles, n_interior, n_boundary, n_vert, n_slots;
  int stage_stride;      // floats per staging plane (>= max tile elems, multiple of 32)
  int max_lv;
  int max_rounds;
  int nc;                // corners per element (dim + 1)
  int prefetch_dist;     // tiles ahead whose element planes are prefetched into L2 (0 = off)
  int grid;              // CTAs of the persistent tile kernels (SM slots)
  int n_halo_tiles;      // multi-GPU: tiles [0, n_halo_tiles) touch a vertex shared with another rank
  const uint8_t* bskip;  // [n_boundary] multi-GPU: 1 = shared vertex counted by ANOTHER rank in global
                         // reductions (dots, norms); nullptr on single-GPU objects
  // multi-GPU, set per launch (run_tile_pass) when the halo exchange runs INSIDE the pass's own kernels:
  const HaloDev* hd;     // nullptr: no exchange in this launch
  int n_send;            // tile kernel: its first n_send CTAs send the halo sums of the (finished) halo tiles
  int send_at;           // -1: the senders are CTAs [0, n_send) of the SECOND launch of a split pass (released by
                         // griddepcontrol.wait); >= 0: one launch over all tiles, senders are CTAs [send_at, send_at +
                         // n_send) and wait until the halo tiles -- all of lower CTA index -- have counted themselves done
  int fuse_recv;         // boundary kernel: wait for the neighbours and add their sums itself
  // in-tile completion of boundary vertices (tile_pass.cuh: itc_arrive); nullptr = off (FEMB_ITC=0)
  int32_t* bcount;       // [n_boundary] arrival counters, armed with -(partial slots); shared-with-a-rank: never reached
  const int32_t* halo_b; // [n_halo_b] boundary indices of the vertices shared with another rank
  int n_halo_b;
  // kernel timeline (femb_trace_enable; diagnostics only): 8 words per launch, set per launch by run_tile_pass --
  // [0] start of CTA 0, [1] latest CTA end, [2] latest end of a wait for the neighbours, [3] halo flags released,
  // [4..6] first sender CTA: released / values stored / fence done
  unsigned long long* trace;
};

// Shared-memory layout of k_tile (byte offsets; every region is 16-byte aligned):
//   sv      [max_lv] float4        staged vertex values
//   lvt     [max_lv] LocalVert     bulk-copied local-vertex table
//   rounds  [max_rounds] uint32    bulk-copied JDS round counts
//   jds     [nc*stage_stride] u16  bulk-copied JDS entries
//   stage   [n_planes][stage_stride] float
//   sv2     [max_lv] float4        (two_sv only) vertex values of the CTA's NEXT tile, filled by cp.async
struct TileSmem {
  int lvt, lvt2, rounds, jds, stage, sv2, total;   // lvt2: second local-vertex table (the CTA's next tile)
};
__host__ __device__ inline TileSmem tile_smem_layout(int max_lv, int max_rounds, int nc, int n_planes,
                                                     bool two_sv = false) {
  TileSmem s;
  int off = max_lv * 16;
  s.lvt = off;    off += (max_lv * 8 + 15) / 16 * 16;
  s.lvt2 = off;   off += (max_lv * 8 + 15) / 16 * 16;
  s.rounds = off; off += (max_rounds * 4 + 15) / 16 * 16;
  s.jds = off;    off += (nc * kStride * 2 + 15) / 16 * 16;
  s.stage = off;  off += n_planes * kStride * 4;
  s.sv2 = off;    off += two_sv ? max_lv * 16 : 0;   // last: the other offsets do not depend on it
  s.total = off;
  return s;
}

// Element planes, one float4 per slot per plane.
//   3D  e0 = {idx01, idx23, V,   r00}  e1 = {r01, r02, r10, r11}  e2 = {r12, r20, r21, r22}
//   2D  e0 = {idx01, idx2,  V,   r00}  e1 = {r01, r10, r11, 0}
//   3D  h0 = {idx01, idx23, h00, h01}  h1 = {h02, h10, h11, h12}  h2 = {h20, h21, h22, 0}
//   2D  h0 = {idx01, idx2,  h00, h01}  h1 = {h10, h11, 0,   0}
struct Planes {
  float4* p0;
  float4* p1;
  float4* p2;
};

struct Blocks {            // circle_blocks.py:6-25
  const float4* data;      // {cx, cy, cz, radius}
  int count;
};

struct Reduce {            // deterministic two-stage reductions
  double* partial;         // [n_partial_blocks] per-CTA partial sums
  int n;
};

template <int D>
struct Mat {
  float m[D][D];
};

template <int D>
__device__ __forceinline__ float det(const Mat<D>& a);

template <>
__device__ __forceinline__ float det<2>(const Mat<2>& a) {
  return a.m[0][0] * a.m[1][1] - a.m[0][1] * a.m[1][0];
}

// cofactor matrix C with inverse(A) = C^T / det(A), i.e. A^-T = C / det
template <int D>
__device__ __forceinline__ void cofactor(const Mat<D>& a, Mat<D>& c);

template <>
__device__ __forceinline__ void cofactor<2>(const Mat<2>& a, Mat<2>& c) {
  c.m[0][0] = a.m[1][1];
  c.m[0][1] = -a.m[1][0];
  c.m[1][0] = -a.m[0][1];
  c.m[1][1] = a.m[0][0];
}

template <>
__device__ __forceinline__ void cofactor<3>(const Mat<3>& a, Mat<3>& c) {
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
      c.m[i][j] = a.m[i1][j1] * a.m[i2][j2] - a.m[i2][j1] * a.m[i1][j2];
    }
}

template <>
__device__ __forceinline__ float det<3>(const Mat<3>& a) {
  return a.m[0][0] * (a.m[1][1] * a.m[2][2] - a.m[2][1] * a.m[1][2]) -
         a.m[1][0] * (a.m[0][1] * a.m[2][2] - a.m[2][1] * a.m[0][2]) +
         a.m[2][0] * (a.m[0][1] * a.m[1][2] - a.m[1][1] * a.m[0][2]);
}

// c = a @ b, k ascending
template <int D>
__device__ __forceinline__ void matmul(const Mat<D>& a, const Mat<D>& b, Mat<D>& c) {
#pragma unroll
  for (int i = 0; i < D; ++i)
#pragma unroll
    for (int j = 0; j < D; ++j) {
      float s = a.m[i][0] * b.m[0][j];
#pragma unroll
      for (int k = 1; k < D; ++k) s += a.m[i][k] * b.m[k][j];
      c.m[i][j] = s;
    }
}

// c = a @ b^T
template <int D>
__device__ __forceinline__ void matmul_nt(const Mat<D>& a, const Mat<D>& b, Mat<D>& c) {
#pragma unroll
  for (int i = 0; i < D; ++i)
#pragma unroll
    for (int j = 0; j < D; ++j) {
      float s = a.m[i][0] * b.m[j][0];
#pragma unroll
      for (int k = 1; k < D; ++k) s += a.m[i][k] * b.m[j][k];
      c.m[i][j] = s;
    }
}

// c = a^T @ b
template <int D>
__device__ __forceinline__ void matmul_tn(const Mat<D>& a, const Mat<D>& b, Mat<D>& c) {
#pragma unroll
  for (int i = 0; i < D; ++i)
#pragma unroll
    for (int j = 0; j < D; ++j) {
      float s = a.m[0][i] * b.m[0][j];
#pragma unroll
      for (int k = 1; k < D; ++k) s += a.m[k][i] * b.m[k][j];
      c.m[i][j] = s;
    }
}

// ---------------------------------------------------------------------------
// Two elements per thread in packed fp32 (sm_100: FFMA2 / FMUL2 / FADD2 work on a 64-bit register pair, one issue
// slot for two lanes; negated and scalar-broadcast operands are free modifiers).  Lane x is the thread's element
// e, lane y its element e + blockDim: the per-element algebra of the explicit path is issue-bound (61 % issue
// utilisation, FMA pipe 32 %), so halving its instruction count is worth more than the warps it costs.
// ---------------------------------------------------------------------------
struct f2 {
  float2 v;
};
__device__ __forceinline__ f2 pk(float a, float b) { f2 r; r.v = make_float2(a, b); return r; }
__device__ __forceinline__ f2 bc(float s) { return pk(s, s); }
__device__ __forceinline__ f2 operator-(f2 a) { return pk(-a.v.x, -a.v.y); }
__device__ __forceinline__ f2 operator+(f2 a, f2 b) { f2 r; r.v = __fadd2_rn(a.v, b.v); return r; }
__device__ __forceinline__ f2 operator-(f2 a, f2 b) { return a + (-b); }
__device__ __forceinline__ f2 operator*(f2 a, f2 b) { f2 r; r.v = __fmul2_rn(a.v, b.v); return r; }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) { f2 r; r.v = __ffma2_rn(a.v, b.v, c.v); return r; }

template <int D>
struct Mat2 {
  f2 m[D][D];
};
template <int D>
__device__ __forceinline__ void pack(const Mat<D>& a, const Mat<D>& b, Mat2<D>& c) {
#pragma unroll
  for (int i = 0; i < D; ++i)
#pragma unroll
    for (int j = 0; j < D; ++j) c.m[i][j] = pk(a.m[i][j], b.m[i][j]);
}
// c = a @ b, k ascending
template <int D>
__device__ __forceinline__ void matmul(const Mat2<D>& a, const Mat2<D>& b, Mat2<D>& c) {
#pragma unroll
  for (int i = 0; i < D; ++i)
#pragma unroll
    for (int j = 0; j < D; ++j) {
      f2 s = a.m[i][0] * b.m[0][j];
#pragma unroll
      for (int k = 1; k < D; ++k) s = fma2(a.m[i][k], b.m[k][j], s);
      c.m[i][j] = s;
    }
}
// c = a @ b^T
template <int D>
__device__ __forceinline__ void matmul_nt(const Mat2<D>& a, const Mat2<D>& b, Mat2<D>& c) {
#pragma unroll
  for (int i = 0; i < D; ++i)
#pragma unroll
    for (int j = 0; j < D; ++j) {
      f2 s = a.m[i][0] * b.m[j][0];
#pragma unroll
      for (int k = 1; k < D; ++k) s = fma2(a.m[i][k], b.m[j][k], s);
      c.m[i][j] = s;
    }
}
// cofactor matrix and determinant (expansion along column 0 of the cofactors)
template <int D>
__device__ __forceinline__ f2 cofactor_det(const Mat2<D>& a, Mat2<D>& c) {
  if constexpr (D == 2) {
    c.m[0][0] = a.m[1][1];
    c.m[0][1] = -a.m[1][0];
    c.m[1][0] = -a.m[0][1];
    c.m[1][1] = a.m[0][0];
    return fma2(a.m[0][0], a.m[1][1], -(a.m[0][1] * a.m[1][0]));
  } else {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
        c.m[i][j] = fma2(a.m[i1][j1], a.m[i2][j2], -(a.m[i2][j1] * a.m[i1][j2]));
      }
    return fma2(a.m[2][0], c.m[2][0], fma2(a.m[1][0], c.m[1][0], a.m[0][0] * c.m[0][0]));
  }
}

__device__ __forceinline__ float comp(const float4& v, int k) {
  return k == 0 ? v.x : (k == 1 ? v.y : (k == 2 ? v.z : v.w));
}

template <int D>
__device__ __forceinline__ void unpack_ref(const float4& e0, const float4& e1, const float4& e2, Mat<D>& r);

template <>
__device__ __forceinline__ void unpack_ref<3>(const float4& e0, const float4& e1, const float4& e2, Mat<3>& r) {
  r.m[0][0] = e0.w; r.m[0][1] = e1.x; r.m[0][2] = e1.y;
  r.m[1][0] = e1.z; r.m[1][1] = e1.w; r.m[1][2] = e2.x;
  r.m[2][0] = e2.y; r.m[2][1] = e2.z; r.m[2][2] = e2.w;
}

template <>
__device__ __forceinline__ void unpack_ref<2>(const float4& e0, const float4& e1, const float4&, Mat<2>& r) {
  r.m[0][0] = e0.w; r.m[0][1] = e1.x;
  r.m[1][0] = e1.y; r.m[1][1] = e1.z;
}

template <int D>
__device__ __forceinline__ void unpack_h(const float4& h0, const float4& h1, const float4& h2, Mat<D>& h);

template <>
__device__ __forceinline__ void unpack_h<3>(const float4& h0, const float4& h1, const float4& h2, Mat<3>& h) {
  h.m[0][0] = h0.z; h.m[0][1] = h0.w; h.m[0][2] = h1.x;
  h.m[1][0] = h1.y; h.m[1][1] = h1.z; h.m[1][2] = h1.w;
  h.m[2][0] = h2.x; h.m[2][1] = h2.y; h.m[2][2] = h2.z;
}

template <>
__device__ __forceinline__ void unpack_h<2>(const float4& h0, const float4& h1, const float4&, Mat<2>& h) {
  h.m[0][0] = h0.z; h.m[0][1] = h0.w;
  h.m[1][0] = h1.x; h.m[1][1] = h1.y;
}

// Ds[:, j] = p_{j+1} - p_0 from the staged tile vertices (explicit.py:11-16)
template <int D>
__device__ __forceinline__ void edge_matrix(const float4* sv, uint32_t w0, uint32_t w1, Mat<D>& x) {
  const float4 p0 = sv[w0 & 0xffffu];
  const float4 p1 = sv[w0 >> 16];
  const float4 p2 = sv[w1 & 0xffffu];
  x.m[0][0] = p1.x - p0.x; x.m[1][0] = p1.y - p0.y;
  x.m[0][1] = p2.x - p0.x; x.m[1][1] = p2.y - p0.y;
  if constexpr (D == 3) {
    const float4 p3 = sv[w1 >> 16];
    x.m[2][0] = p1.z - p0.z;
    x.m[2][1] = p2.z - p0.z;
    x.m[0][D - 1] = p3.x - p0.x; x.m[1][D - 1] = p3.y - p0.y; x.m[2][D - 1] = p3.z - p0.z;
  }
}

}  // namespace femb
