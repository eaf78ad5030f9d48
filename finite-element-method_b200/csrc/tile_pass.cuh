// tile_pass.cuh -- the one kernel skeleton every element-parallel operation uses.
//
// k_tile<D, Op>:   one CTA per tile
//   0. one thread issues TMA bulk copies of the tile's lists (local-vertex table,
//      JDS round counts, JDS entries) into shared memory and an L2 prefetch of
//      the element planes of the tile that will run on this SM slot next,
//   1. the tile's vertices (Op::load_vertex) are staged in shared memory,
//   2. every thread computes its elements (loads front-batched two elements deep)
//      and stages the per-element contribution matrix in shared memory
//      (plane-major: conflict-free stores),
//   3. every thread gathers the contributions of one local vertex through the
//      tile's jagged-diagonal incidence list in a FIXED order (deterministic --
//      no float atomics), then either finishes the vertex (Op::vertex, interior
//      vertices: touched by this tile only) or writes one partial-sum slot
//      (boundary vertices).
// k_boundary<D, Op>: one thread per boundary vertex sums its partial slots in
//   slot order and finishes the vertex with the same Op::vertex.
#pragma once
#include "dev.cuh"
#include "comm.cuh"
#include "launch.h"
#include "tma.cuh"

#include <algorithm>
#include <stdexcept>
#include <string>
#include <type_traits>

namespace femb {

enum GatherMode {
  G_STAR = 0,     // corner j>=1 gets column j-1 of the staged D x D matrix, corner 0 gets -(sum of columns)
  G_UNIFORM = 1   // every corner gets the staged vector, corner 0 weighted by Op::w0()
};

#ifndef FEMB_PDL
#define FEMB_PDL 0
#endif
// Programmatic dependent launch (off by default: measured no gain on C4 -- the register file is full of
// tile CTAs until the very tail): the tile kernels let the boundary kernel that follows them be scheduled
// as soon as every tile CTA has started; the boundary kernel waits (grid_dependency_wait) for their
// completion before it touches the partial sums.
__device__ __forceinline__ void grid_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
// end of a tile kernel: a kernel launched as programmatic dependent (second part of a split pass) may not
// retire before the kernel it overlapped; no-op otherwise
__device__ __forceinline__ void grid_dependency_wait_always() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
__device__ __forceinline__ void grid_dependency_wait() {
#if FEMB_PDL
  asm volatile("griddepcontrol.wait;" ::: "memory");
#endif
}

// kernel timeline (DevTiling::trace): the last launch's values survive (plain store of CTA 0's start, monotonic
// maxima for the ends), so after a run of steps the buffer holds the timeline of the final step
// (launches that exit early -- the solver has converged -- leave the words of the last live launch alone)
template <class Op>
__device__ __forceinline__ void trace_start(const DevTiling& tl, const typename Op::Args& a) {
  if (kDiagBuild && tl.trace && blockIdx.x == 0 && threadIdx.x == 0 && !Op::early_exit(a)) tl.trace[0] = gtime_ns();
}
template <class Op>
__device__ __forceinline__ void trace_end(const DevTiling& tl, const typename Op::Args& a) {
  if (kDiagBuild && tl.trace && threadIdx.x == 0 && !Op::early_exit(a)) atomicMax(tl.trace + 1, gtime_ns());
}

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

// the three float4 of one element slot, loaded with streaming (no L1 allocate) loads
struct PlaneElem {
  float4 a, b, c;
};

template <int D>
__device__ __forceinline__ void load_planes(const Planes& p, int slot, PlaneElem& d) {
  d.a = __ldcs(p.p0 + slot);
  d.b = __ldcs(p.p1 + slot);
  if constexpr (D == 3) d.c = __ldcs(p.p2 + slot);
  else d.c = make_float4(0.f, 0.f, 0.f, 0.f);
}

// Number of shared-memory staging planes of an Op: G_STAR ops stage the D x D contribution
// matrix plus its negated column sum (the share of corner 0), so every incidence reads D
// consecutive planes at a precomputed offset -- no branch on the corner in the gather.
// (A float4-per-(corner, element) layout with one 16-byte load per incidence was measured 8 % SLOWER
// on C4: random 16-byte shared loads cost >= 4 wavefronts each and the larger footprint shrinks the tiles.)
template <int D, class Op>
__host__ __device__ constexpr int n_planes() {
  return Op::GATHER == G_STAR ? Op::NSTAGE + D : Op::NSTAGE;
}

// stage one element's results plane-major (conflict-free stores)
template <int D, class Op>
__device__ __forceinline__ void stage_element(float* stage, int e, const float (&out)[Op::NSTAGE]) {
  constexpr int S = kStride;
  float* s = stage + e;
#pragma unroll
  for (int i = 0; i < Op::NSTAGE; ++i) s[i * S] = out[i];
  if constexpr (Op::GATHER == G_STAR) {   // share of corner 0: f0 = -(sum of columns)   (explicit.py:42-49)
#pragma unroll
    for (int i = 0; i < D; ++i) {
      float c = -out[i];                 // -(a + b + c) as (-a - b) - c: the negation rides on the operands
#pragma unroll
      for (int j = 1; j < D; ++j) c -= out[j * D + i];
      s[(D * D + i) * S] = c;
    }
  }
}

// the same for the two elements of a packed pair (lane x -> element e0, lane y -> element e1)
template <int D, class Op>
__device__ __forceinline__ void stage_pair(float* stage, int e0, int e1, const f2 (&out)[Op::NSTAGE]) {
  constexpr int S = kStride;
  float* s0 = stage + e0;
  float* s1 = stage + e1;
#pragma unroll
  for (int i = 0; i < Op::NSTAGE; ++i) {
    s0[i * S] = out[i].v.x;
    s1[i * S] = out[i].v.y;
  }
  if constexpr (Op::GATHER == G_STAR) {
#pragma unroll
    for (int i = 0; i < D; ++i) {
      f2 c = -out[i];
#pragma unroll
      for (int j = 1; j < D; ++j) c = c - out[j * D + i];
      s0[(D * D + i) * S] = c.v.x;
      s1[(D * D + i) * S] = c.v.y;
    }
  }
}

template <int D, int MODE, int NV>
__device__ __forceinline__ void gather_local_vertex(int n_rounds, const uint32_t* __restrict__ rounds,
                                                    const uint16_t* __restrict__ jds, const float* __restrict__ stage,
                                                    int l, float w0, float (&acc)[D]) {
  constexpr int S = kStride;
#pragma unroll
  for (int i = 0; i < D; ++i) acc[i] = 0.f;
  int off = l;
  // Four rounds per trip: the four round counts come in one LDS.128 (the table is padded with
  // zeros to a multiple of 4), the four entries and their 4*D staged values are independent
  // loads, and the adds keep the round order -> same sums as the one-round-at-a-time loop.
  const uint4* rounds4 = reinterpret_cast<const uint4*>(rounds);
  for (int k = 0; k < n_rounds; k += 4) {
    const uint4 c4 = rounds4[k >> 2];
    const int cnt[4] = {(int)c4.x, (int)c4.y, (int)c4.z, (int)c4.w};
    if (l >= cnt[0]) break;  // valences are sorted descending: no later round includes l
    uint32_t ent[4];
    bool on[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      on[j] = l < cnt[j];
      ent[j] = on[j] ? jds[off] : 0u;   // = jds_entry(D, e, c): word offset of the first plane to read
      off += cnt[j];
    }
    float v[4][D];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (MODE == G_STAR) {
#pragma unroll
        for (int i = 0; i < D; ++i) v[j][i] = on[j] ? stage[ent[j] + i * S] : 0.f;
      } else {
        const int e = ent[j] % S;
        const float w = ent[j] >= (uint32_t)(D * D * S) ? w0 : 1.f;   // corner 0 entries live past plane D*D
#pragma unroll
        for (int i = 0; i < D; ++i) v[j][i] = (on[j] && i < NV) ? w * stage[i * S + e] : 0.f;
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int i = 0; i < D; ++i) acc[i] += v[j][i];
  }
}

// The element phase of a tile: every thread works through its elements tid, tid + B, ... (d0 / d1: the planes of
// its first two, already requested by the caller) and stages the results plane-major.
template <int D, class Op>
__device__ __forceinline__ void element_phase(const typename Op::Args& a, int slot_off, int n_elem, const float4* sv,
                                              float* stage, typename Op::Elem& d0, typename Op::Elem& d1, float& red) {
  const int B = blockDim.x;
  auto one = [&](int e, const typename Op::Elem& d) {
    float out[Op::NSTAGE];
    Op::compute(a, slot_off + e, d, sv, out, red);
    stage_element<D, Op>(stage, e, out);
  };
  if constexpr (Op::PAIR) {
    // elements (e, e + BP) of a thread go through the packed arithmetic together; the trip count is a compile-time
    // bound (a tile holds at most kStride elements), so the loop unrolls and the next pair's planes are requested
    // into fresh registers before the current pair is worked on
    constexpr int BP = Op::BLOCK, NIT = (kStride + 2 * BP - 1) / (2 * BP);
#pragma unroll
    for (int it = 0; it < NIT; ++it) {
      const int e = (int)threadIdx.x + it * 2 * BP;
      if (e >= n_elem) break;
      const bool two = e + BP < n_elem;
      typename Op::Elem n0 = d0, n1 = d1;
      if (FEMB_PAIR_PREFETCH) {
        if (it + 1 < NIT) {
          if (e + 2 * BP < n_elem) Op::load(a, slot_off + e + 2 * BP, n0);
          if (e + 3 * BP < n_elem) Op::load(a, slot_off + e + 3 * BP, n1);
        }
      } else if (it > 0) {
        Op::load(a, slot_off + e, d0);
        if (two) Op::load(a, slot_off + e + BP, d1);
      }
      if (two) {
        f2 out[Op::NSTAGE];
        Op::compute2(a, slot_off + e, slot_off + e + BP, d0, d1, sv, out, red);
        stage_pair<D, Op>(stage, e, e + BP, out);
      } else {
        one(e, d0);
      }
      if (FEMB_PAIR_PREFETCH) {
        d0 = n0;
        d1 = n1;
      }
    }
  } else {
    for (int e = threadIdx.x; e < n_elem; e += 2 * B) {
      const bool two = e + B < n_elem;
      if (e != (int)threadIdx.x) {
        Op::load(a, slot_off + e, d0);
        if (two) Op::load(a, slot_off + e + B, d1);
      }
      one(e, d0);
      if (two) one(e + B, d1);
    }
  }
}

// In-tile completion (ITC) of boundary vertices.  Every boundary vertex b carries an arrival counter armed with
// -(number of partial slots).  A tile CTA that has stored its partial sum of b makes it visible (fence) and bumps
// the counter; the CTA whose bump finds -1 is the LAST of the tiles that touch b: it re-arms the counter, sums the
// slots IN SLOT ORDER (so the result does not depend on who arrived last -- bit-identical to k_boundary) and
// finishes the vertex with Op::vertex.  By then every touching tile has read the vertex's operands (they are
// staged at tile start), so writing its new state is safe; no other tile reads it.  The boundary kernel, its
// launch and its tail disappear; vertices shared with another RANK are armed with a count that is never reached
// and stay with the (tiny) halo boundary kernel, which also receives the neighbour's sums.
// Ops that accumulate a reduction in Op::vertex keep k_boundary (the split of the sum over CTAs would depend on
// arrival order); Op::itc(args) says which launches qualify.
template <int D, class Op>
__device__ __forceinline__ void itc_arrive(const DevTiling& tl, const typename Op::Args& a, int gid, const float4& vtx,
                                           float& red) {
  __threadfence();                                   // my partial slot is visible before the counter moves
  const int b = gid - tl.n_interior;
  if (atomicAdd(tl.bcount + b, 1) != -1) return;
  __threadfence();                                   // the other tiles' slots are visible after it has moved
  const int s0 = tl.bptr[b], s1 = tl.bptr[b + 1];
  tl.bcount[b] = s0 - s1;                            // re-armed for the next pass
  const float4* part = Op::partial_buf(a);
  const float4 pre = Op::pre(a, gid);
  float acc[D];
#pragma unroll
  for (int i = 0; i < D; ++i) acc[i] = 0.f;
  for (int s = s0; s < s1; s += 4) {                 // same order of adds as k_boundary
    float4 p[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) p[k] = s + k < s1 ? __ldcg(part + s + k) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (s + k < s1) {
        acc[0] += p[k].x;
        acc[1] += p[k].y;
        if (D == 3) acc[D - 1] += p[k].z;
      }
    }
  }
  Op::vertex(a, gid, vtx, pre, acc, red);
}

// Persistent form: CTA c owns tiles c, c+G, c+2G, ...  While it works on tile t it already has, in flight or
// landed, everything its next tile t' = t+G needs to start without a dependent global round trip:
//   iteration start : TMA {descriptor, local-vertex table} of t' -> second buffer; TMA {JDS lists} of t;
//                     L2 prefetch of the element planes / JDS / interior vertices of t'
//   before the gather of t : vertex values of t' requested into registers (their table has landed)
//   next iteration  : registers -> shared memory, first element loads already issued.
template <int D, class Op>
__global__ void __launch_bounds__(Op::BLOCK, Op::MIN_CTAS) k_tile_persistent(DevTiling tl, typename Op::Args a, int t_begin,
                                                                           int t_end, int red_off) {
  extern __shared__ float4 smem4[];
  __shared__ float s_warp[32];
  __shared__ uint64_t s_bar[3];          // [0],[1]: descriptor + local-vertex table buffers; [2]: JDS lists
  __shared__ __align__(16) DevTile s_td[2];
  grid_launch_dependents();
  trace_start<Op>(tl, a);
  const int n_send = tl.n_send;
  if ((int)blockIdx.x < n_send) {   // sender CTA of a split multi-GPU pass (see run_tile_pass)
    grid_dependency_wait_always();  // the halo tiles (the kernel this one overlaps) are complete and visible
    halo_send_cta(tl.hd->pc, tl.hd->n, tl.hd->hrange, Op::partial_buf(a), blockIdx.x, n_send, tl.trace);
    trace_end<Op>(tl, a);
    return;
  }
  if (Op::early_exit(a)) return;
  const bool itc = kItcBuild && tl.bcount != nullptr && Op::itc(a);
  const int bid = (int)blockIdx.x - n_send;
  int t = t_begin + bid;
  if (t >= t_end) return;
  constexpr int S = kStride;
  constexpr bool AV = Op::ASYNC_VERTEX;   // next tile's vertex values by cp.async into a second buffer
  const TileSmem lay = tile_smem_layout(tl.max_lv, tl.max_rounds, tl.nc, n_planes<D, Op>(), AV);
  char* smem = reinterpret_cast<char*>(smem4);
  float4* const sv0 = smem4;
  float4* const sv1 = AV ? reinterpret_cast<float4*>(smem + lay.sv2) : smem4;
  auto lv_buf = [&](int i) { return reinterpret_cast<LocalVert*>(smem + (i ? lay.lvt2 : lay.lvt)); };
  uint32_t* s_rounds = reinterpret_cast<uint32_t*>(smem + lay.rounds);
  uint16_t* s_jds = reinterpret_cast<uint16_t*>(smem + lay.jds);
  float* stage = reinterpret_cast<float*>(smem + lay.stage);
  const int B = blockDim.x, G = (int)gridDim.x - n_send, tid = threadIdx.x;

  if (tid == 0) {
    mbar_init(&s_bar[0], 1);
    mbar_init(&s_bar[1], 1);
    mbar_init(&s_bar[2], 1);
    mbar_fence_init();
    const DevTile d0 = tl.tiles[t];      // the only exposed descriptor fetch of this CTA
    s_td[0] = d0;
    const uint32_t lv_bytes = (uint32_t)round_up(d0.n_lv * 8, 16);
    mbar_expect_tx(&s_bar[0], lv_bytes);
    bulk_g2s(lv_buf(0), tl.lv + d0.lv_off, lv_bytes, &s_bar[0]);
  }
  __syncthreads();
  mbar_wait(&s_bar[0], 0);
  // vertex values of the tile about to be processed: two per thread in registers, the rest (tiles with
  // more than 2*B local vertices: irregular meshes only) fetched directly at staging time
  float4 vreg0 = make_float4(0.f, 0.f, 0.f, 0.f), vreg1 = vreg0;
  float red = 0.f;
  constexpr bool EARLY = Op::PAIR;       // register budget of the 128-thread packed kernels allows it
  typename Op::Elem d0, d1;
  if constexpr (AV) {
    const int n_lv = s_td[0].n_lv;
    for (int l = tid; l < n_lv; l += B) cp_async16(&sv0[l], Op::vertex_src(a) + lv_buf(0)[l].gid);
    cp_async_commit();
  } else {
    const int n_lv = s_td[0].n_lv;
    if (tid < n_lv) vreg0 = Op::load_vertex(a, lv_buf(0)[tid].gid);
    if (tid + B < n_lv) vreg1 = Op::load_vertex(a, lv_buf(0)[tid + B].gid);
  }

  for (int k = 0;; ++k) {
    const int b = k & 1;
    const DevTile& td = s_td[b];          // read on demand from shared memory (keeps registers for the math)
    const LocalVert* s_lv = lv_buf(b);
    float4* const sv = (AV && b) ? sv1 : sv0;
    const int n_elem = td.n_elem, slot_off = td.slot_off;
    const int tn = t + G;
    const bool has_next = tn < t_end;
    if (tid == 0) {
      const uint32_t r_bytes = (uint32_t)round_up(td.n_rounds * 4, 16);
      const uint32_t j_bytes = (uint32_t)round_up(td.n_elem * tl.nc * 2, 16);
      mbar_expect_tx(&s_bar[2], r_bytes + j_bytes);
      bulk_g2s(s_rounds, tl.rounds + td.rounds_off, r_bytes, &s_bar[2]);
      bulk_g2s(s_jds, tl.jds + td.jds_off, j_bytes, &s_bar[2]);
      if (has_next) {
        const uint32_t lv_bytes = (uint32_t)round_up(td.nx_n_lv * 8, 16);
        mbar_expect_tx(&s_bar[b ^ 1], lv_bytes + (uint32_t)sizeof(DevTile));
        bulk_g2s(&s_td[b ^ 1], tl.tiles + tn, (uint32_t)sizeof(DevTile), &s_bar[b ^ 1]);
        bulk_g2s(lv_buf(b ^ 1), tl.lv + td.nx_lv_off, lv_bytes, &s_bar[b ^ 1]);
        Op::prefetch(a, td.nx_slot_off, (uint32_t)(round_up(td.nx_n_elem, 32) * 16));
        bulk_prefetch_l2(tl.jds + td.nx_jds_off, (uint32_t)round_up(td.nx_n_elem * tl.nc * 2, 16));
        if (td.nx_n_int > 0) Op::prefetch_verts(a, td.nx_int_base, (uint32_t)(td.nx_n_int * 16));
      }
    }
    // element data of the first two elements of this thread: in flight while the vertices are staged
    // (EARLY: already requested before the previous tile's gather)
    if (!EARLY || k == 0) {
      if (tid < n_elem) Op::load(a, slot_off + tid, d0);
      if (tid + B < n_elem) Op::load(a, slot_off + tid + B, d1);
    }

    if constexpr (AV) {
      cp_async_wait_all();               // this thread's copies have landed; the barrier publishes everyone's
    } else {
      if (tid < td.n_lv) sv[tid] = vreg0;
      if (tid + B < td.n_lv) sv[tid + B] = vreg1;
      for (int l = tid + 2 * B; l < td.n_lv; l += B) sv[l] = Op::load_vertex(a, s_lv[l].gid);
    }
    __syncthreads();

    element_phase<D, Op>(a, slot_off, n_elem, sv, stage, d0, d1, red);
    __syncthreads();

    // the next tile's vertex values: its local-vertex table landed during the element phase
    if (has_next) {
      mbar_wait(&s_bar[b ^ 1], (uint32_t)(((k + 1) >> 1) & 1));
      const int n_lv = s_td[b ^ 1].n_lv;
      const LocalVert* nlv = lv_buf(b ^ 1);
      if constexpr (AV) {
        // the other vertex buffer held tile k-1, whose gather ended before the barrier that closed iteration k-1
        float4* const svn = b ? sv0 : sv1;
        for (int l = tid; l < n_lv; l += B) cp_async16(&svn[l], Op::vertex_src(a) + nlv[l].gid);
        cp_async_commit();
      } else {
        if (tid < n_lv) vreg0 = Op::load_vertex(a, nlv[tid].gid);
        if (tid + B < n_lv) vreg1 = Op::load_vertex(a, nlv[tid + B].gid);
      }
      if (EARLY) {   // the next tile's first element planes cross HBM -> registers behind this tile's gather
        if (tid < td.nx_n_elem) Op::load(a, td.nx_slot_off + tid, d0);
        if (tid + B < td.nx_n_elem) Op::load(a, td.nx_slot_off + tid + B, d1);
      }
    }
    mbar_wait(&s_bar[2], (uint32_t)(k & 1));

    for (int l = tid; l < td.n_lv; l += B) {
      const LocalVert v = s_lv[l];
      // per-vertex global operand of the epilogue (e.g. vel): requested before the gather hides its latency
      float4 pre = make_float4(0.f, 0.f, 0.f, 0.f);
      if (v.part < 0) pre = Op::pre(a, v.gid);
      float acc[D];
      gather_local_vertex<D, Op::GATHER, (Op::NSTAGE < D ? Op::NSTAGE : D)>(td.n_rounds, s_rounds, s_jds, stage, l,
                                                                           Op::w0(), acc);
      if (v.part < 0) {
        Op::vertex(a, v.gid, sv[l], pre, acc, red);
      } else {
        float4 p;
        p.x = acc[0]; p.y = acc[1]; p.z = D == 3 ? acc[D - 1] : 0.f; p.w = 0.f;
        Op::partial_buf(a)[v.part] = p;
        if (itc) itc_arrive<D, Op>(tl, a, v.gid, sv[l], red);
      }
    }
    if (!has_next) break;
    t = tn;
    __syncthreads();   // staging planes, vertex values and JDS lists are free for the next tile
  }
  if (Op::REDUCE) {
    const float r = block_sum(red, s_warp);
    if (tid == 0) Op::red_buf(a)[red_off + bid] = (double)r;
  }
  trace_end<Op>(tl, a);
  grid_dependency_wait_always();
}

// One CTA per tile.
template <int D, class Op>
__global__ void __launch_bounds__(Op::BLOCK, Op::MIN_CTAS) k_tile(DevTiling tl, typename Op::Args a, int t_begin, int t_end,
                                                               int red_off) {
  extern __shared__ float4 smem4[];
  __shared__ float s_warp[32];
  __shared__ uint64_t s_bar[2];   // [0]: local-vertex table landed, [1]: JDS lists landed
  grid_launch_dependents();
  trace_start<Op>(tl, a);
  const int n_send = tl.n_send, send_at = tl.send_at;
  const int bx = (int)blockIdx.x;
  // one launch over all tiles, senders in the middle of the grid (diagnostics build only, dev.cuh)
  const bool counted = kDiagBuild && send_at >= 0 && n_send > 0;
  if (n_send > 0 && (counted ? (bx >= send_at && bx < send_at + n_send) : bx < n_send)) {
    // sender CTA of a multi-GPU pass (see run_tile_pass): the halo tiles are complete and visible
    if (counted) halo_tiles_wait(tl.hd->pc, tl.n_halo_tiles);
    else grid_dependency_wait_always();            // (they were the kernel this one overlaps)
    halo_send_cta(tl.hd->pc, tl.hd->n, tl.hd->hrange, Op::partial_buf(a), counted ? bx - send_at : bx, n_send,
                  tl.trace, counted);
    trace_end<Op>(tl, a);
    return;
  }
  const int bid = bx - ((counted ? bx >= send_at : true) ? n_send : 0);
  const int tile = t_begin + bid;
  if (Op::early_exit(a)) {
    if (counted && tile < tl.n_halo_tiles) halo_tile_done(tl.hd->pc);   // the senders still run: every rank keeps counting
    return;
  }
  const bool itc = kItcBuild && tl.bcount != nullptr && Op::itc(a);
  const DevTile td = tl.tiles[tile];
  constexpr int S = kStride;
  const TileSmem lay = tile_smem_layout(tl.max_lv, tl.max_rounds, tl.nc, n_planes<D, Op>());
  char* smem = reinterpret_cast<char*>(smem4);
  float4* sv = smem4;
  LocalVert* s_lv = reinterpret_cast<LocalVert*>(smem + lay.lvt);
  uint32_t* s_rounds = reinterpret_cast<uint32_t*>(smem + lay.rounds);
  uint16_t* s_jds = reinterpret_cast<uint16_t*>(smem + lay.jds);
  float* stage = reinterpret_cast<float*>(smem + lay.stage);

  if (threadIdx.x == 0) {
    mbar_init(&s_bar[0], 1);
    mbar_init(&s_bar[1], 1);
    mbar_fence_init();
    const uint32_t lv_bytes = (uint32_t)round_up(td.n_lv * 8, 16);
    const uint32_t r_bytes = (uint32_t)round_up(td.n_rounds * 4, 16);
    const uint32_t j_bytes = (uint32_t)round_up(td.n_elem * tl.nc * 2, 16);
    mbar_expect_tx(&s_bar[0], lv_bytes);
    bulk_g2s(s_lv, tl.lv + td.lv_off, lv_bytes, &s_bar[0]);
    mbar_expect_tx(&s_bar[1], r_bytes + j_bytes);
    bulk_g2s(s_rounds, tl.rounds + td.rounds_off, r_bytes, &s_bar[1]);
    bulk_g2s(s_jds, tl.jds + td.jds_off, j_bytes, &s_bar[1]);
    // L2 prefetch of the tile that will occupy this SM slot after us
    const int nxt = tile + tl.prefetch_dist;
    if (tl.prefetch_dist > 0 && nxt < t_end) {
      const DevTile tn = tl.tiles[nxt];
      Op::prefetch(a, tn.slot_off, (uint32_t)(round_up(tn.n_elem, 32) * 16));
      bulk_prefetch_l2(tl.lv + tn.lv_off, (uint32_t)round_up(tn.n_lv * 8, 16));
      bulk_prefetch_l2(tl.jds + tn.jds_off, (uint32_t)round_up(tn.n_elem * tl.nc * 2, 16));
      if (tn.n_int > 0) Op::prefetch_verts(a, tn.int_base, (uint32_t)(tn.n_int * 16));
    }
  }
  const int B = blockDim.x;
  // element data of the first two elements of this thread: in flight while the vertices are staged
  typename Op::Elem d0, d1;
  if ((int)threadIdx.x < td.n_elem) Op::load(a, td.slot_off + threadIdx.x, d0);
  if ((int)threadIdx.x + B < td.n_elem) Op::load(a, td.slot_off + threadIdx.x + B, d1);

  __syncthreads();          // barrier words initialised before anyone polls them
  mbar_wait(&s_bar[0], 0);

  float red = 0.f;
  for (int l = threadIdx.x; l < td.n_lv; l += B) sv[l] = Op::load_vertex(a, s_lv[l].gid);
  __syncthreads();

  element_phase<D, Op>(a, td.slot_off, td.n_elem, sv, stage, d0, d1, red);
  __syncthreads();
  mbar_wait(&s_bar[1], 0);

  for (int l = threadIdx.x; l < td.n_lv; l += B) {
    const LocalVert v = s_lv[l];
    // per-vertex global operand of the epilogue (e.g. vel): requested before the gather hides its latency
    float4 pre = make_float4(0.f, 0.f, 0.f, 0.f);
    if (v.part < 0) pre = Op::pre(a, v.gid);
    float acc[D];
    gather_local_vertex<D, Op::GATHER, (Op::NSTAGE < D ? Op::NSTAGE : D)>(td.n_rounds, s_rounds, s_jds, stage, l,
                                                                         Op::w0(), acc);
    if (v.part < 0) {
      Op::vertex(a, v.gid, sv[l], pre, acc, red);
    } else {
      float4 p;
      p.x = acc[0]; p.y = acc[1]; p.z = D == 3 ? acc[D - 1] : 0.f; p.w = 0.f;
      Op::partial_buf(a)[v.part] = p;
      if (itc) itc_arrive<D, Op>(tl, a, v.gid, sv[l], red);
    }
  }
  // (recomputed from the launch parameters: nothing extra stays live across the tile)
  if (kDiagBuild && tl.send_at >= 0 && tl.n_send > 0 && (int)blockIdx.x < tl.n_halo_tiles) halo_tile_done(tl.hd->pc);
  if (Op::REDUCE) {
    const float r = block_sum(red, s_warp);
    if (threadIdx.x == 0) Op::red_buf(a)[red_off + bid] = (double)r;
  }
  trace_end<Op>(tl, a);
  grid_dependency_wait_always();
}

constexpr int kBoundaryBlock = 256;

// FUSE: multi-GPU launches whose halo exchange is received here (the single-GPU instantiation carries none of it)
template <int D, class Op, bool FUSE>
__global__ void __launch_bounds__(kBoundaryBlock) k_boundary(DevTiling tl, typename Op::Args a, int red_base,
                                                             const int32_t* list, int n_list) {
  __shared__ float s_warp[32];
  // multi-GPU, exchange fused into this kernel: the neighbours' sums of the shared vertices are read straight
  // from the receive plane they stored into (no unpack kernel).  The sequence bookkeeping runs even when the
  // solver has already converged, so that all ranks keep counting the same exchanges.
  const HaloDev* hd = FUSE ? tl.hd : nullptr;
  // "late send" arrangement (object.h: peer_pass): the first tl.n_send CTAs of THIS kernel store the shared
  // vertices' local sums into the neighbours' memory (the tile kernel is complete: stream order); the flags travel
  // while the other CTAs finish the ordinary boundary vertices, and only CTAs that own a shared vertex wait for them
  const int n_send = FUSE ? tl.n_send : 0;
  trace_start<Op>(tl, a);
  if ((int)blockIdx.x < n_send) {
    // (also when the solver has converged: the sequence numbers must keep advancing on every rank)
    halo_send_cta(hd->pc, hd->n, hd->hrange, Op::partial_buf(a), blockIdx.x, n_send, tl.trace);
    trace_end<Op>(tl, a);
    return;
  }
  const unsigned n_recv_ctas = gridDim.x - (unsigned)n_send;
  uint32_t rseq = 0, rparity = 0;
  // list != nullptr: only these boundary vertices (the ones shared with another rank -- every other boundary vertex
  // was finished inside the tile kernel, see itc_arrive); their arrival counters are re-armed by their huge count
  const int bid = (int)blockIdx.x - n_send;
  const int idx = bid * blockDim.x + threadIdx.x;
  const int b = list ? (idx < n_list ? list[idx] : tl.n_boundary) : idx;
  if (hd) rparity = halo_recv_begin(hd->pc, rseq);
  if (Op::early_exit(a)) {
    if (hd) halo_recv_end(hd->pc, rseq, n_recv_ctas);
    return;
  }
  float red = 0.f;
  if (b < tl.n_boundary) {
    const float4* part = Op::partial_buf(a);
    const int hk = FUSE ? hd->bhalo[b] : -1;        // >= 0: shared vertex, its last slot is the remote one
    const int s0 = tl.bptr[b], s1 = tl.bptr[b + 1] - (hk >= 0 ? 1 : 0);
    const int gid = tl.n_interior + b;
    // the vertex operands do not depend on the partial sums: request them first
    const float4 vtx = Op::load_vertex(a, gid);
    const float4 pre = Op::pre(a, gid);
    // Launched with programmatic stream serialization: everything above overlaps the tail of the tile
    // kernel; the partial sums it wrote are visible only after this wait (no-op in a plain launch).
    grid_dependency_wait();
    float acc[D];
#pragma unroll
    for (int i = 0; i < D; ++i) acc[i] = 0.f;
    // four slots per trip: independent loads, adds in slot order (same sums as one at a time)
    for (int s = s0; s < s1; s += 4) {
      float4 p[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) p[k] = s + k < s1 ? __ldcs(part + s + k) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (s + k < s1) {
          acc[0] += p[k].x;
          acc[1] += p[k].y;
          if (D == 3) acc[D - 1] += p[k].z;
        }
      }
    }
    if (hk >= 0) {   // local + remote: the same two-term sum on both owners -> identical bits
      const float4 r = ll_recv_vertex(hd->pc, rseq, hk);   // waits for this vertex's own words only
      if (kDiagBuild && tl.trace) {   // one lane per warp: tens of thousands of atomics on one word would distort the timeline
        const unsigned m = __activemask();
        if ((int)(threadIdx.x & 31) == __ffs(m) - 1) atomicMax(tl.trace + 2, gtime_ns());
      }
      acc[0] += r.x;
      acc[1] += r.y;
      if (D == 3) acc[D - 1] += r.z;
    }
    Op::vertex(a, gid, vtx, pre, acc, red);
    if (Op::REDUCE && tl.bskip && tl.bskip[b]) red = 0.f;   // shared vertex owned by the neighbouring rank
  }
  if (Op::REDUCE) {
    const float r = block_sum(red, s_warp);
    if (threadIdx.x == 0) Op::red_buf(a)[red_base + bid] = (double)r;
  }
  if (hd) halo_recv_end(hd->pc, rseq, n_recv_ctas);
  trace_end<Op>(tl, a);
}

// prefetch helper shared by the Ops: `bytes` per plane starting at slot `slot`
__device__ __forceinline__ void prefetch_planes(const Planes& p, int slot, uint32_t bytes, bool three) {
  bulk_prefetch_l2(p.p0 + slot, bytes);
  bulk_prefetch_l2(p.p1 + slot, bytes);
  if (three) bulk_prefetch_l2(p.p2 + slot, bytes);
}

// CTAs the tile kernel of an Op uses for `n` tiles == per-CTA reduction partials it writes
// persist = false forces the one-CTA-per-tile kernel (LaunchCfg::no_persist: fused multi-GPU passes)
template <class Op>
inline int tile_ctas(const DevTiling& tl, int n, bool persist = true) {
  return Op::PERSIST && persist ? (n < tl.grid ? n : tl.grid) : n;
}
// partials of a whole pass; split: the pass is launched as halo tiles + other tiles (multi-GPU overlap)
template <class Op>
inline int tile_grid(const DevTiling& tl, bool split = false, bool persist = true) {
  if (!split) return tile_ctas<Op>(tl, tl.n_tiles, persist);
  return tile_ctas<Op>(tl, tl.n_halo_tiles, persist) + tile_ctas<Op>(tl, tl.n_tiles - tl.n_halo_tiles, persist);
}

template <int D, class Op>
inline void run_tile_pass(const char* name, const LaunchCfg& c, const DevTiling& tl_in, const typename Op::Args& a) {
  DevTiling tl = tl_in;          // per-launch copy: the fused halo exchange is switched on per kernel
  tl.hd = c.hd;
  tl.n_send = 0;
  tl.send_at = -1;
  tl.fuse_recv = 0;
  tl.trace = nullptr;
  // bytes of dynamic shared memory allowed so far, per kernel function (persistent / one CTA per tile) and per
  // device (the attribute lives in the device's context)
  static int attr_set_all[2][64] = {};
  // Persistent CTAs own a fixed share of the tiles, so sender CTAs that hold SM slots at the start would make
  // the tile CTAs that wait for those slots finish late: fused multi-GPU passes use one CTA per tile.
  const bool persist = Op::PERSIST && !c.no_persist;
  const int smem = tile_smem_layout(tl.max_lv, tl.max_rounds, tl.nc, n_planes<D, Op>(), persist && Op::ASYNC_VERTEX).total;
  auto* kern = persist ? k_tile_persistent<D, Op> : k_tile<D, Op>;
  int* attr_set = attr_set_all[persist ? 1 : 0];
  if (smem > attr_set[c.device & 63]) {
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) {
      cudaGetLastError();
      throw std::runtime_error(std::string(name) + ": cannot reserve shared memory: " + cudaGetErrorString(e));
    }
    attr_set[c.device & 63] = smem;
  }
  const bool split = c.part != 0;
  if (c.phase != 2) {
    int t0 = 0, t1 = tl.n_tiles, red_off = 0;
    if (c.part == 1) t1 = tl.n_halo_tiles;
    if (c.part == 2) { t0 = tl.n_halo_tiles; red_off = tile_ctas<Op>(tl, tl.n_halo_tiles, persist); }
    if (t1 > t0 && c.part == 2) {
      // second part of a split pass: independent of the first part (other tiles, other partial slots), so
      // it is launched as a programmatic dependent -- its CTAs fill the SMs while the first part's last
      // wave drains instead of waiting for it (no second tail)
      // With c.n_send, the first n_send CTAs are SENDERS: they wait for the first part (griddepcontrol.wait)
      // and store the halo sums into the neighbours' memory -- no send kernel, no second stream.
      tl.n_send = c.hd ? c.n_send : 0;
      tl.trace = trace_take_slot(c, name, tl.n_send ? "tiles:rest+send" : "tiles:rest");
      cudaLaunchConfig_t lc = {};
      lc.gridDim = dim3((unsigned)(tile_ctas<Op>(tl, t1 - t0, persist) + tl.n_send));
      lc.blockDim = dim3(Op::BLOCK);
      lc.dynamicSmemBytes = (size_t)smem;
      lc.stream = c.stream;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      at[0].val.programmaticStreamSerializationAllowed = 1;
      lc.attrs = at;
      lc.numAttrs = 1;
      cudaLaunchKernelEx(&lc, kern, tl, a, t0, t1, red_off);
      note_launch(name, c.stream);
      tl.n_send = 0;
    } else if (t1 > t0) {
      // c.send_at >= 0 (multi-GPU, all tiles in ONE launch of the one-CTA-per-tile kernel): the halo tiles come
      // first in tile order; the sender CTAs sit one wave of ordinary tiles behind them, so that by the time they
      // get their SM slots the halo tiles have (all but) finished and the rest of the kernel hides the transfer
      const bool one = c.part == 0 && c.hd && c.n_send > 0 && c.send_at >= 0 && !persist;
      if (one) {
        tl.n_send = c.n_send;
        tl.send_at = std::min(c.send_at, t1 - t0);
      }
      tl.trace = trace_take_slot(c, name, c.part == 1 ? "tiles:halo" : (one ? "tiles+send" : "tiles"));
      kern<<<tile_ctas<Op>(tl, t1 - t0, persist) + tl.n_send, Op::BLOCK, smem, c.stream>>>(tl, a, t0, t1, red_off);
      note_launch(name, c.stream);
      tl.n_send = 0;
      tl.send_at = -1;
    }
  }
  if (c.ev_mid) cudaEventRecord(c.ev_mid, c.stream);
  const bool itc = kItcBuild && tl.bcount != nullptr && Op::itc(a);
  const int32_t* list = itc ? tl.halo_b : nullptr;
  const int n_list = itc ? tl.n_halo_b : 0;
  const int n_bnd = itc ? n_list : tl.n_boundary;
  if (n_bnd > 0 && c.phase != 1) {
    tl.fuse_recv = c.hd ? c.fuse_recv : 0;
    tl.n_send = (c.hd && c.fuse_recv) ? c.bnd_send : 0;     // late send: sender CTAs at the front of the boundary kernel
    const int nb = (n_bnd + kBoundaryBlock - 1) / kBoundaryBlock + tl.n_send;
    tl.trace = trace_take_slot(c, name, tl.n_send ? "boundary+send" : "boundary");
#if FEMB_PDL
    // programmatic dependent launch: the boundary kernel's CTAs are scheduled while the tile kernel drains
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((unsigned)nb);
    lc.blockDim = dim3(kBoundaryBlock);
    lc.dynamicSmemBytes = 0;
    lc.stream = c.stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at;
    lc.numAttrs = 1;
    cudaLaunchKernelEx(&lc, k_boundary<D, Op, false>, tl, a, tile_grid<Op>(tl, split, persist), list, n_list);
#else
    if (tl.fuse_recv && tl.hd)
      k_boundary<D, Op, true><<<nb, kBoundaryBlock, 0, c.stream>>>(tl, a, tile_grid<Op>(tl, split, persist), list, n_list);
    else
      k_boundary<D, Op, false><<<nb, kBoundaryBlock, 0, c.stream>>>(tl, a, tile_grid<Op>(tl, split, persist), list, n_list);
#endif
    note_launch(name, c.stream);
  }
}

}  // namespace femb
