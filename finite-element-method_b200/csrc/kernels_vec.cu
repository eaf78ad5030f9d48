// kernels_vec.cu -- vertex-parallel vector kernels of the CG / Jacobi loops
// (implicit.py:227-262, 311-386), the device-resident solver scalars, and the
// host<->device field marshalling in the caller's ordering.
//
// Reductions are deterministic: per-thread value -> warp shuffle tree -> warp 0
// sums warp partials -> one double per CTA; a single-CTA kernel then sums the
// CTA partials in index order and updates the solver scalars ON THE DEVICE, so
// the iteration loop never synchronises with the host (the reference likewise
// evaluates its convergence test in-kernel, implicit.py:252,341).
#include <stdexcept>

#include "comm.cuh"
#include "launch.h"
#include "tile_pass.cuh"

namespace femb {

enum JacobiAction { JA_NONE = 0, JA_UPDATE = 1, JA_ACCEPT = 2, JA_REVERT = 4 };

__device__ __forceinline__ float dot3(const float4& a, const float4& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// multi-GPU: vertex i (library numbering) is counted in global reductions by this rank
__device__ __forceinline__ bool counted(const RedMask& m, int i) {
  return m.bskip == nullptr || i < m.n_interior || m.bskip[i - m.n_interior] == 0;
}

// Sum red[0..n): every thread sums a strided slice with 8 independent accumulators (8 loads in
// flight: the single CTA is latency-bound), then a fixed tree.  The order depends only on n and
// blockDim, never on timing: deterministic.
constexpr int kFinalThreads = 1024;
__device__ double final_sum(const double* red, int n) {
  __shared__ double s[kFinalThreads];
  double a[8] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  const int B = blockDim.x;
  int i = threadIdx.x;
  for (; i + 7 * B < n; i += 8 * B) {
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] += red[i + k * B];
  }
#pragma unroll
  for (int k = 0; k < 8; ++k)
    if (i + k * B < n) a[k] += red[i + k * B];
  const double v = ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + a[7]));
  s[threadIdx.x] = v;
  __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
    __syncthreads();
  }
  return s[0];
}

// The global sum a scalar kernel consumes: single GPU -> the per-CTA partials red[0..n); multi-GPU over peer
// memory (pc.st != nullptr) -> every rank's already-reduced partial, gathered through the comm buffers.
// Result valid in thread 0.  In peer mode a sequence number is consumed on every call, `done` or not.
// n > 0 in peer mode: this kernel also does the local sum and the scatter (no separate reduce kernel).
// two sums in ONE exchange (single-reduction CG); results valid in thread 0
__device__ void global_sum2(const double* red_a, int n_a, const double* red_b, int n_b, const PeerComm& pc, double& a,
                            double& b) {
  __shared__ double s_ab[2];
  const double sa = final_sum(red_a, n_a);
  __syncthreads();                                       // final_sum's shared scratch is reused
  const double sb = final_sum(red_b, n_b);
  if (pc.st == nullptr) {
    a = sa;
    b = sb;
    return;
  }
  if (threadIdx.x == 0) { s_ab[0] = sa; s_ab[1] = sb; }
  __syncthreads();
  a = b = 0.0;
  if (threadIdx.x < 32) {
    allsum_scatter2(pc, s_ab[0], s_ab[1], threadIdx.x);
    allsum_gather2(pc, threadIdx.x, a, b);
  }
}

__device__ double global_sum(const double* red, int n, const PeerComm& pc) {
  if (pc.st == nullptr) return final_sum(red, n);
  if (n > 0) {
    const double s = final_sum(red, n);
    if (threadIdx.x < 32) allsum_scatter(pc, s, threadIdx.x);   // same sequence number as the gather below
  }
  return threadIdx.x < 32 ? allsum_gather(pc, threadIdx.x) : 0.0;
}

__global__ void k_scalars_reset(SolveScalars* sc) {
  SolveScalars z = {};
  z.first = 1;
  *sc = z;
}

// x = b ; r = rhs - op(x) ; d = r ; partial r.r        (implicit.py:313-334)
__global__ void __launch_bounds__(256) k_cg_init(int n, ImplicitVecs iv, const float4* rhs, const float4* opx,
                                                 double* red, RedMask mask, int zero_ps) {
  __shared__ float s_warp[32];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  float v = 0.f;
  if (i < n) {
    const float4 b = iv.b[i], a = rhs[i], o = opx[i];
    const float4 r = make_float4(a.x - o.x, a.y - o.y, a.z - o.z, 0.f);
    iv.x[i] = b;
    iv.r[i] = r;
    if (zero_ps) {   // single-reduction CG: p_{-1} = s_{-1} = 0 (the first update has beta = 0, but 0 * stale inf = NaN)
      iv.d[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      iv.q[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
      iv.d[i] = r;
    }
    v = counted(mask, i) ? dot3(r, r) : 0.f;
  }
  const float t = block_sum(v, s_warp);
  if (threadIdx.x == 0) red[blockIdx.x] = (double)t;
}

// kind 0: delta_new = sum (loop entry test) ; kind 1: alpha = delta_new / (d.q) ;
// kind 2: delta_old = delta_new, delta_new = sum, beta, ++iters, loop test   (implicit.py:341,354,378-386)
__global__ void __launch_bounds__(kFinalThreads) k_cg_scalar(int kind, int n_red, const double* red, SolveScalars* sc, float tol,
                                                   int max_iter, PeerComm pc) {
  if (pc.st == nullptr && sc->done) return;
  const double s = global_sum(red, n_red, pc);
  if (sc->done) return;
  if (threadIdx.x != 0) return;
  if (kind == 0) {
    sc->delta_new = s;
    sc->iters = 0;
    sc->done = !(0 < max_iter && (tol < 0.f || (float)s > tol));   // tol < 0: fixed iteration count
  } else if (kind == 1) {
    sc->dq = s;
    sc->alpha = (float)sc->delta_new / (float)s;
  } else {
    sc->delta_old = sc->delta_new;
    sc->delta_new = s;
    sc->beta = (float)s / (float)sc->delta_old;
    sc->iters += 1;
    sc->done = !(sc->iters < max_iter && (tol < 0.f || (float)s > tol));
  }
}

// Single-reduction CG: gamma = r.r (partials of the first pass), delta = r.w (partials of the A^T pass).
// Loop test of implicit.py:341 on the EXACT gamma of the current residual, then alpha / beta of the update the
// next first pass will apply (Chronopoulos-Gear recurrences; identical to delta_new / d.q in exact arithmetic).
__global__ void __launch_bounds__(kFinalThreads) k_cgcg_scalar(int n_g, const double* red_g, int n_d, const double* red_d,
                                                              SolveScalars* sc, float tol, int max_iter, PeerComm pc) {
  if (pc.st == nullptr && sc->done) return;
  double g, dl;
  global_sum2(red_g, n_g, red_d, n_d, pc, g, dl);        // multi-GPU: one all-sum for both
  if (sc->done) return;
  if (threadIdx.x != 0) return;
  sc->delta_old = sc->delta_new;
  sc->delta_new = g;
  sc->dq = dl;
  const bool go = sc->iters < max_iter && (tol < 0.f || (float)g > tol);
  if (!go) {
    sc->done = 1;
    return;
  }
  float alpha, beta;
  if (!sc->pending) {                                    // first iteration of the solve: p = r, s = w
    beta = 0.f;
    alpha = (float)g / (float)dl;
  } else {
    beta = (float)g / (float)sc->gamma_old;
    alpha = (float)g / ((float)dl - beta * (float)g / sc->alpha);
  }
  sc->alpha = alpha;
  sc->beta = beta;
  sc->gamma_old = g;
  sc->iters += 1;
  sc->pending = 1;
}

// The four vector updates of a single-reduction CG iteration in ONE streaming kernel, with the partials of the new
// r.r:  p = r + beta p ; s = w + beta s ; x += alpha p ; r -= alpha s   (iv.d = p, iv.q = s = Op p)
__global__ void __launch_bounds__(256) k_cgcg_update(int n, ImplicitVecs iv, double* red, const SolveScalars* sc,
                                                     RedMask mask) {
  __shared__ float s_warp[32];
  if (sc->done) return;
  const float alpha = sc->alpha, beta = sc->beta;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  float v = 0.f;
  if (i < n) {
    float4 r = iv.r[i], x = iv.x[i];
    const float4 w = iv.w[i], s0 = iv.q[i], p0 = iv.d[i];
    const float4 p = make_float4(r.x + beta * p0.x, r.y + beta * p0.y, r.z + beta * p0.z, 0.f);
    const float4 s = make_float4(w.x + beta * s0.x, w.y + beta * s0.y, w.z + beta * s0.z, 0.f);
    x.x += alpha * p.x; x.y += alpha * p.y; x.z += alpha * p.z;
    r.x -= alpha * s.x; r.y -= alpha * s.y; r.z -= alpha * s.z;
    iv.d[i] = p;
    iv.q[i] = s;
    iv.x[i] = x;
    iv.r[i] = r;
    v = counted(mask, i) ? dot3(r, r) : 0.f;
  }
  const float t = block_sum(v, s_warp);
  if (threadIdx.x == 0) red[blockIdx.x] = (double)t;
}

// x += alpha d ; r -= alpha q ; partial r.r               (implicit.py:357-381)
__global__ void __launch_bounds__(256) k_cg_update_xr(int n, ImplicitVecs iv, double* red, const SolveScalars* sc,
                                                      RedMask mask) {
  __shared__ float s_warp[32];
  if (sc->done) return;
  const float alpha = sc->alpha;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  float v = 0.f;
  if (i < n) {
    float4 x = iv.x[i], r = iv.r[i];
    const float4 d = iv.d[i], q = iv.q[i];
    x.x += alpha * d.x; x.y += alpha * d.y; x.z += alpha * d.z;
    r.x -= alpha * q.x; r.y -= alpha * q.y; r.z -= alpha * q.z;
    iv.x[i] = x;
    iv.r[i] = r;
    v = counted(mask, i) ? dot3(r, r) : 0.f;
  }
  const float t = block_sum(v, s_warp);
  if (threadIdx.x == 0) red[blockIdx.x] = (double)t;
}

// d = r + beta d                                          (implicit.py:384-385)
// Skipped once the loop test failed: d is dead after the last iteration.
__global__ void __launch_bounds__(256) k_cg_update_d(int n, ImplicitVecs iv, const SolveScalars* sc) {
  if (sc->done) return;
  const float beta = sc->beta;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float4 d = iv.d[i];
  const float4 r = iv.r[i];
  d.x = r.x + beta * d.x; d.y = r.y + beta * d.y; d.z = r.z + beta * d.z;
  iv.d[i] = d;
}

// Jacobi control flow, implicit.py:234-261, from err = ||b - A x_k|| of the iterate the pass just evaluated (x_k sits
// in buffer sc->cur; the pass has also written the candidate x_{k+1} into the next buffer, see OpJacobi).
//   continue : iters += 1, cur advances -- the next pass evaluates x_{k+1}
//   stop     : terminal; the answer is x_k (ans = cur); past_vec_x := x_k unless no iterate was ever accepted
//   revert   : terminal; the answer is x_{k-1} (the buffer before cur), or past_vec_x ITSELF when the very first
//              sweep made things worse (the reference then restores the STALE vector of the previous time step)
__global__ void __launch_bounds__(kFinalThreads) k_jacobi_decide(int n_red, const double* red, SolveScalars* sc, float tol,
                                                       int max_iter, int fixed, PeerComm pc) {
  if (pc.st == nullptr && sc->done) return;
  const double s = global_sum(red, n_red, pc);
  if (sc->done) return;
  if (sc->terminal) {
    if (threadIdx.x == 0) sc->done = 1;
    return;
  }
  if (threadIdx.x != 0) return;
  const float err = sqrtf((float)s);
  const int cur = sc->cur, nxt = cur == 2 ? 0 : cur + 1, prv = cur == 0 ? 2 : cur - 1;
  sc->err = err;
  if (fixed) {                 // benchmark mode: exactly max_iter updates, no revert / early stop
    sc->p_err = err;
    sc->first = 0;
    if (sc->iters < max_iter) {
      sc->iters += 1;
      sc->cur = nxt;
    } else {
      sc->terminal = 1;
      sc->ans = cur;
      sc->keep_past = 0;
    }
    return;
  }
  bool revert = false;
  if (sc->first) {
    sc->p_err = err;           // p_err = err (235); past_vec_x stays stale
    sc->first = 0;
    sc->keep_past = 1;         // nothing cached yet
  } else if (err >= (float)sc->p_err) {
    revert = true;             // recover_past_frame_x ; break (257-259)
    sc->terminal = 1;
    // keep_past still set: x_{k-1} = x_0 was never cached -> the answer is the stale past_vec_x, left as it is;
    // otherwise past_vec_x holds x_{k-1} (cached when it was accepted) = the answer
    sc->ans = sc->keep_past ? 3 : prv;
  } else {
    sc->p_err = err;           // p_err = err ; cache_x (260-261)
    sc->keep_past = 0;
  }
  if (!revert) {
    if (err > tol && sc->iters < max_iter) {   // while condition (252)
      sc->iters += 1;
      sc->cur = nxt;
    } else {
      sc->terminal = 1;
      sc->ans = cur;
    }
  }
}

// end of the solve: vec_x = the answer, past_vec_x = what the reference's cache_x / recover left there
__global__ void __launch_bounds__(256) k_jacobi_finish(int n, ImplicitVecs iv, const SolveScalars* sc) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int ans = sc->terminal ? sc->ans : sc->cur;   // not terminal: the iteration budget of the caller ran out
  const float4* src = ans == 3 ? iv.past_x : (ans == 0 ? iv.x : (ans == 1 ? iv.tmp : iv.r));
  const float4 v = src[i];
  iv.x[i] = v;
  if (!sc->keep_past) iv.past_x[i] = v;
}

__global__ void __launch_bounds__(256) k_copy_xyz(int n, const float4* src, float4* dst) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 s = src[i];
  float4 d = dst[i];
  d.x = s.x; d.y = s.y; d.z = s.z;
  dst[i] = d;
}

__global__ void __launch_bounds__(256) k_scale_xyz(int n, const float4* src, float4* dst, float f) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 s = src[i];
  float4 d = dst[i];
  d.x = f * s.x; d.y = f * s.y; d.z = f * s.z;
  dst[i] = d;
}

__global__ void __launch_bounds__(256) k_fill(int n, float4* dst, float4 v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = v;
}

__global__ void __launch_bounds__(kFinalThreads) k_reduce_final(int n_red, const double* red, double* out) {
  const double s = final_sum(red, n_red);
  if (threadIdx.x == 0) *out = s;
}

// local sum of the per-CTA partials -> slot[rank] of every rank (peer memory), in one kernel
__global__ void __launch_bounds__(kFinalThreads) k_reduce_final_send(int n_red, const double* red, PeerComm pc) {
  __shared__ double s_v;
  const double s = final_sum(red, n_red);
  if (threadIdx.x == 0) s_v = s;
  __syncthreads();
  if (threadIdx.x < 32) allsum_scatter(pc, s_v, threadIdx.x);
}

// dst[old*ncomp + k] = src[new_of_old[old]].comp[comp0 + k]
__global__ void __launch_bounds__(256) k_pack_vertex(int n, const int32_t* new_of_old, const float4* src, int comp0,
                                                     int ncomp, float* dst) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 v = src[new_of_old[i]];
  for (int k = 0; k < ncomp; ++k) dst[(size_t)i * ncomp + k] = comp(v, comp0 + k);
}

__global__ void __launch_bounds__(256) k_unpack_vertex(int n, const int32_t* new_of_old, const float* src, int comp0,
                                                       int ncomp, float4* dst) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float* d = reinterpret_cast<float*>(dst + new_of_old[i]);
  for (int k = 0; k < ncomp; ++k) d[comp0 + k] = src[(size_t)i * ncomp + k];
}

// State diagnostics on the device (main.py --diag-every): elements whose deformation gradient has det F <= 0 or
// NaN (the explicit kernel takes log(det F) unclamped, explicit.py:24-25), elements whose energy density is NaN,
// vertices with a NaN coordinate.  Integer counters, so the atomics are order-independent.  One CTA per tile.
template <int D>
__global__ void __launch_bounds__(256) k_diag_elements(DevTiling tl, Planes e, const float4* pos, const float* phi,
                                                       unsigned long long* out) {
  const DevTile td = tl.tiles[blockIdx.x];
  const LocalVert* lv = tl.lv + td.lv_off;
  unsigned inv = 0, nan_e = 0;
  for (int k = threadIdx.x; k < td.n_elem; k += blockDim.x) {
    const int s = td.slot_off + k;
    const float4 e0 = e.p0[s], e1 = e.p1[s];
    const float4 e2 = D == 3 ? e.p2[s] : make_float4(0.f, 0.f, 0.f, 0.f);
    const uint32_t w0 = __float_as_uint(e0.x), w1 = __float_as_uint(e0.y);
    const float4 p0 = pos[lv[w0 & 0xffffu].gid], p1 = pos[lv[w0 >> 16].gid], p2 = pos[lv[w1 & 0xffffu].gid];
    Mat<D> R, X, F;
    unpack_ref<D>(e0, e1, e2, R);
    X.m[0][0] = p1.x - p0.x; X.m[1][0] = p1.y - p0.y;
    X.m[0][1] = p2.x - p0.x; X.m[1][1] = p2.y - p0.y;
    if constexpr (D == 3) {
      const float4 p3 = pos[lv[w1 >> 16].gid];
      X.m[2][0] = p1.z - p0.z; X.m[2][1] = p2.z - p0.z;
      X.m[0][2] = p3.x - p0.x; X.m[1][2] = p3.y - p0.y; X.m[2][2] = p3.z - p0.z;
    }
    matmul<D>(X, R, F);
    if (!(det<D>(F) > 0.f)) ++inv;
    if (phi[s] != phi[s]) ++nan_e;
  }
  inv = __reduce_add_sync(0xffffffffu, inv);
  nan_e = __reduce_add_sync(0xffffffffu, nan_e);
  if ((threadIdx.x & 31) == 0) {
    if (inv) atomicAdd(out + 0, (unsigned long long)inv);
    if (nan_e) atomicAdd(out + 1, (unsigned long long)nan_e);
  }
}
__global__ void __launch_bounds__(256) k_diag_vertices(int n, int dim, const float4* pos, unsigned long long* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned bad = 0;
  if (i < n) {
    const float4 p = pos[i];
    bad = (p.x != p.x) | (p.y != p.y) | (dim == 3 && p.z != p.z);
  }
  bad = __reduce_add_sync(0xffffffffu, bad);
  if ((threadIdx.x & 31) == 0 && bad) atomicAdd(out + 2, (unsigned long long)bad);
}
void launch_diagnostics(const LaunchCfg& c, int dim, const DevTiling& tl, Planes e, const float4* pos, const float* phi,
                        unsigned long long* out3) {
  cudaMemsetAsync(out3, 0, 3 * sizeof(unsigned long long), c.stream);
  if (dim == 3) k_diag_elements<3><<<tl.n_tiles, 256, 0, c.stream>>>(tl, e, pos, phi, out3);
  else k_diag_elements<2><<<tl.n_tiles, 256, 0, c.stream>>>(tl, e, pos, phi, out3);
  note_launch("diag_elements", c.stream);
  k_diag_vertices<<<(tl.n_vert + 255) / 256, 256, 0, c.stream>>>(tl.n_vert, dim, pos, out3);
  note_launch("diag_vertices", c.stream);
}

// what: 0 volume, 1 ref (row-major dim*dim), 2 phi, 3 H_e (row-major)
__global__ void __launch_bounds__(256) k_pack_elem(int dim, int n, const int32_t* slot_of_elem, Planes e, Planes h,
                                                   const float* phi, int what, float* dst) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int s = slot_of_elem[i];
  if (what == 0) {
    dst[i] = e.p0[s].z;
  } else if (what == 2) {
    dst[i] = phi[s];
  } else if (what == 1) {
    const float4 a = e.p0[s], b = e.p1[s];
    if (dim == 3) {
      const float4 c = e.p2[s];
      float* o = dst + (size_t)i * 9;
      o[0] = a.w; o[1] = b.x; o[2] = b.y; o[3] = b.z; o[4] = b.w; o[5] = c.x; o[6] = c.y; o[7] = c.z; o[8] = c.w;
    } else {
      float* o = dst + (size_t)i * 4;
      o[0] = a.w; o[1] = b.x; o[2] = b.y; o[3] = b.z;
    }
  } else {
    const float4 a = h.p0[s], b = h.p1[s];
    if (dim == 3) {
      const float4 c = h.p2[s];
      float* o = dst + (size_t)i * 9;
      o[0] = a.z; o[1] = a.w; o[2] = b.x; o[3] = b.y; o[4] = b.z; o[5] = b.w; o[6] = c.x; o[7] = c.y; o[8] = c.z;
    } else {
      float* o = dst + (size_t)i * 4;
      o[0] = a.z; o[1] = a.w; o[2] = b.x; o[3] = b.y;
    }
  }
}

// ---------------------------------------------------------------------------
static inline int nblk(int n) { return (n + 255) / 256; }

void launch_scalars_reset(const LaunchCfg& c, SolveScalars* sc) {
  k_scalars_reset<<<1, 1, 0, c.stream>>>(sc);
  note_launch("scalars_reset", c.stream);
}
void launch_cg_init(const LaunchCfg& c, int n_vert, ImplicitVecs iv, const float4* rhs, const float4* opx,
                    double* red, RedMask mask) {
  launch_cg_init_ex(c, n_vert, iv, rhs, opx, red, mask, 0);
}
void launch_cg_init_ex(const LaunchCfg& c, int n_vert, ImplicitVecs iv, const float4* rhs, const float4* opx,
                       double* red, RedMask mask, int zero_ps) {
  k_cg_init<<<nblk(n_vert), 256, 0, c.stream>>>(n_vert, iv, rhs, opx, red, mask, zero_ps);
  note_launch("cg_init", c.stream);
}
void launch_cgcg_scalar(const LaunchCfg& c, int n_g, const double* red_g, int n_d, const double* red_d, SolveScalars* sc,
                        float tol, int max_iter, const PeerComm& pc) {
  k_cgcg_scalar<<<1, kFinalThreads, 0, c.stream>>>(n_g, red_g, n_d, red_d, sc, tol, max_iter, pc);
  note_launch("cgcg_scalar", c.stream);
}
void launch_cgcg_update(const LaunchCfg& c, int n_vert, ImplicitVecs iv, double* red, const SolveScalars* sc, RedMask mask) {
  k_cgcg_update<<<nblk(n_vert), 256, 0, c.stream>>>(n_vert, iv, red, sc, mask);
  note_launch("cgcg_update", c.stream);
}
void launch_cg_delta0(const LaunchCfg& c, int n_red, const double* red, SolveScalars* sc, float tol, int max_iter,
                      const PeerComm& pc) {
  k_cg_scalar<<<1, kFinalThreads, 0, c.stream>>>(0, n_red, red, sc, tol, max_iter, pc);
  note_launch("cg_scalar0", c.stream);
}
void launch_cg_alpha(const LaunchCfg& c, int n_red, const double* red, SolveScalars* sc, const PeerComm& pc) {
  k_cg_scalar<<<1, kFinalThreads, 0, c.stream>>>(1, n_red, red, sc, 0.f, 0, pc);
  note_launch("cg_alpha", c.stream);
}
void launch_cg_update_xr(const LaunchCfg& c, int n_vert, ImplicitVecs iv, double* red, SolveScalars* sc, RedMask mask) {
  k_cg_update_xr<<<nblk(n_vert), 256, 0, c.stream>>>(n_vert, iv, red, sc, mask);
  note_launch("cg_update_xr", c.stream);
}
void launch_cg_beta(const LaunchCfg& c, int n_red, const double* red, SolveScalars* sc, float tol, int max_iter,
                    const PeerComm& pc) {
  k_cg_scalar<<<1, kFinalThreads, 0, c.stream>>>(2, n_red, red, sc, tol, max_iter, pc);
  note_launch("cg_beta", c.stream);
}
void launch_cg_update_d(const LaunchCfg& c, int n_vert, ImplicitVecs iv, const SolveScalars* sc) {
  k_cg_update_d<<<nblk(n_vert), 256, 0, c.stream>>>(n_vert, iv, sc);
  note_launch("cg_update_d", c.stream);
}
void launch_jacobi_decide(const LaunchCfg& c, int n_red, const double* red, SolveScalars* sc, float tol,
                          int max_iter, int fixed, const PeerComm& pc) {
  k_jacobi_decide<<<1, kFinalThreads, 0, c.stream>>>(n_red, red, sc, tol, max_iter, fixed, pc);
  note_launch("jacobi_decide", c.stream);
}
void launch_jacobi_finish(const LaunchCfg& c, int n_vert, ImplicitVecs iv, const SolveScalars* sc) {
  k_jacobi_finish<<<nblk(n_vert), 256, 0, c.stream>>>(n_vert, iv, sc);
  note_launch("jacobi_finish", c.stream);
}
void launch_copy_xyz(const LaunchCfg& c, int n, const float4* src, float4* dst) {
  k_copy_xyz<<<nblk(n), 256, 0, c.stream>>>(n, src, dst);
  note_launch("copy_xyz", c.stream);
}
void launch_scale_xyz(const LaunchCfg& c, int n, const float4* src, float4* dst, float s) {
  k_scale_xyz<<<nblk(n), 256, 0, c.stream>>>(n, src, dst, s);
  note_launch("scale_xyz", c.stream);
}
void launch_fill(const LaunchCfg& c, int n, float4* dst, float4 v) {
  k_fill<<<nblk(n), 256, 0, c.stream>>>(n, dst, v);
  note_launch("fill", c.stream);
}
void launch_reduce_final(const LaunchCfg& c, int n_red, const double* red, double* out) {
  k_reduce_final<<<1, kFinalThreads, 0, c.stream>>>(n_red, red, out);
  note_launch("reduce_final", c.stream);
}
void launch_reduce_final_send(const LaunchCfg& c, int n_red, const double* red, const PeerComm& pc) {
  k_reduce_final_send<<<1, kFinalThreads, 0, c.stream>>>(n_red, red, pc);
  note_launch("reduce_final_send", c.stream);
}
void launch_pack_vertex(const LaunchCfg& c, int, int n_vert, const int32_t* new_of_old, const float4* src, int comp0,
                        int ncomp, float* dst) {
  k_pack_vertex<<<nblk(n_vert), 256, 0, c.stream>>>(n_vert, new_of_old, src, comp0, ncomp, dst);
  note_launch("pack_vertex", c.stream);
}
void launch_unpack_vertex(const LaunchCfg& c, int, int n_vert, const int32_t* new_of_old, const float* src, int comp0,
                          int ncomp, float4* dst) {
  k_unpack_vertex<<<nblk(n_vert), 256, 0, c.stream>>>(n_vert, new_of_old, src, comp0, ncomp, dst);
  note_launch("unpack_vertex", c.stream);
}
void launch_pack_elem(const LaunchCfg& c, int dim, int n_elem, const int32_t* slot_of_elem, Planes e, Planes h,
                      const float* phi, int what, float* dst) {
  k_pack_elem<<<nblk(n_elem), 256, 0, c.stream>>>(dim, n_elem, slot_of_elem, e, h, phi, what, dst);
  note_launch("pack_elem", c.stream);
}

// ---------------------------------------------------------------------------
// halo (multi-GPU slabs): a shared vertex has its local partial slots first and ONE remote slot last.
// Both ranks compute (sum of local slots) and exchange it, so each evaluates local + remote =
// remote + local: the same two-term fp32 sum on both sides (bit-identical shared vertices).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_halo_pack(int n, const int32_t* hb, const int32_t* bptr, const float4* partial,
                                                   float4* send) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int b = hb[k];
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int i = bptr[b]; i < bptr[b + 1] - 1; ++i) {   // same order as k_boundary
    const float4 p = partial[i];
    s.x += p.x; s.y += p.y; s.z += p.z;
  }
  send[k] = s;
}

__global__ void __launch_bounds__(256) k_halo_unpack(int n, const int32_t* hb, const int32_t* bptr, float4* partial,
                                                     const float4* recv, int zero) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  partial[bptr[hb[k] + 1] - 1] = zero ? make_float4(0.f, 0.f, 0.f, 0.f) : recv[k];
}

__global__ void __launch_bounds__(256) k_halo_mass(int n, const int32_t* hb, int n_interior, float4* pos, float4* send,
                                                   const float4* recv, int phase) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int gid = n_interior + hb[k];
  if (phase == 0) send[k] = make_float4(pos[gid].w, 0.f, 0.f, 0.f);
  else pos[gid].w += recv[k].x;
}

void launch_halo_pack(const LaunchCfg& c, const DevTiling& tl, Halo h, const float4* partial) {
  if (h.n == 0) return;
  k_halo_pack<<<nblk(h.n), 256, 0, c.stream>>>(h.n, h.b, tl.bptr, partial, h.send);
  note_launch("halo_pack", c.stream);
}
void launch_halo_unpack(const LaunchCfg& c, const DevTiling& tl, Halo h, float4* partial) {
  if (h.n == 0) return;
  k_halo_unpack<<<nblk(h.n), 256, 0, c.stream>>>(h.n, h.b, tl.bptr, partial, h.recv, 0);
  note_launch("halo_unpack", c.stream);
}
void launch_halo_zero_remote(const LaunchCfg& c, const DevTiling& tl, Halo h, float4* partial) {
  if (h.n == 0) return;
  k_halo_unpack<<<nblk(h.n), 256, 0, c.stream>>>(h.n, h.b, tl.bptr, partial, nullptr, 1);
  note_launch("halo_zero", c.stream);
}
void launch_halo_mass(const LaunchCfg& c, const DevTiling& tl, Halo h, float4* pos, int phase) {
  if (h.n == 0) return;
  k_halo_mass<<<nblk(h.n), 256, 0, c.stream>>>(h.n, h.b, tl.n_interior, pos, h.send, h.recv, phase);
  note_launch("halo_mass", c.stream);
}

}  // namespace femb
