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
                                                 double* red, RedMask mask) {
  __shared__ float s_warp[32];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  float v = 0.f;
  if (i < n) {
    const float4 b = iv.b[i], a = rhs[i], o = opx[i];
    const float4 r = make_float4(a.x - o.x, a.y - o.y, a.z - o.z, 0.f);
    iv.x[i] = b;
    iv.r[i] = r;
    iv.d[i] = r;
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

// Jacobi control flow, implicit.py:234-261, from err = ||b - A x|| of the CURRENT iterate.
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
  int action = JA_NONE;
  sc->err = err;
  if (fixed) {                 // benchmark mode: exactly max_iter updates, no revert / early stop
    sc->p_err = err;
    sc->first = 0;
    if (sc->iters < max_iter) {
      sc->iters += 1;
      sc->action = JA_ACCEPT | JA_UPDATE;
    } else {
      sc->terminal = 1;
      sc->action = JA_ACCEPT;
    }
    return;
  }
  if (sc->first) {
    sc->p_err = err;           // p_err = err (235); past_vec_x stays stale
    sc->first = 0;
  } else if (err >= (float)sc->p_err) {
    action = JA_REVERT;        // recover_past_frame_x ; break (257-259)
    sc->terminal = 1;
  } else {
    sc->p_err = err;           // p_err = err ; cache_x (260-261)
    action = JA_ACCEPT;
  }
  if (action != JA_REVERT) {
    if (err > tol && sc->iters < max_iter) {   // while condition (252)
      action |= JA_UPDATE;
      sc->iters += 1;
    } else {
      sc->terminal = 1;
    }
  }
  sc->action = action;
}

// cache_x / recover / one damped Jacobi update (implicit.py:266-274, 391-404)
__global__ void __launch_bounds__(256) k_jacobi_update(int n, ImplicitVecs iv, const SolveScalars* sc) {
  if (sc->done) return;
  const int action = sc->action;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || action == JA_NONE) return;
  if (action & JA_REVERT) {
    iv.x[i] = iv.past_x[i];
    return;
  }
  float4 x = iv.x[i];
  float4 past = iv.past_x[i];
  if (action & JA_ACCEPT) {
    past = x;
    iv.past_x[i] = x;
  }
  if (action & JA_UPDATE) {
    const float omega = 0.75f;
    const float4 r = iv.r[i], dg = iv.diag[i];
    // b_k - sum_j A_ij x_j + a_ii x_i[k]  ==  r_k + a_ii x_k
    x.x = fabsf(dg.x) < 1e-6f ? 0.f : omega * (r.x + dg.x * x.x) / dg.x + (1.f - omega) * past.x;
    x.y = fabsf(dg.y) < 1e-6f ? 0.f : omega * (r.y + dg.y * x.y) / dg.y + (1.f - omega) * past.y;
    x.z = fabsf(dg.z) < 1e-6f ? 0.f : omega * (r.z + dg.z * x.z) / dg.z + (1.f - omega) * past.z;
    iv.x[i] = x;
  }
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
  k_cg_init<<<nblk(n_vert), 256, 0, c.stream>>>(n_vert, iv, rhs, opx, red, mask);
  note_launch("cg_init", c.stream);
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
void launch_jacobi_update(const LaunchCfg& c, int n_vert, ImplicitVecs iv, SolveScalars* sc) {
  k_jacobi_update<<<nblk(n_vert), 256, 0, c.stream>>>(n_vert, iv, sc);
  note_launch("jacobi_update", c.stream);
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
