// sweep_core.h -- the reference's "Jacobi" exactly as it runs: a SERIAL in-place sweep (implicit.py:391-404, rows in
// ascending vertex order, every row sees the rows before it already updated) under the accept / revert / stop rule of
// implicit.py:234-261.  This header holds the arithmetic of ONE row and the loop control, written once for the device
// kernels (kernels_sweep.cu) and for the host harness that runs the very same functions without a GPU
// (tests/native/sweep_host.cpp) -- so the row formula, the table layout and the level schedule are checked on the
// CPU against the oracle and the reference's own fixtures, and only the launch plumbing needs a GPU.
//
// Matrix-free row: A = I - dt^2 M^-1 K, K = sum_e L_star (x) H_e.  For an element with corners
// v_0..v_D and block H:  (K x)_{v_j} += H (x_{v_j} - x_{v_0})  (j >= 1),   (K x)_{v_0} -= H sum_j (x_{v_j} - x_{v_0}).
// The sweep's row i:  r = b_i - sum_j A_ij x_j  (current, partly updated x);  per component k:
//   x_i[k] = |a_kk| < 1e-6 ? 0 : omega (r[k] + a_kk x_i[k]) / a_kk + (1 - omega) past_x_i[k],   omega = 0.75.
// Rows only couple along star edges (corner 0 <-> corner j of one element), so rows that share no star edge can run
// at the same time without changing a single bit of the serial result: sweep.cpp sorts the rows into dependency
// levels, and a level is what one kernel launch (or one barrier-separated phase of a single CTA) updates.
#pragma once
#include <cstdint>
#include <cmath>

#include <vector_types.h>   // float4, int4: plain structs, usable from g++ as well

#if defined(__CUDACC__)
#define FEMB_HD __host__ __device__ __forceinline__
#else
#define FEMB_HD inline
#endif

namespace femb {

// Incidence tables of the sweep, in the library's vertex numbering (device pointers in the kernels, host pointers in
// the harness).  Built lazily by the first solve that asks for the sweep (object.h: ensure_sweep).
struct SweepTables {
  const int32_t* vptr;     // [n_vert + 1] incidence range of vertex v
  const uint32_t* inc;     // element slot << 2 | corner
  const int4* sv;          // [n_slots] vertex ids of the slot's corners (.w = -1 in 2D; padding slots unused)
  const int32_t* lptr;     // [n_levels + 1] range of each dependency level in lverts
  const int32_t* lverts;   // vertices by level
};

struct SweepVecs {
  const float4 *h0, *h1, *h2;   // H_e planes (kernels_implicit.cu: OpAssemble writes them); h2 == nullptr in 2D
  const float4* pos;            // .w = lumped mass
  const float4* b;
  const float4* diag;           // diag(A) (OpDiag)
  const float4* past;           // past_vec_x
  float4* x;                    // vec_x, updated in place
  float dt2;
};

// x[] is read while other threads of the same level write OTHER entries of it: on the device the loads must not
// sit in a non-coherent cache across the in-CTA barrier between two levels.
#if defined(__CUDA_ARCH__)
#define FEMB_SWEEP_LDX(p) __ldcg(p)
#else
#define FEMB_SWEEP_LDX(p) (*(p))
#endif

// acc = (K x)_v from the current x (xv4 = x[v], already loaded)
template <int D>
FEMB_HD void sweep_row_kx(int v, const SweepTables& t, const SweepVecs& a, const float4& xv4, float (&acc)[3]) {
  acc[0] = acc[1] = acc[2] = 0.f;
  const int p1 = t.vptr[v + 1];
  for (int p = t.vptr[v]; p < p1; ++p) {
    const uint32_t w = t.inc[p];
    const int slot = (int)(w >> 2), c = (int)(w & 3u);
    const int4 g = t.sv[slot];
    float H[3][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};
    const float4 h0 = a.h0[slot], h1 = a.h1[slot];
    if (D == 3) {                                        // dev.cuh: unpack_h
      const float4 h2 = a.h2[slot];
      H[0][0] = h0.z; H[0][1] = h0.w; H[0][2] = h1.x;
      H[1][0] = h1.y; H[1][1] = h1.z; H[1][2] = h1.w;
      H[2][0] = h2.x; H[2][1] = h2.y; H[2][2] = h2.z;
    } else {
      H[0][0] = h0.z; H[0][1] = h0.w;
      H[1][0] = h1.x; H[1][1] = h1.y;
    }
    const float4 x0 = c == 0 ? xv4 : FEMB_SWEEP_LDX(a.x + g.x);
    float e[3];
    float sgn;
    if (c == 0) {                                        // row of corner 0: -H sum_j (x_{v_j} - x_{v_0})
      const float4 q1 = FEMB_SWEEP_LDX(a.x + g.y), q2 = FEMB_SWEEP_LDX(a.x + g.z);
      e[0] = (q1.x - x0.x) + (q2.x - x0.x);
      e[1] = (q1.y - x0.y) + (q2.y - x0.y);
      e[2] = (q1.z - x0.z) + (q2.z - x0.z);
      if (D == 3) {
        const float4 q3 = FEMB_SWEEP_LDX(a.x + g.w);
        e[0] += q3.x - x0.x; e[1] += q3.y - x0.y; e[2] += q3.z - x0.z;
      }
      sgn = -1.f;
    } else {                                             // row of corner j >= 1: +H (x_v - x_{v_0})
      e[0] = xv4.x - x0.x; e[1] = xv4.y - x0.y; e[2] = xv4.z - x0.z;
      sgn = 1.f;
    }
    for (int i = 0; i < D; ++i) {
      float s = H[i][0] * e[0];
      for (int k = 1; k < D; ++k) s += H[i][k] * e[k];
      acc[i] += sgn * s;
    }
  }
}

// one row of the sweep, in place
template <int D>
FEMB_HD void sweep_row(int v, const SweepTables& t, const SweepVecs& a) {
  const float4 xv4 = FEMB_SWEEP_LDX(a.x + v);
  const float xv[3] = {xv4.x, xv4.y, xv4.z};
  float acc[3];
  sweep_row_kx<D>(v, t, a, xv4, acc);
  const float s = a.dt2 * (1.f / a.pos[v].w);            // row scale dt^2 / m_v (implicit.py:183-187)
  const float4 b4 = a.b[v], d4 = a.diag[v], p4 = a.past[v];
  const float bv[3] = {b4.x, b4.y, b4.z}, dv[3] = {d4.x, d4.y, d4.z}, pv[3] = {p4.x, p4.y, p4.z};
  float xn[3] = {0.f, 0.f, 0.f};
  const float omega = 0.75f;                             // implicit.py:392
  for (int i = 0; i < D; ++i) {
    const float r = bv[i] - (xv[i] - s * acc[i]);        // b_i - (A x)_i
    xn[i] = fabsf(dv[i]) < 1e-6f ? 0.f : omega * (r + dv[i] * xv[i]) / dv[i] + (1.f - omega) * pv[i];
  }
  float4 out;
  out.x = xn[0]; out.y = xn[1]; out.z = xn[2]; out.w = 0.f;
  a.x[v] = out;
}

// Loop control of implicit.py:234-261 from err = ||b - A x|| of the iterate that has just been evaluated.
//   SW_ACCEPT: past_vec_x = vec_x (cache_x, 260-261)      SW_REVERT: vec_x = past_vec_x (recover_past_frame_x, 257-259)
// S needs: err, p_err, iters (sweeps issued so far), first, terminal, action.  `fixed` is the benchmark mode: exactly
// max_iter sweeps, every iterate accepted, no early stop.
enum SweepAction { SW_NONE = 0, SW_ACCEPT = 2, SW_REVERT = 4 };

template <class S>
FEMB_HD void sweep_decide(S* sc, float err, float tol, int max_iter, int fixed) {
  sc->err = err;
  int action = SW_NONE;
  bool go;
  if (fixed) {
    if (!sc->first) action = SW_ACCEPT;
    sc->p_err = err;
    sc->first = 0;
    go = sc->iters < max_iter;
  } else if (sc->first) {
    sc->p_err = err;                                     // p_err = err (235); past_vec_x stays STALE
    sc->first = 0;
    go = err > tol && sc->iters < max_iter;              // while (252)
  } else if (err >= (float)sc->p_err) {
    action = SW_REVERT;                                  // recover_past_frame_x ; break
    go = false;
  } else {
    sc->p_err = err;                                     // p_err = err ; cache_x
    action = SW_ACCEPT;
    go = err > tol && sc->iters < max_iter;
  }
  sc->action = action;
  if (go) sc->iters += 1;
  else sc->terminal = 1;
}

}  // namespace femb
