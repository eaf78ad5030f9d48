// kernels_implicit.cu -- the implicit path of solver/implicit.py, matrix-free.
//
// The reference assembles a dense Nv x Nv block matrix A = I - dt^2 M^-1 K
// (implicit.py:114-194) and runs CG / a damped sweep on it.  K is a sum over
// elements of L_star (x) H_e with ONE dim x dim block H_e per element
// (implicit.py:141-178), so A d needs only H_e (36 B/tet) and
// the same tile gather as the explicit gradient.
#include <stdexcept>

#include "launch.h"
#include "tile_pass.cuh"

namespace femb {

// ---------------------------------------------------------------------------
// assemble: H_e -> h planes, b = vel + dt M^-1 f   (implicit.py:65-105, 116-147)
// ---------------------------------------------------------------------------
template <int D>
struct OpAssemble {
  static constexpr bool PERSIST = FEMB_PERSIST_ASSEMBLE != 0;
  static constexpr bool ASYNC_VERTEX = false, PAIR = false;
  static constexpr int BLOCK = kBlock, MIN_CTAS = kMinCtas,NSTAGE = D * D, GATHER = G_STAR;
  static constexpr bool REDUCE = false;
  struct Args {
    Planes e, h;
    VertexState vs;
    Material mat;
    float dt;
    ImplicitVecs iv;
  };
  __device__ static bool early_exit(const Args&) { return false; }
  __host__ __device__ static bool itc(const Args&) { return true; }
  __device__ static float w0() { return 0.f; }
  __device__ static float4* partial_buf(const Args& a) { return a.vs.partial; }
  __device__ static double* red_buf(const Args&) { return nullptr; }
  __device__ static float4 load_vertex(const Args& a, int gid) { return a.vs.pos[gid]; }

  using Elem = PlaneElem;
  __device__ static void prefetch(const Args& a, int slot, uint32_t bytes) { prefetch_planes(a.e, slot, bytes, D == 3); }
  __device__ static void prefetch_verts(const Args& a, int base, uint32_t bytes) {
    bulk_prefetch_l2(a.vs.pos + base, bytes);
    bulk_prefetch_l2(a.vs.vel + base, bytes);
  }
  __device__ static void load(const Args& a, int slot, Elem& d) { load_planes<D>(a.e, slot, d); }
  __device__ static void compute(const Args& a, int slot, const Elem& d, const float4* sv, float (&out)[NSTAGE],
                                 float&) {
    const float4 e0 = d.a, e1 = d.b, e2 = d.c;
    Mat<D> R, X, F, C, FiT, Fi;
    unpack_ref<D>(e0, e1, e2, R);
    edge_matrix<D>(sv, __float_as_uint(e0.x), __float_as_uint(e0.y), X);
    matmul<D>(X, R, F);
    cofactor<D>(F, C);
    const float J = det<D>(F);
    const float inv_j = 1.f / J;
    const float V = e0.z;
    const float mu = a.mat.mu, lam = a.mat.lambda;
#pragma unroll
    for (int i = 0; i < D; ++i)
#pragma unroll
      for (int j = 0; j < D; ++j) {
        FiT.m[i][j] = inv_j * C.m[i][j];
        Fi.m[j][i] = FiT.m[i][j];
      }
    // ---- right-hand side force (implicit.py:87-88): log|J| through det(F^T F)
    Mat<D> FtF, P, G;
    matmul_tn<D>(F, F, FtF);
    const float lg2 = lam / 2.f * logf(det<D>(FtF));
#pragma unroll
    for (int i = 0; i < D; ++i)
#pragma unroll
      for (int j = 0; j < D; ++j) P.m[i][j] = mu * F.m[i][j] - mu * FiT.m[i][j] + lg2 * FiT.m[i][j];
    matmul_nt<D>(P, R, G);
#pragma unroll
    for (int j = 0; j < D; ++j)
#pragma unroll
      for (int i = 0; i < D; ++i) out[j * D + i] = G.m[i][j] * (-V);
    // ---- H_e (implicit.py:131-147): dF = R, log_J clamped at 1e-4
    const float log_j = logf(fmaxf(J, 1e-4f));
    Mat<D> FiR, T1, M1, H;
    matmul<D>(Fi, R, FiR);
    float tr = FiR.m[0][0];
#pragma unroll
    for (int k = 1; k < D; ++k) tr += FiR.m[k][k];
    matmul_nt<D>(FiT, R, T1);       // F^-T @ dF^T
    matmul<D>(T1, FiT, M1);         // ... @ F^-T
    const float c1 = mu - lam * log_j, c2 = lam * tr;
    Mat<D> term;
#pragma unroll
    for (int i = 0; i < D; ++i)
#pragma unroll
      for (int j = 0; j < D; ++j) term.m[i][j] = (-V) * (mu * R.m[i][j] + c1 * M1.m[i][j] + c2 * FiT.m[i][j]);
    matmul_nt<D>(term, R, H);
    float4 h0, h1;
    h0.x = e0.x; h0.y = e0.y; h0.z = H.m[0][0]; h0.w = H.m[0][1];
    if constexpr (D == 3) {
      h1 = make_float4(H.m[0][2], H.m[1][0], H.m[1][1], H.m[1][2]);
      a.h.p2[slot] = make_float4(H.m[2][0], H.m[2][1], H.m[2][2], 0.f);
    } else {
      h1 = make_float4(H.m[1][0], H.m[1][1], 0.f, 0.f);
    }
    a.h.p0[slot] = h0;
    a.h.p1[slot] = h1;
  }

  __device__ static float4 pre(const Args& a, int gid) { return a.vs.vel[gid]; }
  __device__ static void vertex(const Args& a, int gid, float4 p, float4 vel, const float (&acc)[D], float&) {
    // b = vel ; b[v] += dt * (1/m_v) * f_v              (implicit.py:67-68, 98-105)
    const float s = a.dt * (1.f / p.w);
    float4 b = vel;
    b.x += s * acc[0]; b.y += s * acc[1]; if (D == 3) b.z += s * acc[D - 1];
    b.w = 0.f;
    a.iv.b[gid] = b;
  }
};

// diag(A)_{v,k} = 1 - dt^2/m_v sum_e c_{e,v} H_e[k,k]; c = dim at local vertex 0
template <int D>
struct OpDiag {
  static constexpr bool PERSIST = FEMB_PERSIST_APPLY != 0;
  static constexpr bool ASYNC_VERTEX = false, PAIR = false;
  static constexpr int BLOCK = kBlock, MIN_CTAS = kMinCtas,NSTAGE = D, GATHER = G_UNIFORM;
  static constexpr bool REDUCE = false;
  struct Args {
    Planes h;
    VertexState vs;
    float dt2;
    ImplicitVecs iv;
  };
  __device__ static bool early_exit(const Args&) { return false; }
  __host__ __device__ static bool itc(const Args&) { return true; }
  __device__ static float w0() { return (float)D; }
  __device__ static float4* partial_buf(const Args& a) { return a.vs.partial; }
  __device__ static double* red_buf(const Args&) { return nullptr; }
  __device__ static float4 load_vertex(const Args& a, int gid) { return a.vs.pos[gid]; }
  using Elem = PlaneElem;
  __device__ static void prefetch(const Args& a, int slot, uint32_t bytes) { prefetch_planes(a.h, slot, bytes, D == 3); }
  __device__ static void prefetch_verts(const Args& a, int base, uint32_t bytes) { bulk_prefetch_l2(a.vs.pos + base, bytes); }
  __device__ static void load(const Args& a, int slot, Elem& d) { load_planes<D>(a.h, slot, d); }
  __device__ static void compute(const Args&, int, const Elem& d, const float4*, float (&out)[NSTAGE], float&) {
    const float4 h0 = d.a, h1 = d.b;
    if constexpr (D == 3) {
      out[0] = h0.z; out[1] = h1.z; out[2] = d.c.z;
    } else {
      out[0] = h0.z; out[1] = h1.y;
    }
  }
  __device__ static float4 pre(const Args&, int) { return make_float4(0.f, 0.f, 0.f, 0.f); }
  __device__ static void vertex(const Args& a, int gid, float4 p, float4, const float (&acc)[D], float&) {
    const float s = a.dt2 * (1.f / p.w);
    a.iv.diag[gid] = make_float4(1.f - s * acc[0], 1.f - s * acc[1], D == 3 ? 1.f - s * acc[D - 1] : 1.f, 0.f);
  }
};

// ---------------------------------------------------------------------------
// out = A in  or  out = A^T in   (the reference's dense mat-vec, implicit.py:347-349)
// ---------------------------------------------------------------------------
template <int D, int TR>
struct OpApply {
  static constexpr bool PERSIST = FEMB_PERSIST_APPLY != 0;
  static constexpr bool ASYNC_VERTEX = false, PAIR = false;
  static constexpr int BLOCK = kBlock, MIN_CTAS = kMinCtas,NSTAGE = D * D, GATHER = G_STAR;
  static constexpr bool REDUCE = true;
  struct Args {
    Planes h;
    const float4* pos;     // .w = mass
    const float4* in;
    float4* out;
    float4* partial;
    float dt2;
    int epilogue;
    const float4* aux;     // AP_DOT: vector dotted with out ; AP_RESID: b
    float4* resid;         // AP_RESID: r = b - out
    double* red;
    const SolveScalars* sc;
  };
  __device__ static bool early_exit(const Args& a) { return a.sc && (a.sc->done | a.sc->terminal); }
  __host__ __device__ static bool itc(const Args& a) { return a.epilogue == AP_PLAIN; }   // reductions keep k_boundary
  __device__ static float w0() { return 0.f; }
  __device__ static float4* partial_buf(const Args& a) { return a.partial; }
  __device__ static double* red_buf(const Args& a) { return a.red; }
  __device__ static float4 load_vertex(const Args& a, int gid) {
    float4 d = a.in[gid];
    if (TR) {  // A^T d = d - dt^2 K^T (M^-1 d)
      const float im = 1.f / a.pos[gid].w;
      d.x *= im; d.y *= im; d.z *= im;
    }
    return d;
  }
  using Elem = PlaneElem;
  __device__ static void prefetch(const Args& a, int slot, uint32_t bytes) { prefetch_planes(a.h, slot, bytes, D == 3); }
  __device__ static void prefetch_verts(const Args& a, int base, uint32_t bytes) {
    bulk_prefetch_l2(a.in + base, bytes);
    bulk_prefetch_l2(a.pos + base, bytes);
  }
  __device__ static void load(const Args& a, int slot, Elem& d) { load_planes<D>(a.h, slot, d); }
  __device__ static void compute(const Args&, int, const Elem& d, const float4* sv, float (&out)[NSTAGE], float&) {
    const float4 h0 = d.a, h1 = d.b, h2 = d.c;
    Mat<D> H, E;
    unpack_h<D>(h0, h1, h2, H);
    edge_matrix<D>(sv, __float_as_uint(h0.x), __float_as_uint(h0.y), E);   // e_j = d_{v_j} - d_{v_0}
#pragma unroll
    for (int j = 0; j < D; ++j)
#pragma unroll
      for (int i = 0; i < D; ++i) {
        float s = (TR ? H.m[0][i] : H.m[i][0]) * E.m[0][j];
#pragma unroll
        for (int k = 1; k < D; ++k) s += (TR ? H.m[k][i] : H.m[i][k]) * E.m[k][j];
        out[j * D + i] = s;
      }
  }
  // {in.xyz, mass}: the un-scaled input vector and the row scale of the epilogue
  __device__ static float4 pre(const Args& a, int gid) {
    float4 d = a.in[gid];
    d.w = TR ? 1.f : a.pos[gid].w;
    return d;
  }
  __device__ static void vertex(const Args& a, int gid, float4, float4 d, const float (&acc)[D], float& red) {
    const float s = TR ? a.dt2 : a.dt2 * (1.f / d.w);
    float4 q;
    q.x = d.x - s * acc[0];
    q.y = d.y - s * acc[1];
    q.z = D == 3 ? d.z - s * acc[D - 1] : 0.f;
    q.w = 0.f;
    a.out[gid] = q;
    if (a.epilogue == AP_DOT) {
      const float4 u = a.aux[gid];
      red += u.x * q.x + u.y * q.y + (D == 3 ? u.z * q.z : 0.f);
    } else if (a.epilogue == AP_RESID) {
      const float4 b = a.aux[gid];
      float4 r = make_float4(b.x - q.x, b.y - q.y, D == 3 ? b.z - q.z : 0.f, 0.f);
      a.resid[gid] = r;
      red += r.x * r.x + r.y * r.y + r.z * r.z;
    }
  }
};

// ---------------------------------------------------------------------------
// One iteration of the damped Jacobi of implicit.py:227-262 / 391-404 in ONE pass: r = b - A x_k with its norm
// (compute_error, 277-286) and, in the same vertex epilogue, the NEXT iterate
//   x_{k+1} = omega (r + a_ii x_k) / a_ii + (1 - omega) past      (b_k - sum_j A_ij x_j + a_ii x_i == r + a_ii x_i)
// written speculatively into the next of three rotating buffers.  `past` is what the reference's past_vec_x holds
// when that sweep runs: the accepted iterate x_k itself (cache_x, 260-261), or -- for the first sweep of a solve --
// the STALE vector of the previous time step (231-232 initialise only vec_x).  Whether x_{k+1} is kept is decided
// after the pass, from the norm, by k_jacobi_decide (accept / revert / stop, 252-261): a rejected x_{k+1} is simply
// never read, and the revert target x_{k-1} still sits in the third buffer.  No separate update kernel, r and
// A x never touch HBM.
// ---------------------------------------------------------------------------
template <int D>
struct OpJacobi {
  static constexpr bool PERSIST = FEMB_PERSIST_APPLY != 0;
  static constexpr bool ASYNC_VERTEX = false, PAIR = false;
  static constexpr int BLOCK = kBlock, MIN_CTAS = kMinCtas, NSTAGE = D * D, GATHER = G_STAR;
  static constexpr bool REDUCE = true;
  struct Args {
    Planes h;
    const float4* pos;     // .w = mass
    float4 *x0, *x1, *x2;  // the three rotating iterate buffers (selected with compares: a dynamically indexed
                           // pointer array in the kernel arguments would put the whole struct in local memory)
    const float4* b;
    const float4* diag;
    const float4* past;
    float4* partial;
    float dt2;
    double* red;
    const SolveScalars* sc;
  };
  __device__ static bool early_exit(const Args& a) { return (a.sc->done | a.sc->terminal) != 0; }
  __host__ __device__ static bool itc(const Args&) { return false; }
  __device__ static float w0() { return 0.f; }
  __device__ static float4* partial_buf(const Args& a) { return a.partial; }
  __device__ static double* red_buf(const Args& a) { return a.red; }
  __device__ static int cur_of(const Args& a) { return __ldg(&a.sc->cur); }
  __device__ static float4* buf(const Args& a, int i) { return i == 0 ? a.x0 : (i == 1 ? a.x1 : a.x2); }
  __device__ static const float4* in_buf(const Args& a) { return buf(a, cur_of(a)); }
  __device__ static float4 load_vertex(const Args& a, int gid) { return in_buf(a)[gid]; }
  using Elem = PlaneElem;
  __device__ static void prefetch(const Args& a, int slot, uint32_t bytes) { prefetch_planes(a.h, slot, bytes, D == 3); }
  __device__ static void prefetch_verts(const Args& a, int base, uint32_t bytes) {
    bulk_prefetch_l2(in_buf(a) + base, bytes);
    bulk_prefetch_l2(a.b + base, bytes);
    bulk_prefetch_l2(a.diag + base, bytes);
  }
  __device__ static void load(const Args& a, int slot, Elem& d) { load_planes<D>(a.h, slot, d); }
  __device__ static void compute(const Args&, int, const Elem& d, const float4* sv, float (&out)[NSTAGE], float&) {
    const float4 h0 = d.a, h1 = d.b, h2 = d.c;
    Mat<D> H, E;
    unpack_h<D>(h0, h1, h2, H);
    edge_matrix<D>(sv, __float_as_uint(h0.x), __float_as_uint(h0.y), E);   // e_j = x_{v_j} - x_{v_0}
#pragma unroll
    for (int j = 0; j < D; ++j)
#pragma unroll
      for (int i = 0; i < D; ++i) {
        float s = H.m[i][0] * E.m[0][j];
#pragma unroll
        for (int k = 1; k < D; ++k) s += H.m[i][k] * E.m[k][j];
        out[j * D + i] = s;
      }
  }
  // the right-hand side of the vertex: requested before the gather
  __device__ static float4 pre(const Args& a, int gid) { return a.b[gid]; }
  __device__ static void vertex(const Args& a, int gid, float4 x, float4 b, const float (&acc)[D], float& red) {
    const float s = a.dt2 * (1.f / a.pos[gid].w);
    const float4 dg = a.diag[gid];
    const int cur = cur_of(a);
    const float4 past = __ldg(&a.sc->first) ? a.past[gid] : x;
    float r[3] = {0.f, 0.f, 0.f};
    const float xv[3] = {x.x, x.y, x.z}, bv[3] = {b.x, b.y, b.z}, dv[3] = {dg.x, dg.y, dg.z}, pv[3] = {past.x, past.y, past.z};
    float xn[3] = {0.f, 0.f, 0.f};
    const float omega = 0.75f;
#pragma unroll
    for (int i = 0; i < D; ++i) {
      r[i] = bv[i] - (xv[i] - s * acc[i]);                 // b - (A x)_i
      red += r[i] * r[i];
      xn[i] = fabsf(dv[i]) < 1e-6f ? 0.f : omega * (r[i] + dv[i] * xv[i]) / dv[i] + (1.f - omega) * pv[i];
    }
    buf(a, cur == 2 ? 0 : cur + 1)[gid] = make_float4(xn[0], xn[1], xn[2], 0.f);
  }
};

void launch_jacobi_pass(int dim, const LaunchCfg& c, const DevTiling& tl, Planes h, VertexState vs, float dt,
                        ImplicitVecs iv, double* red, const SolveScalars* sc) {
  if (dim == 3) {
    OpJacobi<3>::Args a{h, vs.pos, iv.x, iv.tmp, iv.r, iv.b, iv.diag, iv.past_x, vs.partial, dt * dt, red, sc};
    run_tile_pass<3, OpJacobi<3>>("jacobi_pass", c, tl, a);
  } else {
    OpJacobi<2>::Args a{h, vs.pos, iv.x, iv.tmp, iv.r, iv.b, iv.diag, iv.past_x, vs.partial, dt * dt, red, sc};
    run_tile_pass<2, OpJacobi<2>>("jacobi_pass", c, tl, a);
  }
}

void launch_implicit_assemble(int dim, const LaunchCfg& c, const DevTiling& tl, Planes e, Planes h,
                              VertexState vs, Material mat, float dt, ImplicitVecs iv) {
  if (dim == 3) {
    OpAssemble<3>::Args a{e, h, vs, mat, dt, iv};
    run_tile_pass<3, OpAssemble<3>>("implicit_assemble<3>", c, tl, a);
  } else {
    OpAssemble<2>::Args a{e, h, vs, mat, dt, iv};
    run_tile_pass<2, OpAssemble<2>>("implicit_assemble<2>", c, tl, a);
  }
}

void launch_implicit_diag(int dim, const LaunchCfg& c, const DevTiling& tl, Planes h, VertexState vs, float dt,
                          ImplicitVecs iv) {
  if (dim == 3) {
    OpDiag<3>::Args a{h, vs, dt * dt, iv};
    run_tile_pass<3, OpDiag<3>>("implicit_diag<3>", c, tl, a);
  } else {
    OpDiag<2>::Args a{h, vs, dt * dt, iv};
    run_tile_pass<2, OpDiag<2>>("implicit_diag<2>", c, tl, a);
  }
}

template <int D, int TR>
static void launch_apply_t(const LaunchCfg& c, const DevTiling& tl, Planes h, VertexState vs, float dt,
                           const float4* in, float4* out, int epilogue, const float4* aux, float4* resid,
                           double* red, const SolveScalars* sc) {
  typename OpApply<D, TR>::Args a{h, vs.pos, in, out, vs.partial, dt * dt, epilogue, aux, resid, red, sc};
  run_tile_pass<D, OpApply<D, TR>>("implicit_apply", c, tl, a);
}

void launch_apply(int dim, const LaunchCfg& c, const DevTiling& tl, Planes h, VertexState vs, float dt,
                  int transpose, const float4* in, float4* out, int epilogue, const float4* aux,
                  float4* resid_out, double* red, const SolveScalars* sc) {
  if (in == out) throw std::runtime_error("launch_apply: in-place application is not supported");
  if (dim == 3) {
    if (transpose) launch_apply_t<3, 1>(c, tl, h, vs, dt, in, out, epilogue, aux, resid_out, red, sc);
    else launch_apply_t<3, 0>(c, tl, h, vs, dt, in, out, epilogue, aux, resid_out, red, sc);
  } else {
    if (transpose) launch_apply_t<2, 1>(c, tl, h, vs, dt, in, out, epilogue, aux, resid_out, red, sc);
    else launch_apply_t<2, 0>(c, tl, h, vs, dt, in, out, epilogue, aux, resid_out, red, sc);
  }
}

int apply_tile_grid(const DevTiling& tl, bool split, bool persist) { return tile_grid<OpApply<3, 0>>(tl, split, persist); }

int reduce_slots(const DevTiling& tl, const LaunchCfg&) {
  const int nbb = (tl.n_boundary + kBoundaryBlock - 1) / kBoundaryBlock;
  const int nvb = (tl.n_vert + 255) / 256;
  const int a = tl.n_tiles + nbb;   // upper bound for every Op (persistent ones write fewer)
  return a > nvb ? a : nvb;
}

}  // namespace femb
