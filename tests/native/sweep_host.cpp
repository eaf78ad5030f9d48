// sweep_host.cpp -- TEST INFRASTRUCTURE.  Runs the reference-order sweep of the library on the HOST: the same
// sweep_row / sweep_decide (csrc/sweep_core.h) the CUDA kernels call, on the tables csrc/sweep.cpp builds from the real
// tiling (csrc/tiling.cpp), level by level, so that tests/test_sweep_host.py can check the row formula, the H-plane
// layout, the slot -> vertex tables and the dependency levels against the oracle and the reference's fixtures
// without a GPU.  What this cannot cover is the launch plumbing of kernels_sweep.cu / api_solve.cu.
// Built by the test with g++ (no nvcc, no CUDA runtime); never part of libfemb200.so.
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include "sweep.h"
#include "sweep_core.h"
#include "tiling.h"

using namespace femb;

namespace {
struct Scalars {   // the fields of launch.h: SolveScalars that sweep_decide touches
  double err = 0, p_err = 0;
  int iters = 0, first = 1, terminal = 0, action = 0;
};

struct Problem {
  int dim = 0, nv = 0;
  Tiling til;
  SweepHost sw;
  std::vector<float4> h0, h1, h2, pos, b, diag, past, x;
  float dt2 = 0;
  SweepTables tab() const { return SweepTables{sw.vptr.data(), sw.inc.data(), sw.sv.data(), sw.lptr.data(), sw.lverts.data()}; }
  SweepVecs vecs() { return SweepVecs{h0.data(), h1.data(), dim == 3 ? h2.data() : nullptr, pos.data(), b.data(), diag.data(), past.data(), x.data(), dt2}; }
};

float4 f4(float x, float y, float z, float w) { float4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }

template <int D>
double resid2(Problem& P) {   // ||b - A x||^2 (the device takes it from the operator pass, OpApply + AP_RESID)
  const SweepTables t = P.tab();
  const SweepVecs a = P.vecs();
  double s2 = 0;
  for (int v = 0; v < P.nv; ++v) {
    float acc[3];
    sweep_row_kx<D>(v, t, a, P.x[v], acc);
    const float s = a.dt2 * (1.f / a.pos[v].w);
    const float xv[3] = {P.x[v].x, P.x[v].y, P.x[v].z}, bv[3] = {P.b[v].x, P.b[v].y, P.b[v].z};
    for (int i = 0; i < D; ++i) {
      const float r = bv[i] - (xv[i] - s * acc[i]);
      s2 += (double)(r * r);
    }
  }
  return s2;
}

template <int D>
void one_sweep(Problem& P, int order) {
  const SweepTables t = P.tab();
  const SweepVecs a = P.vecs();
  for (size_t g = 0; g + 1 < P.sw.gptr.size(); ++g)
    for (int l = P.sw.gptr[g]; l < P.sw.gptr[g + 1]; ++l) {
      const int off = P.sw.lptr[l], n = P.sw.lptr[l + 1] - off;
      // inside a level the rows are independent: any order must give the same bits (order 1 runs them backwards)
      for (int i = 0; i < n; ++i) sweep_row<D>(P.sw.lverts[off + (order ? n - 1 - i : i)], t, a);
    }
}
}  // namespace

extern "C" {

// levels_out[old vertex id] = dependency level; returns the number of levels (< 0: error), groups in *n_groups
int sweep_host_levels(int dim, int n_vert, int n_elem, const float* verts, const int32_t* elems, int tile_elems,
                      int32_t* levels_out, int32_t* n_groups, int32_t* max_chain_rows) {
  Tiling til;
  if (tile_elems <= 0) tile_elems = kStride;   // api.cu: femb_object_create
  if (!build_tiling(dim, n_vert, n_elem, verts, elems, tile_elems, til).empty()) return -1;
  SweepHost sw;
  if (!build_sweep(til, sw).empty()) return -2;
  for (int l = 0; l < sw.n_levels; ++l)
    for (int p = sw.lptr[l]; p < sw.lptr[l + 1]; ++p) levels_out[til.old_of_new[sw.lverts[p]]] = l;
  *n_groups = (int)sw.gptr.size() - 1;
  int mc = 0;
  for (size_t g = 0; g + 1 < sw.gptr.size(); ++g)
    if (sw.gptr[g + 1] - sw.gptr[g] > 1)
      for (int l = sw.gptr[g]; l < sw.gptr[g + 1]; ++l) mc = std::max(mc, sw.lptr[l + 1] - sw.lptr[l]);
  *max_chain_rows = mc;
  return sw.n_levels;
}

// The whole solve of method 2 as api_solve.cu sequences it.  All arrays in the CALLER's numbering:
//   H [n_elem, dim, dim], mass [n_vert], b / past_x [n_vert, dim] -> x_out, past_out [n_vert, dim]
//   trace (optional) [max_trace, n_vert, dim]: the iterate after every sweep.  order: 0 / 1 row order inside a level.
int sweep_host_solve(int dim, int n_vert, int n_elem, const float* verts, const int32_t* elems, int tile_elems,
                     const float* H, const float* mass, const float* b, const float* past_x, float dt, float tol,
                     int max_iter, int fixed, int order, float* x_out, float* past_out, int32_t* iters_out,
                     float* err_out, float* trace, float* trace_err, int max_trace) {
  Problem P;
  P.dim = dim;
  P.nv = n_vert;
  if (tile_elems <= 0) tile_elems = kStride;
  if (!build_tiling(dim, n_vert, n_elem, verts, elems, tile_elems, P.til).empty()) return -1;
  if (!build_sweep(P.til, P.sw).empty()) return -2;
  const Tiling& t = P.til;
  const int nc = dim + 1;
  P.dt2 = dt * dt;
  P.h0.assign(t.n_slots, f4(0, 0, 0, 0)); P.h1 = P.h0; P.h2 = P.h0;
  for (int e = 0; e < n_elem; ++e) {       // kernels_implicit.cu: OpAssemble's plane layout
    const int s = t.slot_of_elem[e];
    const float* h = H + (size_t)e * dim * dim;
    if (dim == 3) {
      P.h0[s] = f4(0, 0, h[0], h[1]);
      P.h1[s] = f4(h[2], h[3], h[4], h[5]);
      P.h2[s] = f4(h[6], h[7], h[8], 0);
    } else {
      P.h0[s] = f4(0, 0, h[0], h[1]);
      P.h1[s] = f4(h[2], h[3], 0, 0);
    }
  }
  auto vec = [&](const float* src, std::vector<float4>& dst) {
    dst.assign(n_vert, f4(0, 0, 0, 0));
    for (int o = 0; o < n_vert; ++o) {
      const int v = t.new_of_old[o];
      dst[v].x = src[(size_t)o * dim]; dst[v].y = src[(size_t)o * dim + 1];
      if (dim == 3) dst[v].z = src[(size_t)o * dim + 2];
    }
  };
  vec(b, P.b);
  vec(past_x, P.past);
  P.pos.assign(n_vert, f4(0, 0, 0, 0));
  for (int o = 0; o < n_vert; ++o) P.pos[t.new_of_old[o]].w = mass[o];
  // diag(A) as OpDiag: 1 - dt^2/m_v sum_e c H_e[k,k], c = dim at corner 0
  std::vector<float> acc((size_t)n_vert * 3, 0.f);
  for (int e = 0; e < n_elem; ++e)
    for (int c = 0; c < nc; ++c) {
      const int v = t.new_of_old[elems[(size_t)e * nc + c]];
      const float w = c == 0 ? (float)dim : 1.f;
      for (int k = 0; k < dim; ++k) acc[(size_t)v * 3 + k] += w * H[(size_t)e * dim * dim + k * dim + k];
    }
  P.diag.assign(n_vert, f4(1, 1, 1, 0));
  for (int v = 0; v < n_vert; ++v) {
    const float s = P.dt2 * (1.f / P.pos[v].w);
    P.diag[v] = f4(1.f - s * acc[(size_t)v * 3], 1.f - s * acc[(size_t)v * 3 + 1], dim == 3 ? 1.f - s * acc[(size_t)v * 3 + 2] : 1.f, 0);
  }
  P.x.assign(n_vert, f4(0, 0, 0, 0));
  for (int v = 0; v < n_vert; ++v) P.x[v] = f4(0.5f * P.b[v].x, 0.5f * P.b[v].y, 0.5f * P.b[v].z, 0);   // implicit.py:231-232

  if (tol < 0.f) tol = 1e-5f;
  if (max_iter <= 0) max_iter = 20000;
  Scalars sc;
  int n_trace = 0;
  for (int it = 0; it < max_iter + 2 && !sc.terminal; ++it) {    // the iteration of api_solve.cu: build_program
    const double s2 = dim == 3 ? resid2<3>(P) : resid2<2>(P);
    sweep_decide(&sc, std::sqrt((float)s2), tol, max_iter, fixed);
    if (trace_err && it > 0 && it - 1 < max_trace) trace_err[it - 1] = (float)sc.err;
    if (sc.action == SW_ACCEPT) P.past = P.x;
    else if (sc.action == SW_REVERT) P.x = P.past;
    if (sc.terminal) break;
    if (dim == 3) one_sweep<3>(P, order); else one_sweep<2>(P, order);
    if (trace && n_trace < max_trace) {
      for (int o = 0; o < n_vert; ++o) {
        const float4 q = P.x[t.new_of_old[o]];
        float* d = trace + ((size_t)n_trace * n_vert + o) * dim;
        d[0] = q.x; d[1] = q.y; if (dim == 3) d[2] = q.z;
      }
      ++n_trace;
    }
  }
  for (int o = 0; o < n_vert; ++o) {
    const float4 q = P.x[t.new_of_old[o]], p = P.past[t.new_of_old[o]];
    x_out[(size_t)o * dim] = q.x; x_out[(size_t)o * dim + 1] = q.y;
    past_out[(size_t)o * dim] = p.x; past_out[(size_t)o * dim + 1] = p.y;
    if (dim == 3) { x_out[(size_t)o * dim + 2] = q.z; past_out[(size_t)o * dim + 2] = p.z; }
  }
  *iters_out = sc.iters;
  *err_out = (float)sc.err;
  return P.sw.n_levels;
}

}  // extern "C"
