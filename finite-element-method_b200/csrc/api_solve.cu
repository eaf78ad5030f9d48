// api_solve.cu -- implicit_solver_neo_hookean (implicit.py:201-223) driven from the
// host: assemble once per time step, then matrix-free CG / A^T A-CG / Jacobi
// iterations whose convergence state lives on the device.  The host only polls
// the `done` flag every few iterations.
#include <algorithm>
#include <cstring>

#include "object.h"

using namespace femb;

namespace femb {
int fail(const std::exception& ex);
}

#define FEMB_TRY try {
#define FEMB_CATCH                   \
  }                                  \
  catch (const std::exception& ex) { \
    return femb::fail(ex);           \
  }                                  \
  return FEMB_OK;

static int apply_red_slots(const femb_object* o) { return o->dtl.n_tiles + (o->dtl.n_boundary + 255) / 256; }

static void poll(femb_object* o) {
  FEMB_CUDA(cudaMemcpyAsync(o->h_sc, o->d_sc, sizeof(SolveScalars), cudaMemcpyDeviceToHost, o->stream));
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
}

static void assemble(femb_object* o, float dt, bool with_diag) {
  o->prep_full_pass();
  launch_implicit_assemble(o->dim, o->cfg, o->dtl, o->e, o->h, o->vs, o->mat, dt, o->iv);
  if (with_diag) launch_implicit_diag(o->dim, o->cfg, o->dtl, o->h, o->vs, dt, o->iv);
}

extern "C" {

int femb_implicit_assemble(femb_object* o, float dt) {
  FEMB_TRY
  if (!o) throw ArgError("NULL object");
  if (!o->implicit) throw StateError("object was created without implicit buffers");
  FEMB_CUDA(cudaSetDevice(o->device));
  assemble(o, dt, true);
  FEMB_CATCH
}

int femb_implicit_apply(femb_object* o, float dt, int32_t transpose, const float* in, float* out) {
  FEMB_TRY
  if (!o || !in || !out) throw ArgError("NULL argument");
  if (!o->implicit) throw StateError("object was created without implicit buffers");
  FEMB_CUDA(cudaSetDevice(o->device));
  const size_t n = (size_t)o->n_vert * o->dim;
  o->ensure_stage(n);
  FEMB_CUDA(cudaMemsetAsync(o->iv.d, 0, (size_t)o->n_vert * sizeof(float4), o->stream));
  FEMB_CUDA(cudaMemcpyAsync(o->d_stage, in, n * sizeof(float), cudaMemcpyHostToDevice, o->stream));
  launch_unpack_vertex(o->cfg, o->dim, o->n_vert, o->d_new_of_old, o->d_stage, 0, o->dim, o->iv.d);
  launch_apply(o->dim, o->cfg, o->dtl, o->h, o->vs, dt, transpose, o->iv.d, o->iv.q, AP_PLAIN, nullptr, nullptr,
               o->d_red, nullptr);
  launch_pack_vertex(o->cfg, o->dim, o->n_vert, o->d_new_of_old, o->iv.q, 0, o->dim, o->d_stage);
  FEMB_CUDA(cudaMemcpyAsync(out, o->d_stage, n * sizeof(float), cudaMemcpyDeviceToHost, o->stream));
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
  FEMB_CATCH
}

int femb_implicit_solve(femb_object* o, float dt, int32_t method, int32_t preconditioned, int32_t max_iter,
                        float tol, int32_t* iters_out, float* resid_out) {
  FEMB_TRY
  if (!o) throw ArgError("NULL object");
  if (!o->implicit) throw StateError("object was created without implicit buffers");
  if (method != 0 && method != 1) throw ArgError("implicit method must be 0 (Jacobi) or 1 (CG)");
  FEMB_CUDA(cudaSetDevice(o->device));
  const bool fixed_count = (tol == 0.f);            // benchmark mode: exactly max_iter iterations, no polling
  if (tol < 0.f) tol = 1e-5f;                        // implicit.py:236,341
  if (max_iter <= 0) max_iter = method == 1 ? 500 : 20000;  // implicit.py:337,237
  const LaunchCfg& c = o->cfg;
  const DevTiling& tl = o->dtl;
  ImplicitVecs& iv = o->iv;
  const int nv = o->n_vert, ared = apply_red_slots(o), vred = (nv + 255) / 256;

  launch_scalars_reset(c, o->d_sc);
  assemble(o, dt, method == 0);
  int chunk = 4, since_poll = 0;
  bool finished = false;

  if (method == 1) {
    // ---- conjugate gradient, implicit.py:311-386
    const float4* rhs = iv.b;
    if (preconditioned == 1) {                       // ATb = A^T b, ATA = A^T A (289-299)
      launch_apply(o->dim, c, tl, o->h, o->vs, dt, 1, iv.b, iv.atb, AP_PLAIN, nullptr, nullptr, o->d_red, nullptr);
      launch_apply(o->dim, c, tl, o->h, o->vs, dt, 0, iv.b, iv.tmp, AP_PLAIN, nullptr, nullptr, o->d_red, nullptr);
      launch_apply(o->dim, c, tl, o->h, o->vs, dt, 1, iv.tmp, iv.q, AP_PLAIN, nullptr, nullptr, o->d_red, nullptr);
      rhs = iv.atb;
    } else {
      launch_apply(o->dim, c, tl, o->h, o->vs, dt, 0, iv.b, iv.q, AP_PLAIN, nullptr, nullptr, o->d_red, nullptr);
    }
    launch_cg_init(c, nv, iv, rhs, iv.q, o->d_red, o->d_sc, tol, max_iter);   // x = b, r = rhs - op(x), d = r
    for (int it = 0; it < max_iter && !finished; ++it) {
      if (preconditioned == 1) {
        launch_apply(o->dim, c, tl, o->h, o->vs, dt, 0, iv.d, iv.tmp, AP_PLAIN, nullptr, nullptr, o->d_red, o->d_sc);
        launch_apply(o->dim, c, tl, o->h, o->vs, dt, 1, iv.tmp, iv.q, AP_DOT, iv.d, nullptr, o->d_red, o->d_sc);
      } else {
        launch_apply(o->dim, c, tl, o->h, o->vs, dt, 0, iv.d, iv.q, AP_DOT, iv.d, nullptr, o->d_red, o->d_sc);
      }
      launch_cg_alpha(c, ared, o->d_red, o->d_sc);
      launch_cg_update_xr(c, nv, iv, o->d_red, o->d_sc);
      launch_cg_beta(c, vred, o->d_red, o->d_sc, tol, max_iter);
      launch_cg_update_d(c, nv, iv, o->d_sc);
      if (!fixed_count && ++since_poll >= chunk) {
        poll(o);
        finished = o->h_sc->done != 0;
        since_poll = 0;
        chunk = std::min(chunk * 2, 32);
      }
    }
  } else {
    // ---- the reference's damped "Jacobi" (227-262, 391-404) as a true parallel Jacobi
    launch_scale_xyz(c, nv, iv.b, iv.x, 0.5f);      // x0 = 0.5 b (231-232)
    for (int it = 0; it < max_iter + 2 && !finished; ++it) {
      launch_apply(o->dim, c, tl, o->h, o->vs, dt, 0, iv.x, iv.tmp, AP_RESID, iv.b, iv.r, o->d_red, o->d_sc);
      launch_jacobi_decide(c, ared, o->d_red, o->d_sc, tol, max_iter);
      launch_jacobi_update(c, nv, iv, o->d_sc);
      if (!fixed_count && ++since_poll >= chunk) {
        poll(o);
        finished = (o->h_sc->done | o->h_sc->terminal) != 0;
        since_poll = 0;
        chunk = std::min(chunk * 2, 32);
      }
    }
  }
  launch_copy_xyz(c, nv, iv.x, o->vs.vel);          // vel = x (222-223)
  if (iters_out || resid_out) {
    poll(o);
    if (iters_out) *iters_out = o->h_sc->iters;
    if (resid_out) *resid_out = method == 1 ? (float)o->h_sc->delta_new : (float)o->h_sc->err;
  }
  FEMB_CATCH
}

}  // extern "C"
