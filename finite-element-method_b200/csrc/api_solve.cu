// api_solve.cu -- implicit_solver_neo_hookean (implicit.py:201-223) driven from the
// host: assemble once per time step, then matrix-free CG / A^T A-CG / Jacobi
// iterations whose convergence state lives on the device.  The host only polls
// the `done` flag every few iterations.
//
// The solve is a SolveProgram (object.h): device-work segments separated by
// communication points.  femb_implicit_solve runs it straight through on one GPU;
// femb_implicit_solve_begin / _advance / _finish hand the communication points to a
// caller that owns the collectives (one process per GPU, torch.distributed / NCCL).
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "object.h"
#include "sweep.h"

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

constexpr bool kApplyPersist = FEMB_PERSIST_APPLY != 0;   // the operator passes use the persistent tile kernel

static int apply_red_slots(const femb_object* o) {
  const bool fused = o->fused_pass(kApplyPersist);   // fused exchange: one CTA per tile (object.h: peer_pass)
  return apply_tile_grid(o->dtl, o->split_reds(kApplyPersist), !fused) + (o->dtl.n_boundary + 255) / 256;
}

static void poll(femb_object* o) {
  FEMB_CUDA(cudaMemcpyAsync(o->h_sc, o->d_sc, sizeof(SolveScalars), cudaMemcpyDeviceToHost, o->stream));
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
  o->check_comm_error();
}

// tables of the reference-order sweep: built from the host copy of the tiling and uploaded on first use
void femb_object::ensure_sweep() {
  if (sweep_tab.vptr) return;
  SweepHost h;
  const std::string err = build_sweep(til, h);
  if (!err.empty()) throw LimitError(err);
  auto up = [&](auto*& dst, const auto& vec) {
    using T_ = typename std::remove_reference<decltype(vec)>::type::value_type;
    T_* p = alloc<T_>(vec.size(), false);
    if (!vec.empty()) FEMB_CUDA(cudaMemcpyAsync(p, vec.data(), vec.size() * sizeof(T_), cudaMemcpyHostToDevice, stream));
    dst = p;
  };
  SweepTables t{};
  up(t.inc, h.inc);
  up(t.sv, h.sv);
  up(t.lptr, h.lptr);
  up(t.lverts, h.lverts);
  up(t.vptr, h.vptr);
  FEMB_CUDA(cudaStreamSynchronize(stream));   // the host vectors go out of scope
  sweep_gptr = h.gptr;
  sweep_gmax.assign(h.gptr.size() - 1, 0);
  for (size_t g = 0; g + 1 < h.gptr.size(); ++g)
    for (int l = h.gptr[g]; l < h.gptr[g + 1]; ++l) sweep_gmax[g] = std::max(sweep_gmax[g], h.lptr[l + 1] - h.lptr[l]);
  sweep_levels = h.n_levels;
  sweep_tab = t;
}

static void assemble(femb_object* o, float dt, bool with_diag) {
  o->prep_full_pass();
  launch_implicit_assemble(o->dim, o->cfg, o->dtl, o->e, o->h, o->vs, o->mat, dt, o->iv);
  if (with_diag) launch_implicit_diag(o->dim, o->cfg, o->dtl, o->h, o->vs, dt, o->iv);
}

// ---------------------------------------------------------------------------
// program construction
// ---------------------------------------------------------------------------
namespace {

using PassFn = std::function<void(const LaunchCfg&)>;
using SumFn = std::function<void(int, const double*, const PeerComm&)>;

struct Builder {
  femb_object* o;
  bool dist;
  std::vector<SolveOp>* L;

  void add(std::function<void()> f, int comm = FEMB_COMM_NONE) { L->push_back(SolveOp{std::move(f), comm}); }

  // one tile pass (tile kernel + boundary kernel); multi-GPU: split around the halo exchange
  void pass(PassFn f, bool persistent_op = false) {
    femb_object* ob = o;
    if (!dist) {
      add([ob, f] { ob->prep_full_pass(); f(ob->cfg); });
      return;
    }
    if (ob->peer_on) {   // halo sums through peer memory, inside our own kernels: no communication point
      add([ob, f, persistent_op] { ob->peer_pass(f, persistent_op); });
      return;
    }
    add([ob, f] {
      LaunchCfg c = ob->cfg;
      c.phase = 1;
      f(c);
      launch_halo_pack(c, ob->dtl, ob->halo, ob->vs.partial);
    }, FEMB_COMM_HALO);
    add([ob, f] {
      LaunchCfg c = ob->cfg;
      c.phase = 2;
      launch_halo_unpack(c, ob->dtl, ob->halo, ob->vs.partial);
      ob->halo_dirty = true;
      f(c);
    });
  }

  // global sum of the per-CTA partials red[0..n_red) consumed by a scalar kernel; multi-GPU: the
  // local sum goes to the bound allreduce buffer, the caller sum-allreduces it, the consumer reads it
  void gsum(int n_red, SumFn consume) {
    femb_object* ob = o;
    if (!dist) {
      add([ob, n_red, consume] { consume(n_red, ob->d_red, PeerComm{}); });
      return;
    }
    if (ob->peer_on) {
      add([ob, n_red, consume] {
        if (ob->d_halo_dev) {   // fused: the scalar kernel sums locally, scatters, gathers and does the math
          consume(n_red, ob->d_red, ob->peer);
          return;
        }
        launch_reduce_final_send(ob->cfg, n_red, ob->d_red, ob->peer);   // my partial -> every rank's slot
        consume(0, nullptr, ob->peer);                                     // gathers the slots, then the scalar math
      });
      return;
    }
    add([ob, n_red] { launch_reduce_final(ob->cfg, n_red, ob->d_red, ob->d_allred); }, FEMB_COMM_ALLREDUCE);
    add([ob, consume] { consume(1, ob->d_allred, PeerComm{}); });
  }

  // two global sums consumed by ONE scalar kernel (single-reduction CG): partials red_a[0..n_a), red_b[0..n_b)
  using Sum2Fn = std::function<void(int, const double*, int, const double*, const PeerComm&)>;
  void gsum2(int n_a, const double* red_a, int n_b, const double* red_b, Sum2Fn consume) {
    femb_object* ob = o;
    if (!dist) {
      add([=] { consume(n_a, red_a, n_b, red_b, PeerComm{}); });
      return;
    }
    if (ob->peer_on) {   // the scalar kernel sums locally, scatters, gathers (twice) and does the math
      add([=] { consume(n_a, red_a, n_b, red_b, ob->peer); });
      return;
    }
    add([=] {            // caller-side collective: both local sums into the bound buffer, ONE all-reduce
      launch_reduce_final(ob->cfg, n_a, red_a, ob->d_allred);
      launch_reduce_final(ob->cfg, n_b, red_b, ob->d_allred + 1);
    }, FEMB_COMM_ALLREDUCE);
    add([=] { consume(1, ob->d_allred, 1, ob->d_allred + 1, PeerComm{}); });
  }
};

void build_program(femb_object* o, SolveProgram& P, float dt, int method, int preconditioned, int max_iter, float tol) {
  const bool dist = o->halo.n > 0;
  if (dist && !o->peer_on && (!o->halo.send || !o->halo.recv || !o->d_allred))
    throw StateError("multi-GPU solve: bind the halo buffers (femb_halo_bind) and the allreduce buffer (femb_reduce_bind) first");
  P.fixed = (tol == 0.f);                              // benchmark mode: exactly max_iter iterations
  if (tol < 0.f) tol = 1e-5f;                          // implicit.py:236,341
  if (max_iter <= 0) max_iter = method == 1 ? 500 : 20000;  // implicit.py:337,237
  P.method = method;
  P.pre = 0;   // set below: 1 when the single-reduction CG arrangement is chosen
  o->last_method = method;
  P.max_iter = max_iter;
  const float ktol = P.fixed ? -1.f : tol;
  const int fixed = P.fixed ? 1 : 0;
  const int dim = o->dim, nv = o->n_vert, ared = apply_red_slots(o), vred = (nv + 255) / 256;
  const RedMask mask{o->dtl.n_interior, o->dtl.bskip};
  ImplicitVecs iv = o->iv;
  SolveScalars* sc = o->d_sc;

  auto apply = [o, dim, dt](int tr, const float4* in, float4* out, int epi, const float4* aux, float4* resid,
                            const SolveScalars* s) -> PassFn {
    return [=](const LaunchCfg& c) {
      launch_apply(dim, c, o->dtl, o->h, o->vs, dt, tr, in, out, epi, aux, resid, o->d_red, s);
    };
  };

  Builder pro{o, dist, &P.prologue}, it{o, dist, &P.iteration};
  pro.add([o, sc] { launch_scalars_reset(o->cfg, sc); });
  pro.pass([o, dim, dt](const LaunchCfg& c) {
    launch_implicit_assemble(dim, c, o->dtl, o->e, o->h, o->vs, o->mat, dt, o->iv);
  });

  if (method == 1) {
    // ---- conjugate gradient, implicit.py:311-386
    const float4* rhs = iv.b;
    if (preconditioned == 1) {                         // ATb = A^T b, ATA = A^T A (289-299)
      pro.pass(apply(1, iv.b, iv.atb, AP_PLAIN, nullptr, nullptr, nullptr), kApplyPersist);
      pro.pass(apply(0, iv.b, iv.tmp, AP_PLAIN, nullptr, nullptr, nullptr), kApplyPersist);
      pro.pass(apply(1, iv.tmp, iv.q, AP_PLAIN, nullptr, nullptr, nullptr), kApplyPersist);
      rhs = iv.atb;
    } else {
      pro.pass(apply(0, iv.b, iv.q, AP_PLAIN, nullptr, nullptr, nullptr), kApplyPersist);
    }
    // A^T A variant: textbook loop or the single-reduction arrangement (below).  Measured on C4, one GPU: 0.583 vs
    // 0.624 ms per iteration -- the merged vector kernel streams nine 43 MB vectors from HBM where the textbook
    // kernels find most of theirs in L2 -- so the single-reduction form is used only where a rank's vectors are
    // L2-resident and the launch / all-sum count is what costs (thin slabs).  FEMB_CG_SINGLE_REDUCTION=0|1 overrides.
    int cgcg = 0;
    if (preconditioned == 1) {
      cgcg = dist && (size_t)nv * sizeof(float4) * 9 <= (size_t)64 << 20;
      if (const char* e = std::getenv("FEMB_CG_SINGLE_REDUCTION")) cgcg = e[0] == '1';
    }
    pro.add([=] { launch_cg_init_ex(o->cfg, nv, iv, rhs, iv.q, o->d_red, mask, cgcg); });   // x = b, r = rhs - op(x), d = r
    pro.gsum(vred, [=](int n, const double* red, const PeerComm& pc) { launch_cg_delta0(o->cfg, n, red, sc, ktol, max_iter, pc); });
    if (cgcg)   // the first iteration's gamma: the same partials, where the iteration's scalar kernel looks for them
      pro.add([=] { FEMB_CUDA(cudaMemcpyAsync(o->d_red2, o->d_red, (size_t)vred * sizeof(double), cudaMemcpyDeviceToDevice, o->stream)); });

    if (preconditioned == 1 && !cgcg) {
      it.pass(apply(0, iv.d, iv.tmp, AP_PLAIN, nullptr, nullptr, sc), kApplyPersist);
      it.pass(apply(1, iv.tmp, iv.q, AP_DOT, iv.d, nullptr, sc), kApplyPersist);
      it.gsum(ared, [=](int n, const double* red, const PeerComm& pc) { launch_cg_alpha(o->cfg, n, red, sc, pc); });
      it.add([=] { launch_cg_update_xr(o->cfg, nv, iv, o->d_red, sc, mask); });
      it.gsum(vred, [=](int n, const double* red, const PeerComm& pc) { launch_cg_beta(o->cfg, n, red, sc, ktol, max_iter, pc); });
      it.add([=] { launch_cg_update_d(o->cfg, nv, iv, sc); });
    } else if (cgcg) {
      P.pre = 1;
      // Single-reduction CG (Chronopoulos & Gear) -- algebraically the loop of implicit.py:341-386:
      //   w = Op r ; gamma = r.r ; delta = r.w ; beta = gamma / gamma_old ; alpha = gamma / (delta - beta gamma / alpha_old)
      //   p = r + beta p ; s = w + beta s ; x += alpha p ; r -= alpha s                 (s = Op p by linearity)
      // Both dot products are consumed at ONE point, so an iteration is: two operator passes, one scalar kernel
      // (one all-sum on a slab instead of two), one streaming vector kernel -- 6 launches instead of 8.
      // gamma's partials come from the vector kernel of the previous iteration (cg_init for the first).
      // (Folding the vector updates into the first pass's vertex loads was measured: 0.649 vs 0.583 ms per
      // iteration on C4 -- the tile kernels are latency-bound, the streaming kernel runs at the HBM roof.)
      it.pass(apply(0, iv.r, iv.tmp, AP_PLAIN, nullptr, nullptr, sc), kApplyPersist);
      it.pass(apply(1, iv.tmp, iv.w, AP_DOT, iv.r, nullptr, sc), kApplyPersist);
      it.gsum2(vred, o->d_red2, ared, o->d_red,
               [=](int na, const double* ra, int nb, const double* rb, const PeerComm& pc) {
                 launch_cgcg_scalar(o->cfg, na, ra, nb, rb, sc, ktol, max_iter, pc);
               });
      it.add([=] { launch_cgcg_update(o->cfg, nv, iv, o->d_red2, sc, mask); });
    } else {
      it.pass(apply(0, iv.d, iv.q, AP_DOT, iv.d, nullptr, sc), kApplyPersist);
      it.gsum(ared, [=](int n, const double* red, const PeerComm& pc) { launch_cg_alpha(o->cfg, n, red, sc, pc); });
      it.add([=] { launch_cg_update_xr(o->cfg, nv, iv, o->d_red, sc, mask); });
      it.gsum(vred, [=](int n, const double* red, const PeerComm& pc) { launch_cg_beta(o->cfg, n, red, sc, ktol, max_iter, pc); });
      it.add([=] { launch_cg_update_d(o->cfg, nv, iv, sc); });
    }
  } else if (method == FEMB_METHOD_SWEEP) {
    // ---- the reference's damped "Jacobi" in its own iterate order: the serial in-place sweep (391-404) run level by
    // dependency level (kernels_sweep.cu), loop control of 234-261 on the device
    if (dist)
      throw StateError("the reference-order sweep (method 2) orders rows by global vertex id and runs on one GPU; "
                       "use method 0 or 1 on slab-partitioned objects");
    o->ensure_sweep();
    pro.pass([o, dim, dt](const LaunchCfg& c) { launch_implicit_diag(dim, c, o->dtl, o->h, o->vs, dt, o->iv); });
    pro.add([=] { launch_scale_xyz(o->cfg, nv, iv.b, iv.x, 0.5f); });               // x0 = 0.5 b (231-232)
    it.pass(apply(0, iv.x, iv.q, AP_RESID, iv.b, iv.r, sc), kApplyPersist);         // compute_error (277-286)
    it.add([=] { launch_sweep_decide(o->cfg, ared, o->d_red, sc, tol, max_iter, fixed); });
    it.add([=] { launch_sweep_action(o->cfg, nv, iv.x, iv.past_x, sc); });
    it.add([o, dim, dt, iv, sc] {
      const SweepVecs a{o->h.p0, o->h.p1, o->h.p2, o->vs.pos, iv.b, iv.diag, iv.past_x, iv.x, dt * dt};
      for (size_t g = 0; g + 1 < o->sweep_gptr.size(); ++g)
        launch_sweep_levels(dim, o->cfg, o->sweep_tab, a, o->sweep_gptr[g], o->sweep_gptr[g + 1] - o->sweep_gptr[g],
                            o->sweep_gmax[g], sc);
    });
  } else {
    // ---- the reference's damped "Jacobi" (227-262, 391-404) as a true parallel Jacobi
    pro.pass([o, dim, dt](const LaunchCfg& c) { launch_implicit_diag(dim, c, o->dtl, o->h, o->vs, dt, o->iv); });
    pro.add([=] { launch_scale_xyz(o->cfg, nv, iv.b, iv.x, 0.5f); });               // x0 = 0.5 b (231-232)
    // one fused pass per iteration (residual norm + the next iterate, kernels_implicit.cu: OpJacobi) + the decision
    it.pass([o, dim, dt, sc](const LaunchCfg& c) {
      launch_jacobi_pass(dim, c, o->dtl, o->h, o->vs, dt, o->iv, o->d_red, sc);
    }, kApplyPersist);
    it.gsum(ared, [=](int n, const double* red, const PeerComm& pc) { launch_jacobi_decide(o->cfg, n, red, sc, tol, max_iter, fixed, pc); });
  }
}

// Runs segments up to the next communication point; returns that femb_comm.
int advance(femb_object* o) {
  SolveProgram& P = *o->prog;
  std::vector<SolveOp>& L = P.in_prologue ? P.prologue : P.iteration;
  while (P.pc < L.size()) {
    const SolveOp& op = L[P.pc++];
    op.run();
    if (op.comm != FEMB_COMM_NONE) return op.comm;
  }
  P.pc = 0;
  if (!P.in_prologue) P.iters_issued++;
  P.in_prologue = false;
  return FEMB_COMM_ITER;
}

void check_solve_args(femb_object* o, int method) {
  if (!o) throw ArgError("NULL object");
  if (!o->implicit) throw StateError("object was created without implicit buffers");
  if (method != FEMB_METHOD_JACOBI && method != FEMB_METHOD_CG && method != FEMB_METHOD_SWEEP)
    throw ArgError("implicit method must be 0 (Jacobi), 1 (CG) or 2 (reference-order sweep)");
}

void finish(femb_object* o, int32_t* iters_out, float* resid_out) {
  if (o->prog->method == 0) launch_jacobi_finish(o->cfg, o->n_vert, o->iv, o->d_sc);   // answer buffer -> vec_x, past_vec_x
  launch_copy_xyz(o->cfg, o->n_vert, o->iv.x, o->vs.vel);          // vel = x (222-223)
  if (iters_out || resid_out) {
    poll(o);
    if (iters_out) *iters_out = o->h_sc->iters;
    if (resid_out) *resid_out = o->prog->method == 1 ? (float)o->h_sc->delta_new : (float)o->h_sc->err;
  }
  o->prog.reset();
}

}  // namespace

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
  o->prep_full_pass();
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

int femb_implicit_apply_timed(femb_object* o, float dt, int32_t transpose, int32_t n, float* tile_ms, float* boundary_ms) {
  FEMB_TRY
  if (!o || n <= 0 || !tile_ms || !boundary_ms) throw ArgError("femb_implicit_apply_timed: bad argument");
  if (!o->implicit) throw StateError("object was created without implicit buffers");
  FEMB_CUDA(cudaSetDevice(o->device));
  o->prep_full_pass();
  std::vector<cudaEvent_t> ev((size_t)n * 3);
  for (auto& e : ev) FEMB_CUDA(cudaEventCreate(&e));
  LaunchCfg cfg = o->cfg;
  for (int s = 0; s < n; ++s) {      // q = A d (or A^T d) on the vectors of the last solve: same traffic as in the solve
    FEMB_CUDA(cudaEventRecord(ev[3 * s], o->stream));
    cfg.ev_mid = ev[3 * s + 1];
    launch_apply(o->dim, cfg, o->dtl, o->h, o->vs, dt, transpose, o->iv.d, o->iv.q, AP_DOT, o->iv.d, nullptr, o->d_red,
                 nullptr);
    FEMB_CUDA(cudaEventRecord(ev[3 * s + 2], o->stream));
  }
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
  double t = 0.0, b = 0.0;
  for (int s = 0; s < n; ++s) {
    float ms = 0.f;
    FEMB_CUDA(cudaEventElapsedTime(&ms, ev[3 * s], ev[3 * s + 1])); t += ms;
    FEMB_CUDA(cudaEventElapsedTime(&ms, ev[3 * s + 1], ev[3 * s + 2])); b += ms;
  }
  for (auto& e : ev) cudaEventDestroy(e);
  *tile_ms = (float)t;
  *boundary_ms = (float)b;
  FEMB_CATCH
}

int femb_implicit_solve(femb_object* o, float dt, int32_t method, int32_t preconditioned, int32_t max_iter,
                        float tol, int32_t* iters_out, float* resid_out) {
  FEMB_TRY
  check_solve_args(o, method);
  if (o->halo.n > 0 && !o->peer_on)
    throw StateError("object has halo vertices: bind peers, or drive the solve with femb_implicit_solve_begin/_advance/_finish");
  FEMB_CUDA(cudaSetDevice(o->device));
  o->prog.reset(new SolveProgram());
  SolveProgram& P = *o->prog;
  build_program(o, P, dt, method, preconditioned, max_iter, tol);
  advance(o);                                          // prologue (no communication on one GPU)
  // Jacobi needs one more residual evaluation than updates, plus the pass that sets `done`
  // single-reduction CG evaluates r.r of the iterate of update k in iteration k + 1: one more pass to see the last one
  const int n_it = method == 1 ? P.max_iter + ((P.pre && !P.fixed) ? 1 : 0) : P.max_iter + 2;
  // one iteration (all its kernels, peer-memory communication included) as a CUDA graph: the argument
  // values of an iteration never change, so the graph is cached per (dt, method, tolerance ...)
  SolveGraph& g = o->solve_graph;
  const bool use_graph = n_it >= 4 && o->stream != nullptr && o->stream != cudaStreamLegacy;
  if (use_graph) {
    const int arrangement = preconditioned | (P.pre << 4);   // the single-reduction loop is a different graph
    const bool hit = g.exec && g.dt == dt && g.tol == tol && g.method == method && g.pre == arrangement &&
                     g.max_iter == P.max_iter;
    if (!hit) {
      if (g.exec) { cudaGraphExecDestroy(g.exec); g = SolveGraph(); }
      cudaGraph_t graph = nullptr;
      const uint64_t before = launch_count();
      FEMB_CUDA(cudaStreamBeginCapture(o->stream, cudaStreamCaptureModeThreadLocal));
      try {
        advance(o);
      } catch (...) {
        cudaStreamEndCapture(o->stream, &graph);
        if (graph) cudaGraphDestroy(graph);
        throw;
      }
      FEMB_CUDA(cudaStreamEndCapture(o->stream, &graph));
      const cudaError_t ie = cudaGraphInstantiate(&g.exec, graph, 0);
      cudaGraphDestroy(graph);
      cuda_check(ie, "cudaGraphInstantiate");
      g.launches = launch_count() - before;
      g.dt = dt; g.tol = tol; g.method = method; g.pre = arrangement; g.max_iter = P.max_iter;
    }
  }
  int chunk = 4, since_poll = 0;
  bool finished = false;
  for (int it = 0; it < n_it && !finished; ++it) {
    if (use_graph) {
      FEMB_CUDA(cudaGraphLaunch(g.exec, o->stream));
      for (uint64_t i = 0; i < g.launches; ++i) note_launch("graph", o->stream);
    } else {
      advance(o);
    }
    if (!P.fixed && ++since_poll >= chunk) {
      poll(o);
      finished = (o->h_sc->done | o->h_sc->terminal) != 0;
      since_poll = 0;
      chunk = std::min(chunk * 2, 32);
    }
  }
  finish(o, iters_out, resid_out);
  FEMB_CATCH
}

int femb_reduce_bind(femb_object* o, void* dev_doubles) {
  FEMB_TRY
  if (!o || !dev_doubles) throw ArgError("femb_reduce_bind: NULL argument");
  o->d_allred = static_cast<double*>(dev_doubles);
  FEMB_CATCH
}

int femb_implicit_solve_begin(femb_object* o, float dt, int32_t method, int32_t preconditioned, int32_t max_iter,
                              float tol) {
  FEMB_TRY
  check_solve_args(o, method);
  FEMB_CUDA(cudaSetDevice(o->device));
  o->prog.reset(new SolveProgram());
  try {
    build_program(o, *o->prog, dt, method, preconditioned, max_iter, tol);
  } catch (...) {
    o->prog.reset();
    throw;
  }
  FEMB_CATCH
}

int femb_implicit_solve_advance(femb_object* o) {
  try {
    if (!o) throw ArgError("NULL object");
    if (!o->prog) throw StateError("no solve in progress: call femb_implicit_solve_begin first");
    FEMB_CUDA(cudaSetDevice(o->device));
    return advance(o);
  } catch (const std::exception& ex) {
    return femb::fail(ex);
  }
}

int femb_implicit_solve_poll(femb_object* o, int32_t* done_out, int32_t* iters_out, float* resid_out) {
  FEMB_TRY
  if (!o) throw ArgError("NULL object");
  if (!o->implicit) throw StateError("object was created without implicit buffers");
  FEMB_CUDA(cudaSetDevice(o->device));
  poll(o);
  if (done_out) *done_out = (o->h_sc->done | o->h_sc->terminal) != 0;
  if (iters_out) *iters_out = o->h_sc->iters;
  if (resid_out) *resid_out = o->last_method == 1 ? (float)o->h_sc->delta_new : (float)o->h_sc->err;
  FEMB_CATCH
}

int femb_implicit_solve_finish(femb_object* o, int32_t* iters_out, float* resid_out) {
  FEMB_TRY
  if (!o) throw ArgError("NULL object");
  if (!o->prog) throw StateError("no solve in progress");
  FEMB_CUDA(cudaSetDevice(o->device));
  finish(o, iters_out, resid_out);
  FEMB_CATCH
}

}  // extern "C"
