// kernels_sweep.cu -- the reference's "Jacobi" in its own iterate order (implicit method FEMB_METHOD_SWEEP):
// the serial in-place sweep of implicit.py:391-404, parallelised over dependency levels without changing its
// result (sweep_core.h, sweep.cpp), under the loop of implicit.py:234-261.  Opt-in: implicit_method 0 keeps the
// simultaneous weighted Jacobi of kernels_implicit.cu (OpJacobi), which is 5-10x faster per iteration on large meshes
// but follows a different iterate path.
//
// Per iteration of the solve (api_solve.cu): residual pass (OpApply + AP_RESID: ||b - A x||^2 partials) ->
// k_sweep_decide (norm, accept / revert / stop, on the device) -> k_sweep_action (past_vec_x = vec_x or
// vec_x = past_vec_x) -> one k_sweep_levels launch per level group.  Everything early-exits once the loop has ended,
// so an iteration is a fixed launch sequence and replays as a CUDA graph.
#include <stdexcept>

#include "launch.h"
#include "sweep.h"
#include "sweep_core.h"

namespace femb {

template <int D>
__global__ void __launch_bounds__(kSweepBlock) k_sweep_levels(SweepTables t, SweepVecs a, int lvl0, int n_lvl,
                                                                     const SolveScalars* sc) {
  if (sc->done | sc->terminal) return;
  for (int l = lvl0; l < lvl0 + n_lvl; ++l) {
    const int off = t.lptr[l], n = t.lptr[l + 1] - off;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
      sweep_row<D>(t.lverts[off + i], t, a);
    // a chain of small levels runs in ONE CTA (the launcher guarantees gridDim.x == 1): the barrier orders this
    // level's stores to x before the next level's loads (which bypass L1, sweep_core.h)
    if (n_lvl > 1) __syncthreads();
  }
}

// sum of red[0..n) in a fixed order (thread-strided partial sums, then a tree): deterministic
__global__ void __launch_bounds__(1024) k_sweep_decide(int n_red, const double* red, SolveScalars* sc, float tol,
                                                       int max_iter, int fixed) {
  __shared__ double s[1024];
  if (sc->done) return;
  if (sc->terminal) {                      // the loop ended in the previous iteration: nothing left to apply
    if (threadIdx.x == 0) {
      sc->done = 1;
      sc->action = SW_NONE;
    }
    return;
  }
  double v = 0.0;
  for (int i = threadIdx.x; i < n_red; i += blockDim.x) v += red[i];
  s[threadIdx.x] = v;
  __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) sweep_decide(sc, sqrtf((float)s[0]), tol, max_iter, fixed);
}

__global__ void __launch_bounds__(256) k_sweep_action(int n, float4* x, float4* past, const SolveScalars* sc) {
  if (sc->done) return;
  const int action = sc->action;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || action == SW_NONE) return;
  if (action == SW_ACCEPT) past[i] = x[i];
  else x[i] = past[i];
}

void launch_sweep_decide(const LaunchCfg& c, int n_red, const double* red, SolveScalars* sc, float tol, int max_iter,
                         int fixed) {
  k_sweep_decide<<<1, 1024, 0, c.stream>>>(n_red, red, sc, tol, max_iter, fixed);
  note_launch("sweep_decide", c.stream);
}

void launch_sweep_action(const LaunchCfg& c, int n_vert, float4* x, float4* past, const SolveScalars* sc) {
  k_sweep_action<<<(n_vert + 255) / 256, 256, 0, c.stream>>>(n_vert, x, past, sc);
  note_launch("sweep_action", c.stream);
}

void launch_sweep_levels(int dim, const LaunchCfg& c, const SweepTables& t, const SweepVecs& a, int lvl0, int n_lvl,
                         int max_level_rows, const SolveScalars* sc) {
  if (n_lvl <= 0) return;
  if (n_lvl > 1 && max_level_rows > kSweepBlock) throw std::runtime_error("launch_sweep_levels: a chained level does not fit one CTA");
  const int grid = n_lvl > 1 ? 1 : (max_level_rows + kSweepBlock - 1) / kSweepBlock;
  if (dim == 3) k_sweep_levels<3><<<grid, kSweepBlock, 0, c.stream>>>(t, a, lvl0, n_lvl, sc);
  else k_sweep_levels<2><<<grid, kSweepBlock, 0, c.stream>>>(t, a, lvl0, n_lvl, sc);
  note_launch("sweep_levels", c.stream);
}

}  // namespace femb
