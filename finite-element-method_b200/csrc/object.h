// object.h -- the host-side object behind the opaque femb_object handle: what the reference keeps in the Taichi
// fields of object.Object (object.py:52-93) plus the launch state (streams, CUDA graphs, multi-GPU exchange).
#pragma once
#include <cuda_runtime.h>

#include <functional>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/femb200.h"
#include "launch.h"
#include "sweep_core.h"
#include "tiling.h"

namespace femb {

struct CudaError : std::runtime_error {
  using std::runtime_error::runtime_error;
};
struct ArgError : std::runtime_error {
  using std::runtime_error::runtime_error;
};
struct StateError : std::runtime_error {
  using std::runtime_error::runtime_error;
};
struct LimitError : std::runtime_error {
  using std::runtime_error::runtime_error;
};

inline void cuda_check(cudaError_t e, const char* what) {
  if (e != cudaSuccess) throw CudaError(std::string(what) + ": " + cudaGetErrorString(e));
}
#define FEMB_CUDA(x) ::femb::cuda_check((x), #x)

struct SolveGraph {         // cached CUDA graph of ONE solver iteration
  cudaGraphExec_t exec = nullptr;
  float dt = 0.f, tol = 0.f;
  int method = -1, pre = -1, max_iter = 0;
  uint64_t launches = 0;
};

struct StepGraph {          // cached CUDA graph of n fused explicit steps
  cudaGraphExec_t exec = nullptr;
  float dt = 0.f, g[3] = {0, 0, 0}, damp = 0.f;
  int n_steps = 0, blocks_version = -1;
  uint64_t launches = 0;
};

// The implicit solve as a list of device-work segments separated by communication points, so that
// the same code drives the single-GPU solve (no communication) and the slab-partitioned solve, where
// the caller (one process per GPU) performs the halo exchange / allreduce between segments.
struct SolveOp {
  std::function<void()> run;
  int comm;                 // femb_comm the caller must perform AFTER this op (0 = none)
};
struct SolveProgram {
  std::vector<SolveOp> prologue, iteration;
  size_t pc = 0;
  bool in_prologue = true;
  int iters_issued = 0;
  int method = 0, max_iter = 0, pre = 0;
  bool fixed = false;
};

}  // namespace femb

struct femb_object {
  int dim = 0, n_vert = 0, n_elem = 0, device = 0;
  bool implicit = false;
  femb::Material mat{};
  femb::Tiling til;                 // host copy (maps kept for diagnostics)
  femb::DevTiling dtl{};
  femb::Planes e{}, h{};
  femb::Planes ex{};                // what the explicit kernels read: e, or its pair-interleaved copy (launch.h)
  uint32_t* d_lidx = nullptr;
  femb::VertexState vs{};
  float4* ref_pos = nullptr;
  femb::ImplicitVecs iv{};
  float* d_phi = nullptr;           // per slot
  double* d_red = nullptr;
  double* d_red2 = nullptr;         // second partial buffer (single-reduction CG: r.r of the first pass)
  int red_slots = 0;
  double* d_U = nullptr;
  unsigned long long* d_diag = nullptr;   // femb_diagnostics counters
  unsigned long long* d_trace = nullptr;  // femb_trace_enable: kernel timeline words
  femb::SolveScalars* d_sc = nullptr;
  femb::SolveScalars* h_sc = nullptr;  // pinned
  int32_t* d_new_of_old = nullptr;
  int32_t* d_slot_of_elem = nullptr;
  float* d_stage = nullptr;         // marshalling buffer
  // asynchronous field transfers (femb_set_async / femb_get_async): one staging buffer per field and direction,
  // copies on two dedicated streams (PCIe is full duplex), ordered against the compute stream by events
  struct Xfer {
    float* buf = nullptr;
    size_t floats = 0;
    cudaEvent_t copied = nullptr;   // the PCIe copy of this buffer has finished
    cudaEvent_t consumed = nullptr; // the unpack kernel has read it / the pack kernel may overwrite it
    bool used = false;
  };
  Xfer xin[48], xout[48];
  cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
  size_t stage_floats = 0;
  float4* d_blocks = nullptr;
  int n_blocks = 0, blocks_version = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  femb::LaunchCfg cfg{};
  femb::StepGraph graph;
  femb::SolveGraph solve_graph;
  std::unique_ptr<femb::SolveProgram> prog;   // implicit solve in progress (begin/advance/finish)
  int last_method = 1;              // method of the most recent solve (what `resid` means)
  double* d_allred = nullptr;       // multi-GPU: caller-owned device doubles the caller sum-allreduces
  femb::PeerComm peer{};            // multi-GPU over peer memory (femb_peer_bind); peer_on: no caller-side collectives
  bool peer_on = false;
  double* d_allred_own = nullptr;   // library-owned scalar for the in-kernel all-sum
  femb::Halo halo;                  // multi-GPU: vertices shared with neighbouring ranks
  bool halo_dirty = false;          // remote partial slots hold data of an earlier exchange
  // every un-phased pass on a halo object must see zeros in the remote slots
  void prep_full_pass() {
    if (halo.n && halo_dirty) {
      femb::launch_halo_zero_remote(cfg, dtl, halo, vs.partial);
      halo_dirty = false;
    }
  }
  // headless frames (femb_render): primitives in the caller's numbering + key / RGB frame buffers
  std::vector<int32_t> elems2d;     // dim 2: the element triangles, kept for the default primitive list
  int32_t* d_prims = nullptr;
  int n_prims = 0, prims_version = -1;
  unsigned long long* d_fb = nullptr;
  uint8_t* d_rgb = nullptr;
  int fb_pixels = 0;
  std::vector<void*> allocs;
  int64_t device_bytes = 0;

  template <class T>
  T* alloc(size_t count, bool zero = true) {
    void* p = nullptr;
    const size_t bytes = (count ? count : 1) * sizeof(T);
    FEMB_CUDA(cudaMalloc(&p, bytes));
    allocs.push_back(p);
    device_bytes += (int64_t)bytes;
    if (zero) FEMB_CUDA(cudaMemsetAsync(p, 0, bytes, stream));
    return static_cast<T*>(p);
  }
  const femb::HaloDev* d_halo_dev = nullptr;   // peer mode, fused exchange (default; FEMB_PEER_FUSED=0 disables)
  const int2* d_hrange = nullptr;   // local partial slots of every shared vertex (sender side)
  int n_send_ctas = 0;
  cudaStream_t side = nullptr;      // peer mode: the halo send runs here, overlapping the non-halo tiles
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  bool pass_split() const { return peer_on && side && dtl.n_halo_tiles > 0 && dtl.n_halo_tiles < dtl.n_tiles; }
  // Does a pass of this kind run with the exchange fused into its own kernels and one CTA per tile?
  // Ops whose tile kernel is persistent keep it (plus the send kernel on the side stream) while a rank still
  // owns many tiles per SM slot -- there the persistent kernel is worth more than the kernel boundaries the
  // fusion removes (N=2 on C4: 0.342 vs 0.370 ms per A^T A-CG iteration); on thin slabs it is the other way
  // round (the slab of N=8: 0.137 vs 0.123).  FEMB_PEER_FUSED_WAVES moves the threshold.
  int fuse_max_waves = 8;
  // "late send" (FEMB_PEER_MODE=late): ONE tile kernel over all tiles (persistent where the Op wants it), then the
  // boundary kernel whose first CTAs send the shared vertices' sums while the rest finish the ordinary boundary
  // vertices; only CTAs that own a shared vertex wait for the neighbours.  Two kernels per pass, no split, no
  // programmatic dependent launch; the NVLink latency hides behind the boundary kernel's own work.
  bool late_send = false;
  // "one tile kernel" (FEMB_PEER_MODE=one): halo tiles first, the sender CTAs in the middle of the SAME launch (they
  // wait for a counter of finished halo tiles), then the boundary kernel that receives -- two kernels per pass like
  // on one GPU.  Measured against the split form (default) on C4, same box, after the sender and the fence-free
  // exchange had been rewritten: N = 8 explicit 0.04733 vs 0.04671 ms, A^T A-CG 0.1337 vs 0.1302, Jacobi 0.0732 vs
  // 0.0704; N = 2 0.1508 vs 0.1484.  The split form's second kernel overlaps the first one's last wave
  // (programmatic dependent launch) and its senders need no counter, so it stays the default.
  int one_kernel = 0;
  bool one_pass(bool persistent_op) const { return one_kernel != 0 && fused_pass(persistent_op); }
  bool fused_pass(bool persistent_op) const {
    if (late_send) return false;
    if (!pass_split() || !d_halo_dev) return false;
    return !persistent_op || (dtl.n_tiles - dtl.n_halo_tiles) <= (long long)fuse_max_waves * dtl.grid;
  }
  bool late_pass() const { return late_send && peer_on && d_halo_dev; }
  // tile partials laid out as two launches?
  bool split_reds(bool persistent_op = false) const { return pass_split() && !late_pass() && !one_pass(persistent_op); }
  // One tile pass of a slab rank over peer memory.  f(cfg) launches the pass's kernels for the phase / part
  // in cfg.
  //   fused:   halo tiles | [sender CTAs + other tiles] (programmatic dependent launch) | boundary kernel that
  //            waits for the neighbours itself.  One stream, three kernels.
  //   else:    halo tiles | send kernel (side stream) || other tiles | boundary kernel (receives itself when
  //            the fused description exists, else after the receive kernel)
  template <class F>
  void peer_pass(F f, bool persistent_op = false) {
    using namespace femb;
    LaunchCfg c1 = cfg, c2 = cfg;
    c1.phase = 1;
    c2.phase = 2;
    if (late_pass()) {
      f(c1);                 // every tile, one launch
      c2.hd = d_halo_dev;
      c2.fuse_recv = 1;
      c2.bnd_send = n_send_ctas;
      f(c2);
      return;
    }
    if (one_pass(persistent_op)) {
      c1.no_persist = c2.no_persist = 1;
      c1.hd = d_halo_dev;
      c1.n_send = n_send_ctas;
      c1.send_at = dtl.n_halo_tiles + cfg.sms * femb::kMinCtas;   // one wave of ordinary tiles after the last halo tile
      f(c1);
      c2.hd = d_halo_dev;
      c2.fuse_recv = 1;
      f(c2);
      return;
    }
    if (fused_pass(persistent_op)) {
      c1.no_persist = c2.no_persist = 1;   // sender CTAs and persistent tile CTAs do not mix (run_tile_pass)
      c1.part = 1;
      f(c1);
      c1.part = 2;
      c1.hd = d_halo_dev;
      c1.n_send = n_send_ctas;
      f(c1);
      c2.part = 1;
      c2.hd = d_halo_dev;
      c2.fuse_recv = 1;
      f(c2);   // (the remote partial slots are not touched: halo_dirty keeps its value)
      return;
    }
    if (pass_split()) {
      c1.part = 1;
      f(c1);
      FEMB_CUDA(cudaEventRecord(ev_fork, stream));
      FEMB_CUDA(cudaStreamWaitEvent(side, ev_fork, 0));
      LaunchCfg cs = cfg;
      cs.stream = side;
      launch_halo_send(cs, dtl, peer, halo, vs.partial, vs.pos, d_halo_dev ? 2 : 0, d_hrange);
      FEMB_CUDA(cudaEventRecord(ev_join, side));
      c1.part = 2;
      f(c1);
      FEMB_CUDA(cudaStreamWaitEvent(stream, ev_join, 0));
      c2.part = 1;   // tells the boundary kernel where the tile partials of a split pass end
    } else {
      f(c1);
      launch_halo_send(c1, dtl, peer, halo, vs.partial, vs.pos, d_halo_dev ? 2 : 0, d_hrange);
    }
    if (d_halo_dev) {
      c2.hd = d_halo_dev;
      c2.fuse_recv = 1;
    } else {
      launch_halo_recv(c2, dtl, peer, halo, vs.partial, vs.pos, 0);
      halo_dirty = true;
    }
    f(c2);
  }
  // reference-order sweep (FEMB_METHOD_SWEEP, kernels_sweep.cu): incidence tables and level schedule, built and
  // uploaded by the first solve that asks for it (api_solve.cu) -- objects that never do pay nothing
  femb::SweepTables sweep_tab{};
  std::vector<int32_t> sweep_gptr;   // launch plan: group g = levels [gptr[g], gptr[g+1])
  std::vector<int32_t> sweep_gmax;   // rows of the largest level of group g
  int sweep_levels = 0;
  void ensure_sweep();
  // a peer wait that timed out leaves CommState.error set: surface it (call after a stream sync)
  void check_comm_error() {
    if (!peer_on) return;
    uint32_t err = 0;
    FEMB_CUDA(cudaMemcpy(&err, &peer.st->error, sizeof(err), cudaMemcpyDeviceToHost));
    if (err) throw femb::CudaError("multi-GPU: timed out waiting for a peer rank (halo flag / dot-product slot)");
  }
  void ensure_stage(size_t floats);
  femb::KinArgs kin_args(float dt, const float* g_dir) const;
};
