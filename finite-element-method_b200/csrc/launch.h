// launch.h -- host-callable launchers implemented in kernels.cu (one per kernel family).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

#include "dev.cuh"

namespace femb {

struct VertexState {       // SoA-of-float4 vertex arrays in the library's numbering
  float4* pos;             // {x, y, z, mass}           particles.pos / .mass
  float4* vel;             // particles.vel
  float4* vel_g;           // particles.vel_g  (implicit objects only)
  float4* force;           // particles.force  (+dE/dx)
  float4* grad;            // particles.pos.grad
  float4* partial;         // [n_partials] boundary partial sums
  float4* partial2;        // second partial buffer (implicit assemble)
};

struct Material {
  float mu, lambda, rho, damping;
};

struct KinArgs {           // kinematic.py:14-45 / implicit.py:407-437 constants
  float dt;
  float g98[3];            // 9.8f * g_dir
  float damp;              // expf(-dt * damping)
  Blocks blocks;
};

struct ImplicitVecs {      // object.py:83-93 vectors, float4 per vertex
  float4 *b, *x, *past_x, *d, *r, *q, *atb, *tmp, *diag;
  float4 *w;               // single-reduction CG: w = Op r (d doubles as p, q as s = Op p)
};

struct SolveScalars {      // device-resident solver state (one per object)
  double delta_new, delta_old, dq, err, p_err;
  float alpha, beta;
  int iters, done, terminal, action, first;
  // fused Jacobi (kernels_implicit.cu: OpJacobi): the iterate rotates through three buffers
  int cur;                 // buffer holding the iterate the current pass evaluates
  int ans;                 // terminal: buffer holding the answer (3 = past_x itself: revert before any accept)
  int keep_past;           // terminal: past_vec_x keeps its (stale) value
  // single-reduction CG (api_solve.cu): the vector update of iteration i is applied by the first pass of i+1
  double gamma_old;        // r.r of the previous iteration
  int pending;             // 1 after the first iteration of a single-reduction CG solve (beta is defined)
};

enum ExplicitMode { EX_GRAD = 0, EX_STEP = 1, EX_ENERGY = 2 };

struct LaunchCfg {
  cudaStream_t stream;
  int block;               // threads per CTA of the tile kernels
  int smem;                // dynamic shared memory of the tile kernels
  int sms;
  int device;
  cudaEvent_t ev_mid = nullptr;  // if set, recorded between the tile and the boundary kernel of a pass
  int phase = 0;                 // 0: tile + boundary kernel; 1: tile kernel only; 2: boundary kernel only
                                 // (multi-GPU: the halo exchange of the partial sums runs between 1 and 2)
  int part = 0;                  // phase 1 only -- 0: all tiles; 1: the halo tiles [0, n_halo_tiles);
                                 // 2: the other tiles (the exchange overlaps part 2)
  const struct HaloDev* hd = nullptr;   // peer-memory exchange fused into the pass's kernels:
  int n_send = 0;                //   part 2: sender CTAs at the front of the tile kernel
  int fuse_recv = 0;             //   phase 2: the boundary kernel receives
  int no_persist = 0;            //   one CTA per tile even for Ops that default to the persistent kernel
  int bnd_send = 0;              //   phase 2 with fuse_recv: sender CTAs at the front of the BOUNDARY kernel ("late send")
  int send_at = -1;              //   phase 1, part 0, n_send > 0: ONE tile kernel over all tiles whose CTAs
                                 //   [send_at, send_at + n_send) are the senders; they wait for the halo-tile counter
  unsigned long long* trace = nullptr;   // kernel timeline buffer (femb_trace_enable); nullptr = off
};

struct Halo {                    // vertices shared with a neighbouring rank (slab interface planes)
  int n = 0;
  const int32_t* b = nullptr;    // [n] boundary index of each halo vertex
  float4* send = nullptr;        // [n] caller-owned device buffers (torch tensors / peer memory)
  float4* recv = nullptr;
};
// ---- peer-memory communication (kernels_comm.cu): the comm buffer every rank exposes to all ranks
constexpr int kMaxRanks = 16;
struct CommHeader {              // at offset 0 of a comm buffer; written by REMOTE ranks
  uint32_t halo_flag[2];         // [0]: sequence released by the LEFT neighbour, [1]: by the RIGHT one
  uint32_t ar_flag[kMaxRanks];   // [r]: dot-product sequence released by rank r
  uint32_t pad[46];              // header = 256 bytes
};
struct CommState {               // LOCAL device memory: sequence counters (graph-replayable) and the error flag
  uint32_t send_seq, recv_seq, ar_seq, pack_count, unpack_count, error, pad0, pad1;
  // "one tile kernel" passes (LaunchCfg::send_at): halo tiles finished since the object was bound, and the number
  // of such passes completed -- a sender CTA of pass p waits for halo_done >= (p + 1) * n_halo_tiles
  unsigned long long halo_done;
  uint32_t count_passes, pad2;
};
struct PeerComm {                // kernel argument
  char* self;                    // my comm buffer
  char* left;                    // neighbours' comm buffers (peer-mapped), nullptr at the ends of the slab chain
  char* right;
  char* all[kMaxRanks];          // every rank's comm buffer, all[rank] == self
  int rank, world;
  int n_left;                    // halo entries [0, n_left) go to / come from the left neighbour
  int left_dst_off, right_dst_off;   // where my planes start in the neighbours' receive arrays
  size_t recv_off, slot_off, recv_stride;
  size_t ll_slot_off, ll_off;    // LL all-sum slots / LL halo planes (comm.cuh), behind the float4 receive planes
  CommState* st;
};
struct HaloDev {                 // device-resident: what a tile / boundary kernel needs to run the exchange itself
  PeerComm pc;
  int n;                         // halo vertices
  const int32_t* hb;             // [n] boundary index of halo vertex k
  const int2* hrange;            // [n] local partial slots [x, y) of halo vertex k (the remote slot excluded)
  const int32_t* bhalo;          // [n_boundary] halo index of boundary vertex b, -1: not shared
};
size_t comm_layout(int n_halo, size_t* recv_off, size_t* slot_off, size_t* recv_stride);
// offsets of the LL areas for a given (common) receive stride
void comm_ll_offsets(size_t recv_off, size_t recv_stride, size_t* ll_slot_off, size_t* ll_off);
// what: 0 partial sums, 1 lumped masses (flag protocol, received by launch_halo_recv); 2 partial sums as LL words
// (received by the boundary kernel itself)
void launch_halo_send(const LaunchCfg& c, const DevTiling& tl, const PeerComm& pc, Halo h, const float4* partial,
                      const float4* pos, int what, const int2* hrange = nullptr);   // what == 2: LL words (needs hrange)
void launch_halo_recv(const LaunchCfg& c, const DevTiling& tl, const PeerComm& pc, Halo h, float4* partial,
                      float4* pos, int what);
// local sum of red[0..n_red) -> slot[rank] of every rank; the consuming scalar kernel gathers the slots
void launch_reduce_final_send(const LaunchCfg& c, int n_red, const double* red, const PeerComm& pc);

// send[k] = sum of the local partial slots of halo vertex k ; partial[remote slot of k] = recv[k]
void launch_halo_pack(const LaunchCfg& c, const DevTiling& tl, Halo h, const float4* partial);
void launch_halo_unpack(const LaunchCfg& c, const DevTiling& tl, Halo h, float4* partial);
// lumped mass of shared vertices: phase 0 send[k].x = local mass ; phase 1 mass += recv[k].x
void launch_halo_mass(const LaunchCfg& c, const DevTiling& tl, Halo h, float4* pos, int phase);
void launch_halo_zero_remote(const LaunchCfg& c, const DevTiling& tl, Halo h, float4* partial);

// setup: Dm^-1, volume -> element planes; lumped mass -> pos.w   (object.py:345-362)
// kernel timeline: every launch of run_tile_pass takes the next 8-word slot of LaunchCfg::trace (dev.cuh)
constexpr int kTraceSlots = 512, kTraceWords = 8;
void trace_restart();                                   // slot counter back to 0 (new capture)
int trace_slots_used();
const char* trace_slot_name(int slot);
unsigned long long* trace_take_slot(const LaunchCfg& c, const char* name, const char* what);   // nullptr when off / full
// inverted elements / NaN energy densities / NaN vertices -> out3[0..2] (device)
void launch_diagnostics(const LaunchCfg& c, int dim, const DevTiling& tl, Planes e, const float4* pos, const float* phi,
                        unsigned long long* out3);
// packed explicit kernels (FEMB_PAIR_EXPLICIT): they read a pair-interleaved copy of the element planes
bool explicit_pair_layout();
void launch_pair_interleave(int dim, const LaunchCfg& c, const DevTiling& tl, Planes src, Planes dst);
void launch_setup(int dim, const LaunchCfg& c, const DevTiling& tl, Planes e, const uint32_t* lidx,
                  VertexState vs, Material mat);
// explicit tile + boundary pass
void launch_explicit(int dim, int mode, const LaunchCfg& c, const DevTiling& tl, Planes e, VertexState vs,
                     Material mat, KinArgs kin, float* phi, double* red);
void launch_kinematic(int dim, const LaunchCfg& c, int n_vert, VertexState vs, KinArgs kin, int use_grad);
void launch_advect_implicit(int dim, const LaunchCfg& c, int n_vert, VertexState vs, KinArgs kin);
// implicit: H_e planes + b ; diag(A)
void launch_implicit_assemble(int dim, const LaunchCfg& c, const DevTiling& tl, Planes e, Planes h,
                              VertexState vs, Material mat, float dt, ImplicitVecs iv);
void launch_implicit_diag(int dim, const LaunchCfg& c, const DevTiling& tl, Planes h, VertexState vs,
                          float dt, ImplicitVecs iv);
// out = A in / A^T in ; optional fused dot <dotv, out> or residual r = b - out with ||r||^2
enum ApplyEpilogue { AP_PLAIN = 0, AP_DOT = 1, AP_RESID = 2 };
void launch_apply(int dim, const LaunchCfg& c, const DevTiling& tl, Planes h, VertexState vs, float dt,
                  int transpose, const float4* in, float4* out, int epilogue, const float4* aux,
                  float4* resid_out, double* red, const SolveScalars* sc);
int reduce_slots(const DevTiling& tl, const LaunchCfg& c);
int apply_tile_grid(const DevTiling& tl, bool split = false, bool persist = true);   // per-CTA reduction partials written by an operator pass
int energy_tile_grid(const DevTiling& tl);  // ... by the energy pass  // doubles needed in a reduction buffer

// CG / Jacobi vector kernels (all early-exit on sc->done)
struct RedMask {           // which vertices this rank counts in global reductions (multi-GPU owner mask)
  int n_interior;
  const uint8_t* bskip;    // nullptr: count every vertex
};
// x = b, r = rhs - opx, d = r, per-CTA partials of r.r -> red[0 .. ceil(n_vert/256))
void launch_cg_init(const LaunchCfg& c, int n_vert, ImplicitVecs iv, const float4* rhs, const float4* opx,
                    double* red, RedMask mask);
// delta_new = sum red[0..n_red), loop entry test
void launch_cg_delta0(const LaunchCfg& c, int n_red, const double* red, SolveScalars* sc, float tol, int max_iter,
                      const PeerComm& pc);
void launch_cg_alpha(const LaunchCfg& c, int n_red, const double* red, SolveScalars* sc, const PeerComm& pc);
void launch_cg_update_xr(const LaunchCfg& c, int n_vert, ImplicitVecs iv, double* red, SolveScalars* sc, RedMask mask);
void launch_cg_beta(const LaunchCfg& c, int n_red, const double* red, SolveScalars* sc, float tol, int max_iter,
                    const PeerComm& pc);
void launch_cg_update_d(const LaunchCfg& c, int n_vert, ImplicitVecs iv, const SolveScalars* sc);
// single-reduction (Chronopoulos-Gear) CG on A^T A (api_solve.cu)
//   gamma = sum red_g, delta = sum red_d -> loop test, alpha, beta
void launch_cgcg_scalar(const LaunchCfg& c, int n_g, const double* red_g, int n_d, const double* red_d, SolveScalars* sc,
                        float tol, int max_iter, const PeerComm& pc);
//   p = r + beta p ; s = w + beta s ; x += alpha p ; r -= alpha s ; partials of r.r -> red
void launch_cgcg_update(const LaunchCfg& c, int n_vert, ImplicitVecs iv, double* red, const SolveScalars* sc, RedMask mask);
void launch_cg_init_ex(const LaunchCfg& c, int n_vert, ImplicitVecs iv, const float4* rhs, const float4* opx,
                       double* red, RedMask mask, int zero_ps);
void launch_jacobi_decide(const LaunchCfg& c, int n_red, const double* red, SolveScalars* sc, float tol,
                          int max_iter, int fixed, const PeerComm& pc);
void launch_jacobi_finish(const LaunchCfg& c, int n_vert, ImplicitVecs iv, const SolveScalars* sc);
// one fused Jacobi pass: r = b - A x_k (norm^2 -> red) and, speculatively, x_{k+1} into the next of the three
// rotating buffers {iv.x, iv.tmp, iv.r} (sc->cur says which holds x_k)
void launch_jacobi_pass(int dim, const LaunchCfg& c, const DevTiling& tl, Planes h, VertexState vs, float dt,
                        ImplicitVecs iv, double* red, const SolveScalars* sc);
void launch_scalars_reset(const LaunchCfg& c, SolveScalars* sc);
// reference-order sweep (kernels_sweep.cu, sweep_core.h): loop control, the accept / revert copy, one group of levels
struct SweepTables;
struct SweepVecs;
void launch_sweep_decide(const LaunchCfg& c, int n_red, const double* red, SolveScalars* sc, float tol, int max_iter,
                         int fixed);
void launch_sweep_action(const LaunchCfg& c, int n_vert, float4* x, float4* past, const SolveScalars* sc);
void launch_sweep_levels(int dim, const LaunchCfg& c, const SweepTables& t, const SweepVecs& a, int lvl0, int n_lvl,
                         int max_level_rows, const SolveScalars* sc);

// misc vector ops
void launch_copy_xyz(const LaunchCfg& c, int n, const float4* src, float4* dst);           // dst.xyz = src.xyz
void launch_scale_xyz(const LaunchCfg& c, int n, const float4* src, float4* dst, float s); // dst.xyz = s*src.xyz
void launch_fill(const LaunchCfg& c, int n, float4* dst, float4 v);
void launch_reduce_final(const LaunchCfg& c, int n_red, const double* red, double* out);
// host <-> device field marshalling in the caller's ordering
void launch_pack_vertex(const LaunchCfg& c, int dim, int n_vert, const int32_t* new_of_old, const float4* src,
                        int comp0, int ncomp, float* dst);
void launch_unpack_vertex(const LaunchCfg& c, int dim, int n_vert, const int32_t* new_of_old, const float* src,
                          int comp0, int ncomp, float4* dst);
void launch_pack_elem(const LaunchCfg& c, int dim, int n_elem, const int32_t* slot_of_elem, Planes e, Planes h,
                      const float* phi, int what, float* dst);

// headless rasteriser (kernels_render.cu): what render/render.py:64-106 draws, into a device frame buffer
void launch_render2d(const LaunchCfg& c, int n_elem, const int32_t* elems, const int32_t* new_of_old,
                     const int32_t* slot_of_elem, const float4* pos, const float* phi, int n_vert, const float4* blocks,
                     int n_blocks, unsigned long long* fb, uint8_t* rgb, int W, int H);
void launch_render3d(const LaunchCfg& c, const float* eye, const float* lookat, const float* up, float fov_deg,
                     int n_faces, const int32_t* faces, const int32_t* new_of_old, const float4* pos,
                     unsigned long long* fb, uint8_t* rgb, int W, int H);

uint64_t launch_count();
void note_launch(const char* what, cudaStream_t s);  // counts + checks cudaGetLastError (throws std::runtime_error)

}  // namespace femb
