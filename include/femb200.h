/* femb200.h -- C ABI of libfemb200.so, the B200-native Neo-Hookean FEM hot path.
 *
 * The reference (Jukgei/Finite-Element-Method) has NO FFI layer: its boundary
 * is Python calling Taichi @ti.kernel functions that mutate Taichi fields
 * (main.py:101-113).  Each entry point below replaces one of those kernels /
 * field operations; the citation names the reference interface it stands in
 * for.  Plain C types only: pointers, sizes, scalars.  Host pointers are
 * borrowed for the duration of the call.  All device memory is owned by the
 * library behind the opaque handle.
 *
 * Every function returns 0 on success or a negative femb_status; the message
 * is available from femb_last_error() (thread-local).  There is no CPU
 * fallback: without a CUDA device object creation fails with FEMB_E_CUDA.
 *
 * Calls enqueue work on the object's stream and return; femb_get / femb_sync
 * synchronise.  A handle is not thread-safe.
 */
#ifndef FEMB200_H
#define FEMB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct femb_object femb_object; /* one soft body: object.py:41-93 `Object` */
typedef struct femb_tiling femb_tiling; /* host-only view of the element tiling (tests, tools) */

typedef enum femb_status {
  FEMB_OK = 0,
  FEMB_E_ARG = -1,   /* bad argument / bad mesh */
  FEMB_E_CUDA = -2,  /* CUDA runtime error, or no device */
  FEMB_E_STATE = -3, /* call not valid for this object (e.g. implicit buffers not allocated) */
  FEMB_E_LIMIT = -4  /* mesh exceeds a tiling limit */
} femb_status;

/* Field ids for femb_get / femb_set.  Vertex fields are [n_vert*dim] float32 in
 * the caller's ORIGINAL vertex order; element fields are in original element
 * order.  (The library renumbers both internally.) */
typedef enum femb_field {
  FEMB_F_POS = 0,      /* particles.pos            object.py:13  */
  FEMB_F_VEL = 1,      /* particles.vel            object.py:14  */
  FEMB_F_VEL_G = 2,    /* particles.vel_g          object.py:15  */
  FEMB_F_FORCE = 3,    /* particles.force (= +dE/dx) object.py:19, explicit.py:42-49 */
  FEMB_F_GRAD = 4,     /* particles.pos.grad       object.py:59 (needs_grad), main.py:107 */
  FEMB_F_MASS = 5,     /* particles.mass  [n_vert] object.py:18,360 */
  FEMB_F_REF_POS = 6,  /* particles.ref_pos        object.py:20,342 */
  FEMB_F_VEC_B = 7,    /* obj.vec_b                object.py:83  */
  FEMB_F_VEC_X = 8,    /* obj.vec_x                object.py:84  */
  FEMB_F_PAST_X = 9,   /* obj.past_vec_x           object.py:85  */
  FEMB_F_DIAG_A = 10,  /* diagonal of matrix_A (matrix-free; implicit.py:399) */
  FEMB_F_ELEM_VOLUME = 32, /* elements.volume [n_elem]          object.py:34,356 */
  FEMB_F_ELEM_REF = 33,    /* elements.ref = Dm^-1 [n_elem*dim*dim] row-major, object.py:35,362 */
  FEMB_F_ELEM_PHI = 34,    /* obj.phi [n_elem] (sized by elements; fixes object.py:68) */
  FEMB_F_ELEM_H = 35       /* per-element block H_e [n_elem*dim*dim] (implicit.py:141-147) */
} femb_field;

typedef struct femb_desc {
  int32_t dim;          /* 2 (triangles) or 3 (tets)      constants.dim, utils.py:10 */
  int32_t n_vert;
  int32_t n_elem;
  const float *verts;   /* [n_vert*dim] raw vertices (object.py:62 ti_vertices) */
  const int32_t *elems; /* [n_elem*(dim+1)] (object.py:66-67 ti_element) */
  float center[3];      /* offset added to every vertex, object.py:341 */
  float rho;            /* object.py:45 */
  float mu, lambda;     /* object.py:48 */
  float damping;        /* object.py:50 */
  int32_t implicit;     /* !=0: allocate the implicit-solver vectors, object.py:78-93 */
  int32_t tile_elems;   /* elements per tile (<= femb_info.stage_stride); 0 = library default */
  int32_t device;       /* CUDA device ordinal; -1 = current device */
  /* Multi-GPU (one process per GPU, mesh partitioned by the caller): vertices of THIS rank's
   * mesh that another rank also holds.  Each may appear once; the element gradient / operator
   * output on them is completed by a halo exchange (femb_halo_* below).  Not in the reference
   * (single device, main.py:40). */
  int32_t n_halo;
  const int32_t *halo_verts; /* [n_halo] original vertex ids */
  /* [n_halo] 1 = THIS rank counts the shared vertex in global reductions (CG dots, residual norms),
   * 0 = the neighbour does.  NULL = count all (single rank). */
  const uint8_t *halo_owned;
} femb_desc;

typedef struct femb_info {
  int32_t dim, n_vert, n_elem;
  int32_t n_tiles, tile_elems, max_tile_verts, max_rounds;
  int32_t n_interior, n_boundary, n_partials;
  int32_t smem_bytes, block_threads;
  int32_t stage_stride;  /* compile-time tile capacity == staging-plane stride (JDS entry encoding) */
  int32_t n_halo_tiles;  /* multi-GPU: tiles [0, n_halo_tiles) touch a shared vertex and are launched first */
  int64_t device_bytes; /* total device memory owned by the object */
} femb_info;

/* ---- lifetime: Object.__init__ + particles_init + elements_init (object.py:44-93,337-362) */
int femb_object_create(const femb_desc *desc, femb_object **out);
int femb_object_destroy(femb_object *obj);
int femb_object_info(const femb_object *obj, femb_info *out);
/* Run subsequent work on a caller-provided cudaStream_t (e.g. torch's current stream). */
int femb_set_stream(femb_object *obj, void *cuda_stream);

/* ---- CircleBlocks (circle_blocks.py:15-25): centers [n*dim], radii [n]; radius<=0 disables */
int femb_blocks_set(femb_object *obj, int32_t n, const float *centers, const float *radii);

/* ---- explicit path */
/* neo_hookean_1_grad(obj), explicit.py:8-49: force += dE/dx */
int femb_explicit_grad(femb_object *obj);
/* compute_energy(obj) under ti.ad.Tape, explicit_auto_diff.py:8-30 + main.py:107-108:
 * phi[e], U, and pos.grad = dU/dx (overwritten).  U_out may be NULL (no sync). */
int femb_energy_grad(femb_object *obj, float *U_out);
/* kinematic(obj, circle_blocks), kinematic.py:9-45.  use_grad!=0 reads pos.grad
 * (auto_diff==1) and leaves it; otherwise reads force and zeroes it. */
int femb_kinematic(femb_object *obj, float dt, const float *g_dir, int32_t use_grad);
/* n_steps x [neo_hookean_1_grad ; kinematic] (main.py:103-110 with auto_diff==0),
 * fused so the gradient never round-trips through HBM.  Results equal the
 * two-call sequence bit for bit. */
int femb_explicit_step(femb_object *obj, float dt, const float *g_dir, int32_t n_steps);
/* Same n_steps fused steps, launched one by one with CUDA events around each kernel:
 * returns the summed device time (ms) of the tile kernel and of the boundary kernel.
 * Measurement aid for bench.py's per-kernel roofline; synchronises. */
int femb_explicit_step_timed(femb_object *obj, float dt, const float *g_dir, int32_t n_steps,
                             float *tile_ms, float *boundary_ms);

/* ---- multi-GPU halo (new: the reference is single-device).  send/recv are caller-owned DEVICE
 * buffers of n_halo float4 each (torch tensors, NCCL or peer-mapped memory), in halo_verts order.
 * A phased pass is: phase 1 (tile kernel + pack into send) ; caller exchanges send -> the
 * neighbour's recv ; phase 2 (unpack recv + boundary kernel).  Shared vertices come out
 * bit-identical on both ranks (two-term commutative sum). */
int femb_halo_bind(femb_object *obj, void *send_dev, void *recv_dev);
/* lumped mass of shared vertices, once after creation: phase 1 packs, phase 2 adds the remote part */
int femb_halo_mass(femb_object *obj, int32_t phase);
/* one fused explicit step split around the exchange (phase 1 | exchange | phase 2) */
int femb_explicit_step_phase(femb_object *obj, float dt, const float *g_dir, int32_t phase);

/* ---- multi-GPU over peer memory (new).  Every rank allocates one comm buffer of femb_comm_bytes(max over
 * ranks of n_halo) ZEROED bytes that all ranks can address (torch symmetric memory / CUDA IPC) and binds
 * the table of mapped pointers.  From then on femb_explicit_step and femb_implicit_solve run the halo sums
 * and the dot-product sums inside the library's own kernels over NVLink (remote stores + release/acquire
 * flags): no collective call on the data path, and the whole step / iteration stays one CUDA graph.  Halo
 * entries [0, n_left) belong to rank `left`, the rest to rank `right` (-1: none); *_dst_off = index of my
 * plane inside that neighbour's halo list. */
typedef struct femb_peer_desc {
  int32_t rank, world;
  int32_t left, right;
  int32_t n_left;
  int32_t left_dst_off, right_dst_off;
  int32_t recv_stride;      /* from femb_comm_bytes, identical on all ranks */
  void *const *bufs;        /* [world] comm buffer of every rank as mapped in THIS process */
} femb_peer_desc;
int femb_comm_bytes(int32_t max_halo, size_t *bytes_out, int32_t *recv_stride_out);
int femb_peer_bind(femb_object *obj, const femb_peer_desc *desc);
int femb_peer_mass(femb_object *obj); /* lumped masses of shared vertices, once after binding (synchronises) */

/* ---- implicit path */
/* implicit_solver_neo_hookean(obj), implicit.py:201-223, matrix-free.
 * method: 0 Jacobi (implicit.py:227-262) as a simultaneous weighted Jacobi, 1 CG (implicit.py:311-386),
 * 2 (FEMB_METHOD_SWEEP) the reference's "Jacobi" in ITS iterate order -- the serial in-place sweep of
 * implicit.py:391-404 (rows in ascending vertex id, each row sees the rows before it updated), run level by
 * dependency level so that the result is the serial one; same start vector, stale past_vec_x, accept / revert /
 * stop rule as method 0; single-GPU objects only;
 * preconditioned: CG on A^T A x = A^T b (implicit.py:289-299).
 * max_iter<=0 / tol<0 select the reference's constants (500 / 1e-5 for CG,
 * 20000 / 1e-5 for Jacobi and the sweep).  tol == 0: exactly max_iter iterations (benchmark / test mode).
 * Writes vel = x. */
#define FEMB_METHOD_JACOBI 0
#define FEMB_METHOD_CG 1
#define FEMB_METHOD_SWEEP 2
int femb_implicit_solve(femb_object *obj, float dt, int32_t method, int32_t preconditioned,
                        int32_t max_iter, float tol, int32_t *iters_out, float *resid_out);
/* advect_implicit(obj, circle_blocks), implicit.py:407-437 */
int femb_advect_implicit(femb_object *obj, float dt, const float *g_dir);

/* The same solve for a caller that owns the collectives (multi-GPU slabs, one process per GPU; new --
 * the reference is single-device).  _begin builds the solve; every _advance enqueues device work up to
 * the next communication point and returns (>= 0) what the caller must do before advancing again:
 *   FEMB_COMM_HALO       exchange the bound halo buffers (send -> the neighbour's recv),
 *   FEMB_COMM_ALLREDUCE  sum-allreduce the double bound with femb_reduce_bind,
 *   FEMB_COMM_ITER       nothing: the prologue / one solver iteration is fully enqueued; the caller may
 *                        _poll the convergence state, then _advance into the next iteration or _finish.
 * The sequence of return values of one iteration never changes, so a caller may capture one iteration
 * (kernels + collectives) in a CUDA graph and replay it.  Negative return = femb_status error. */
typedef enum femb_comm { FEMB_COMM_NONE = 0, FEMB_COMM_HALO = 1, FEMB_COMM_ALLREDUCE = 2, FEMB_COMM_ITER = 3 } femb_comm;
int femb_reduce_bind(femb_object *obj, void *dev_double); /* caller-owned DEVICE double[1] */
int femb_implicit_solve_begin(femb_object *obj, float dt, int32_t method, int32_t preconditioned,
                              int32_t max_iter, float tol);
int femb_implicit_solve_advance(femb_object *obj);
/* synchronises; done != 0 once the loop test of implicit.py:252 / :341 has failed.  Also valid after the
 * solve has ended (state of the most recent solve). */
int femb_implicit_solve_poll(femb_object *obj, int32_t *done_out, int32_t *iters_out, float *resid_out);
/* vel = x (implicit.py:222-223); ends the solve */
int femb_implicit_solve_finish(femb_object *obj, int32_t *iters_out, float *resid_out);
/* Pieces of the solve, exposed for parity tests: assemble H_e, b, diag(A) */
int femb_implicit_assemble(femb_object *obj, float dt);
/* out = A in (transpose==0) or A^T in (transpose!=0); host vectors [n_vert*dim] */
int femb_implicit_apply(femb_object *obj, float dt, int32_t transpose, const float *in, float *out);

/* n operator passes q = A d (transpose != 0: A^T d) on the vectors of the last solve, launched one by one
 * with CUDA events around each kernel: summed device time (ms) of the tile kernel and of the boundary
 * kernel.  Measurement aid for bench.py's per-kernel roofline; synchronises. */
int femb_implicit_apply_timed(femb_object *obj, float dt, int32_t transpose, int32_t n, float *tile_ms,
                              float *boundary_ms);

/* ---- headless frame (render/render.py: Render.render2d :64-75, render3d :86-98).  The reference draws
 * through a Taichi GUI / GGUI (Vulkan) window, unavailable on this box; the same picture is rasterised by
 * CUDA kernels into a device frame buffer and returned as width*height*3 RGB8 bytes, row 0 = top.
 *   dim 2: element triangles coloured (phi+0.13, 0.13, 0.13), vertex dots (radius 2 px, 0xAAAA00), blocks
 *          (radius r*width, 0x343434); unit square -> image, y up.  prims = NULL (the object's elements).
 *   dim 3: the (0..5)^3 box (colour (0.99,0.68,0.28), width 2) and the surface mesh prims[n_prims*3]
 *          (obj.indices, ORIGINAL vertex numbering) as a wireframe shaded with ambient 0.8 + the point light
 *          at (3.5,3.5,3.5); look-at perspective camera.  prims are uploaded when prims_version changes. */
typedef struct femb_view {
  int32_t width, height;      /* render.py:11-12: 640 x 640 */
  float eye[3], lookat[3], up[3];   /* render.py:37-39 */
  float fov_deg;              /* GGUI default 45 */
  int32_t prims_version;      /* bump to re-upload prims (a new surface) */
} femb_view;
int femb_render(femb_object *obj, const femb_view *view, const int32_t *prims, int32_t n_prims, uint8_t *rgb_out);

/* ---- field access: field.to_numpy() / field.from_numpy() */
int femb_get(femb_object *obj, int32_t field, void *host, size_t bytes);
int femb_set(femb_object *obj, int32_t field, const void *host, size_t bytes);
/* Asynchronous variants for pipelined host traffic (new; the reference's to_numpy/from_numpy, object.py:62-67,
 * render.py:67, are synchronous copies).  `host` should be page-locked and must stay valid until
 * femb_transfers_wait() (or femb_sync after it) returns.  femb_set_async copies on a dedicated H2D stream and orders
 * the scatter into the field BEFORE any later call on the object; femb_get_async packs the field AFTER all earlier
 * calls and copies on a dedicated D2H stream -- so H2D of step k+1, the kernels of step k and D2H of step k-1
 * overlap (PCIe is full duplex).  One staging buffer per field and direction: re-use is ordered by events. */
int femb_set_async(femb_object *obj, int32_t field, const void *host, size_t bytes);
int femb_get_async(femb_object *obj, int32_t field, void *host, size_t bytes);
int femb_transfers_wait(femb_object *obj);
/* obj.U[None] (object.py:69) */
/* Kernel timeline of the tile passes (diagnostics; nothing in the reference corresponds).  While enabled, every
 * tile / boundary kernel launch records, in its own 8-word slot, %globaltimer nanoseconds of: [0] the start of its
 * CTA 0, [1] its latest CTA end, [2] the latest end of a wait for a neighbouring rank's halo flags, [3] the release
 * of this rank's halo flags, [4..6] its first sender CTA: released / values stored / fence done (0 where not
 * applicable; [7] unused).  Slots are handed out in launch order from 0 after each femb_trace_enable (a cached CUDA
 * graph keeps the slots it was captured with, so replays overwrite them: after a run of steps the buffer holds the
 * timeline of the last one).  femb_trace_get synchronises and returns the first n slots (8 words each) and their
 * newline-separated names ("<pass>:<kernel>").  Serves tools/trace_steps.py -> profiles/r02_multi_gpu.txt. */
int femb_trace_enable(femb_object *obj, int32_t on);
int femb_trace_get(femb_object *obj, uint64_t *words, int32_t max_slots, int32_t *n_out, char *names, size_t names_bytes);

/* State diagnostics counted on the device (replaces nothing in the reference -- it has no such check; serves
 * main.py --diag-every): out3 = { elements with det F <= 0 or NaN (explicit.py:24-25 takes
 * log(det F) unclamped), elements whose energy density phi is NaN (as of the last femb_energy_grad), vertices with
 * a NaN coordinate }.  Synchronises. */
int femb_diagnostics(femb_object *obj, int64_t *out3);
int femb_get_energy(femb_object *obj, float *U_out);
int femb_sync(femb_object *obj);

/* ---- diagnostics */
const char *femb_last_error(void);
/* number of CUDA kernels this library has launched in this process */
uint64_t femb_launch_count(void);
const char *femb_version(void);

/* OpenMP threads of the host-side set-up code (tiling builder); returns the count in effect.  n <= 0 only queries. */
int femb_set_host_threads(int32_t n);

/* ---- host-only tiling (no CUDA calls): used by tests to check the layout invariants */
int femb_tiling_build(int32_t dim, int32_t n_vert, int32_t n_elem, const float *verts,
                      const int32_t *elems, int32_t tile_elems, femb_tiling **out);
/* same with shared (halo) vertices and the device's wave size (CTAs resident at once; 0 = do not round) */
int femb_tiling_build_ex(int32_t dim, int32_t n_vert, int32_t n_elem, const float *verts, const int32_t *elems,
                         int32_t tile_elems, int32_t n_halo, const int32_t *halo_verts, int32_t wave,
                         femb_tiling **out);
int femb_tiling_destroy(femb_tiling *t);
int femb_tiling_info(const femb_tiling *t, femb_info *out);
/* what: 0 new_of_old[n_vert], 1 slot_of_elem[n_elem], 2 tile descriptors [n_tiles*12],
 * 3 local vertex table [sum n_lv * 2], 4 jds entries (uint16), 5 round counts (uint32),
 * 6 packed local indices [n_slots*2] (uint32), 7 boundary ptr [n_boundary+1].
 * Returns the element count; copies min(count, cap) elements when out != NULL. */
int64_t femb_tiling_get(const femb_tiling *t, int32_t what, void *out, int64_t cap);

#ifdef __cplusplus
}
#endif
#endif /* FEMB200_H */
