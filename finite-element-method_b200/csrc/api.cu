// api.cu -- the extern "C" surface of include/femb200.h: object lifetime, field
// marshalling, explicit path.  (The implicit solve lives in api_solve.cu.)
#include <cmath>
#include <cstdlib>
#include <cstring>

#include <omp.h>
#include "object.h"

using namespace femb;

namespace femb {
thread_local std::string g_last_error;

int fail(const std::exception& ex) {
  g_last_error = ex.what();
  if (dynamic_cast<const ArgError*>(&ex)) return FEMB_E_ARG;
  if (dynamic_cast<const StateError*>(&ex)) return FEMB_E_STATE;
  if (dynamic_cast<const LimitError*>(&ex)) return FEMB_E_LIMIT;
  return FEMB_E_CUDA;
}
}  // namespace femb

#define FEMB_TRY try {
#define FEMB_CATCH                   \
  }                                  \
  catch (const std::exception& ex) { \
    return femb::fail(ex);           \
  }                                  \
  return FEMB_OK;

void femb_object::ensure_stage(size_t floats) {
  if (floats <= stage_floats) return;
  // grow-only; the old buffer stays in `allocs` and is freed with the object
  d_stage = alloc<float>(floats, false);
  stage_floats = floats;
}

KinArgs femb_object::kin_args(float dt, const float* g_dir) const {
  if (!g_dir) throw ArgError("g_dir is NULL");
  KinArgs k{};
  k.dt = dt;
  for (int i = 0; i < 3; ++i) k.g98[i] = i < dim ? 9.8f * g_dir[i] : 0.f;
  k.damp = expf(-dt * mat.damping);  // ti.exp(-dt * damping), kinematic.py:23
  k.blocks.data = d_blocks;
  k.blocks.count = n_blocks;
  return k;
}

static const int kMaxSmem = 226 * 1024;  // 227 KB per CTA minus the kernels' static shared memory

static int tile_smem_bytes(const Tiling& t) {
  // largest layout any Op uses: D*D + D staging planes; the second vertex buffer only when an Op with
  // ASYNC_VERTEX is built persistent (FEMB_PERSIST_EXPLICIT, off by default)
  return tile_smem_layout(t.max_lv, (t.max_rounds + 3) / 4 * 4, t.nc, (t.dim + 1) * t.dim, FEMB_PERSIST_EXPLICIT != 0).total;
}

extern "C" {

const char* femb_last_error(void) { return g_last_error.c_str(); }
uint64_t femb_launch_count(void) { return launch_count(); }
const char* femb_version(void) { return "femb200 0.1 (sm_100a)"; }

int femb_object_create(const femb_desc* d, femb_object** out) {
  femb_object* o = nullptr;
  try {
    if (!d || !out) throw ArgError("femb_object_create: NULL argument");
    *out = nullptr;
    if (d->dim != 2 && d->dim != 3) throw ArgError("dim must be 2 or 3");
    if (d->n_vert <= 0 || d->n_elem <= 0 || !d->verts || !d->elems) throw ArgError("empty mesh");
    if (!(d->rho > 0.f)) throw ArgError("rho must be positive");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
      cudaGetLastError();
      throw CudaError("no CUDA device: libfemb200 has no CPU fallback");
    }
    o = new femb_object();
    if (d->device >= 0) FEMB_CUDA(cudaSetDevice(d->device));
    FEMB_CUDA(cudaGetDevice(&o->device));
    o->dim = d->dim; o->n_vert = d->n_vert; o->n_elem = d->n_elem;
    o->implicit = d->implicit != 0;
    o->mat = Material{d->mu, d->lambda, d->rho, d->damping};
    const int dim = d->dim;

    // pos = raw + center in fp32                       (object.py:341)
    std::vector<float> pos((size_t)d->n_vert * dim);
    for (int v = 0; v < d->n_vert; ++v)
      for (int k = 0; k < dim; ++k) pos[(size_t)v * dim + k] = d->verts[(size_t)v * dim + k] + d->center[k];

    int T = d->tile_elems > 0 ? std::min(d->tile_elems, kStride) : kStride;
    cudaDeviceProp prop{};
    FEMB_CUDA(cudaGetDeviceProperties(&prop, o->device));
    const int wave = prop.multiProcessorCount * kMinCtas;   // tile CTAs resident at once
    for (;;) {
      if (d->n_halo < 0 || (d->n_halo > 0 && !d->halo_verts)) throw ArgError("bad halo vertex list");
      const std::string err = build_tiling(dim, d->n_vert, d->n_elem, pos.data(), d->elems, T, o->til, d->n_halo,
                                           d->halo_verts, wave);
      if (!err.empty()) throw ArgError(err);
      if (tile_smem_bytes(o->til) <= kMaxSmem) break;
      if (T <= 64) throw LimitError("tiling does not fit shared memory even with 64-element tiles");
      T /= 2;  // irregular mesh with very many vertices per tile: shrink the tiles
    }
    const Tiling& t = o->til;
    if (dim == 2) o->elems2d.assign(d->elems, d->elems + (size_t)d->n_elem * 3);

    FEMB_CUDA(cudaStreamCreateWithFlags(&o->stream, cudaStreamNonBlocking));
    o->own_stream = true;
    o->cfg = LaunchCfg{o->stream, 256, tile_smem_bytes(t), prop.multiProcessorCount, o->device};

    auto upload = [&](auto*& dst, const auto& vec) {
      using T_ = typename std::remove_reference<decltype(vec)>::type::value_type;
      dst = o->alloc<T_>(vec.size(), false);
      if (!vec.empty())
        FEMB_CUDA(cudaMemcpyAsync(dst, vec.data(), vec.size() * sizeof(T_), cudaMemcpyHostToDevice, o->stream));
    };
    DevTile* d_tiles; LocalVert* d_lv; uint16_t* d_jds; uint32_t* d_rounds; int32_t* d_bptr;
    // CTAs of the tile kernels resident at once
    const int occ = std::max(1, std::min(std::min(2048 / kBlock, kMinCtas), (227 * 1024) / (tile_smem_bytes(t) + 1024)));
    const int grid = std::min(t.n_tiles, prop.multiProcessorCount * occ);
    std::vector<DevTile> dev_tiles((size_t)t.n_tiles);
    for (int i = 0; i < t.n_tiles; ++i) {
      const TileDesc& s = t.tiles[i];
      DevTile& q = dev_tiles[i];
      q = DevTile{s.slot_off, s.n_elem, s.lv_off, s.n_lv, s.jds_off, s.rounds_off, s.n_rounds, s.n_int, s.int_base,
                  0, 0, 0, 0, 0, 0, 0};
      if (i + grid < t.n_tiles) {
        const TileDesc& n = t.tiles[i + grid];
        q.nx_slot_off = n.slot_off; q.nx_n_elem = n.n_elem; q.nx_lv_off = n.lv_off; q.nx_n_lv = n.n_lv;
        q.nx_jds_off = n.jds_off; q.nx_int_base = n.int_base; q.nx_n_int = n.n_int;
      }
    }
    upload(d_tiles, dev_tiles);
    upload(d_lv, t.lv);
    upload(d_jds, t.jds);
    upload(d_rounds, t.rounds);
    upload(d_bptr, t.bptr);
    upload(o->d_lidx, t.lidx);
    upload(o->d_new_of_old, t.new_of_old);
    upload(o->d_slot_of_elem, t.slot_of_elem);
    if (!t.halo_b.empty()) {
      int32_t* d_hb;
      upload(d_hb, t.halo_b);
      o->halo.n = (int)t.halo_b.size();
      o->halo.b = d_hb;
    }
    uint8_t* d_bskip = nullptr;
    if (d->halo_owned && d->n_halo > 0) {
      std::vector<uint8_t> bskip((size_t)t.n_boundary, 0);
      for (int k = 0; k < d->n_halo; ++k) bskip[t.halo_b[k]] = d->halo_owned[k] ? 0 : 1;
      upload(d_bskip, bskip);
    }
    o->dtl = DevTiling{d_tiles, d_lv, d_jds, d_rounds, d_bptr, t.n_tiles, t.n_interior, t.n_boundary, t.n_vert,
                       t.n_slots, (t.max_tile_elems + 31) / 32 * 32, t.max_lv, (t.max_rounds + 3) / 4 * 4, t.nc,
                       prop.multiProcessorCount * occ, grid, t.n_halo_tiles, d_bskip};

    // in-tile completion of boundary vertices (tile_pass.cuh: itc_arrive): arrival counters armed with -(slots);
    // vertices shared with another rank get a count that is never reached (the halo boundary kernel owns them)
    // Measured (profiles/r02_ab_itc.log): correct, but SLOWER -- C4 explicit 0.274 -> 0.355 ms, C3 22.7 -> 27.1 us: the
    // fence -> atomic -> dependent loads chain sits at the end of every tile CTA's life and idles its SM slot.
    // Compiled in only with -DFEMB_ITC_BUILD=1 (dev.cuh), and then off unless FEMB_ITC=1.
    const char* itc_env = std::getenv("FEMB_ITC");
    if (kItcBuild && t.n_boundary > 0 && itc_env && itc_env[0] == '1') {
      std::vector<int32_t> cnt((size_t)t.n_boundary);
      for (int b = 0; b < t.n_boundary; ++b) cnt[b] = t.bptr[b] - t.bptr[b + 1];
      for (int32_t b : t.halo_b) cnt[b] = -(1 << 30);
      int32_t* d_cnt;
      upload(d_cnt, cnt);
      o->dtl.bcount = d_cnt;
      o->dtl.halo_b = o->halo.b;
      o->dtl.n_halo_b = o->halo.n;
    }

    const size_t ns = (size_t)t.n_slots, nv = (size_t)t.n_vert;
    o->e.p0 = o->alloc<float4>(ns);
    o->e.p1 = o->alloc<float4>(ns);
    o->e.p2 = dim == 3 ? o->alloc<float4>(ns) : nullptr;
    o->ex = o->e;
    if (explicit_pair_layout()) {
      o->ex.p0 = o->alloc<float4>(ns);
      o->ex.p1 = o->alloc<float4>(ns);
      o->ex.p2 = dim == 3 ? o->alloc<float4>(ns) : nullptr;
    }
    o->vs.pos = o->alloc<float4>(nv);
    o->vs.vel = o->alloc<float4>(nv);
    o->vs.force = o->alloc<float4>(nv);
    o->vs.grad = o->alloc<float4>(nv);
    o->vs.partial = o->alloc<float4>((size_t)t.n_partials);
    o->ref_pos = o->alloc<float4>(nv);
    o->d_phi = o->alloc<float>(ns);
    o->red_slots = reduce_slots(o->dtl, o->cfg);
    o->d_red = o->alloc<double>((size_t)o->red_slots);
    o->d_U = o->alloc<double>(1);
    if (o->implicit) {
      o->h.p0 = o->alloc<float4>(ns);
      o->h.p1 = o->alloc<float4>(ns);
      o->h.p2 = dim == 3 ? o->alloc<float4>(ns) : nullptr;
      o->vs.vel_g = o->alloc<float4>(nv);
      float4** vecs[] = {&o->iv.b, &o->iv.x, &o->iv.past_x, &o->iv.d, &o->iv.r,
                         &o->iv.q, &o->iv.atb, &o->iv.tmp, &o->iv.diag};
      for (float4** p : vecs) *p = o->alloc<float4>(nv);
      o->iv.w = o->alloc<float4>(nv);
      o->d_red2 = o->alloc<double>((size_t)o->red_slots);
      o->d_sc = o->alloc<SolveScalars>(1);
      FEMB_CUDA(cudaMallocHost(&o->h_sc, sizeof(SolveScalars)));
    }

    // positions in, then Dm^-1 / volume / mass on the device (object.py:71-73)
    o->ensure_stage(nv * 3);
    FEMB_CUDA(cudaMemcpyAsync(o->d_stage, pos.data(), pos.size() * sizeof(float), cudaMemcpyHostToDevice, o->stream));
    launch_unpack_vertex(o->cfg, dim, t.n_vert, o->d_new_of_old, o->d_stage, 0, dim, o->vs.pos);
    launch_setup(dim, o->cfg, o->dtl, o->e, o->d_lidx, o->vs, o->mat);
    if (explicit_pair_layout()) launch_pair_interleave(dim, o->cfg, o->dtl, o->e, o->ex);
    FEMB_CUDA(cudaMemcpyAsync(o->ref_pos, o->vs.pos, nv * sizeof(float4), cudaMemcpyDeviceToDevice, o->stream));
    FEMB_CUDA(cudaStreamSynchronize(o->stream));
    *out = o;
  } catch (const std::exception& ex) {
    if (o) femb_object_destroy(o);
    return femb::fail(ex);
  }
  return FEMB_OK;
}

int femb_object_destroy(femb_object* o) {
  if (!o) return FEMB_OK;
  cudaSetDevice(o->device);
  if (o->stream) cudaStreamSynchronize(o->stream);
  if (o->graph.exec) cudaGraphExecDestroy(o->graph.exec);
  if (o->solve_graph.exec) cudaGraphExecDestroy(o->solve_graph.exec);
  for (void* p : o->allocs) cudaFree(p);
  if (o->h_sc) cudaFreeHost(o->h_sc);
  if (o->side) { cudaStreamSynchronize(o->side); cudaStreamDestroy(o->side); }
  if (o->s_h2d) { cudaStreamSynchronize(o->s_h2d); cudaStreamDestroy(o->s_h2d); }
  if (o->s_d2h) { cudaStreamSynchronize(o->s_d2h); cudaStreamDestroy(o->s_d2h); }
  for (auto* arr : {o->xin, o->xout})
    for (int i = 0; i < 48; ++i) {
      if (arr[i].buf) cudaFree(arr[i].buf);
      if (arr[i].copied) cudaEventDestroy(arr[i].copied);
      if (arr[i].consumed) cudaEventDestroy(arr[i].consumed);
    }
  if (o->ev_fork) cudaEventDestroy(o->ev_fork);
  if (o->ev_join) cudaEventDestroy(o->ev_join);
  if (o->own_stream && o->stream) cudaStreamDestroy(o->stream);
  cudaGetLastError();
  delete o;
  return FEMB_OK;
}

static void fill_info(const Tiling& t, femb_info* i) {
  std::memset(i, 0, sizeof(*i));
  i->dim = t.dim; i->n_vert = t.n_vert; i->n_elem = t.n_elem;
  i->n_tiles = t.n_tiles; i->tile_elems = t.T; i->max_tile_verts = t.max_lv; i->max_rounds = t.max_rounds;
  i->n_interior = t.n_interior; i->n_boundary = t.n_boundary; i->n_partials = t.n_partials;
  i->smem_bytes = tile_smem_bytes(t);
  i->block_threads = kBlock;
  i->stage_stride = kStride;
  i->n_halo_tiles = t.n_halo_tiles;
}

int femb_object_info(const femb_object* o, femb_info* out) {
  FEMB_TRY
  if (!o || !out) throw ArgError("femb_object_info: NULL argument");
  fill_info(o->til, out);
  out->device_bytes = o->device_bytes;
  FEMB_CATCH
}

int femb_set_stream(femb_object* o, void* s) {
  FEMB_TRY
  if (!o) throw ArgError("NULL object");
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
  if (o->own_stream) cudaStreamDestroy(o->stream);
  o->own_stream = false;
  o->stream = static_cast<cudaStream_t>(s);
  o->cfg.stream = o->stream;
  if (o->graph.exec) { cudaGraphExecDestroy(o->graph.exec); o->graph = StepGraph(); }
  if (o->solve_graph.exec) { cudaGraphExecDestroy(o->solve_graph.exec); o->solve_graph = SolveGraph(); }
  FEMB_CATCH
}

int femb_blocks_set(femb_object* o, int32_t n, const float* centers, const float* radii) {
  FEMB_TRY
  if (!o || n < 0 || (n > 0 && (!centers || !radii))) throw ArgError("femb_blocks_set: bad argument");
  std::vector<float4> b((size_t)n);
  for (int i = 0; i < n; ++i) {
    b[i].x = centers[(size_t)i * o->dim + 0];
    b[i].y = centers[(size_t)i * o->dim + 1];
    b[i].z = o->dim == 3 ? centers[(size_t)i * 3 + 2] : 0.f;
    b[i].w = radii[i];
  }
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
  o->d_blocks = n ? o->alloc<float4>((size_t)n, false) : nullptr;
  if (n) FEMB_CUDA(cudaMemcpyAsync(o->d_blocks, b.data(), b.size() * sizeof(float4), cudaMemcpyHostToDevice, o->stream));
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
  o->n_blocks = n;
  o->blocks_version++;
  FEMB_CATCH
}

int femb_explicit_grad(femb_object* o) {
  FEMB_TRY
  if (!o) throw ArgError("NULL object");
  FEMB_CUDA(cudaSetDevice(o->device));
  o->prep_full_pass();
  launch_explicit(o->dim, EX_GRAD, o->cfg, o->dtl, o->ex, o->vs, o->mat, KinArgs{}, nullptr, nullptr);
  FEMB_CATCH
}

int femb_energy_grad(femb_object* o, float* U_out) {
  FEMB_TRY
  if (!o) throw ArgError("NULL object");
  FEMB_CUDA(cudaSetDevice(o->device));
  o->prep_full_pass();
  launch_explicit(o->dim, EX_ENERGY, o->cfg, o->dtl, o->ex, o->vs, o->mat, KinArgs{}, o->d_phi, o->d_red);
  const int nbb = (o->dtl.n_boundary + 255) / 256;
  launch_reduce_final(o->cfg, energy_tile_grid(o->dtl) + nbb, o->d_red, o->d_U);
  if (U_out) return femb_get_energy(o, U_out);
  FEMB_CATCH
}

int femb_render(femb_object* o, const femb_view* v, const int32_t* prims, int32_t n_prims, uint8_t* rgb_out) {
  FEMB_TRY
  if (!o || !v || !rgb_out) throw ArgError("femb_render: NULL argument");
  if (v->width <= 0 || v->height <= 0 || (int64_t)v->width * v->height > (1 << 26)) throw ArgError("femb_render: bad image size");
  if (n_prims < 0 || (n_prims > 0 && !prims)) throw ArgError("femb_render: bad primitive list");
  FEMB_CUDA(cudaSetDevice(o->device));
  if (o->dim == 2 && !prims) { prims = o->elems2d.data(); n_prims = o->n_elem; }
  if (o->dim == 2 && n_prims != o->n_elem) throw ArgError("femb_render: 2D primitives are the object's elements (phi is per element)");
  if (o->prims_version != v->prims_version || o->n_prims != n_prims) {
    for (int64_t i = 0; i < (int64_t)n_prims * 3; ++i)
      if (prims[i] < 0 || prims[i] >= o->n_vert) throw ArgError("femb_render: vertex index out of range");
    FEMB_CUDA(cudaStreamSynchronize(o->stream));
    o->d_prims = n_prims ? o->alloc<int32_t>((size_t)n_prims * 3, false) : nullptr;
    if (n_prims)
      FEMB_CUDA(cudaMemcpyAsync(o->d_prims, prims, (size_t)n_prims * 3 * sizeof(int32_t), cudaMemcpyHostToDevice, o->stream));
    FEMB_CUDA(cudaStreamSynchronize(o->stream));
    o->n_prims = n_prims;
    o->prims_version = v->prims_version;
  }
  const int npx = v->width * v->height;
  if (npx > o->fb_pixels) {
    o->d_fb = o->alloc<unsigned long long>((size_t)npx, false);
    o->d_rgb = o->alloc<uint8_t>((size_t)npx * 3, false);
    o->fb_pixels = npx;
  }
  if (o->dim == 2)
    launch_render2d(o->cfg, o->n_prims, o->d_prims, o->d_new_of_old, o->d_slot_of_elem, o->vs.pos, o->d_phi, o->n_vert,
                    o->d_blocks, o->n_blocks, o->d_fb, o->d_rgb, v->width, v->height);
  else
    launch_render3d(o->cfg, v->eye, v->lookat, v->up, v->fov_deg > 0.f ? v->fov_deg : 45.f, o->n_prims, o->d_prims,
                    o->d_new_of_old, o->vs.pos, o->d_fb, o->d_rgb, v->width, v->height);
  FEMB_CUDA(cudaMemcpyAsync(rgb_out, o->d_rgb, (size_t)npx * 3, cudaMemcpyDeviceToHost, o->stream));
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
  FEMB_CATCH
}

int femb_trace_enable(femb_object* o, int32_t on) {
  FEMB_TRY
  if (!o) throw ArgError("NULL object");
  if (on && !kDiagBuild)
    throw StateError("femb_trace_enable: this build carries no timeline code; load the diagnostics build "
                     "(make -C csrc diag -> femb200/libfemb200_diag.so, FEMB200_LIB=<path>)");
  FEMB_CUDA(cudaSetDevice(o->device));
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
  // cached graphs carry the trace pointers of the launches they were captured from
  if (o->graph.exec) { cudaGraphExecDestroy(o->graph.exec); o->graph = StepGraph(); }
  if (o->solve_graph.exec) { cudaGraphExecDestroy(o->solve_graph.exec); o->solve_graph = SolveGraph(); }
  const size_t words = (size_t)kTraceSlots * kTraceWords;
  if (on && !o->d_trace) o->d_trace = o->alloc<unsigned long long>(words);
  if (on) FEMB_CUDA(cudaMemsetAsync(o->d_trace, 0, words * sizeof(unsigned long long), o->stream));
  o->cfg.trace = on ? o->d_trace : nullptr;
  trace_restart();
  FEMB_CATCH
}

int femb_trace_get(femb_object* o, uint64_t* words, int32_t max_slots, int32_t* n_out, char* names, size_t names_bytes) {
  FEMB_TRY
  if (!o || !words || !n_out) throw ArgError("NULL argument");
  if (!o->d_trace) throw StateError("femb_trace_get: tracing was never enabled");
  FEMB_CUDA(cudaSetDevice(o->device));
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
  const int n = std::min((int)max_slots, trace_slots_used());
  FEMB_CUDA(cudaMemcpy(words, o->d_trace, (size_t)n * kTraceWords * sizeof(uint64_t), cudaMemcpyDeviceToHost));
  *n_out = n;
  if (names && names_bytes) {
    std::string all;
    for (int i = 0; i < n; ++i) { all += trace_slot_name(i); all += '\n'; }
    std::snprintf(names, names_bytes, "%s", all.c_str());
  }
  FEMB_CATCH
}

int femb_diagnostics(femb_object* o, int64_t* out3) {
  FEMB_TRY
  if (!o || !out3) throw ArgError("NULL argument");
  FEMB_CUDA(cudaSetDevice(o->device));
  if (!o->d_diag) o->d_diag = o->alloc<unsigned long long>(3);
  launch_diagnostics(o->cfg, o->dim, o->dtl, o->e, o->vs.pos, o->d_phi, o->d_diag);
  unsigned long long h[3];
  FEMB_CUDA(cudaMemcpyAsync(h, o->d_diag, sizeof(h), cudaMemcpyDeviceToHost, o->stream));
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
  for (int i = 0; i < 3; ++i) out3[i] = (int64_t)h[i];
  FEMB_CATCH
}

int femb_get_energy(femb_object* o, float* U_out) {
  FEMB_TRY
  if (!o || !U_out) throw ArgError("NULL argument");
  double u = 0.0;
  FEMB_CUDA(cudaMemcpyAsync(&u, o->d_U, sizeof(double), cudaMemcpyDeviceToHost, o->stream));
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
  *U_out = (float)u;
  FEMB_CATCH
}

int femb_kinematic(femb_object* o, float dt, const float* g_dir, int32_t use_grad) {
  FEMB_TRY
  if (!o) throw ArgError("NULL object");
  FEMB_CUDA(cudaSetDevice(o->device));
  launch_kinematic(o->dim, o->cfg, o->n_vert, o->vs, o->kin_args(dt, g_dir), use_grad);
  FEMB_CATCH
}

int femb_explicit_step(femb_object* o, float dt, const float* g_dir, int32_t n_steps) {
  FEMB_TRY
  if (!o || n_steps < 0) throw ArgError("femb_explicit_step: bad argument");
  FEMB_CUDA(cudaSetDevice(o->device));
  const KinArgs kin = o->kin_args(dt, g_dir);
  o->prep_full_pass();
  if (o->halo.n > 0 && !o->peer_on)
    throw StateError("object has halo vertices: bind peers (femb_peer_bind) or drive femb_explicit_step_phase");
  auto run = [&]() {
    for (int s = 0; s < n_steps; ++s) {
      if (!o->peer_on) {
        launch_explicit(o->dim, EX_STEP, o->cfg, o->dtl, o->ex, o->vs, o->mat, kin, nullptr, nullptr);
        continue;
      }
      // slab rank: tile kernel | shared-vertex sums straight into the neighbours' memory | boundary kernel
      o->peer_pass([&](const LaunchCfg& c) {
        launch_explicit(o->dim, EX_STEP, c, o->dtl, o->ex, o->vs, o->mat, kin, nullptr, nullptr);
      });
    }
  };
  // (the legacy default stream cannot be captured)
  if (n_steps < 4 || o->stream == nullptr || o->stream == cudaStreamLegacy) {
    run();
  } else {
    StepGraph& g = o->graph;
    const bool hit = g.exec && g.n_steps == n_steps && g.dt == dt && g.damp == kin.damp &&
                     g.blocks_version == o->blocks_version && std::memcmp(g.g, kin.g98, sizeof(g.g)) == 0;
    if (!hit) {
      if (g.exec) { cudaGraphExecDestroy(g.exec); g = StepGraph(); }
      // warm the function attributes outside the capture
      cudaGraph_t graph = nullptr;
      const uint64_t before = launch_count();
      FEMB_CUDA(cudaStreamBeginCapture(o->stream, cudaStreamCaptureModeThreadLocal));
      try {
        run();
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
      g.n_steps = n_steps; g.dt = dt; g.damp = kin.damp; g.blocks_version = o->blocks_version;
      std::memcpy(g.g, kin.g98, sizeof(g.g));
    } else {
      for (uint64_t i = 0; i < g.launches; ++i) note_launch("graph", o->stream);
    }
    FEMB_CUDA(cudaGraphLaunch(g.exec, o->stream));
  }
  FEMB_CATCH
}

int femb_explicit_step_timed(femb_object* o, float dt, const float* g_dir, int32_t n_steps, float* tile_ms,
                             float* boundary_ms) {
  FEMB_TRY
  if (!o || n_steps <= 0 || !tile_ms || !boundary_ms) throw ArgError("femb_explicit_step_timed: bad argument");
  FEMB_CUDA(cudaSetDevice(o->device));
  const KinArgs kin = o->kin_args(dt, g_dir);
  std::vector<cudaEvent_t> ev((size_t)n_steps * 3);
  for (auto& e : ev) FEMB_CUDA(cudaEventCreate(&e));
  LaunchCfg cfg = o->cfg;
  for (int s = 0; s < n_steps; ++s) {
    FEMB_CUDA(cudaEventRecord(ev[3 * s], o->stream));
    cfg.ev_mid = ev[3 * s + 1];
    launch_explicit(o->dim, EX_STEP, cfg, o->dtl, o->ex, o->vs, o->mat, kin, nullptr, nullptr);
    FEMB_CUDA(cudaEventRecord(ev[3 * s + 2], o->stream));
  }
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
  double t = 0.0, b = 0.0;
  for (int s = 0; s < n_steps; ++s) {
    float ms = 0.f;
    FEMB_CUDA(cudaEventElapsedTime(&ms, ev[3 * s], ev[3 * s + 1])); t += ms;
    FEMB_CUDA(cudaEventElapsedTime(&ms, ev[3 * s + 1], ev[3 * s + 2])); b += ms;
  }
  for (auto& e : ev) cudaEventDestroy(e);
  *tile_ms = (float)t;
  *boundary_ms = (float)b;
  FEMB_CATCH
}

int femb_halo_bind(femb_object* o, void* send_dev, void* recv_dev) {
  FEMB_TRY
  if (!o) throw ArgError("NULL object");
  if (o->halo.n == 0) throw StateError("object was created without halo vertices");
  if (!send_dev || !recv_dev) throw ArgError("femb_halo_bind: NULL buffer");
  o->halo.send = static_cast<float4*>(send_dev);
  o->halo.recv = static_cast<float4*>(recv_dev);
  FEMB_CATCH
}

int femb_comm_bytes(int32_t max_halo, size_t* bytes_out, int32_t* recv_stride_out) {
  FEMB_TRY
  if (max_halo <= 0 || !bytes_out || !recv_stride_out) throw ArgError("femb_comm_bytes: bad argument");
  size_t a, b, stride;
  *bytes_out = comm_layout(max_halo, &a, &b, &stride);
  *recv_stride_out = (int32_t)stride;
  FEMB_CATCH
}

int femb_peer_bind(femb_object* o, const femb_peer_desc* d) {
  FEMB_TRY
  if (!o || !d || !d->bufs) throw ArgError("femb_peer_bind: NULL argument");
  if (o->halo.n == 0) throw StateError("object was created without halo vertices");
  if (d->world < 2 || d->world > kMaxRanks || d->rank < 0 || d->rank >= d->world) throw ArgError("bad rank / world size");
  if (d->n_left < 0 || d->n_left > o->halo.n) throw ArgError("bad n_left");
  if ((d->n_left > 0) != (d->left >= 0) || (d->n_left < o->halo.n) != (d->right >= 0))
    throw ArgError("neighbour ranks do not match the halo split");
  FEMB_CUDA(cudaSetDevice(o->device));
  PeerComm pc{};
  for (int r = 0; r < d->world; ++r) {
    if (!d->bufs[r]) throw ArgError("NULL comm buffer");
    pc.all[r] = static_cast<char*>(d->bufs[r]);
  }
  pc.self = pc.all[d->rank];
  pc.left = d->left >= 0 ? pc.all[d->left] : nullptr;
  pc.right = d->right >= 0 ? pc.all[d->right] : nullptr;
  pc.rank = d->rank; pc.world = d->world; pc.n_left = d->n_left;
  pc.left_dst_off = d->left_dst_off; pc.right_dst_off = d->right_dst_off;
  comm_layout(o->halo.n, &pc.recv_off, &pc.slot_off, &pc.recv_stride);
  if (d->recv_stride < o->halo.n) throw ArgError("recv_stride must be the value femb_comm_bytes returned for the largest halo");
  pc.recv_stride = (size_t)d->recv_stride;   // common to all ranks: remote addresses are computed with it
  comm_ll_offsets(pc.recv_off, pc.recv_stride, &pc.ll_slot_off, &pc.ll_off);
  pc.st = o->alloc<CommState>(1);
  if (!o->side) {
    // highest priority: when the halo tiles finish, the (small) send kernel gets SM slots before the
    // CTAs of the other tiles fill the device
    int prio_lo = 0, prio_hi = 0;
    FEMB_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    FEMB_CUDA(cudaStreamCreateWithPriority(&o->side, cudaStreamNonBlocking, prio_hi));
    FEMB_CUDA(cudaEventCreateWithFlags(&o->ev_fork, cudaEventDisableTiming));
    FEMB_CUDA(cudaEventCreateWithFlags(&o->ev_join, cudaEventDisableTiming));
  }
  o->d_allred_own = o->alloc<double>(1);
  o->d_allred = o->d_allred_own;
  {
    // device-resident exchange description for the kernels that send / receive themselves
    const char* env = std::getenv("FEMB_PEER_FUSED");
    const bool fused = !(env && env[0] == '0');
    std::vector<int32_t> bhalo((size_t)o->dtl.n_boundary, -1);
    for (size_t k = 0; k < o->til.halo_b.size(); ++k) bhalo[(size_t)o->til.halo_b[k]] = (int32_t)k;
    int32_t* d_bhalo = o->alloc<int32_t>(bhalo.size(), false);
    FEMB_CUDA(cudaMemcpyAsync(d_bhalo, bhalo.data(), bhalo.size() * sizeof(int32_t), cudaMemcpyHostToDevice, o->stream));
    // local partial slots of every shared vertex (its last slot is the remote one)
    std::vector<int2> hrange(o->til.halo_b.size());
    for (size_t k = 0; k < hrange.size(); ++k) {
      const int32_t b = o->til.halo_b[k];
      hrange[k] = make_int2(o->til.bptr[(size_t)b], o->til.bptr[(size_t)b + 1] - 1);
    }
    int2* d_hrange = o->alloc<int2>(hrange.size(), false);
    FEMB_CUDA(cudaMemcpyAsync(d_hrange, hrange.data(), hrange.size() * sizeof(int2), cudaMemcpyHostToDevice, o->stream));
    o->d_hrange = d_hrange;
    HaloDev hd{pc, o->halo.n, o->halo.b, d_hrange, d_bhalo};
    HaloDev* d_hd = o->alloc<HaloDev>(1, false);
    FEMB_CUDA(cudaMemcpyAsync(d_hd, &hd, sizeof(HaloDev), cudaMemcpyHostToDevice, o->stream));
    FEMB_CUDA(cudaStreamSynchronize(o->stream));   // bhalo / hd are stack / local storage
    o->d_halo_dev = fused ? d_hd : nullptr;
    o->n_send_ctas = std::max(1, std::min(o->cfg.sms, (o->halo.n + 255) / 256));   // one vertex per sender thread
    if (const char* w = std::getenv("FEMB_PEER_FUSED_WAVES")) o->fuse_max_waves = std::atoi(w);
    if (const char* m = std::getenv("FEMB_PEER_MODE")) {
      o->late_send = fused && std::string(m) == "late";
      if (std::string(m) == "one") {
        if (!kDiagBuild) throw StateError("FEMB_PEER_MODE=one needs the diagnostics build (libfemb200_diag.so)");
        o->one_kernel = 1;
      }
      if (std::string(m) == "split") o->one_kernel = 0;
    }
  }
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
  o->peer = pc;
  o->peer_on = true;
  if (o->graph.exec) { cudaGraphExecDestroy(o->graph.exec); o->graph = StepGraph(); }
  if (o->solve_graph.exec) { cudaGraphExecDestroy(o->solve_graph.exec); o->solve_graph = SolveGraph(); }
  FEMB_CATCH
}

// completes the lumped masses of shared vertices through the peer buffers (once, after femb_peer_bind)
int femb_peer_mass(femb_object* o) {
  FEMB_TRY
  if (!o || !o->peer_on) throw StateError("femb_peer_mass: peers are not bound");
  FEMB_CUDA(cudaSetDevice(o->device));
  launch_halo_send(o->cfg, o->dtl, o->peer, o->halo, nullptr, o->vs.pos, 1);
  launch_halo_recv(o->cfg, o->dtl, o->peer, o->halo, nullptr, o->vs.pos, 1);
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
  o->check_comm_error();
  FEMB_CATCH
}

static void need_halo(const femb_object* o, int phase) {
  if (!o) throw ArgError("NULL object");
  if (phase != 1 && phase != 2) throw ArgError("phase must be 1 or 2");
  if (o->halo.n == 0 || !o->halo.send || !o->halo.recv) throw StateError("halo buffers are not bound");
}

int femb_halo_mass(femb_object* o, int32_t phase) {
  FEMB_TRY
  need_halo(o, phase);
  FEMB_CUDA(cudaSetDevice(o->device));
  launch_halo_mass(o->cfg, o->dtl, o->halo, o->vs.pos, phase - 1);
  FEMB_CATCH
}

int femb_explicit_step_phase(femb_object* o, float dt, const float* g_dir, int32_t phase) {
  FEMB_TRY
  need_halo(o, phase);
  FEMB_CUDA(cudaSetDevice(o->device));
  LaunchCfg cfg = o->cfg;
  cfg.phase = phase;
  if (phase == 2) {
    launch_halo_unpack(cfg, o->dtl, o->halo, o->vs.partial);
    o->halo_dirty = true;
  }
  launch_explicit(o->dim, EX_STEP, cfg, o->dtl, o->ex, o->vs, o->mat, o->kin_args(dt, g_dir), nullptr, nullptr);
  if (phase == 1) launch_halo_pack(cfg, o->dtl, o->halo, o->vs.partial);
  FEMB_CATCH
}

int femb_advect_implicit(femb_object* o, float dt, const float* g_dir) {
  FEMB_TRY
  if (!o) throw ArgError("NULL object");
  if (!o->implicit) throw StateError("object was created without implicit buffers");
  FEMB_CUDA(cudaSetDevice(o->device));
  launch_advect_implicit(o->dim, o->cfg, o->n_vert, o->vs, o->kin_args(dt, g_dir));
  FEMB_CATCH
}

int femb_sync(femb_object* o) {
  FEMB_TRY
  if (!o) throw ArgError("NULL object");
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
  o->check_comm_error();
  FEMB_CATCH
}

// ---- field access ---------------------------------------------------------
struct FieldRef {
  float4* vec = nullptr;   // vertex field
  int comp0 = 0, ncomp = 0;
  int elem_what = -1;      // element field selector for k_pack_elem
  size_t count = 0;        // floats on the host side
};

static FieldRef resolve(femb_object* o, int field) {
  FieldRef r;
  const size_t nv = (size_t)o->n_vert, ne = (size_t)o->n_elem, d = (size_t)o->dim;
  auto vec = [&](float4* p) { r.vec = p; r.comp0 = 0; r.ncomp = o->dim; r.count = nv * d; };
  auto need_implicit = [&]() { if (!o->implicit) throw StateError("field needs an implicit object"); };
  switch (field) {
    case FEMB_F_POS: vec(o->vs.pos); break;
    case FEMB_F_VEL: vec(o->vs.vel); break;
    case FEMB_F_VEL_G: need_implicit(); vec(o->vs.vel_g); break;
    case FEMB_F_FORCE: vec(o->vs.force); break;
    case FEMB_F_GRAD: vec(o->vs.grad); break;
    case FEMB_F_MASS: r.vec = o->vs.pos; r.comp0 = 3; r.ncomp = 1; r.count = nv; break;
    case FEMB_F_REF_POS: vec(o->ref_pos); break;
    case FEMB_F_VEC_B: need_implicit(); vec(o->iv.b); break;
    case FEMB_F_VEC_X: need_implicit(); vec(o->iv.x); break;
    case FEMB_F_PAST_X: need_implicit(); vec(o->iv.past_x); break;
    case FEMB_F_DIAG_A: need_implicit(); vec(o->iv.diag); break;
    case FEMB_F_ELEM_VOLUME: r.elem_what = 0; r.count = ne; break;
    case FEMB_F_ELEM_REF: r.elem_what = 1; r.count = ne * d * d; break;
    case FEMB_F_ELEM_PHI: r.elem_what = 2; r.count = ne; break;
    case FEMB_F_ELEM_H: need_implicit(); r.elem_what = 3; r.count = ne * d * d; break;
    default: throw ArgError("unknown field id");
  }
  return r;
}

int femb_get(femb_object* o, int32_t field, void* host, size_t bytes) {
  FEMB_TRY
  if (!o || !host) throw ArgError("femb_get: NULL argument");
  FEMB_CUDA(cudaSetDevice(o->device));
  const FieldRef r = resolve(o, field);
  if (bytes != r.count * sizeof(float)) throw ArgError("femb_get: byte count does not match the field size");
  o->ensure_stage(r.count);
  if (r.vec)
    launch_pack_vertex(o->cfg, o->dim, o->n_vert, o->d_new_of_old, r.vec, r.comp0, r.ncomp, o->d_stage);
  else
    launch_pack_elem(o->cfg, o->dim, o->n_elem, o->d_slot_of_elem, o->e, o->h, o->d_phi, r.elem_what, o->d_stage);
  FEMB_CUDA(cudaMemcpyAsync(host, o->d_stage, bytes, cudaMemcpyDeviceToHost, o->stream));
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
  o->check_comm_error();   // a peer wait that timed out must not hand out stale data silently
  FEMB_CATCH
}

int femb_set(femb_object* o, int32_t field, const void* host, size_t bytes) {
  FEMB_TRY
  if (!o || !host) throw ArgError("femb_set: NULL argument");
  FEMB_CUDA(cudaSetDevice(o->device));
  const FieldRef r = resolve(o, field);
  if (!r.vec) throw ArgError("femb_set: element fields are read-only");
  if (field == FEMB_F_REF_POS) throw ArgError("femb_set: ref_pos is read-only");
  if (bytes != r.count * sizeof(float)) throw ArgError("femb_set: byte count does not match the field size");
  o->ensure_stage(r.count);
  FEMB_CUDA(cudaMemcpyAsync(o->d_stage, host, bytes, cudaMemcpyHostToDevice, o->stream));
  launch_unpack_vertex(o->cfg, o->dim, o->n_vert, o->d_new_of_old, o->d_stage, r.comp0, r.ncomp, r.vec);
  // the staging buffer is reused by the next call: make the copy-in complete before returning
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
  FEMB_CATCH
}

// ---- asynchronous field transfers (pipelined host <-> device traffic) ------
static femb_object::Xfer& xfer_slot(femb_object* o, femb_object::Xfer* arr, int field, size_t floats) {
  if (field < 0 || field >= 48) throw ArgError("unknown field id");
  femb_object::Xfer& x = arr[field];
  if (!o->s_h2d) {
    FEMB_CUDA(cudaStreamCreateWithFlags(&o->s_h2d, cudaStreamNonBlocking));
    FEMB_CUDA(cudaStreamCreateWithFlags(&o->s_d2h, cudaStreamNonBlocking));
  }
  if (x.floats < floats) {
    if (x.buf) { FEMB_CUDA(cudaDeviceSynchronize()); FEMB_CUDA(cudaFree(x.buf)); }
    FEMB_CUDA(cudaMalloc(&x.buf, floats * sizeof(float)));
    x.floats = floats;
  }
  if (!x.copied) {
    FEMB_CUDA(cudaEventCreateWithFlags(&x.copied, cudaEventDisableTiming));
    FEMB_CUDA(cudaEventCreateWithFlags(&x.consumed, cudaEventDisableTiming));
  }
  return x;
}

int femb_set_async(femb_object* o, int32_t field, const void* host, size_t bytes) {
  FEMB_TRY
  if (!o || !host) throw ArgError("femb_set_async: NULL argument");
  FEMB_CUDA(cudaSetDevice(o->device));
  const FieldRef r = resolve(o, field);
  if (!r.vec) throw ArgError("femb_set_async: element fields are read-only");
  if (field == FEMB_F_REF_POS) throw ArgError("femb_set_async: ref_pos is read-only");
  if (bytes != r.count * sizeof(float)) throw ArgError("femb_set_async: byte count does not match the field size");
  femb_object::Xfer& x = xfer_slot(o, o->xin, field, r.count);
  // the previous unpack of this staging buffer must have read it before the next copy lands in it
  if (x.used) FEMB_CUDA(cudaStreamWaitEvent(o->s_h2d, x.consumed, 0));
  FEMB_CUDA(cudaMemcpyAsync(x.buf, host, bytes, cudaMemcpyHostToDevice, o->s_h2d));
  FEMB_CUDA(cudaEventRecord(x.copied, o->s_h2d));
  FEMB_CUDA(cudaStreamWaitEvent(o->stream, x.copied, 0));
  launch_unpack_vertex(o->cfg, o->dim, o->n_vert, o->d_new_of_old, x.buf, r.comp0, r.ncomp, r.vec);
  FEMB_CUDA(cudaEventRecord(x.consumed, o->stream));
  x.used = true;
  FEMB_CATCH
}

int femb_get_async(femb_object* o, int32_t field, void* host, size_t bytes) {
  FEMB_TRY
  if (!o || !host) throw ArgError("femb_get_async: NULL argument");
  FEMB_CUDA(cudaSetDevice(o->device));
  const FieldRef r = resolve(o, field);
  if (bytes != r.count * sizeof(float)) throw ArgError("femb_get_async: byte count does not match the field size");
  femb_object::Xfer& x = xfer_slot(o, o->xout, field, r.count);
  // the previous device->host copy out of this staging buffer must have finished before it is packed again
  if (x.used) FEMB_CUDA(cudaStreamWaitEvent(o->stream, x.copied, 0));
  if (r.vec)
    launch_pack_vertex(o->cfg, o->dim, o->n_vert, o->d_new_of_old, r.vec, r.comp0, r.ncomp, x.buf);
  else
    launch_pack_elem(o->cfg, o->dim, o->n_elem, o->d_slot_of_elem, o->e, o->h, o->d_phi, r.elem_what, x.buf);
  FEMB_CUDA(cudaEventRecord(x.consumed, o->stream));
  FEMB_CUDA(cudaStreamWaitEvent(o->s_d2h, x.consumed, 0));
  FEMB_CUDA(cudaMemcpyAsync(host, x.buf, bytes, cudaMemcpyDeviceToHost, o->s_d2h));
  FEMB_CUDA(cudaEventRecord(x.copied, o->s_d2h));
  x.used = true;
  FEMB_CATCH
}

int femb_transfers_wait(femb_object* o) {
  FEMB_TRY
  if (!o) throw ArgError("NULL object");
  FEMB_CUDA(cudaSetDevice(o->device));
  if (o->s_h2d) FEMB_CUDA(cudaStreamSynchronize(o->s_h2d));
  FEMB_CUDA(cudaStreamSynchronize(o->stream));
  if (o->s_d2h) FEMB_CUDA(cudaStreamSynchronize(o->s_d2h));
  FEMB_CATCH
}

int femb_set_host_threads(int32_t n) {
  // launchers such as torchrun export OMP_NUM_THREADS=1; the host-side tiling builder is OpenMP code
  if (n > 0) omp_set_num_threads(n);
  return omp_get_max_threads();
}

// ---- host-only tiling ------------------------------------------------------
struct femb_tiling {
  Tiling t;
};

int femb_tiling_build(int32_t dim, int32_t n_vert, int32_t n_elem, const float* verts, const int32_t* elems,
                      int32_t tile_elems, femb_tiling** out) {
  FEMB_TRY
  if (!out) throw ArgError("NULL out");
  *out = nullptr;
  femb_tiling* h = new femb_tiling();
  const std::string err = build_tiling(dim, n_vert, n_elem, verts, elems, tile_elems, h->t);
  if (!err.empty()) { delete h; throw ArgError(err); }
  *out = h;
  FEMB_CATCH
}

int femb_tiling_build_ex(int32_t dim, int32_t n_vert, int32_t n_elem, const float* verts, const int32_t* elems,
                         int32_t tile_elems, int32_t n_halo, const int32_t* halo_verts, int32_t wave,
                         femb_tiling** out) {
  FEMB_TRY
  if (!out) throw ArgError("NULL out");
  *out = nullptr;
  femb_tiling* h = new femb_tiling();
  const std::string err = build_tiling(dim, n_vert, n_elem, verts, elems, tile_elems, h->t, n_halo, halo_verts, wave);
  if (!err.empty()) { delete h; throw ArgError(err); }
  *out = h;
  FEMB_CATCH
}

int femb_tiling_destroy(femb_tiling* t) {
  delete t;
  return FEMB_OK;
}

int femb_tiling_info(const femb_tiling* t, femb_info* out) {
  FEMB_TRY
  if (!t || !out) throw ArgError("NULL argument");
  fill_info(t->t, out);
  FEMB_CATCH
}

int64_t femb_tiling_get(const femb_tiling* h, int32_t what, void* out, int64_t cap) {
  if (!h) return -1;
  const Tiling& t = h->t;
  const void* src = nullptr;
  int64_t count = 0;
  size_t esz = 4;
  switch (what) {
    case 0: src = t.new_of_old.data(); count = (int64_t)t.new_of_old.size(); break;
    case 1: src = t.slot_of_elem.data(); count = (int64_t)t.slot_of_elem.size(); break;
    case 2: src = t.tiles.data(); count = (int64_t)t.tiles.size() * 12; break;
    case 3: src = t.lv.data(); count = (int64_t)t.lv.size() * 2; break;
    case 4: src = t.jds.data(); count = (int64_t)t.jds.size(); esz = 2; break;
    case 5: src = t.rounds.data(); count = (int64_t)t.rounds.size(); break;
    case 6: src = t.lidx.data(); count = (int64_t)t.lidx.size(); break;
    case 7: src = t.bptr.data(); count = (int64_t)t.bptr.size(); break;
    default: return -1;
  }
  if (out && cap > 0) std::memcpy(out, src, (size_t)std::min(cap, count) * esz);
  return count;
}

}  // extern "C"
