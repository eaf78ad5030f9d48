// kernels_explicit.cu -- setup (object.py:345-362), explicit element gradient
// (explicit.py:8-49), energy (explicit_auto_diff.py:8-30) and the symplectic-Euler
// vertex update with walls and circle blocks (kinematic.py:9-45), hand-written
// for sm_100a on the tile layout of tiling.h.
#include <atomic>
#include <stdexcept>
#include <string>

#include "launch.h"
#include "tile_pass.cuh"

namespace femb {

static std::atomic<uint64_t> g_launches{0};
uint64_t launch_count() { return g_launches.load(); }
void note_launch(const char* what, cudaStream_t) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) throw std::runtime_error(std::string(what) + ": " + cudaGetErrorString(err));
}

static int g_trace_next = 0;
static std::string g_trace_names[kTraceSlots];
void trace_restart() { g_trace_next = 0; }
int trace_slots_used() { return g_trace_next; }
const char* trace_slot_name(int slot) { return slot >= 0 && slot < g_trace_next ? g_trace_names[slot].c_str() : ""; }
unsigned long long* trace_take_slot(const LaunchCfg& c, const char* name, const char* what) {
  if (!c.trace || g_trace_next >= kTraceSlots) return nullptr;
  g_trace_names[g_trace_next] = std::string(name) + ":" + what;
  return c.trace + kTraceWords * (g_trace_next++);
}

// ---------------------------------------------------------------------------
// vertex update shared by the fused step, the stand-alone kinematic kernel
// ---------------------------------------------------------------------------
template <int D>
__device__ __forceinline__ void kinematic_update(float4& p, float4& v, const float (&g)[D], const KinArgs& k) {
  float pos[3] = {p.x, p.y, p.z};
  float vel[3] = {v.x, v.y, v.z};
  const float m = p.w;
#pragma unroll
  for (int i = 0; i < D; ++i) {
    // vel += (acc + 9.8*g_dir - grad/mass) * dt, acc == 0     (kinematic.py:15-20)
    vel[i] += (k.g98[i] - g[i] / m) * k.dt;
    vel[i] *= k.damp;                                    // kinematic.py:23
    if (pos[i] < 0.f && vel[i] < 0.f) vel[i] = 0.f;      // kinematic.py:25-30 (old positions)
    if (pos[i] > 1.f && vel[i] > 0.f) vel[i] = 0.f;
  }
  for (int b = 0; b < k.blocks.count; ++b) {             // kinematic.py:32-41
    const float4 blk = k.blocks.data[b];
    if (blk.w <= 0.f) continue;
    const float c[3] = {blk.x, blk.y, blk.z};
    float disp[3], nsq = 0.f, vdc = 0.f, vdd = 0.f;
#pragma unroll
    for (int i = 0; i < D; ++i) {
      disp[i] = pos[i] - c[i];
      nsq += disp[i] * disp[i];
      vdc += vel[i] * (c[i] - pos[i]);
      vdd += vel[i] * disp[i];
    }
    if (sqrtf(nsq) < blk.w && vdc > 0.f) {
#pragma unroll
      for (int i = 0; i < D; ++i) vel[i] -= vdd * disp[i] / nsq;
    }
  }
#pragma unroll
  for (int i = 0; i < D; ++i) pos[i] += vel[i] * k.dt;   // kinematic.py:43
  p.x = pos[0]; p.y = pos[1]; if (D == 3) p.z = pos[2];
  v.x = vel[0]; v.y = vel[1]; if (D == 3) v.z = vel[2];
}

// ---------------------------------------------------------------------------
// setup: Dm^-1, volume, lumped mass.  Contraction-free arithmetic (__fmul_rn /
// __fsub_rn ...) so ref and volume are bit-identical to the fp32 oracle.
// ---------------------------------------------------------------------------
template <int D>
struct OpSetup {
  static constexpr bool PERSIST = false;
  static constexpr bool ASYNC_VERTEX = false, PAIR = false;
  static constexpr int BLOCK = kBlock, MIN_CTAS = kMinCtas,NSTAGE = 1, GATHER = G_UNIFORM;
  static constexpr bool REDUCE = false;
  struct Args {
    Planes e;
    const uint32_t* lidx;
    VertexState vs;
    Material mat;
  };
  __device__ static bool early_exit(const Args&) { return false; }
  __host__ __device__ static bool itc(const Args&) { return true; }
  __device__ static float w0() { return 1.f; }
  __device__ static float4* partial_buf(const Args& a) { return a.vs.partial; }
  __device__ static double* red_buf(const Args&) { return nullptr; }
  __device__ static float4 load_vertex(const Args& a, int gid) { return a.vs.pos[gid]; }

  struct Elem {
    uint32_t w0, w1;
  };
  __device__ static void prefetch(const Args&, int, uint32_t) {}
  __device__ static void prefetch_verts(const Args&, int, uint32_t) {}
  __device__ static void load(const Args& a, int slot, Elem& d) {
    d.w0 = a.lidx[2 * slot];
    d.w1 = a.lidx[2 * slot + 1];
  }
  __device__ static void compute(const Args& a, int slot, const Elem& d, const float4* sv, float (&out)[NSTAGE],
                                 float&) {
    const uint32_t w0 = d.w0, w1 = d.w1;
    Mat<D> r;
    edge_matrix<D>(sv, w0, w1, r);  // plain subtractions: nothing to contract
    float vol;
    Mat<D> inv;
    if constexpr (D == 2) {
      // object.py:388-389 ; 2x2 adjugate inverse
      const float cross = __fsub_rn(__fmul_rn(r.m[0][0], r.m[1][1]), __fmul_rn(r.m[1][0], r.m[0][1]));
      vol = fabsf(cross) / 2.f;
      const float dt = __fsub_rn(__fmul_rn(r.m[0][0], r.m[1][1]), __fmul_rn(r.m[0][1], r.m[1][0]));
      const float id = __fdiv_rn(1.f, dt);
      inv.m[0][0] = __fmul_rn(r.m[1][1], id);
      inv.m[0][1] = __fmul_rn(-r.m[0][1], id);
      inv.m[1][0] = __fmul_rn(-r.m[1][0], id);
      inv.m[1][1] = __fmul_rn(r.m[0][0], id);
    } else {
      // object.py:390-391: (1/6)|p0 . (p1 x p2)| with p_j = column j
      const float cx = __fsub_rn(__fmul_rn(r.m[1][1], r.m[2][2]), __fmul_rn(r.m[2][1], r.m[1][2]));
      const float cy = __fsub_rn(__fmul_rn(r.m[2][1], r.m[0][2]), __fmul_rn(r.m[0][1], r.m[2][2]));
      const float cz = __fsub_rn(__fmul_rn(r.m[0][1], r.m[1][2]), __fmul_rn(r.m[1][1], r.m[0][2]));
      const float dot = __fadd_rn(__fadd_rn(__fmul_rn(r.m[0][0], cx), __fmul_rn(r.m[1][0], cy)), __fmul_rn(r.m[2][0], cz));
      vol = __fmul_rn((float)(1.0 / 6.0), fabsf(dot));
      const float t0 = __fsub_rn(__fmul_rn(r.m[1][1], r.m[2][2]), __fmul_rn(r.m[2][1], r.m[1][2]));
      const float t1 = __fsub_rn(__fmul_rn(r.m[0][1], r.m[2][2]), __fmul_rn(r.m[2][1], r.m[0][2]));
      const float t2 = __fsub_rn(__fmul_rn(r.m[0][1], r.m[1][2]), __fmul_rn(r.m[1][1], r.m[0][2]));
      const float dt = __fadd_rn(__fsub_rn(__fmul_rn(r.m[0][0], t0), __fmul_rn(r.m[1][0], t1)), __fmul_rn(r.m[2][0], t2));
      const float id = __fdiv_rn(1.f, dt);
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
          inv.m[j][i] = __fmul_rn(id, __fsub_rn(__fmul_rn(r.m[i1][j1], r.m[i2][j2]), __fmul_rn(r.m[i2][j1], r.m[i1][j2])));
        }
    }
    float4 e0, e1, e2;
    e0.x = __uint_as_float(w0); e0.y = __uint_as_float(w1); e0.z = vol; e0.w = inv.m[0][0];
    if constexpr (D == 3) {
      e1 = make_float4(inv.m[0][1], inv.m[0][2], inv.m[1][0], inv.m[1][1]);
      e2 = make_float4(inv.m[1][2], inv.m[2][0], inv.m[2][1], inv.m[2][2]);
      a.e.p2[slot] = e2;
    } else {
      e1 = make_float4(inv.m[0][1], inv.m[1][0], inv.m[1][1], 0.f);
    }
    a.e.p0[slot] = e0;
    a.e.p1[slot] = e1;
    // mass[v] += 1/(dim+1) * V * rho          (object.py:358-360)
    out[0] = __fmul_rn(__fmul_rn((float)(1.0 / (D + 1)), vol), a.mat.rho);
  }
  __device__ static float4 pre(const Args&, int) { return make_float4(0.f, 0.f, 0.f, 0.f); }
  __device__ static void vertex(const Args& a, int gid, float4, float4, const float (&acc)[D], float&) {
    a.vs.pos[gid].w = acc[0];
  }
};

// ---------------------------------------------------------------------------
// explicit element gradient / fused step / energy
// ---------------------------------------------------------------------------
template <int D, int MODE>
struct OpExplicit {
  static constexpr bool PERSIST = FEMB_PERSIST_EXPLICIT != 0;
  // the tile's vertex values are pos[gid] untransformed: the persistent kernel fetches the NEXT tile's values
  // global -> shared with cp.async into a second vertex buffer (no registers held across the element phase)
  static constexpr bool ASYNC_VERTEX = true;
  // PAIR: every thread works on two elements at once in packed fp32 (compute2; dev.cuh f2).  Half the threads per
  // CTA, twice the registers per thread, the same 4 CTAs/SM.
  static constexpr bool PAIR = FEMB_PAIR_EXPLICIT != 0;
  static constexpr int BLOCK = PAIR ? kBlock / 2 : kBlock, MIN_CTAS = kMinCtas, NSTAGE = D * D, GATHER = G_STAR;
  static constexpr bool REDUCE = (MODE == EX_ENERGY);
  struct Args {
    Planes e;
    VertexState vs;
    Material mat;
    KinArgs kin;
    float* phi;
    double* red;
  };
  __device__ static bool early_exit(const Args&) { return false; }
  __host__ __device__ static bool itc(const Args&) { return !REDUCE; }
  __device__ static float w0() { return 0.f; }
  __device__ static float4* partial_buf(const Args& a) { return a.vs.partial; }
  __device__ static double* red_buf(const Args& a) { return a.red; }
  __device__ static float4 load_vertex(const Args& a, int gid) { return a.vs.pos[gid]; }
  __device__ static const float4* vertex_src(const Args& a) { return a.vs.pos; }   // ASYNC_VERTEX source array

  using Elem = PlaneElem;
  __device__ static void prefetch(const Args& a, int slot, uint32_t bytes) { prefetch_planes(a.e, slot, bytes, D == 3); }
  // interior vertices of a tile are one contiguous id range: pos (+ vel / force) of the next tile
  __device__ static void prefetch_verts(const Args& a, int base, uint32_t bytes) {
    bulk_prefetch_l2(a.vs.pos + base, bytes);
    if (MODE == EX_STEP) bulk_prefetch_l2(a.vs.vel + base, bytes);
    if (MODE == EX_GRAD) bulk_prefetch_l2(a.vs.force + base, bytes);
  }
  __device__ static void load(const Args& a, int slot, Elem& d) { load_planes<D>(a.e, slot, d); }
  __device__ static void compute(const Args& a, int slot, const Elem& d, const float4* sv, float (&out)[NSTAGE],
                                 float& red) {
    const float4 e0 = d.a, e1 = d.b, e2 = d.c;
    Mat<D> R, X, F, C;
    unpack_ref<D>(e0, e1, e2, R);
    edge_matrix<D>(sv, __float_as_uint(e0.x), __float_as_uint(e0.y), X);  // explicit.py:11-16
    matmul<D>(X, R, F);                                                   // F = X @ R_inv (19)
    cofactor<D>(F, C);                                                    // F^-T = C / det
    const float J = det<D>(F);
    const float log_j = logf(J);                                          // explicit.py:24 (no clamp)
    const float V = e0.z;
    // V * P = V * (mu F - mu F^-T + lambda log(J) F^-T), F^-T = C / J   (explicit.py:25, 28): the volume is
    // folded into the two scalar coefficients (18 instead of 27 multiplies); the division by J is the
    // 2-ulp fast division -- far inside the 1e-5 force tolerance, and J <= 0 gives NaN/inf either way
    const float muV = a.mat.mu * V;
    const float cfV = __fdividef((a.mat.lambda * log_j - a.mat.mu) * V, J);
    Mat<D> P, G;
#pragma unroll
    for (int i = 0; i < D; ++i)
#pragma unroll
      for (int j = 0; j < D; ++j) P.m[i][j] = muV * F.m[i][j] + cfV * C.m[i][j];
    matmul_nt<D>(P, R, G);                                                // @ R_inv^T
#pragma unroll
    for (int j = 0; j < D; ++j)
#pragma unroll
      for (int i = 0; i < D; ++i) out[j * D + i] = G.m[i][j];
    if (MODE == EX_ENERGY) {
      // explicit_auto_diff.py:25-30
      float tr = 0.f;
#pragma unroll
      for (int i = 0; i < D; ++i)
#pragma unroll
        for (int j = 0; j < D; ++j) tr += F.m[i][j] * F.m[i][j];
      float phi = a.mat.mu / 2.f * (tr - (float)D);
      phi -= a.mat.mu * log_j;
      phi += a.mat.lambda / 2.f * log_j * log_j;
      const float pv = phi * V;
      a.phi[slot] = pv;
      red += pv;
    }
  }

  // the same for the elements in slot0 (lane x) and slot1 (lane y)
  __device__ static void compute2(const Args& a, int slot0, int slot1, const Elem& d0, const Elem& d1,
                                  const float4* sv, f2 (&out)[NSTAGE], float& red) {
    // d0 / d1 are the two slots of a pair in the pair-interleaved layout (k_pair_interleave below):
    //   3D  d0 = {i01a i01b i23a i23b | r01 r02 | r12 r20}   d1 = {Va Vb r00a r00b | r10 r11 | r21 r22}
    //   2D  d0 = {i01a i01b i2a  i2b  | r01 r10}             d1 = {Va Vb r00a r00b | r11 0}
    // (each r.. an (a, b) pair) -- every operand of the packed arithmetic is an aligned register pair as loaded
    Mat2<D> R, X, F, C;
    const f2 V = pk(d1.a.x, d1.a.y);
    R.m[0][0] = pk(d1.a.z, d1.a.w);
    R.m[0][1] = pk(d0.b.x, d0.b.y);
    if constexpr (D == 3) {
      R.m[0][2] = pk(d0.b.z, d0.b.w);
      R.m[1][0] = pk(d1.b.x, d1.b.y); R.m[1][1] = pk(d1.b.z, d1.b.w); R.m[1][2] = pk(d0.c.x, d0.c.y);
      R.m[2][0] = pk(d0.c.z, d0.c.w); R.m[2][1] = pk(d1.c.x, d1.c.y); R.m[2][2] = pk(d1.c.z, d1.c.w);
    } else {
      R.m[1][0] = pk(d0.b.z, d0.b.w); R.m[1][1] = pk(d1.b.x, d1.b.y);
    }
    {
      Mat<D> X0, X1;
      edge_matrix<D>(sv, __float_as_uint(d0.a.x), __float_as_uint(d0.a.z), X0);
      edge_matrix<D>(sv, __float_as_uint(d0.a.y), __float_as_uint(d0.a.w), X1);
      pack<D>(X0, X1, X);
    }
    matmul<D>(X, R, F);
    const f2 J = cofactor_det<D>(F, C);
    const f2 log_j = pk(logf(J.v.x), logf(J.v.y));
    const f2 muV = bc(a.mat.mu) * V;
    const f2 num = fma2(bc(a.mat.lambda), log_j, bc(-a.mat.mu)) * V;
    const f2 cfV = pk(__fdividef(num.v.x, J.v.x), __fdividef(num.v.y, J.v.y));
    Mat2<D> P, G;
#pragma unroll
    for (int i = 0; i < D; ++i)
#pragma unroll
      for (int j = 0; j < D; ++j) P.m[i][j] = fma2(muV, F.m[i][j], cfV * C.m[i][j]);
    matmul_nt<D>(P, R, G);
#pragma unroll
    for (int j = 0; j < D; ++j)
#pragma unroll
      for (int i = 0; i < D; ++i) out[j * D + i] = G.m[i][j];
    if (MODE == EX_ENERGY) {
      f2 tr = F.m[0][0] * F.m[0][0];
#pragma unroll
      for (int i = 0; i < D; ++i)
#pragma unroll
        for (int j = 0; j < D; ++j)
          if (i + j > 0) tr = fma2(F.m[i][j], F.m[i][j], tr);
      const float lj[2] = {log_j.v.x, log_j.v.y}, t[2] = {tr.v.x, tr.v.y}, vol[2] = {V.v.x, V.v.y};
      const int slot[2] = {slot0, slot1};
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        float phi = a.mat.mu / 2.f * (t[q] - (float)D);
        phi -= a.mat.mu * lj[q];
        phi += a.mat.lambda / 2.f * lj[q] * lj[q];
        const float pv = phi * vol[q];
        a.phi[slot[q]] = pv;
        red += pv;
      }
    }
  }

  // the vertex operand the epilogue updates: force (EX_GRAD) or vel (EX_STEP)
  __device__ static float4 pre(const Args& a, int gid) {
    if (MODE == EX_GRAD) return a.vs.force[gid];
    if (MODE == EX_STEP) return a.vs.vel[gid];
    return make_float4(0.f, 0.f, 0.f, 0.f);
  }
  __device__ static void vertex(const Args& a, int gid, float4 p, float4 pre, const float (&acc)[D], float&) {
    if (MODE == EX_GRAD) {
      float4 f = pre;                                                     // force[v] += f (explicit.py:42-49)
      f.x += acc[0]; f.y += acc[1]; if (D == 3) f.z += acc[D - 1];
      a.vs.force[gid] = f;
    } else if (MODE == EX_ENERGY) {
      a.vs.grad[gid] = make_float4(acc[0], acc[1], D == 3 ? acc[D - 1] : 0.f, 0.f);
    } else {
      float4 v = pre;
      kinematic_update<D>(p, v, acc, a.kin);
      a.vs.pos[gid] = p;
      a.vs.vel[gid] = v;
    }
  }
};

template <int D>
__global__ void __launch_bounds__(256) k_kinematic(int n, VertexState vs, KinArgs kin, int use_grad) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float4 p = vs.pos[i], v = vs.vel[i];
  const float4 gsrc = use_grad ? vs.grad[i] : vs.force[i];
  float g[D];
  g[0] = gsrc.x; g[1] = gsrc.y; if (D == 3) g[D - 1] = gsrc.z;
  kinematic_update<D>(p, v, g, kin);
  vs.pos[i] = p;
  vs.vel[i] = v;
  if (!use_grad) vs.force[i] = make_float4(0.f, 0.f, 0.f, 0.f);           // kinematic.py:44-45
}

// advect_implicit, implicit.py:407-437
template <int D>
__global__ void __launch_bounds__(256) k_advect_implicit(int n, VertexState vs, KinArgs k) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  float4 p4 = vs.pos[idx], v4 = vs.vel[idx], g4 = vs.vel_g[idx];
  float pos[3] = {p4.x, p4.y, p4.z}, vel[3] = {v4.x, v4.y, v4.z}, vg[3] = {g4.x, g4.y, g4.z}, v[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int i = 0; i < D; ++i) {
    vg[i] += k.g98[i] * k.dt;      // 409
    vel[i] *= k.damp;              // 410
    vg[i] *= k.damp;               // 411
    v[i] = vel[i] + vg[i];         // 412
    if (pos[i] < 0.f && v[i] < 0.f) { vel[i] = 0.f; vg[i] = 0.f; v[i] = 0.f; }   // 415-418
    if (pos[i] > 1.f && v[i] > 0.f) { vel[i] = 0.f; v[i] = 0.f; }                 // 420-423 (vel_g kept)
  }
  for (int b = 0; b < k.blocks.count; ++b) {   // 424-435
    const float4 blk = k.blocks.data[b];
    if (blk.w <= 0.f) continue;
    const float c[3] = {blk.x, blk.y, blk.z};
    float disp[3], nsq = 0.f, vdc = 0.f, d_v = 0.f, d_vel = 0.f, d_vg = 0.f;
#pragma unroll
    for (int i = 0; i < D; ++i) {
      disp[i] = pos[i] - c[i];
      nsq += disp[i] * disp[i];
      vdc += v[i] * (c[i] - pos[i]);
      d_v += v[i] * disp[i];
      d_vel += vel[i] * disp[i];
      d_vg += vg[i] * disp[i];
    }
    if (sqrtf(nsq) < blk.w && vdc > 0.f) {
#pragma unroll
      for (int i = 0; i < D; ++i) {
        v[i] -= d_v * disp[i] / nsq;
        vel[i] -= d_vel * disp[i] / nsq;
        vg[i] -= d_vg * disp[i] / nsq;
      }
    }
  }
#pragma unroll
  for (int i = 0; i < D; ++i) pos[i] += v[i] * k.dt;   // 437
  p4.x = pos[0]; p4.y = pos[1]; if (D == 3) p4.z = pos[2];
  v4.x = vel[0]; v4.y = vel[1]; if (D == 3) v4.z = vel[2];
  g4.x = vg[0]; g4.y = vg[1]; if (D == 3) g4.z = vg[2];
  vs.pos[idx] = p4; vs.vel[idx] = v4; vs.vel_g[idx] = g4;
}

// Pair-interleaved copy of the element planes for the packed (two elements per thread) kernels: the two slots
// (e, e + BP) a thread of a BP-thread CTA works on together exchange halves, plane by plane,
//   q[e]      = {p[e].x, p[e+BP].x, p[e].y, p[e+BP].y}        q[e+BP] = {p[e].z, p[e+BP].z, p[e].w, p[e+BP].w}
// so each 16-byte load lands (element a, element b) operand pairs in aligned register pairs.  Elements without
// a partner (ragged tile end) keep the plain layout.  Same addresses, same bytes as the plain planes.
__global__ void __launch_bounds__(128) k_pair_interleave(DevTiling tl, Planes src, Planes dst, int bp, int three) {
  const DevTile td = tl.tiles[blockIdx.x];
  for (int e = threadIdx.x; e < td.n_elem; e += blockDim.x) {
    const int half = (e / bp) & 1, e0 = half ? e - bp : e, e1 = e0 + bp;
    const int s = td.slot_off + e;
    const bool paired = e1 < td.n_elem;
    const float4* sp[3] = {src.p0, src.p1, src.p2};
    float4* dp[3] = {dst.p0, dst.p1, dst.p2};
    for (int k = 0; k < (three ? 3 : 2); ++k) {
      if (!paired) {
        dp[k][s] = sp[k][s];
        continue;
      }
      const float4 a = sp[k][td.slot_off + e0], b = sp[k][td.slot_off + e1];
      dp[k][s] = half ? make_float4(a.z, b.z, a.w, b.w) : make_float4(a.x, b.x, a.y, b.y);
    }
  }
}

// ---------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------
bool explicit_pair_layout() { return FEMB_PAIR_EXPLICIT != 0; }
void launch_pair_interleave(int dim, const LaunchCfg& c, const DevTiling& tl, Planes src, Planes dst) {
  k_pair_interleave<<<tl.n_tiles, 128, 0, c.stream>>>(tl, src, dst, OpExplicit<3, EX_STEP>::BLOCK, dim == 3);
  note_launch("pair_interleave", c.stream);
}

void launch_setup(int dim, const LaunchCfg& c, const DevTiling& tl, Planes e, const uint32_t* lidx,
                  VertexState vs, Material mat) {
  if (dim == 3) {
    OpSetup<3>::Args a{e, lidx, vs, mat};
    run_tile_pass<3, OpSetup<3>>("setup<3>", c, tl, a);
  } else {
    OpSetup<2>::Args a{e, lidx, vs, mat};
    run_tile_pass<2, OpSetup<2>>("setup<2>", c, tl, a);
  }
}

template <int D, int MODE>
static void launch_explicit_t(const LaunchCfg& c, const DevTiling& tl, Planes e, VertexState vs, Material mat,
                              KinArgs kin, float* phi, double* red) {
  typename OpExplicit<D, MODE>::Args a{e, vs, mat, kin, phi, red};
  run_tile_pass<D, OpExplicit<D, MODE>>("explicit", c, tl, a);
}

void launch_explicit(int dim, int mode, const LaunchCfg& c, const DevTiling& tl, Planes e, VertexState vs,
                     Material mat, KinArgs kin, float* phi, double* red) {
#define FEMB_CASE(D, M) \
  if (dim == D && mode == M) return launch_explicit_t<D, M>(c, tl, e, vs, mat, kin, phi, red);
  FEMB_CASE(3, EX_GRAD) FEMB_CASE(3, EX_STEP) FEMB_CASE(3, EX_ENERGY)
  FEMB_CASE(2, EX_GRAD) FEMB_CASE(2, EX_STEP) FEMB_CASE(2, EX_ENERGY)
#undef FEMB_CASE
  throw std::runtime_error("launch_explicit: bad dim/mode");
}

int energy_tile_grid(const DevTiling& tl) { return tile_grid<OpExplicit<3, EX_ENERGY>>(tl); }

void launch_kinematic(int dim, const LaunchCfg& c, int n_vert, VertexState vs, KinArgs kin, int use_grad) {
  const int nb = (n_vert + 255) / 256;
  if (dim == 3) k_kinematic<3><<<nb, 256, 0, c.stream>>>(n_vert, vs, kin, use_grad);
  else k_kinematic<2><<<nb, 256, 0, c.stream>>>(n_vert, vs, kin, use_grad);
  note_launch("kinematic", c.stream);
}

void launch_advect_implicit(int dim, const LaunchCfg& c, int n_vert, VertexState vs, KinArgs kin) {
  const int nb = (n_vert + 255) / 256;
  if (dim == 3) k_advect_implicit<3><<<nb, 256, 0, c.stream>>>(n_vert, vs, kin);
  else k_advect_implicit<2><<<nb, 256, 0, c.stream>>>(n_vert, vs, kin);
  note_launch("advect_implicit", c.stream);
}

}  // namespace femb
