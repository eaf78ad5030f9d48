// comm.cuh -- device-side primitives of the peer-memory communication (see kernels_comm.cu).
#pragma once
#include "launch.h"

namespace femb {

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

constexpr long long kSpinLimit = 20000000000ll;   // ~10 s at 2 GHz

__device__ __forceinline__ unsigned long long gtime_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// wait until *flag >= seq (flags only grow); false on time-out
__device__ __forceinline__ bool spin_until(const uint32_t* flag, uint32_t seq, CommState* st) {
  const long long t0 = clock64();
  while ((int32_t)(ld_acquire_sys(flag) - seq) < 0) {
    if (clock64() - t0 > kSpinLimit) {
      st->error = 1;
      return false;
    }
    __nanosleep(64);
  }
  return true;
}

// ---- "LL" words: 32 bits of payload + the 32-bit sequence number of the exchange in ONE 64-bit word, written and
// read with single (single-copy atomic) 64-bit accesses -- a word whose upper half carries the expected sequence
// number has arrived, whatever the order in which the words travel.  No system-scope fence, no separate flag, no
// second round trip: the fence of the flag protocol cost 3-10 us per sender CTA while the SMs stream at full rate
// (profiles/r02_multi_gpu.txt).  Same idea as NCCL's low-latency protocol.
__device__ __forceinline__ void ll_store(unsigned long long* p, uint32_t data, uint32_t seq) {
  const unsigned long long v = ((unsigned long long)seq << 32) | (unsigned long long)data;
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ bool ll_try(const unsigned long long* p, uint32_t seq, uint32_t& data) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  data = (uint32_t)v;
  return (uint32_t)(v >> 32) == seq;
}
// wait for n (<= 4) consecutive words; false on time-out (error flag set)
template <int N>
__device__ __forceinline__ bool ll_wait(const unsigned long long* p, uint32_t seq, uint32_t (&data)[N], CommState* st) {
  const long long t0 = clock64();
  for (;;) {
    bool ok = true;
#pragma unroll
    for (int i = 0; i < N; ++i) ok = ll_try(p + i, seq, data[i]) && ok;
    if (ok) return true;
    if (clock64() - t0 > kSpinLimit) {
      st->error = 1;
      return false;
    }
  }
}

__device__ __forceinline__ CommHeader* hdr_of(char* base) { return reinterpret_cast<CommHeader*>(base); }
// LL halo planes: four words per shared vertex (x, y, z, unused), double-buffered by sequence parity
__device__ __forceinline__ unsigned long long* ll_plane(char* base, const PeerComm& pc, uint32_t parity) {
  return reinterpret_cast<unsigned long long*>(base + pc.ll_off) + (size_t)parity * pc.recv_stride * 4;
}
__device__ __forceinline__ float4* recv_of(char* base, const PeerComm& pc, uint32_t parity) {
  return reinterpret_cast<float4*>(base + pc.recv_off) + (size_t)parity * pc.recv_stride;
}

// One CTA's share of a halo send (what == 0 of k_halo_send): send[k] = sum of the local partial slots of halo
// vertex k, stored straight into the neighbour's receive plane; the last of the n_ctas CTAs releases the flags.
// partial is read through L2 (__ldcg): it was written by other CTAs, possibly of an overlapped kernel.
// The sender sits on the critical path of the exchange while the memory system is saturated by the tile CTAs
// (loaded round trips of 1-2 us; the first form -- hb -> bptr -> one slot after the other, four vertices per thread
// in sequence -- took 8 us to its stores and 18 us to the flags on C4 at N = 2), so it is written for few DEPENDENT
// round trips: hrange[k] = {first slot, end} is precomputed, one vertex per thread and trip, and the slot loads of
// a vertex are issued eight at a time before the (slot-ordered) adds.
__device__ __forceinline__ void halo_send_cta(const PeerComm& pc, int n, const int2* __restrict__ hrange,
                                              const float4* partial, int cta, int n_ctas,
                                              unsigned long long* trace = nullptr, bool count_pass = false) {
  CommState* st = pc.st;
  if (trace && cta == 0 && threadIdx.x == 0) trace[4] = gtime_ns();
  const uint32_t seq = st->send_seq + 1, parity = seq & 1u;
  unsigned long long* const lplane = pc.left ? ll_plane(pc.left, pc, parity) + (size_t)pc.left_dst_off * 4 : nullptr;
  unsigned long long* const rplane = pc.right ? ll_plane(pc.right, pc, parity) + (size_t)pc.right_dst_off * 4 : nullptr;
  for (int k = cta * blockDim.x + threadIdx.x; k < n; k += n_ctas * blockDim.x) {
    const int2 r = hrange[k];
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i0 = r.x; i0 < r.y; i0 += 8) {
      float4 p[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) p[j] = i0 + j < r.y ? __ldcg(partial + i0 + j) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (i0 + j < r.y) { s.x += p[j].x; s.y += p[j].y; s.z += p[j].z; }
    }
    unsigned long long* const dst = k < pc.n_left ? lplane + (size_t)k * 4 : rplane + (size_t)(k - pc.n_left) * 4;
    ll_store(dst + 0, __float_as_uint(s.x), seq);
    ll_store(dst + 1, __float_as_uint(s.y), seq);
    ll_store(dst + 2, __float_as_uint(s.z), seq);
  }
  __syncthreads();
  if (threadIdx.x == 0) {   // local bookkeeping only: the words carry their own sequence number
    if (trace && cta == 0) trace[5] = gtime_ns();
    const unsigned prev = atomicAdd(&st->pack_count, 1u);
    if (prev == (unsigned)n_ctas - 1) {
      st->pack_count = 0;
      st->send_seq = seq;
      if (count_pass) st->count_passes = st->count_passes + 1;
      if (trace) trace[3] = gtime_ns();
    }
  }
}

// the received sum of shared vertex hk of exchange `seq` (k_boundary): waits for its three words
__device__ __forceinline__ float4 ll_recv_vertex(const PeerComm& pc, uint32_t seq, int hk) {
  uint32_t w[3];
  ll_wait<3>(ll_plane(pc.self, pc, seq & 1u) + (size_t)hk * 4, seq, w, pc.st);
  return make_float4(__uint_as_float(w[0]), __uint_as_float(w[1]), __uint_as_float(w[2]), 0.f);
}

// "One tile kernel" passes: a halo tile's CTA counts itself done once its partial sums are stored (call with the
// whole CTA); the sender CTAs -- all of higher CTA index, so every halo tile is resident or finished when they
// start -- wait for the count of this pass.  No dispatch-order assumption is needed for progress: the senders are
// far fewer than the resident CTA slots, so tiles keep retiring and every halo tile gets a slot.
__device__ __forceinline__ void halo_tile_done(const PeerComm& pc) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(&pc.st->halo_done, 1ull);
  }
}
__device__ __forceinline__ void halo_tiles_wait(const PeerComm& pc, int n_halo_tiles) {
  CommState* st = pc.st;
  if (threadIdx.x == 0) {
    const unsigned long long need = (unsigned long long)(st->count_passes + 1u) * (unsigned long long)n_halo_tiles;
    const long long t0 = clock64();
    while (atomicAdd(&st->halo_done, 0ull) < need) {
      if (clock64() - t0 > kSpinLimit) {
        st->error = 1;
        break;
      }
      __nanosleep(200);
    }
    __threadfence();
  }
  __syncthreads();
}

// Receive side inside a consumer kernel: every CTA waits for both neighbours' flags (begin) and the last CTA to
// finish advances the sequence (end).  Returns the parity of the receive plane to read.
// `need`: this thread will read a received value.  Only CTAs that own a shared vertex wait for the neighbours'
// flags -- the (many) CTAs of ordinary boundary vertices never see the NVLink latency.
// (the LL words are waited for vertex by vertex, ll_recv_vertex: nothing to do here but to read the sequence number)
__device__ __forceinline__ uint32_t halo_recv_begin(const PeerComm& pc, uint32_t& seq) {
  seq = pc.st->recv_seq + 1;
  return seq & 1u;
}
// n_ctas: CTAs of the grid that call this (the boundary kernel's sender CTAs do not)
__device__ __forceinline__ void halo_recv_end(const PeerComm& pc, uint32_t seq, unsigned n_ctas) {
  CommState* st = pc.st;
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(&st->unpack_count, 1u);
    if (prev == n_ctas - 1) {
      st->unpack_count = 0;
      st->recv_seq = seq;
    }
  }
}

// All-sum of up to two doubles over all ranks: this rank's values -> slot[rank] of EVERY rank as LL words (four per
// value pair: lo / hi halves), then every rank sums the slots in rank order (identical bits on all ranks).  Called
// by one full warp (lanes 0..world-1 talk to rank `lane`); results valid in lane 0.  Always consumes one sequence
// number.  Slots are double-buffered by sequence parity.
__device__ __forceinline__ unsigned long long* ll_slot(char* base, const PeerComm& pc, uint32_t parity, int rank) {
  return reinterpret_cast<unsigned long long*>(base + pc.ll_slot_off) + ((size_t)parity * kMaxRanks + rank) * 4;
}
__device__ __forceinline__ void allsum_scatter2(const PeerComm& pc, double v0, double v1, int lane) {
  const uint32_t seq = pc.st->ar_seq + 1, parity = seq & 1u;
  if (lane < pc.world) {
    unsigned long long* slot = ll_slot(pc.all[lane], pc, parity, pc.rank);
    const unsigned long long b0 = (unsigned long long)__double_as_longlong(v0), b1 = (unsigned long long)__double_as_longlong(v1);
    ll_store(slot + 0, (uint32_t)b0, seq);
    ll_store(slot + 1, (uint32_t)(b0 >> 32), seq);
    ll_store(slot + 2, (uint32_t)b1, seq);
    ll_store(slot + 3, (uint32_t)(b1 >> 32), seq);
  }
}
__device__ __forceinline__ void allsum_gather2(const PeerComm& pc, int lane, double& s0, double& s1) {
  CommState* st = pc.st;
  const uint32_t seq = st->ar_seq + 1, parity = seq & 1u;
  double v0 = 0.0, v1 = 0.0;
  if (lane < pc.world) {
    uint32_t w[4] = {0u, 0u, 0u, 0u};
    ll_wait<4>(ll_slot(pc.self, pc, parity, lane), seq, w, st);
    v0 = __longlong_as_double((long long)(((unsigned long long)w[1] << 32) | w[0]));
    v1 = __longlong_as_double((long long)(((unsigned long long)w[3] << 32) | w[2]));
  }
  s0 = 0.0;
  s1 = 0.0;
  for (int i = 0; i < pc.world; ++i) {   // rank order, the same on every rank
    s0 += __shfl_sync(0xffffffffu, v0, i);
    s1 += __shfl_sync(0xffffffffu, v1, i);
  }
  if (lane == 0) st->ar_seq = seq;
}
__device__ __forceinline__ void allsum_scatter(const PeerComm& pc, double v, int lane) { allsum_scatter2(pc, v, 0.0, lane); }
__device__ __forceinline__ double allsum_gather(const PeerComm& pc, int lane) {
  double s0, s1;
  allsum_gather2(pc, lane, s0, s1);
  return s0;
}

}  // namespace femb
