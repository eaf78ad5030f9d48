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

__device__ __forceinline__ CommHeader* hdr_of(char* base) { return reinterpret_cast<CommHeader*>(base); }
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
  float4* const lplane = pc.left ? recv_of(pc.left, pc, parity) + pc.left_dst_off : nullptr;
  float4* const rplane = pc.right ? recv_of(pc.right, pc, parity) + pc.right_dst_off : nullptr;
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
    if (k < pc.n_left) lplane[k] = s;
    else rplane[k - pc.n_left] = s;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (trace && cta == 0) trace[5] = gtime_ns();
    __threadfence_system();
    if (trace && cta == 0) trace[6] = gtime_ns();
    const unsigned prev = atomicAdd(&st->pack_count, 1u);
    if (prev == (unsigned)n_ctas - 1) {
      st->pack_count = 0;
      st->send_seq = seq;
      if (count_pass) st->count_passes = st->count_passes + 1;
      __threadfence_system();
      if (pc.left) st_release_sys(&hdr_of(pc.left)->halo_flag[1], seq);
      if (pc.right) st_release_sys(&hdr_of(pc.right)->halo_flag[0], seq);
      if (trace) trace[3] = gtime_ns();
    }
  }
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
__device__ __forceinline__ uint32_t halo_recv_begin(const PeerComm& pc, uint32_t& seq, bool need = true,
                                                    unsigned long long* trace = nullptr) {
  CommState* st = pc.st;
  seq = st->recv_seq + 1;
  if (__syncthreads_or(need ? 1 : 0)) {
    if (threadIdx.x == 0) {
      if (pc.left) spin_until(&hdr_of(pc.self)->halo_flag[0], seq, st);
      if (pc.right) spin_until(&hdr_of(pc.self)->halo_flag[1], seq, st);
      if (trace) atomicMax(trace + 2, gtime_ns());
    }
    __syncthreads();
  }
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

// this rank's partial `v` -> slot[rank] of every rank, then the flag (lanes 0..world-1 of one warp)
__device__ __forceinline__ void allsum_scatter(const PeerComm& pc, double v, int lane) {
  const uint32_t seq = pc.st->ar_seq + 1, parity = seq & 1u;
  if (lane < pc.world) {
    double* slot = reinterpret_cast<double*>(pc.all[lane] + pc.slot_off) + parity * kMaxRanks + pc.rank;
    *slot = v;
    __threadfence_system();
    st_release_sys(&hdr_of(pc.all[lane])->ar_flag[pc.rank], seq);
  }
}

// wait for every rank's slot and sum them in rank order (identical bits on all ranks).  Called by one full
// warp; the result is valid in lane 0.  Always consumes one sequence number.
__device__ __forceinline__ double allsum_gather(const PeerComm& pc, int lane) {
  CommState* st = pc.st;
  const uint32_t seq = st->ar_seq + 1, parity = seq & 1u;
  if (lane < pc.world) spin_until(&hdr_of(pc.self)->ar_flag[lane], seq, st);
  __syncwarp();
  double s = 0.0;
  if (lane == 0) {
    const double* slot = reinterpret_cast<const double*>(pc.self + pc.slot_off) + parity * kMaxRanks;
    for (int i = 0; i < pc.world; ++i) s += __ldcv(slot + i);
    st->ar_seq = seq;
  }
  return s;
}

}  // namespace femb
