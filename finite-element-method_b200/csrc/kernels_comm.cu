// kernels_comm.cu -- multi-GPU communication INSIDE our kernels, over NVLink peer memory.
// (New: the reference runs on one device, main.py:40 / utils.py:12-17; SURVEY.md 8e defines the slab partition.)
//
// One process per GPU; every rank owns one symmetric "comm buffer" that all ranks have mapped
// (torch symmetric memory / CUDA IPC on the host side, femb_peer_bind).  No NCCL call sits on the
// data path of a step or a solver iteration:
//   halo sum      the pack kernel stores each shared vertex's local sum straight into the neighbour's
//                 receive plane (remote st over NVLink) and its last CTA releases a sequence flag there;
//                 the neighbour's unpack kernel acquires the flag and reads its own memory.
//   dot products  every rank stores its partial into slot[rank] of EVERY rank and releases a flag; each
//                 rank sums the slots in rank order -> the same bits on all ranks (the device-side
//                 convergence flags can never diverge), one NVLink write latency instead of an allreduce.
// Sequence numbers live in device memory (CommState), so the kernels are CUDA-graph replayable.  Receive
// planes and slots are double-buffered by sequence parity: a rank can run at most one exchange ahead of
// its neighbour (its next phase needs the neighbour's data of this one), so parity never collides.
// Every spin is bounded (kSpinLimit cycles): on time-out CommState.error is set and the kernel goes on,
// so a lost peer cannot hang the GPU; the host turns the flag into an error at the next sync.
#include <stdexcept>

#include "comm.cuh"
#include "launch.h"

namespace femb {

// what == 0: send[k] = sum of the local partial slots of halo vertex k (same order as k_boundary)
// what == 1: send[k] = {lumped mass of vertex k, 0, 0, 0}
// what == 2: as 0, but as LL words for a pass whose boundary kernel receives itself (comm.cuh: halo_send_cta)
__global__ void __launch_bounds__(256) k_halo_send(PeerComm pc, int n, const int32_t* hb, const int32_t* bptr,
                                                   const int2* hrange, const float4* partial, const float4* pos,
                                                   int n_interior, int what) {
  if (what == 2) {
    halo_send_cta(pc, n, hrange, partial, blockIdx.x, gridDim.x);
    return;
  }
  CommState* st = pc.st;
  const uint32_t seq = st->send_seq + 1, parity = seq & 1u;
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) {
    const int b = hb[k];
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (what == 0) {
      for (int i = bptr[b]; i < bptr[b + 1] - 1; ++i) {
        const float4 p = partial[i];
        s.x += p.x; s.y += p.y; s.z += p.z;
      }
    } else {
      s.x = pos[n_interior + b].w;
    }
    // my left plane is the left neighbour's RIGHT plane and vice versa
    if (k < pc.n_left) recv_of(pc.left, pc, parity)[pc.left_dst_off + k] = s;
    else recv_of(pc.right, pc, parity)[pc.right_dst_off + (k - pc.n_left)] = s;
  }
  // release pattern of a grid barrier: CTA barrier, then ONE system-scope fence (cumulative over the stores
  // the barrier ordered before it) -- a fence per thread would wait for an NVLink round trip 256 times
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    const unsigned prev = atomicAdd(&st->pack_count, 1u);
    if (prev == gridDim.x - 1) {            // last CTA: every remote store of this kernel is fenced
      st->pack_count = 0;
      st->send_seq = seq;
      __threadfence_system();
      if (pc.left) st_release_sys(&hdr_of(pc.left)->halo_flag[1], seq);
      if (pc.right) st_release_sys(&hdr_of(pc.right)->halo_flag[0], seq);
    }
  }
}

// what == 0: partial[remote slot of k] = recv[k] ; what == 1: mass of vertex k += recv[k].x
__global__ void __launch_bounds__(256) k_halo_recv(PeerComm pc, int n, const int32_t* hb, const int32_t* bptr,
                                                   float4* partial, float4* pos, int n_interior, int what) {
  CommState* st = pc.st;
  const uint32_t seq = st->recv_seq + 1, parity = seq & 1u;
  if (threadIdx.x == 0) {
    if (pc.left) spin_until(&hdr_of(pc.self)->halo_flag[0], seq, st);
    if (pc.right) spin_until(&hdr_of(pc.self)->halo_flag[1], seq, st);
  }
  __syncthreads();
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) {
    const float4 r = __ldcv(recv_of(pc.self, pc, parity) + k);   // written by the neighbour: never from a stale L1 line
    const int b = hb[k];
    if (what == 0) partial[bptr[b + 1] - 1] = r;
    else pos[n_interior + b].w += r.x;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned prev = atomicAdd(&st->unpack_count, 1u);
    if (prev == gridDim.x - 1) {
      st->unpack_count = 0;
      st->recv_seq = seq;
    }
  }
}

static inline int nblk(int n) { return (n + 255) / 256; }

void launch_halo_send(const LaunchCfg& c, const DevTiling& tl, const PeerComm& pc, Halo h, const float4* partial,
                      const float4* pos, int what, const int2* hrange) {
  if (h.n == 0) return;
  if (what == 2 && !hrange) throw std::runtime_error("halo_send: LL send without slot ranges");
  k_halo_send<<<nblk(h.n), 256, 0, c.stream>>>(pc, h.n, h.b, tl.bptr, hrange, partial, pos, tl.n_interior, what);
  note_launch("halo_send", c.stream);
}
void launch_halo_recv(const LaunchCfg& c, const DevTiling& tl, const PeerComm& pc, Halo h, float4* partial,
                      float4* pos, int what) {
  if (h.n == 0) return;
  k_halo_recv<<<nblk(h.n), 256, 0, c.stream>>>(pc, h.n, h.b, tl.bptr, partial, pos, tl.n_interior, what);
  note_launch("halo_recv", c.stream);
}
size_t comm_layout(int n_halo, size_t* recv_off, size_t* slot_off, size_t* recv_stride) {
  size_t off = sizeof(CommHeader);
  *slot_off = off;
  off += 2 * kMaxRanks * sizeof(double);
  off = (off + 255) / 256 * 256;
  *recv_off = off;
  *recv_stride = ((size_t)n_halo + 15) / 16 * 16;   // float4 elements per parity
  size_t ll_slot, ll;
  comm_ll_offsets(*recv_off, *recv_stride, &ll_slot, &ll);
  return ll + 2 * *recv_stride * 4 * sizeof(unsigned long long);
}
void comm_ll_offsets(size_t recv_off, size_t recv_stride, size_t* ll_slot_off, size_t* ll_off) {
  size_t off = recv_off + 2 * recv_stride * sizeof(float4);
  off = (off + 255) / 256 * 256;
  *ll_slot_off = off;
  off += 2 * kMaxRanks * 4 * sizeof(unsigned long long);   // [parity][rank][4 words]
  off = (off + 255) / 256 * 256;
  *ll_off = off;                                           // [parity][recv_stride][4 words]
}

}  // namespace femb
