// tiling.h -- host-side element tiling: the HBM layout every kernel streams.
//
// Elements are grouped into spatially compact tiles (recursive coordinate
// bisection of rest centroids) of at most T elements.  A CTA owns one tile:
// it stages the tile's vertices in shared memory, computes one dim x dim
// contribution matrix per element, stages those in shared memory, and gathers
// them per local vertex in a fixed order through a jagged-diagonal (JDS)
// incidence list -- a deterministic replacement for the reference's float
// atomics (explicit.py:42-49).  Vertices touched by a single tile ("interior")
// are renumbered contiguously per tile so their state is read/written
// coalesced and finished inside the tile kernel; vertices shared by several
// tiles ("boundary") get one partial-sum slot per touching tile, laid out
// vertex-major, and are finished by a small boundary kernel.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace femb {

// Capacity of a tile in elements == stride (in floats) of one shared-memory staging plane.
// Compile-time so that every staging access is base register + immediate offset.
#ifndef FEMB_TILE_CAP
#define FEMB_TILE_CAP 768
#endif
constexpr int kStride = FEMB_TILE_CAP;

// Threads per CTA of the tile kernels and the CTAs per SM the register allocation aims at.
#ifndef FEMB_BLOCK
#define FEMB_BLOCK 256
#endif
#ifndef FEMB_MIN_CTAS
#define FEMB_MIN_CTAS 4
#endif
constexpr int kBlock = FEMB_BLOCK;
constexpr int kMinCtas = FEMB_MIN_CTAS;

// JDS entry of (element slot e in tile, corner c): the staging-plane word offset the gather
// reads first.  Corner j>=1 owns planes (j-1)*dim .. +dim-1 (column j-1 of the element's
// contribution matrix), corner 0 owns planes dim*dim .. +dim-1 (minus the column sum).
constexpr int jds_entry(int dim, int e, int c) { return (c == 0 ? dim * dim : (c - 1) * dim) * kStride + e; }

struct TileDesc {       // 12 x int32, mirrored on the device
  int32_t slot_off;     // first element slot of the tile in the element planes
  int32_t n_elem;
  int32_t lv_off;       // first entry in the local-vertex table
  int32_t n_lv;
  int32_t jds_off;      // first JDS entry
  int32_t rounds_off;   // first round count
  int32_t n_rounds;     // = max in-tile valence
  int32_t n_int;        // number of interior local vertices
  int32_t int_base;     // new global id of the tile's first interior vertex
  int32_t pad0, pad1, pad2;
};

struct LocalVert {      // local-vertex table entry
  int32_t gid;          // vertex id in the library's (new) numbering
  int32_t part;         // partial-sum slot for boundary vertices, -1 for interior
};

struct Tiling {
  int dim = 0, nc = 0;  // nc = dim + 1 corners
  int n_vert = 0, n_elem = 0;
  int T = 0;            // requested max elements per tile
  int n_tiles = 0, n_slots = 0;
  int n_halo_tiles = 0;   // multi-GPU: tiles [0, n_halo_tiles) touch a halo vertex
  int n_interior = 0, n_boundary = 0, n_partials = 0;
  int max_tile_elems = 0, max_lv = 0, max_rounds = 0;
  std::vector<int32_t> new_of_old, old_of_new;     // vertex renumbering
  std::vector<int32_t> slot_of_elem, elem_of_slot; // element -> slot (-1 for padding slots)
  std::vector<TileDesc> tiles;
  std::vector<LocalVert> lv;
  std::vector<uint16_t> jds;      // entry = slot_in_tile | corner << 12
  std::vector<uint32_t> rounds;   // per tile, per round k: #local vertices with valence > k
  std::vector<uint32_t> lidx;     // per slot: 2 x uint32 = 4 x uint16 local corner indices
  std::vector<int32_t> bptr;      // [n_boundary+1] partial-slot range of each boundary vertex
  std::vector<int32_t> halo_b;    // boundary index of each halo vertex; its remote slot is bptr[b+1]-1
};

// verts: [n_vert*dim] positions used only for spatial grouping.  Returns an
// empty string on success, else an error message.
std::string build_tiling(int dim, int n_vert, int n_elem, const float* verts,
                         const int32_t* elems, int tile_elems, Tiling& out, int n_halo = 0,
                         const int32_t* halo_verts = nullptr, int wave = 0);

}  // namespace femb
