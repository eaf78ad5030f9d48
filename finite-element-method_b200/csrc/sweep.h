// sweep.h -- host-side tables of the reference-order sweep (sweep_core.h): vertex -> (element slot, corner)
// incidences, the corner vertex ids of every slot, and the dependency levels of the rows.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include <vector_types.h>

#include "tiling.h"

namespace femb {

constexpr int kSweepBlock = 256;   // threads per CTA of the level kernels == largest level a single CTA chains

struct SweepHost {
  std::vector<int32_t> vptr;       // [n_vert + 1]
  std::vector<uint32_t> inc;       // slot << 2 | corner, per vertex in slot order
  std::vector<int4> sv;            // [n_slots] corner vertex ids (library numbering); {-1,-1,-1,-1} for padding slots
  std::vector<int32_t> lptr;       // [n_levels + 1]
  std::vector<int32_t> lverts;     // vertices by level (ascending library id inside a level)
  // launch plan: group g covers levels [gptr[g], gptr[g+1]).  A group of several levels has only levels of at most
  // kSweepBlock rows and runs in ONE CTA with a barrier between levels; a single large level runs as a grid.
  std::vector<int32_t> gptr;
  int n_levels = 0;
};

// Levels follow the REFERENCE's row order, i.e. ascending caller-side vertex id (Tiling::old_of_new):
//   level(v) = 1 + max{ level(u) : u star-adjacent to v, old(u) < old(v) }   (0 without such a neighbour).
// Two star-adjacent rows never share a level, every lower-numbered neighbour sits in an earlier level and every
// higher-numbered one in a later level -- running the levels in order reproduces the serial sweep exactly.
std::string build_sweep(const Tiling& t, SweepHost& out);

}  // namespace femb
