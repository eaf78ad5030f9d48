// sweep.cpp -- builds the tables of the reference-order sweep from the element tiling (sweep.h).
#include "sweep.h"

#include <algorithm>

namespace femb {

std::string build_sweep(const Tiling& t, SweepHost& out) {
  const int nc = t.nc, nv = t.n_vert;
  if (nc != 3 && nc != 4) return "build_sweep: tiling has no elements";
  if ((int64_t)t.n_slots >= ((int64_t)1 << 30)) return "build_sweep: too many element slots for 30-bit incidence words";
  out = SweepHost();

  // corner vertex ids of every slot: packed local indices (Tiling::lidx) through the tile's local-vertex table
  int4 none;
  none.x = none.y = none.z = none.w = -1;
  out.sv.assign((size_t)t.n_slots, none);
  out.vptr.assign((size_t)nv + 1, 0);
  for (const TileDesc& d : t.tiles)
    for (int i = 0; i < d.n_elem; ++i) {
      const size_t s = (size_t)d.slot_off + i;
      const uint32_t w0 = t.lidx[2 * s], w1 = t.lidx[2 * s + 1];
      const uint32_t l[4] = {w0 & 0xffffu, w0 >> 16, w1 & 0xffffu, w1 >> 16};
      int32_t g[4] = {-1, -1, -1, -1};
      for (int c = 0; c < nc; ++c) {
        if ((int)l[c] >= d.n_lv) return "build_sweep: local vertex index out of range";
        g[c] = t.lv[(size_t)d.lv_off + l[c]].gid;
        if (g[c] < 0 || g[c] >= nv) return "build_sweep: vertex id out of range";
        out.vptr[(size_t)g[c] + 1]++;
      }
      int4 q;
      q.x = g[0]; q.y = g[1]; q.z = g[2]; q.w = g[3];
      out.sv[s] = q;
    }
  for (int v = 0; v < nv; ++v) {
    if ((int64_t)out.vptr[v] + out.vptr[v + 1] > INT32_MAX) return "build_sweep: too many incidences";
    out.vptr[v + 1] += out.vptr[v];
  }
  out.inc.resize((size_t)out.vptr[nv]);
  {
    std::vector<int32_t> cur(out.vptr.begin(), out.vptr.end() - 1);
    for (const TileDesc& d : t.tiles)
      for (int i = 0; i < d.n_elem; ++i) {
        const size_t s = (size_t)d.slot_off + i;
        const int4 q = out.sv[s];
        const int32_t g[4] = {q.x, q.y, q.z, q.w};
        for (int c = 0; c < nc; ++c) out.inc[(size_t)cur[g[c]]++] = (uint32_t)s << 2 | (uint32_t)c;
      }
  }

  // dependency levels in the reference's row order (ascending caller-side id)
  std::vector<int32_t> level((size_t)nv, 0);
  int n_levels = 0;
  for (int old = 0; old < nv; ++old) {
    const int v = t.new_of_old[old];
    int lv = 0;
    for (int p = out.vptr[v]; p < out.vptr[v + 1]; ++p) {
      const uint32_t w = out.inc[p];
      const int4 q = out.sv[w >> 2];
      const int32_t g[4] = {q.x, q.y, q.z, q.w};
      const int c = (int)(w & 3u);
      // star edges only: corner 0 couples to every other corner, corner j >= 1 to corner 0
      for (int k = (c == 0 ? 1 : 0); k < (c == 0 ? nc : 1); ++k) {
        const int u = g[k];
        if (u != v && t.old_of_new[u] < old) lv = std::max(lv, level[u] + 1);
      }
    }
    level[v] = lv;
    n_levels = std::max(n_levels, lv + 1);
  }
  out.n_levels = n_levels;
  out.lptr.assign((size_t)n_levels + 1, 0);
  for (int v = 0; v < nv; ++v) out.lptr[(size_t)level[v] + 1]++;
  for (int l = 0; l < n_levels; ++l) out.lptr[l + 1] += out.lptr[l];
  out.lverts.resize((size_t)nv);
  {
    std::vector<int32_t> cur(out.lptr.begin(), out.lptr.end() - 1);
    for (int v = 0; v < nv; ++v) out.lverts[(size_t)cur[level[v]]++] = v;
  }

  // launch plan
  out.gptr.clear();
  out.gptr.push_back(0);
  bool open_small = false;   // the group being built is a chain of small levels
  for (int l = 0; l < n_levels; ++l) {
    const bool small = out.lptr[l + 1] - out.lptr[l] <= kSweepBlock;
    if (l > 0 && !(small && open_small)) out.gptr.push_back(l);
    open_small = small;
  }
  out.gptr.push_back(n_levels);
  return std::string();
}

}  // namespace femb
