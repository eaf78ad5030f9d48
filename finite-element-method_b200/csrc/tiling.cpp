// tiling.cpp -- builds the tile layout described in tiling.h (host only, OpenMP).
//
// What it replaces: the reference scatters every element's contribution with float atomics into AoS fields
// (explicit.py:42-49, implicit.py:92-105, object.py:358-360).  Here the scatter pattern is resolved ONCE on the
// host into per-tile gather lists, so the device sums in a fixed order without atomics.
#include "tiling.h"

#include <algorithm>
#include <cstdlib>
#include <atomic>
#include <cmath>
#include <cstring>
#include <numeric>

namespace femb {
namespace {

struct Rcb {
  const float* c[3];
  int dim;
  int T;
  int32_t* idx;
  std::vector<char>* is_start;

  // elements [lo, hi) become exactly k tiles of (nearly) equal size
  void run(int lo, int hi, int k) {
    const int n = hi - lo;
    (*is_start)[lo] = 1;
    if (k <= 1) return;
    const int kl = k / 2;
    int m = lo + (int)((int64_t)n * kl / k);
    // longest axis of the centroid bounding box
    float mn[3], mx[3];
    for (int a = 0; a < dim; ++a) { mn[a] = INFINITY; mx[a] = -INFINITY; }
    for (int i = lo; i < hi; ++i)
      for (int a = 0; a < dim; ++a) {
        const float v = c[a][idx[i]];
        mn[a] = std::min(mn[a], v);
        mx[a] = std::max(mx[a], v);
      }
    int axis = 0;
    for (int a = 1; a < dim; ++a)
      if (mx[a] - mn[a] > mx[axis] - mn[axis]) axis = a;
    const float* ca = c[axis];
    auto less = [ca](int32_t x, int32_t y) { return ca[x] < ca[y] || (ca[x] == ca[y] && x < y); };
    std::nth_element(idx + lo, idx + m, idx + hi, less);
    // Snap the cut to a gap between distinct coordinates when the balance
    // allows it: structured meshes then get planar tile faces.
    const float pivot = ca[idx[m]];
    int cl = (int)(std::partition(idx + lo, idx + m, [&](int32_t x) { return ca[x] < pivot; }) - idx);
    int cr = (int)(std::partition(idx + m, idx + hi, [&](int32_t x) { return ca[x] == pivot; }) - idx);
    auto ok = [&](int cut) {
      const int64_t nl = cut - lo, nr = hi - cut;
      return nl > 0 && nr > 0 && nl <= (int64_t)kl * T && nr <= (int64_t)(k - kl) * T;
    };
    int best = -1;
    if (ok(cl)) best = cl;
    if (ok(cr) && (best < 0 || std::abs(cr - m) < std::abs(best - m))) best = cr;
    if (best >= 0) m = best;
    if (n > (1 << 16)) {
#pragma omp task firstprivate(lo, m, kl)
      run(lo, m, kl);
#pragma omp task firstprivate(m, hi, k, kl)
      run(m, hi, k - kl);
#pragma omp taskwait
    } else {
      run(lo, m, kl);
      run(m, hi, k - kl);
    }
  }
};

}  // namespace

// Bank-aware order of a tile's LOCAL VERTICES (default; FEMB_VERTEX_ORDER=id turns it off).  In the element phase the
// 8 threads of a quarter warp read, corner by corner, the float4 of their element's vertex from the staged vertices
// sv[local index] with one 128-bit load: threads that read the same vertex share a wavefront, two DIFFERENT vertices
// whose local indices agree mod 8 conflict (a float4 spans four banks: bank group = index mod 8).  The local order is
// fixed only up to ties -- valence descending, for the jagged-diagonal lists -- so inside every run of equal valence
// the vertices are exchanged until the vertices read together differ mod 8: pairwise-swap descent on
// sum over (quarter warp, corner, residue) of count^2.  Modelled (tools/bank_model.py): 1.39 -> 1.0x wavefronts per
// quarter-warp load on C4.  ev[i * nc + c]: vertex (index into the tile's sorted-unique list) of corner c of the
// tile's i-th element; val: valence by the same index; p: local order -> that index, already sorted by valence.
static void corner_residue_order(int ne, int nc, const std::vector<int32_t>& ev, const std::vector<int32_t>& val,
                                 std::vector<int32_t>& p) {
  const int nl = (int)p.size();
  const int ns = (ne + 7) / 8 * nc;                  // one set per (quarter warp, corner)
  std::vector<int32_t> mem((size_t)ns * 8, -1);      // distinct vertices of a set
  std::vector<uint8_t> nmem(ns, 0), cnt((size_t)ns * 8, 0);
  std::vector<int32_t> sptr(nl + 1, 0), sets;
  for (int i = 0; i < ne; ++i)
    for (int c = 0; c < nc; ++c) {
      const int si = (i / 8) * nc + c, u = ev[(size_t)i * nc + c];
      int32_t* m = mem.data() + (size_t)si * 8;
      int k = 0;
      while (k < nmem[si] && m[k] != u) ++k;
      if (k == nmem[si]) { m[nmem[si]++] = u; sptr[u + 1]++; }
    }
  for (int u = 0; u < nl; ++u) sptr[u + 1] += sptr[u];
  sets.resize(sptr[nl]);
  {
    std::vector<int32_t> cur(sptr.begin(), sptr.end() - 1);
    for (int si = 0; si < ns; ++si)
      for (int k = 0; k < nmem[si]; ++k) sets[cur[mem[(size_t)si * 8 + k]]++] = si;
  }
  std::vector<int32_t> pos(nl);
  for (int l = 0; l < nl; ++l) pos[p[l]] = l;
  for (int si = 0; si < ns; ++si)
    for (int k = 0; k < nmem[si]; ++k) cnt[(size_t)si * 8 + (pos[mem[(size_t)si * 8 + k]] & 7)]++;
  auto conflicted = [&](int u) {
    const int r = pos[u] & 7;
    for (int q = sptr[u]; q < sptr[u + 1]; ++q)
      if (cnt[(size_t)sets[q] * 8 + r] > 1) return true;
    return false;
  };
  // Exchange of u (residue ru) and v (residue rv): every set that holds exactly one of them moves one count from the
  // old residue to the new one, d(objective) = 2 (count[new] - count[old]) + 2; sets that hold both keep their counts.
  std::vector<uint8_t> mark(ns, 0);                  // sets of the current u
  long du[8];                                        // u alone to residue r, over ALL its sets
  auto set_u = [&](int u, int ru, bool on) {
    for (int q = sptr[u]; q < sptr[u + 1]; ++q) mark[sets[q]] = on;
    if (!on) return;
    for (int r = 0; r < 8; ++r) {
      long d = 0;
      for (int q = sptr[u]; q < sptr[u + 1]; ++q)
        d += 2 * ((long)cnt[(size_t)sets[q] * 8 + r] - cnt[(size_t)sets[q] * 8 + ru]) + 2;
      du[r] = d;
    }
  };
  auto delta_pair = [&](int v, int ru, int rv) {     // needs set_u(u, ru, true)
    long d = du[rv];
    for (int q = sptr[v]; q < sptr[v + 1]; ++q) {
      const int si = sets[q];
      if (mark[si]) d -= 2 * ((long)cnt[(size_t)si * 8 + rv] - cnt[(size_t)si * 8 + ru]) + 2;   // shared: u's term out
      else d += 2 * ((long)cnt[(size_t)si * 8 + ru] - cnt[(size_t)si * 8 + rv]) + 2;
    }
    return d;
  };
  auto in_set = [&](int si, int u) {
    const int32_t* m = mem.data() + (size_t)si * 8;
    for (int k = 0; k < nmem[si]; ++k)
      if (m[k] == u) return true;
    return false;
  };
  auto apply = [&](int u, int ru, int rv, int other) {
    for (int q = sptr[u]; q < sptr[u + 1]; ++q) {
      const int si = sets[q];
      if (in_set(si, other)) continue;
      cnt[(size_t)si * 8 + ru]--;
      cnt[(size_t)si * 8 + rv]++;
    }
  };
  for (int pass = 0; pass < 4; ++pass) {
    bool any = false;
    for (int a = 0; a < nl;) {
      int b = a;
      while (b < nl && val[p[b]] == val[p[a]]) ++b;
      for (int i = a; i < b; ++i) {
        if (!conflicted(p[i])) continue;
        set_u(p[i], i & 7, true);
        for (int j = a; j < b; ++j) {
          if (((i ^ j) & 7) == 0) continue;
          const int u = p[i], v = p[j], ri = i & 7, rj = j & 7;
          if (delta_pair(v, ri, rj) >= 0) continue;
          set_u(u, ri, false);
          apply(u, ri, rj, v);
          apply(v, rj, ri, u);
          std::swap(p[i], p[j]);
          pos[u] = j;
          pos[v] = i;
          any = true;
          if (!conflicted(p[i])) break;
          set_u(p[i], i & 7, true);                  // position i now holds v
        }
        set_u(p[i], i & 7, false);
      }
      a = b;
    }
    if (!any) break;
  }
}

// Bank-aware SLOT order (default; FEMB_SLOT_ORDER=id turns it off).  The list order below can only choose AMONG the banks a
// vertex's incident elements happen to have (bank = slot mod 32); this step chooses those banks.  The elements of
// every 8-slot block -- one quarter warp of the element phase, the unit a 128-bit shared-memory load is served in, so
// the vertex reads of the element phase keep their wavefronts, and coalescing is untouched -- are permuted so that for
// every group of 32 local vertices the incident elements spread evenly over the banks: greedy assignment, then
// pairwise-swap descent on  sum over (vertex group, bank) of count^2.
// li[i * nc + c]: sorted-local vertex of corner c of the tile's i-th element.  perm[i] = old position of new slot i.
static void bank_slot_order(int n_elem, int nc, int nl, const std::vector<uint16_t>& li, std::vector<int>& perm) {
  constexpr int B = 8;
  const int ng = (nl + 31) / 32;
  std::vector<int32_t> cnt((size_t)ng * 32, 0);
  std::vector<int> bank(n_elem, -1);
  auto move = [&](int e, int from, int to) {   // returns the change of the objective
    long d = 0;
    for (int c = 0; c < nc; ++c) {
      int32_t* row = cnt.data() + (size_t)(li[(size_t)e * nc + c] / 32) * 32;
      if (from >= 0) { d -= 2 * row[from] - 1; row[from]--; }
      if (to >= 0) { d += 2 * row[to] + 1; row[to]++; }
    }
    return d;
  };
  for (int w0 = 0; w0 < n_elem; w0 += B) {
    const int nb = std::min(B, n_elem - w0);   // the last block may be short
    const int b0 = w0 % 32;
    bool used[B] = {false};
    for (int i = w0; i < w0 + nb; ++i) {
      int best = -1;
      long best_cost = 0;
      for (int b = 0; b < nb; ++b) {
        if (used[b]) continue;
        long cost = 0;
        for (int c = 0; c < nc; ++c) cost += cnt[(size_t)(li[(size_t)i * nc + c] / 32) * 32 + b0 + b];
        if (best < 0 || cost < best_cost) { best = b; best_cost = cost; }
      }
      used[best] = true;
      bank[i] = b0 + best;
      move(i, -1, bank[i]);
    }
  }
  for (int pass = 0; pass < 6; ++pass) {
    bool any = false;
    for (int w0 = 0; w0 < n_elem; w0 += B) {
      const int nb = std::min(B, n_elem - w0);
      for (int i = w0; i < w0 + nb; ++i)
        for (int j = i + 1; j < w0 + nb; ++j) {
          const int bi = bank[i], bj = bank[j];
          const long d = move(i, bi, bj) + move(j, bj, bi);
          if (d < 0) { bank[i] = bj; bank[j] = bi; any = true; }
          else { move(i, bj, bi); move(j, bi, bj); }
        }
    }
    if (!any) break;
  }
  perm.assign(n_elem, 0);
  for (int i = 0; i < n_elem; ++i) perm[i / 32 * 32 + bank[i]] = i;
}

// Bank-aware order of the incidence lists of ONE group of up to 32 local vertices (= one warp of the gather), by
// matching.  lst[l] points at lane l's list (val[l] entries, bank of an entry = entry & 31); on return entry k of
// every list is what the lane reads in round k.  A round is conflict-free when its active lanes (val > k: a prefix,
// the lanes are sorted by valence) read 32 different banks.  Rounds are fixed from the LAST to the first: at round k
// every active lane has exactly k + 1 entries left, lanes that are not active yet still have all of theirs, and a bank
// with d entries left in the group needs d of the remaining k + 1 rounds -- so the banks with the most entries left
// are served first (augmenting paths from the bank side, which never un-match a bank already served), then every
// still unmatched lane gets a bank of its own if one is reachable (augmenting paths from the lane side), and only
// what is left over takes a conflict.  This is the constructive proof of Koenig's edge-colouring theorem, bent to the
// prefix structure of the jagged-diagonal layout.  Measured by tools/bank_model.py on a 48^3-cell box: 1.91 -> 1.1x
// wavefronts per gather load (the greedy order it replaces: FEMB_LIST_ORDER=greedy).
namespace {
struct RoundMatcher {
  uint32_t adj_l[32];      // adj_l[l]: banks lane l has an entry of left
  uint32_t adj_b[32];      // adj_b[b]: active lanes with an entry of bank b left
  int match_l[32], match_b[32];
  uint32_t seen_l = 0, seen_b = 0;
  bool from_bank(int b) {
    uint32_t cand = adj_b[b] & ~seen_l;
    while (cand) {
      const int l = __builtin_ctz(cand);
      seen_l |= 1u << l;
      if (match_l[l] < 0 || from_bank(match_l[l])) { match_l[l] = b; match_b[b] = l; return true; }
      cand = adj_b[b] & ~seen_l;
    }
    return false;
  }
  bool from_lane(int l) {
    uint32_t cand = adj_l[l] & ~seen_b;
    while (cand) {
      const int b = __builtin_ctz(cand);
      seen_b |= 1u << b;
      if (match_b[b] < 0 || from_lane(match_b[b])) { match_l[l] = b; match_b[b] = l; return true; }
      cand = adj_l[l] & ~seen_b;
    }
    return false;
  }
};
}  // namespace

static void order_group_by_matching(int n_lanes, uint16_t* const* lst, const int32_t* val) {
  if (n_lanes <= 0) return;
  int deg[32] = {0};
  for (int l = 0; l < n_lanes; ++l)
    for (int j = 0; j < val[l]; ++j) deg[lst[l][j] & 31]++;
  RoundMatcher m;
  for (int k = val[0] - 1; k >= 0; --k) {
    int n_act = 0;
    while (n_act < n_lanes && val[n_act] > k) ++n_act;
    std::memset(m.adj_l, 0, sizeof(m.adj_l));
    std::memset(m.adj_b, 0, sizeof(m.adj_b));
    for (int l = 0; l < n_act; ++l)
      for (int j = 0; j <= k; ++j) {
        const int b = lst[l][j] & 31;
        m.adj_l[l] |= 1u << b;
        m.adj_b[b] |= 1u << l;
      }
    std::fill(m.match_l, m.match_l + 32, -1);
    std::fill(m.match_b, m.match_b + 32, -1);
    int border[32];
    std::iota(border, border + 32, 0);
    std::stable_sort(border, border + 32, [&](int a, int b) { return deg[a] > deg[b]; });
    int matched = 0;
    for (int q = 0; q < 32 && matched < n_act; ++q) {
      const int b = border[q];
      if (deg[b] == 0) break;
      m.seen_l = 0;
      if (m.from_bank(b)) ++matched;
    }
    for (int l = 0; l < n_act && matched < n_act; ++l) {
      if (m.match_l[l] >= 0) continue;
      m.seen_b = 0;
      if (m.from_lane(l)) ++matched;
    }
    int occ[32] = {0};
    for (int l = 0; l < n_act; ++l)
      if (m.match_l[l] >= 0) occ[m.match_l[l]]++;
    for (int l = 0; l < n_act; ++l) {
      int b = m.match_l[l];
      if (b < 0) {   // no bank of its own this round: the least occupied one, then the one with most entries left
        for (int j = 0; j <= k; ++j) {
          const int c = lst[l][j] & 31;
          if (b < 0 || occ[c] < occ[b] || (occ[c] == occ[b] && deg[c] > deg[b])) b = c;
        }
        occ[b]++;
      }
      int j = 0;
      while ((lst[l][j] & 31) != b) ++j;
      std::swap(lst[l][j], lst[l][k]);
      deg[b]--;
    }
  }
}

std::string build_tiling(int dim, int n_vert, int n_elem, const float* verts,
                         const int32_t* elems, int tile_elems, Tiling& t, int n_halo, const int32_t* halo_verts,
                         int wave) {
  if (dim != 2 && dim != 3) return "dim must be 2 or 3";
  if (n_vert <= 0 || n_elem <= 0) return "empty mesh";
  if (!verts || !elems) return "null mesh arrays";
  if (tile_elems < 32 || tile_elems > kStride) return "tile_elems must be in [32, " + std::to_string(kStride) + "]";
  const int nc = dim + 1;
  for (int64_t i = 0; i < (int64_t)n_elem * nc; ++i)
    if (elems[i] < 0 || elems[i] >= n_vert) return "element index out of range";

  t = Tiling();
  t.dim = dim; t.nc = nc; t.n_vert = n_vert; t.n_elem = n_elem; t.T = tile_elems;

  // 1. spatial key per element = lower corner of its bounding box.  All elements of one
  //    structured cell share the key, so snapped cuts never split a cell.
  std::vector<float> cen[3];
  for (int a = 0; a < dim; ++a) cen[a].resize(n_elem);
#pragma omp parallel for schedule(static)
  for (int e = 0; e < n_elem; ++e)
    for (int a = 0; a < dim; ++a) {
      float s = INFINITY;
      for (int c = 0; c < nc; ++c) s = std::min(s, verts[(int64_t)elems[(int64_t)e * nc + c] * dim + a]);
      cen[a][e] = s;
    }

  // 2. recursive coordinate bisection -> element order + tile starts
  std::vector<int32_t> order(n_elem);
  std::iota(order.begin(), order.end(), 0);
  std::vector<char> is_start(n_elem + 1, 0);
  {
    Rcb r;
    for (int a = 0; a < 3; ++a) r.c[a] = a < dim ? cen[a].data() : nullptr;
    r.dim = dim; r.T = tile_elems; r.idx = order.data(); r.is_start = &is_start;
    // Tile count: ~88 % fill, so that cuts can snap to coordinate gaps (planar tile faces).  When the
    // tile kernels run only a few waves of CTAs (wave = CTAs resident on the device; multi-GPU slabs),
    // round the count up to whole waves of equal tiles: a 6.05-wave launch costs 7 tile times otherwise.
#ifndef FEMB_TILE_SLACK
#define FEMB_TILE_SLACK 64
#endif
    const int Tf = tile_elems - tile_elems / FEMB_TILE_SLACK;
    int k0 = std::max(1, (int)(((int64_t)n_elem + Tf - 1) / Tf));
    if (wave > 0 && k0 >= wave && k0 < 64 * wave) k0 = (k0 + wave - 1) / wave * wave;
    k0 = std::min(k0, n_elem);
    auto rcb = [&](int k) {
      std::iota(order.begin(), order.end(), 0);
      std::fill(is_start.begin(), is_start.end(), 0);
#pragma omp parallel
#pragma omp single
      r.run(0, n_elem, k);
    };
    rcb(k0);
  }
  for (int a = 0; a < dim; ++a) std::vector<float>().swap(cen[a]);
  std::vector<int> tbeg;
  for (int i = 0; i < n_elem; ++i)
    if (is_start[i]) tbeg.push_back(i);
  tbeg.push_back(n_elem);
  const int nt = (int)tbeg.size() - 1;
  t.n_tiles = nt;
  // 2b. multi-GPU: tiles that touch a halo vertex go FIRST (stable), so that the tile kernel can be
  //     launched in two parts and the halo exchange overlaps the interior part.
  t.n_halo_tiles = 0;
  if (n_halo > 0) {
    std::vector<char> is_halo(n_vert, 0);
    for (int k = 0; k < n_halo; ++k) {
      if (halo_verts[k] < 0 || halo_verts[k] >= n_vert) return "halo vertex out of range";
      is_halo[halo_verts[k]] = 1;
    }
    std::vector<char> th(nt, 0);
#pragma omp parallel for schedule(dynamic, 8)
    for (int ti = 0; ti < nt; ++ti) {
      char f = 0;
      for (int i = tbeg[ti]; i < tbeg[ti + 1] && !f; ++i)
        for (int c = 0; c < nc; ++c) f |= is_halo[elems[(int64_t)order[i] * nc + c]];
      th[ti] = f;
    }
    std::vector<int32_t> order2;
    order2.reserve(n_elem);
    std::vector<int> tbeg2;
    for (int pass = 1; pass >= 0; --pass)
      for (int ti = 0; ti < nt; ++ti)
        if (th[ti] == pass) {
          tbeg2.push_back((int)order2.size());
          order2.insert(order2.end(), order.begin() + tbeg[ti], order.begin() + tbeg[ti + 1]);
          t.n_halo_tiles += pass;
        }
    tbeg2.push_back(n_elem);
    order.swap(order2);
    tbeg.swap(tbeg2);
  }
  t.tiles.assign(nt, TileDesc{});

  // 3. per tile: sorted element list, unique vertices, tile-incidence count per vertex
  std::vector<std::vector<int32_t>> tverts(nt);
  std::vector<std::atomic<int32_t>> tcount(n_vert);
  for (auto& a : tcount) a.store(0, std::memory_order_relaxed);
#pragma omp parallel for schedule(dynamic, 8)
  for (int ti = 0; ti < nt; ++ti) {
    std::sort(order.begin() + tbeg[ti], order.begin() + tbeg[ti + 1]);
    const int ne = tbeg[ti + 1] - tbeg[ti];
    std::vector<int32_t>& v = tverts[ti];
    v.resize((size_t)ne * nc);
    for (int i = 0; i < ne; ++i)
      for (int c = 0; c < nc; ++c) v[(size_t)i * nc + c] = elems[(int64_t)order[tbeg[ti] + i] * nc + c];
    std::sort(v.begin(), v.end());
    v.erase(std::unique(v.begin(), v.end()), v.end());
    for (int32_t g : v) tcount[g].fetch_add(1, std::memory_order_relaxed);
  }
  for (int v = 0; v < n_vert; ++v)
    if (tcount[v].load(std::memory_order_relaxed) == 0) return "mesh has a vertex that no element references";
  // halo vertices (shared with another rank) are never finished inside a tile: force them into the
  // boundary class; they also get one extra, LAST partial slot that receives the remote rank's sum
  for (int k = 0; k < n_halo; ++k) {
    if (halo_verts[k] < 0 || halo_verts[k] >= n_vert) return "halo vertex index out of range";
    tcount[halo_verts[k]].fetch_add(1 << 20, std::memory_order_relaxed);
  }

  // 4. per tile: in-tile valence, local order (valence descending; ties: bank-aware, see corner_residue_order)
  const char* vo_env = std::getenv("FEMB_VERTEX_ORDER");   // "id": ties by old id (round-1 tables)
  const bool vertex_order_bank = !(vo_env && std::string(vo_env) == "id");
  std::vector<std::vector<int32_t>> tval(nt);   // valence in local order
  std::vector<std::vector<int32_t>> tperm(nt);  // local order -> position in sorted-unique list
#pragma omp parallel for schedule(dynamic, 8)
  for (int ti = 0; ti < nt; ++ti) {
    const std::vector<int32_t>& v = tverts[ti];
    const int ne = tbeg[ti + 1] - tbeg[ti];
    const int nl = (int)v.size();
    std::vector<int32_t> val(nl, 0);
    for (int i = 0; i < ne; ++i)
      for (int c = 0; c < nc; ++c) {
        const int32_t g = elems[(int64_t)order[tbeg[ti] + i] * nc + c];
        val[std::lower_bound(v.begin(), v.end(), g) - v.begin()]++;
      }
    std::vector<int32_t>& p = tperm[ti];
    p.resize(nl);
    std::iota(p.begin(), p.end(), 0);
    std::stable_sort(p.begin(), p.end(), [&](int32_t a, int32_t b) { return val[a] > val[b]; });
    if (vertex_order_bank) {
      std::vector<int32_t> ev((size_t)ne * nc);
      for (int i = 0; i < ne; ++i)
        for (int c = 0; c < nc; ++c) {
          const int32_t g = elems[(int64_t)order[tbeg[ti] + i] * nc + c];
          ev[(size_t)i * nc + c] = (int32_t)(std::lower_bound(v.begin(), v.end(), g) - v.begin());
        }
      corner_residue_order(ne, nc, ev, val, p);
    }
    tval[ti].resize(nl);
    int n_int = 0;
    for (int l = 0; l < nl; ++l) {
      tval[ti][l] = val[p[l]];
      if (tcount[v[p[l]]].load(std::memory_order_relaxed) == 1) ++n_int;
    }
    TileDesc& d = t.tiles[ti];
    d.n_elem = ne;
    d.n_lv = nl;
    d.n_rounds = nl ? tval[ti][0] : 0;
    d.n_int = n_int;
  }

  // 5. offsets + vertex renumbering (serial, deterministic)
  {
    int64_t slot = 0, lvo = 0, jo = 0, ro = 0;
    int ib = 0;
    for (int ti = 0; ti < nt; ++ti) {
      TileDesc& d = t.tiles[ti];
      if (d.n_lv > 65535) return "tile has more than 65535 local vertices";
      d.slot_off = (int32_t)slot; slot += (d.n_elem + 31) / 32 * 32;
      // every per-tile list starts 16-byte aligned and owns a multiple of 16 bytes: the
      // kernels fetch them with bulk (TMA) copies
      d.lv_off = (int32_t)lvo;   lvo += (d.n_lv + 1) / 2 * 2;
      d.jds_off = (int32_t)jo;   jo += ((int64_t)d.n_elem * nc + 7) / 8 * 8;
      d.rounds_off = (int32_t)ro; ro += (d.n_rounds + 3) / 4 * 4;
      d.int_base = ib;           ib += d.n_int;
      t.max_tile_elems = std::max(t.max_tile_elems, d.n_elem);
      t.max_lv = std::max(t.max_lv, d.n_lv);
      t.max_rounds = std::max(t.max_rounds, d.n_rounds);
      if (slot > INT32_MAX || jo > INT32_MAX) return "mesh too large for 32-bit slot offsets";
    }
    t.n_slots = (int)slot;
    t.n_interior = ib;
    t.n_boundary = n_vert - ib;
    t.lv.resize(lvo);
    t.jds.assign(jo, 0);
    t.rounds.resize(ro);
    t.lidx.assign((size_t)t.n_slots * 2, 0);
    t.slot_of_elem.assign(n_elem, -1);
    t.elem_of_slot.assign(t.n_slots, -1);
  }
  t.new_of_old.assign(n_vert, -1);
  t.old_of_new.assign(n_vert, -1);
  std::vector<int32_t> bcount(t.n_boundary, 0);
  {
    int nb = 0;
    for (int ti = 0; ti < nt; ++ti) {
      const TileDesc& d = t.tiles[ti];
      int ni = 0;
      for (int l = 0; l < d.n_lv; ++l) {
        const int32_t g = tverts[ti][tperm[ti][l]];
        if (tcount[g].load(std::memory_order_relaxed) == 1) {
          t.new_of_old[g] = d.int_base + ni++;
        } else {
          if (t.new_of_old[g] < 0) t.new_of_old[g] = t.n_interior + nb++;
          bcount[t.new_of_old[g] - t.n_interior]++;
        }
      }
    }
    if (nb != t.n_boundary) return "internal error: boundary count mismatch";
  }
  for (int v = 0; v < n_vert; ++v) t.old_of_new[t.new_of_old[v]] = v;
  t.halo_b.resize(n_halo);
  for (int k = 0; k < n_halo; ++k) {
    const int b = t.new_of_old[halo_verts[k]] - t.n_interior;
    if ((tcount[halo_verts[k]].load(std::memory_order_relaxed) >> 20) != 1) return "duplicate halo vertex";
    t.halo_b[k] = b;
    bcount[b]++;   // the remote slot
  }
  t.bptr.assign(t.n_boundary + 1, 0);
  for (int b = 0; b < t.n_boundary; ++b) t.bptr[b + 1] = t.bptr[b] + bcount[b];
  t.n_partials = t.bptr[t.n_boundary];
  {
    std::vector<int32_t> cursor(t.bptr.begin(), t.bptr.end() - 1);
    for (int ti = 0; ti < nt; ++ti) {
      const TileDesc& d = t.tiles[ti];
      for (int l = 0; l < d.n_lv; ++l) {
        const int32_t g = tverts[ti][tperm[ti][l]];
        LocalVert& e = t.lv[(size_t)d.lv_off + l];
        e.gid = t.new_of_old[g];
        e.part = e.gid >= t.n_interior ? cursor[e.gid - t.n_interior]++ : -1;
      }
    }
  }

  // 6. per tile: packed local indices, slots, JDS incidence lists
  const char* so_env = std::getenv("FEMB_SLOT_ORDER");    // "id": elements of a tile in ascending id (round-1 tables)
  const bool slot_order_bank = !(so_env && std::string(so_env) == "id");
  const char* lo_env = std::getenv("FEMB_LIST_ORDER");   // "greedy": the round-1 order of the incidence lists (A/B aid)
  const bool list_order_greedy = lo_env && std::string(lo_env) == "greedy";
#pragma omp parallel for schedule(dynamic, 8)
  for (int ti = 0; ti < nt; ++ti) {
    const TileDesc& d = t.tiles[ti];
    const std::vector<int32_t>& v = tverts[ti];
    const std::vector<int32_t>& p = tperm[ti];
    const int nl = d.n_lv;
    std::vector<int32_t> local_of_sorted(nl);
    for (int l = 0; l < nl; ++l) local_of_sorted[p[l]] = l;
    // incidence lists in local order; entries appended corner-major so that
    // each list is sorted by (corner, slot)
    std::vector<int32_t> start(nl + 1, 0);
    for (int l = 0; l < nl; ++l) start[l + 1] = start[l] + tval[ti][l];
    std::vector<int32_t> fill(start.begin(), start.end() - 1);
    std::vector<uint16_t> inc((size_t)d.n_elem * nc);
    std::vector<uint16_t> li((size_t)d.n_elem * nc);
    if (slot_order_bank) {   // permute the tile's elements inside their 8-slot blocks (each tile owns its segment of `order`)
      for (int i = 0; i < d.n_elem; ++i)
        for (int c = 0; c < nc; ++c) {
          const int32_t g = elems[(int64_t)order[tbeg[ti] + i] * nc + c];
          li[(size_t)i * nc + c] = (uint16_t)local_of_sorted[std::lower_bound(v.begin(), v.end(), g) - v.begin()];
        }
      std::vector<int> perm;
      bank_slot_order(d.n_elem, nc, nl, li, perm);
      std::vector<int32_t> seg(order.begin() + tbeg[ti], order.begin() + tbeg[ti] + d.n_elem);
      for (int i = 0; i < d.n_elem; ++i) order[tbeg[ti] + i] = seg[perm[i]];
    }
    for (int i = 0; i < d.n_elem; ++i) {
      const int32_t e = order[tbeg[ti] + i];
      t.slot_of_elem[e] = d.slot_off + i;
      t.elem_of_slot[d.slot_off + i] = e;
      for (int c = 0; c < nc; ++c) {
        const int32_t g = elems[(int64_t)e * nc + c];
        li[(size_t)i * nc + c] = (uint16_t)local_of_sorted[std::lower_bound(v.begin(), v.end(), g) - v.begin()];
      }
      uint32_t w0 = li[(size_t)i * nc + 0] | ((uint32_t)li[(size_t)i * nc + 1] << 16);
      uint32_t w1 = li[(size_t)i * nc + 2] | ((uint32_t)(nc == 4 ? li[(size_t)i * nc + 3] : 0) << 16);
      t.lidx[(size_t)(d.slot_off + i) * 2 + 0] = w0;
      t.lidx[(size_t)(d.slot_off + i) * 2 + 1] = w1;
    }
    for (int c = 0; c < nc; ++c)
      for (int i = 0; i < d.n_elem; ++i) {
        const int l = li[(size_t)i * nc + c];
        inc[fill[l]++] = (uint16_t)jds_entry(dim, i, c);
      }
    // Bank-aware order of every vertex's list.  In round k the lanes of a warp (32 consecutive local
    // vertices) read staging word (plane * kStride + slot) of their k-th incidence: bank = slot % 32 for
    // every plane.  Pick, round by round, for each lane an unused incidence whose bank is the least
    // occupied in this round -- most rounds become conflict-free (random order: ~3.5 wavefronts per load).
    // Only the (fixed) summation order changes.
    static_assert(kStride % 32 == 0, "staging planes must start on bank 0");
#ifndef FEMB_NO_BANK_ORDER
    for (int w0 = 0; w0 < nl; w0 += 32) {
      const int w1 = std::min(nl, w0 + 32);
      if (!list_order_greedy) {                     // default: rounds by matching (order_group_by_matching)
        uint16_t* lp[32];
        for (int l = w0; l < w1; ++l) lp[l - w0] = inc.data() + start[l];
        order_group_by_matching(w1 - w0, lp, tval[ti].data() + w0);
        continue;
      }
      const int vmax = tval[ti][w0];                // sorted descending: the group's largest valence
      for (int k = 0; k < vmax; ++k) {
        int occ[32] = {0};
        for (int l = w1 - 1; l >= w0; --l) {         // fewest remaining choices first
          const int val = tval[ti][l];
          if (val <= k) continue;
          uint16_t* lst = inc.data() + start[l];
          int best = k, best_occ = occ[lst[k] & 31];
          for (int j = k + 1; j < val && best_occ > 0; ++j) {
            const int oc = occ[lst[j] & 31];
            if (oc < best_occ) { best = j; best_occ = oc; }
          }
          std::swap(lst[k], lst[best]);
          occ[lst[k] & 31]++;
        }
      }
    }
#endif
    // jagged-diagonal layout: round k holds entry k of every vertex with valence > k
    int64_t off = d.jds_off;
    for (int k = 0; k < d.n_rounds; ++k) {
      int cnt = 0;
      while (cnt < nl && tval[ti][cnt] > k) ++cnt;
      t.rounds[(size_t)d.rounds_off + k] = (uint32_t)cnt;
      for (int l = 0; l < cnt; ++l) t.jds[off + l] = inc[start[l] + k];
      off += cnt;
    }
  }
  return std::string();
}

}  // namespace femb
