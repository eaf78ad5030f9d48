// kernels_render.cu -- headless CUDA rasteriser for what the reference draws (render/render.py:64-106).
//
// The reference renders through a Taichi GUI (2D) or a GGUI/Vulkan window (3D); neither can exist on this
// box.  The IMAGE is plain arithmetic, so it is produced here by CUDA kernels into a device frame buffer:
//   2D (render2d, render.py:64-75): every triangle filled with (phi + 0.13, 0.13, 0.13), every vertex a
//       dot of radius 2 px colour 0xAAAA00, every block a disc of radius r * width colour 0x343434, drawn in
//       that order (later primitives win), unit square -> width x height pixels, y up.
//   3D (render3d, render.py:86-98): the (0..5)^3 box as lines of colour (0.99, 0.68, 0.28) and the surface
//       mesh as a WIREFRAME (show_wireframe=True), Lambert-shaded per triangle with ambient 0.8 and a point
//       light at (3.5, 3.5, 3.5), through a look-at perspective camera; nearest fragment wins.
// Pixels are 64-bit keys resolved with atomicMax / atomicMin: the result does not depend on thread timing.
#include <stdexcept>

#include "launch.h"

namespace femb {

__device__ __forceinline__ uint32_t pack_rgb(float r, float g, float b) {
  const uint32_t R = (uint32_t)(fminf(fmaxf(r, 0.f), 1.f) * 255.f + 0.5f);
  const uint32_t G = (uint32_t)(fminf(fmaxf(g, 0.f), 1.f) * 255.f + 0.5f);
  const uint32_t B = (uint32_t)(fminf(fmaxf(b, 0.f), 1.f) * 255.f + 0.5f);
  return (R << 16) | (G << 8) | B;
}

__device__ __forceinline__ uint32_t pack_rgb_trunc(float r, float g, float b) {
  const uint32_t R = (uint32_t)fminf(fmaxf(r * 255.f, 0.f), 255.f);
  const uint32_t G = (uint32_t)fminf(fmaxf(g * 255.f, 0.f), 255.f);
  const uint32_t B = (uint32_t)fminf(fmaxf(b * 255.f, 0.f), 255.f);
  return (R << 16) | (G << 8) | B;
}

__global__ void __launch_bounds__(256) k_fb_fill(int n, unsigned long long* fb, unsigned long long v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) fb[i] = v;
}

// ---- 2D: painter's order through the key  (layer << 56) | (primitive << 24) | rgb, atomicMax
__device__ __forceinline__ void put2d(unsigned long long* fb, int W, int H, int px, int py, int layer, uint32_t prim,
                                      uint32_t rgb) {
  if (px < 0 || py < 0 || px >= W || py >= H) return;
  const unsigned long long key = ((unsigned long long)layer << 56) | ((unsigned long long)prim << 24) | rgb;
  atomicMax(&fb[(size_t)py * W + px], key);
}

// one thread per triangle (element, ORIGINAL numbering): fill the pixels whose centre lies inside
__global__ void __launch_bounds__(128) k_raster_tri2d(int n_elem, const int32_t* elems, const int32_t* new_of_old,
                                                      const int32_t* slot_of_elem, const float4* pos, const float* phi,
                                                      unsigned long long* fb, int W, int H) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_elem) return;
  const float4 a = pos[new_of_old[elems[3 * s]]], b = pos[new_of_old[elems[3 * s + 1]]], c = pos[new_of_old[elems[3 * s + 2]]];
  const float ph = phi[slot_of_elem[s]];
  const uint32_t rgb = pack_rgb_trunc(ph + 0.13f, 0.13f, 0.13f);    // ti.rgb_to_hex truncates (render.py:65-71)
  // pixel space: x * W, (1 - y) * H
  const float ax = a.x * W, ay = (1.f - a.y) * H, bx = b.x * W, by = (1.f - b.y) * H, cx = c.x * W, cy = (1.f - c.y) * H;
  const int x0 = max(0, (int)floorf(fminf(ax, fminf(bx, cx)))), x1 = min(W - 1, (int)ceilf(fmaxf(ax, fmaxf(bx, cx))));
  const int y0 = max(0, (int)floorf(fminf(ay, fminf(by, cy)))), y1 = min(H - 1, (int)ceilf(fmaxf(ay, fmaxf(by, cy))));
  const float area = (bx - ax) * (cy - ay) - (by - ay) * (cx - ax);
  if (area == 0.f || x1 - x0 > 4096 || y1 - y0 > 4096) return;
  const float sgn = area > 0.f ? 1.f : -1.f;
  for (int py = y0; py <= y1; ++py)
    for (int px = x0; px <= x1; ++px) {
      const float x = px + 0.5f, y = py + 0.5f;
      const float e_ab = ((bx - ax) * (y - ay) - (by - ay) * (x - ax)) * sgn;
      const float e_bc = ((cx - bx) * (y - by) - (cy - by) * (x - bx)) * sgn;
      const float e_ca = ((ax - cx) * (y - cy) - (ay - cy) * (x - cx)) * sgn;
      if (e_ab >= 0.f && e_bc >= 0.f && e_ca >= 0.f) put2d(fb, W, H, px, py, 1, (uint32_t)s, rgb);
    }
}

// discs: vertices (radius in pixels) and blocks (radius = r * W), one thread per disc
__global__ void __launch_bounds__(128) k_raster_disc2d(int n, const float4* centers, int use_w_radius, float radius_px,
                                                       int layer, uint32_t rgb, unsigned long long* fb, int W, int H) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 c = centers[i];
  const float r = use_w_radius ? c.w * W : radius_px;                // render.py:72-75
  if (!(r > 0.f) || r > 8192.f) return;
  const float cx = c.x * W, cy = (1.f - c.y) * H;
  const int x0 = max(0, (int)floorf(cx - r)), x1 = min(W - 1, (int)ceilf(cx + r));
  const int y0 = max(0, (int)floorf(cy - r)), y1 = min(H - 1, (int)ceilf(cy + r));
  for (int py = y0; py <= y1; ++py)
    for (int px = x0; px <= x1; ++px) {
      const float dx = px + 0.5f - cx, dy = py + 0.5f - cy;
      if (dx * dx + dy * dy <= r * r) put2d(fb, W, H, px, py, layer, (uint32_t)i, rgb);
    }
}

// ---- 3D: depth-tested lines through the key  (depth bits << 32) | rgb, atomicMin
struct Camera {
  float eye[3], fwd[3], right[3], up[3];
  float f;          // focal length in pixels
  float cx, cy;
};

__device__ __forceinline__ bool project(const Camera& cam, float x, float y, float z, float& sx, float& sy, float& depth) {
  const float d[3] = {x - cam.eye[0], y - cam.eye[1], z - cam.eye[2]};
  depth = d[0] * cam.fwd[0] + d[1] * cam.fwd[1] + d[2] * cam.fwd[2];
  if (depth < 0.05f) return false;
  const float u = d[0] * cam.right[0] + d[1] * cam.right[1] + d[2] * cam.right[2];
  const float v = d[0] * cam.up[0] + d[1] * cam.up[1] + d[2] * cam.up[2];
  sx = cam.cx + cam.f * u / depth;
  sy = cam.cy - cam.f * v / depth;
  return true;
}

__device__ void draw_line(const Camera& cam, const float* p, const float* q, uint32_t rgb, int half_width,
                          unsigned long long* fb, int W, int H) {
  float ax, ay, az, bx, by, bz;
  if (!project(cam, p[0], p[1], p[2], ax, ay, az) || !project(cam, q[0], q[1], q[2], bx, by, bz)) return;
  const float len = fmaxf(fabsf(bx - ax), fabsf(by - ay));
  if (!(len < 16384.f)) return;
  const int n = max(1, (int)ceilf(len));
  for (int i = 0; i <= n; ++i) {
    const float t = (float)i / n;
    const int px = (int)floorf(ax + (bx - ax) * t), py = (int)floorf(ay + (by - ay) * t);
    const float depth = 1.f / ((1.f - t) / az + t / bz);           // perspective-correct
    const unsigned long long key = ((unsigned long long)__float_as_uint(depth) << 32) | rgb;
    for (int oy = -half_width; oy <= half_width; ++oy)
      for (int ox = -half_width; ox <= half_width; ++ox) {
        const int x = px + ox, y = py + oy;
        if (x >= 0 && y >= 0 && x < W && y < H) atomicMin(&fb[(size_t)y * W + x], key);
      }
  }
}

// box edges (render.py:47-60): 12 lines, width 2
__global__ void k_raster_box3d(Camera cam, float lo, float hi, unsigned long long* fb, int W, int H) {
  const int i = threadIdx.x;
  if (i >= 12) return;
  const int idx[24] = {0, 1, 0, 2, 1, 3, 2, 3, 4, 5, 4, 6, 5, 7, 6, 7, 0, 4, 1, 5, 2, 6, 3, 7};
  float v[2][3];
  for (int k = 0; k < 2; ++k) {
    const int c = idx[2 * i + k];                                   // bit 0: y, bit 1: x, bit 2: z  (render.py:51-58)
    v[k][0] = (c & 2) ? hi : lo;
    v[k][1] = (c & 1) ? hi : lo;
    v[k][2] = (c & 4) ? hi : lo;
  }
  draw_line(cam, v[0], v[1], pack_rgb(0.99f, 0.68f, 0.28f), 1, fb, W, H);
}

// surface mesh as a wireframe, one thread per triangle, Lambert shading per triangle (ambient 0.8 + point light)
__global__ void __launch_bounds__(128) k_raster_wire3d(Camera cam, int n_faces, const int32_t* faces, const int32_t* new_of_old,
                                                       const float4* pos, unsigned long long* fb, int W, int H) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n_faces) return;
  float p[3][3];
  for (int k = 0; k < 3; ++k) {
    const float4 v = pos[new_of_old[faces[3 * f + k]]];
    p[k][0] = v.x; p[k][1] = v.y; p[k][2] = v.z;
  }
  const float e1[3] = {p[1][0] - p[0][0], p[1][1] - p[0][1], p[1][2] - p[0][2]};
  const float e2[3] = {p[2][0] - p[0][0], p[2][1] - p[0][1], p[2][2] - p[0][2]};
  float n[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
  const float nl = sqrtf(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
  float diff = 0.f;
  if (nl > 0.f) {
    const float c[3] = {(p[0][0] + p[1][0] + p[2][0]) / 3.f, (p[0][1] + p[1][1] + p[2][1]) / 3.f, (p[0][2] + p[1][2] + p[2][2]) / 3.f};
    float l[3] = {3.5f - c[0], 3.5f - c[1], 3.5f - c[2]};             // render.py:94
    const float ll = sqrtf(l[0] * l[0] + l[1] * l[1] + l[2] * l[2]);
    if (ll > 0.f) diff = fmaxf(0.f, (n[0] * l[0] + n[1] * l[1] + n[2] * l[2]) / (nl * ll));
  }
  const float shade = 0.5f * (0.8f + diff);                           // default mesh colour 0.5, ambient 0.8 (render.py:93)
  const uint32_t rgb = pack_rgb(shade, shade, shade);
  draw_line(cam, p[0], p[1], rgb, 0, fb, W, H);
  draw_line(cam, p[1], p[2], rgb, 0, fb, W, H);
  draw_line(cam, p[2], p[0], rgb, 0, fb, W, H);
}

// key buffer -> RGB8; untouched pixels get the background colour
__global__ void __launch_bounds__(256) k_fb_resolve(int n, const unsigned long long* fb, unsigned long long empty,
                                                    uint32_t background, uint8_t* rgb) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long k = fb[i];
  const uint32_t c = k == empty ? background : (uint32_t)(k & 0xffffffu);
  rgb[3 * i + 0] = (uint8_t)(c >> 16);
  rgb[3 * i + 1] = (uint8_t)(c >> 8);
  rgb[3 * i + 2] = (uint8_t)c;
}

static inline int nb(int n, int b) { return (n + b - 1) / b; }

void launch_render2d(const LaunchCfg& c, int n_elem, const int32_t* elems, const int32_t* new_of_old,
                     const int32_t* slot_of_elem, const float4* pos, const float* phi, int n_vert, const float4* blocks,
                     int n_blocks, unsigned long long* fb, uint8_t* rgb, int W, int H) {
  const int n = W * H;
  k_fb_fill<<<nb(n, 256), 256, 0, c.stream>>>(n, fb, 0ull);
  note_launch("fb_fill", c.stream);
  if (n_elem > 0) {
    k_raster_tri2d<<<nb(n_elem, 128), 128, 0, c.stream>>>(n_elem, elems, new_of_old, slot_of_elem, pos, phi, fb, W, H);
    note_launch("raster_tri2d", c.stream);
  }
  k_raster_disc2d<<<nb(n_vert, 128), 128, 0, c.stream>>>(n_vert, pos, 0, 2.f, 2, 0xAAAA00u, fb, W, H);   // render.py:72
  note_launch("raster_dots2d", c.stream);
  if (n_blocks > 0) {
    k_raster_disc2d<<<nb(n_blocks, 128), 128, 0, c.stream>>>(n_blocks, blocks, 1, 0.f, 3, 0x343434u, fb, W, H);  // 73-75
    note_launch("raster_blocks2d", c.stream);
  }
  k_fb_resolve<<<nb(n, 256), 256, 0, c.stream>>>(n, fb, 0ull, 0x000000u, rgb);
  note_launch("fb_resolve", c.stream);
}

void launch_render3d(const LaunchCfg& c, const float* eye, const float* lookat, const float* up, float fov_deg,
                     int n_faces, const int32_t* faces, const int32_t* new_of_old, const float4* pos,
                     unsigned long long* fb, uint8_t* rgb, int W, int H) {
  Camera cam;
  float f[3] = {lookat[0] - eye[0], lookat[1] - eye[1], lookat[2] - eye[2]};
  const float fl = sqrtf(f[0] * f[0] + f[1] * f[1] + f[2] * f[2]);
  if (!(fl > 0.f)) throw std::runtime_error("render: camera position equals its look-at point");
  for (int i = 0; i < 3; ++i) f[i] /= fl;
  float r[3] = {f[1] * up[2] - f[2] * up[1], f[2] * up[0] - f[0] * up[2], f[0] * up[1] - f[1] * up[0]};
  const float rl = sqrtf(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  if (!(rl > 0.f)) throw std::runtime_error("render: camera up vector is parallel to the view direction");
  for (int i = 0; i < 3; ++i) r[i] /= rl;
  const float u[3] = {r[1] * f[2] - r[2] * f[1], r[2] * f[0] - r[0] * f[2], r[0] * f[1] - r[1] * f[0]};
  for (int i = 0; i < 3; ++i) { cam.eye[i] = eye[i]; cam.fwd[i] = f[i]; cam.right[i] = r[i]; cam.up[i] = u[i]; }
  cam.f = 0.5f * H / tanf(0.5f * fov_deg * 3.14159265358979f / 180.f);
  cam.cx = 0.5f * W;
  cam.cy = 0.5f * H;
  const int n = W * H;
  const unsigned long long empty = ~0ull;
  k_fb_fill<<<nb(n, 256), 256, 0, c.stream>>>(n, fb, empty);
  note_launch("fb_fill", c.stream);
  k_raster_box3d<<<1, 32, 0, c.stream>>>(cam, 0.f, 5.f, fb, W, H);
  note_launch("raster_box3d", c.stream);
  if (n_faces > 0) {
    k_raster_wire3d<<<nb(n_faces, 128), 128, 0, c.stream>>>(cam, n_faces, faces, new_of_old, pos, fb, W, H);
    note_launch("raster_wire3d", c.stream);
  }
  k_fb_resolve<<<nb(n, 256), 256, 0, c.stream>>>(n, fb, empty, 0x000000u, rgb);
  note_launch("fb_resolve", c.stream);
}

}  // namespace femb
