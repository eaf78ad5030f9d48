"""CPU restatement of what the reference's Render draws (render/render.py:64-98) -- TEST INFRASTRUCTURE ONLY.

Only tests/ may import this; the product path is csrc/kernels_render.cu.  parity unpinned: the reference
draws through Taichi's GUI / GGUI (Vulkan), which cannot run on this box, so there are no reference pixels to
pin against; this file restates the DRAW LIST of render.py (what is drawn, in which order, with which
colours, camera and lights) with a plain scan-conversion rule:

  2D  pixel (px, py) covers the unit-square point ((px + .5) / W, 1 - (py + .5) / H); a triangle / disc owns
      the pixels whose CENTRE it contains; painter's order = triangles in element order (render.py:70), then
      the vertex dots (render.py:72), then the blocks (render.py:73-75).
  3D  look-at pinhole camera, vertical field of view `fov`; lines are sampled once per pixel of their longer
      screen extent with perspective-correct depth; the nearest sample wins (ties: smaller RGB key).
"""
import numpy as np

F = np.float32


def _rgb_trunc(r, g, b):          # ti.rgb_to_hex (render.py:71): int(x * 255), clipped
    c = [int(min(max(F(x) * F(255.0), F(0.0)), F(255.0))) for x in (r, g, b)]
    return (c[0] << 16) | (c[1] << 8) | c[2]


def _rgb_round(r, g, b):
    c = [int(min(max(F(x), F(0.0)), F(1.0)) * F(255.0) + F(0.5)) for x in (r, g, b)]
    return (c[0] << 16) | (c[1] << 8) | c[2]


def _unpack(img_u32):
    out = np.empty(img_u32.shape + (3,), np.uint8)
    out[..., 0] = (img_u32 >> 16) & 255
    out[..., 1] = (img_u32 >> 8) & 255
    out[..., 2] = img_u32 & 255
    return out


def render2d(pos, elems, phi, centers, radii, W=640, H=640):
    """render.py:64-75 -> uint8 [H, W, 3]."""
    pos = np.asarray(pos, F)
    img = np.zeros((H, W), np.uint32)
    for e, (i, j, k) in enumerate(np.asarray(elems)):
        ax, ay = pos[i, 0] * F(W), (F(1) - pos[i, 1]) * F(H)
        bx, by = pos[j, 0] * F(W), (F(1) - pos[j, 1]) * F(H)
        cx, cy = pos[k, 0] * F(W), (F(1) - pos[k, 1]) * F(H)
        x0, x1 = max(0, int(np.floor(min(ax, bx, cx)))), min(W - 1, int(np.ceil(max(ax, bx, cx))))
        y0, y1 = max(0, int(np.floor(min(ay, by, cy)))), min(H - 1, int(np.ceil(max(ay, by, cy))))
        area = (bx - ax) * (cy - ay) - (by - ay) * (cx - ax)
        if area == 0 or x1 < x0 or y1 < y0:
            continue
        sgn = F(1) if area > 0 else F(-1)
        x = (np.arange(x0, x1 + 1, dtype=F) + F(0.5))[None, :]
        y = (np.arange(y0, y1 + 1, dtype=F) + F(0.5))[:, None]
        e_ab = ((bx - ax) * (y - ay) - (by - ay) * (x - ax)) * sgn
        e_bc = ((cx - bx) * (y - by) - (cy - by) * (x - bx)) * sgn
        e_ca = ((ax - cx) * (y - cy) - (ay - cy) * (x - cx)) * sgn
        inside = (e_ab >= 0) & (e_bc >= 0) & (e_ca >= 0)
        img[y0:y1 + 1, x0:x1 + 1][inside] = _rgb_trunc(F(phi[e]) + F(0.13), 0.13, 0.13)

    def disc(cx, cy, r, colour):
        cx, cy, r = F(cx) * F(W), (F(1) - F(cy)) * F(H), F(r)
        x0, x1 = max(0, int(np.floor(cx - r))), min(W - 1, int(np.ceil(cx + r)))
        y0, y1 = max(0, int(np.floor(cy - r))), min(H - 1, int(np.ceil(cy + r)))
        if x1 < x0 or y1 < y0:
            return
        dx = (np.arange(x0, x1 + 1, dtype=F) + F(0.5) - cx)[None, :]
        dy = (np.arange(y0, y1 + 1, dtype=F) + F(0.5) - cy)[:, None]
        img[y0:y1 + 1, x0:x1 + 1][dx * dx + dy * dy <= r * r] = colour

    for p in pos:
        disc(p[0], p[1], 2.0, 0xAAAA00)                       # render.py:72
    for c, r in zip(np.asarray(centers, F).reshape(-1, 2), np.asarray(radii, F).reshape(-1)):
        disc(c[0], c[1], r * F(W), 0x343434)                  # render.py:73-75
    return _unpack(img)


class Camera:
    def __init__(self, eye, lookat, up, fov_deg, W, H):
        eye, lookat, up = (np.asarray(v, F) for v in (eye, lookat, up))
        f = lookat - eye
        f = f / F(np.sqrt((f * f).sum(dtype=F)))
        r = np.cross(f, up).astype(F)
        r = r / F(np.sqrt((r * r).sum(dtype=F)))
        self.eye, self.fwd, self.right, self.up = eye, f, r, np.cross(r, f).astype(F)
        self.f = F(0.5) * F(H) / F(np.tan(F(0.5) * F(fov_deg) * F(3.14159265358979) / F(180)))
        self.cx, self.cy = F(0.5) * F(W), F(0.5) * F(H)

    def project(self, p):
        d = np.asarray(p, F) - self.eye
        depth = F(d @ self.fwd)
        if depth < F(0.05):
            return None
        return self.cx + self.f * F(d @ self.right) / depth, self.cy - self.f * F(d @ self.up) / depth, depth


def _line(key, cam, p, q, rgb, half, W, H):
    a, b = cam.project(p), cam.project(q)
    if a is None or b is None:
        return
    (ax, ay, az), (bx, by, bz) = a, b
    n = max(1, int(np.ceil(max(abs(bx - ax), abs(by - ay)))))
    t = np.arange(n + 1, dtype=F) / F(n)
    px = np.floor(ax + (bx - ax) * t).astype(np.int64)
    py = np.floor(ay + (by - ay) * t).astype(np.int64)
    depth = (F(1) / ((F(1) - t) / az + t / bz)).astype(F)
    k = (depth.view(np.uint32).astype(np.uint64) << np.uint64(32)) | np.uint64(rgb)
    for oy in range(-half, half + 1):
        for ox in range(-half, half + 1):
            x, y = px + ox, py + oy
            ok = (x >= 0) & (y >= 0) & (x < W) & (y < H)
            np.minimum.at(key, (y[ok], x[ok]), k[ok])


def render3d(pos, faces, eye=(-6.36, 3.49, 2.44), lookat=(-5.40, 3.19, 2.43), up=(0, 1, 0), fov_deg=45.0, W=640, H=640):
    """render.py:86-98 -> uint8 [H, W, 3]."""
    pos = np.asarray(pos, F)
    cam = Camera(eye, lookat, up, fov_deg, W, H)
    empty = np.uint64(0xFFFFFFFFFFFFFFFF)
    key = np.full((H, W), empty, np.uint64)
    box = [[(5.0 if c & 2 else 0.0), (5.0 if c & 1 else 0.0), (5.0 if c & 4 else 0.0)] for c in range(8)]  # render.py:47-58
    idx = [0, 1, 0, 2, 1, 3, 2, 3, 4, 5, 4, 6, 5, 7, 6, 7, 0, 4, 1, 5, 2, 6, 3, 7]                    # render.py:59
    for i in range(12):
        _line(key, cam, box[idx[2 * i]], box[idx[2 * i + 1]], _rgb_round(0.99, 0.68, 0.28), 1, W, H)     # render.py:95
    light = np.array([3.5, 3.5, 3.5], F)                                                            # render.py:94
    for tri in np.asarray(faces).reshape(-1, 3):
        p = pos[tri]
        n = np.cross(p[1] - p[0], p[2] - p[0]).astype(F)
        nl = F(np.sqrt((n * n).sum(dtype=F)))
        diff = F(0)
        if nl > 0:
            l = light - (p[0] + p[1] + p[2]) / F(3)
            ll = F(np.sqrt((l * l).sum(dtype=F)))
            if ll > 0:
                diff = max(F(0), F(n @ l) / (nl * ll))
        shade = F(0.5) * (F(0.8) + diff)                                                            # render.py:93, 98
        rgb = _rgb_round(shade, shade, shade)
        for a, b in ((0, 1), (1, 2), (2, 0)):
            _line(key, cam, p[a], p[b], rgb, 0, W, H)
    img = (key & np.uint64(0xFFFFFF)).astype(np.uint32)
    img[key == empty] = 0
    return _unpack(img)
