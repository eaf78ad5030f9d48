/* fem_ref.c -- C/OpenMP restatement of the reference's hot path.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/fem_oracle.py header): used by tests as a
 * fast large-size checker and by bench.py as the timed CPU baseline
 * ("cpu_baseline.kind" = "port": the reference itself is Taichi-JIT Python and
 * taichi is not installable here, so there is nothing to compile from
 * /root/reference).  PARITY UNPINNED BY THE REFERENCE (no golden vectors exist).
 *
 * Same algorithm and parallel structure as the reference's Taichi kernels run
 * with ti.cpu: an element-parallel loop with float atomic adds
 * (solver/explicit.py:8-49), a vertex-parallel integrator
 * (solver/kinematic.py:9-45), and -- for the implicit path -- the matrix-free
 * form of the dense operator of solver/implicit.py:114-194 (the
 * dense Nv x Nv form cannot be instantiated beyond a few thousand vertices)
 * driving the CG loop of implicit.py:311-386.
 */
#include <math.h>
#include <omp.h>
#include <stdint.h>
#include <string.h>

#define MAXD 3

static inline float det_m(int d, const float m[MAXD][MAXD]) {
  if (d == 2) return m[0][0] * m[1][1] - m[0][1] * m[1][0];
  return m[0][0] * (m[1][1] * m[2][2] - m[2][1] * m[1][2]) - m[1][0] * (m[0][1] * m[2][2] - m[2][1] * m[0][2]) +
         m[2][0] * (m[0][1] * m[1][2] - m[1][1] * m[0][2]);
}

/* inv = adj / det (ti.math.inverse for n <= 3) */
static inline void inv_m(int d, const float m[MAXD][MAXD], float out[MAXD][MAXD]) {
  const float id = 1.0f / det_m(d, m);
  if (d == 2) {
    out[0][0] = m[1][1] * id; out[0][1] = -m[0][1] * id;
    out[1][0] = -m[1][0] * id; out[1][1] = m[0][0] * id;
    return;
  }
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
      out[j][i] = id * (m[i1][j1] * m[i2][j2] - m[i2][j1] * m[i1][j2]);
    }
}

static inline void mul_m(int d, const float a[MAXD][MAXD], const float b[MAXD][MAXD], float c[MAXD][MAXD]) {
  for (int i = 0; i < d; ++i)
    for (int j = 0; j < d; ++j) {
      float s = a[i][0] * b[0][j];
      for (int k = 1; k < d; ++k) s += a[i][k] * b[k][j];
      c[i][j] = s;
    }
}

static inline void load_F(int d, const int32_t* el, const float* pos, const float* ref, float R[MAXD][MAXD],
                          float F[MAXD][MAXD]) {
  float X[MAXD][MAXD];
  const float* p0 = pos + (size_t)el[0] * d;
  for (int j = 0; j < d; ++j) {
    const float* pj = pos + (size_t)el[j + 1] * d;
    for (int i = 0; i < d; ++i) X[i][j] = pj[i] - p0[i];
  }
  for (int i = 0; i < d; ++i)
    for (int j = 0; j < d; ++j) R[i][j] = ref[i * d + j];
  mul_m(d, X, R, F);
}

int fem_ref_threads(void) { return omp_get_max_threads(); }

/* solver/explicit.py:8-49 : force[v] += dE/dx, element-parallel with atomics */
void fem_ref_explicit_grad(int d, int64_t ne, const int32_t* elems, const float* pos, const float* ref,
                           const float* vol, float mu, float lam, float* force) {
#pragma omp parallel for schedule(static)
  for (int64_t e = 0; e < ne; ++e) {
    const int32_t* el = elems + e * (d + 1);
    float R[MAXD][MAXD], F[MAXD][MAXD], Fi[MAXD][MAXD], P[MAXD][MAXD], G[MAXD][MAXD];
    load_F(d, el, pos, ref + e * d * d, R, F);
    inv_m(d, F, Fi);
    const float lj = logf(det_m(d, F));
    for (int i = 0; i < d; ++i)
      for (int j = 0; j < d; ++j) P[i][j] = mu * F[i][j] - mu * Fi[j][i] + lam * lj * Fi[j][i];
    for (int i = 0; i < d; ++i)
      for (int j = 0; j < d; ++j) {
        float s = P[i][0] * R[j][0];
        for (int k = 1; k < d; ++k) s += P[i][k] * R[j][k];
        G[i][j] = s * vol[e];
      }
    float f0[MAXD] = {0, 0, 0};
    for (int j = 0; j < d; ++j) {
      float* fv = force + (size_t)el[j + 1] * d;
      for (int i = 0; i < d; ++i) {
#pragma omp atomic
        fv[i] += G[i][j];
        f0[i] -= G[i][j];
      }
    }
    float* fv = force + (size_t)el[0] * d;
    for (int i = 0; i < d; ++i) {
#pragma omp atomic
      fv[i] += f0[i];
    }
  }
}

/* solver/kinematic.py:9-45 (auto_diff == 0 flavour: zeroes force) */
void fem_ref_kinematic(int d, int64_t nv, float* pos, float* vel, float* force, const float* mass, float dt,
                       const float* g_dir, float damping, int n_blocks, const float* centers, const float* radii) {
  const float damp = expf(-dt * damping);
#pragma omp parallel for schedule(static)
  for (int64_t v = 0; v < nv; ++v) {
    float* p = pos + v * d;
    float* u = vel + v * d;
    float* f = force + v * d;
    for (int i = 0; i < d; ++i) {
      u[i] += (9.8f * g_dir[i] - f[i] / mass[v]) * dt;
      u[i] *= damp;
    }
    for (int i = 0; i < d; ++i) {
      if (p[i] < 0.f && u[i] < 0.f) u[i] = 0.f;
      if (p[i] > 1.f && u[i] > 0.f) u[i] = 0.f;
    }
    for (int b = 0; b < n_blocks; ++b) {
      if (radii[b] <= 0.f) continue;
      const float* c = centers + (size_t)b * d;
      float disp[MAXD], nsq = 0.f, vdc = 0.f, vdd = 0.f;
      for (int i = 0; i < d; ++i) {
        disp[i] = p[i] - c[i];
        nsq += disp[i] * disp[i];
        vdc += u[i] * (c[i] - p[i]);
        vdd += u[i] * disp[i];
      }
      if (sqrtf(nsq) < radii[b] && vdc > 0.f)
        for (int i = 0; i < d; ++i) u[i] -= vdd * disp[i] / nsq;
    }
    for (int i = 0; i < d; ++i) {
      p[i] += u[i] * dt;
      f[i] = 0.f;
    }
  }
}

/* n explicit steps: main.py:103-110 with auto_diff == 0 */
void fem_ref_explicit_steps(int d, int64_t nv, int64_t ne, const int32_t* elems, float* pos, float* vel, float* force,
                            const float* mass, const float* ref, const float* vol, float mu, float lam, float dt,
                            const float* g_dir, float damping, int n_blocks, const float* centers, const float* radii,
                            int n_steps) {
  for (int s = 0; s < n_steps; ++s) {
    fem_ref_explicit_grad(d, ne, elems, pos, ref, vol, mu, lam, force);
    fem_ref_kinematic(d, nv, pos, vel, force, mass, dt, g_dir, damping, n_blocks, centers, radii);
  }
}

/* implicit.py:116-147 : the one block H_e per element */
void fem_ref_hessian_blocks(int d, int64_t ne, const int32_t* elems, const float* pos, const float* ref,
                            const float* vol, float mu, float lam, float* H) {
#pragma omp parallel for schedule(static)
  for (int64_t e = 0; e < ne; ++e) {
    float R[MAXD][MAXD], F[MAXD][MAXD], Fi[MAXD][MAXD], FiT[MAXD][MAXD], RT[MAXD][MAXD];
    float FiR[MAXD][MAXD], T1[MAXD][MAXD], M1[MAXD][MAXD], term[MAXD][MAXD], He[MAXD][MAXD];
    load_F(d, elems + e * (d + 1), pos, ref + e * d * d, R, F);
    inv_m(d, F, Fi);
    const float lj = logf(fmaxf(det_m(d, F), 1e-4f));
    for (int i = 0; i < d; ++i)
      for (int j = 0; j < d; ++j) { FiT[i][j] = Fi[j][i]; RT[i][j] = R[j][i]; }
    mul_m(d, Fi, R, FiR);
    float tr = 0.f;
    for (int k = 0; k < d; ++k) tr += FiR[k][k];
    mul_m(d, FiT, RT, T1);
    mul_m(d, T1, FiT, M1);
    for (int i = 0; i < d; ++i)
      for (int j = 0; j < d; ++j)
        term[i][j] = -vol[e] * (mu * R[i][j] + (mu - lam * lj) * M1[i][j] + lam * tr * FiT[i][j]);
    mul_m(d, term, RT, He);
    for (int i = 0; i < d; ++i)
      for (int j = 0; j < d; ++j) H[e * d * d + i * d + j] = He[i][j];
  }
}

/* y = A x (transpose == 0) or A^T x: A = I - dt^2 M^-1 K, K = sum_e L_star (x) H_e */
void fem_ref_apply(int d, int64_t nv, int64_t ne, const int32_t* elems, const float* H, const float* mass, float dt,
                   int transpose, const float* x, float* y, float* scratch) {
  const float dt2 = dt * dt;
  float* in = scratch;         /* nv*d : x or M^-1 x */
  float* kx = scratch + nv * d; /* nv*d : K in */
#pragma omp parallel for schedule(static)
  for (int64_t v = 0; v < nv; ++v)
    for (int i = 0; i < d; ++i) {
      in[v * d + i] = transpose ? x[v * d + i] * (1.0f / mass[v]) : x[v * d + i];
      kx[v * d + i] = 0.f;
    }
#pragma omp parallel for schedule(static)
  for (int64_t e = 0; e < ne; ++e) {
    const int32_t* el = elems + e * (d + 1);
    const float* He = H + e * d * d;
    float s[MAXD] = {0, 0, 0};
    for (int j = 0; j < d; ++j) {
      float ej[MAXD], he[MAXD];
      for (int i = 0; i < d; ++i) ej[i] = in[(size_t)el[j + 1] * d + i] - in[(size_t)el[0] * d + i];
      for (int i = 0; i < d; ++i) {
        float a = 0.f;
        for (int k = 0; k < d; ++k) a += (transpose ? He[k * d + i] : He[i * d + k]) * ej[k];
        he[i] = a;
        s[i] += a;
      }
      for (int i = 0; i < d; ++i) {
#pragma omp atomic
        kx[(size_t)el[j + 1] * d + i] += he[i];
      }
    }
    for (int i = 0; i < d; ++i) {
#pragma omp atomic
      kx[(size_t)el[0] * d + i] -= s[i];
    }
  }
#pragma omp parallel for schedule(static)
  for (int64_t v = 0; v < nv; ++v)
    for (int i = 0; i < d; ++i)
      y[v * d + i] = x[v * d + i] - (transpose ? dt2 : dt2 * (1.0f / mass[v])) * kx[v * d + i];
}

static double dot_v(int64_t n, const float* a, const float* b) {
  double s = 0.0;
#pragma omp parallel for reduction(+ : s) schedule(static)
  for (int64_t i = 0; i < n; ++i) s += (double)a[i] * b[i];
  return s;
}

/* implicit.py:311-386 with matrix-free products.  work: 8*nv*d floats.  Returns iterations. */
int fem_ref_cg(int d, int64_t nv, int64_t ne, const int32_t* elems, const float* H, const float* mass, float dt,
               int preconditioned, const float* b, float* x, int iter_max, float tol, float* work, float* delta_out) {
  const int64_t n = nv * d;
  float *r = work, *dd = work + n, *q = work + 2 * n, *rhs = work + 3 * n, *tmp = work + 4 * n, *scr = work + 5 * n;
  if (preconditioned) fem_ref_apply(d, nv, ne, elems, H, mass, dt, 1, b, rhs, scr);
  else memcpy(rhs, b, n * sizeof(float));
  memcpy(x, b, n * sizeof(float));
  if (preconditioned) {
    fem_ref_apply(d, nv, ne, elems, H, mass, dt, 0, x, tmp, scr);
    fem_ref_apply(d, nv, ne, elems, H, mass, dt, 1, tmp, q, scr);
  } else {
    fem_ref_apply(d, nv, ne, elems, H, mass, dt, 0, x, q, scr);
  }
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) { r[i] = rhs[i] - q[i]; dd[i] = r[i]; }
  float delta_new = (float)dot_v(n, r, r);
  int it = 0;
  while (it < iter_max && delta_new > tol) {
    if (preconditioned) {
      fem_ref_apply(d, nv, ne, elems, H, mass, dt, 0, dd, tmp, scr);
      fem_ref_apply(d, nv, ne, elems, H, mass, dt, 1, tmp, q, scr);
    } else {
      fem_ref_apply(d, nv, ne, elems, H, mass, dt, 0, dd, q, scr);
    }
    const float alpha = delta_new / (float)dot_v(n, dd, q);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) { x[i] += alpha * dd[i]; r[i] -= alpha * q[i]; }
    const float delta_old = delta_new;
    delta_new = (float)dot_v(n, r, r);
    const float beta = delta_new / delta_old;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) dd[i] = r[i] + beta * dd[i];
    ++it;
  }
  if (delta_out) *delta_out = delta_new;
  return it;
}

/* bench.py sets the thread count itself: launchers such as torchrun export OMP_NUM_THREADS=1 */
void fem_ref_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }

/* object.py:345-362, 375-392 : Dm^-1, volume, lumped mass (atomic adds) */
void fem_ref_elements_init(int d, int64_t nv, int64_t ne, const int32_t* elems, const float* pos, float rho,
                           float* ref, float* vol, float* mass) {
  memset(mass, 0, (size_t)nv * sizeof(float));
  const float share = (float)(1.0 / (d + 1));
#pragma omp parallel for schedule(static)
  for (int64_t e = 0; e < ne; ++e) {
    const int32_t* el = elems + e * (d + 1);
    float X[MAXD][MAXD], R[MAXD][MAXD];
    const float* p0 = pos + (size_t)el[0] * d;
    for (int j = 0; j < d; ++j) {
      const float* pj = pos + (size_t)el[j + 1] * d;
      for (int i = 0; i < d; ++i) X[i][j] = pj[i] - p0[i];
    }
    float V;
    if (d == 2) {
      V = fabsf(X[0][0] * X[1][1] - X[1][0] * X[0][1]) / 2.0f;
    } else {
      const float cx = X[1][1] * X[2][2] - X[2][1] * X[1][2];
      const float cy = X[2][1] * X[0][2] - X[0][1] * X[2][2];
      const float cz = X[0][1] * X[1][2] - X[1][1] * X[0][2];
      V = (float)(1.0 / 6.0) * fabsf(X[0][0] * cx + X[1][0] * cy + X[2][0] * cz);
    }
    vol[e] = V;
    const float m = share * V * rho;
    for (int j = 0; j <= d; ++j) {
#pragma omp atomic
      mass[el[j]] += m;
    }
    inv_m(d, X, R);
    for (int i = 0; i < d; ++i)
      for (int j = 0; j < d; ++j) ref[(size_t)e * d * d + i * d + j] = R[i][j];
  }
}

/* diag(A)_{v,k} = 1 - dt^2/m_v * sum_e c_{e,v} H_e[k][k]  (c = d for local vertex 0, else 1) */
void fem_ref_diag(int d, int64_t nv, int64_t ne, const int32_t* elems, const float* H, const float* mass, float dt,
                  float* diag) {
  memset(diag, 0, (size_t)nv * d * sizeof(float));
#pragma omp parallel for schedule(static)
  for (int64_t e = 0; e < ne; ++e) {
    const int32_t* el = elems + e * (d + 1);
    for (int j = 0; j <= d; ++j)
      for (int k = 0; k < d; ++k) {
        const float c = (j == 0 ? (float)d : 1.0f) * H[(size_t)e * d * d + k * d + k];
#pragma omp atomic
        diag[(size_t)el[j] * d + k] += c;
      }
  }
#pragma omp parallel for schedule(static)
  for (int64_t v = 0; v < nv; ++v)
    for (int k = 0; k < d; ++k) diag[v * d + k] = 1.0f - dt * dt * (1.0f / mass[v]) * diag[v * d + k];
}

/* Weighted Jacobi iterations with the matrix-free operator (the loop of implicit.py:227-262 with the sweep of
 * 391-404 evaluated simultaneously -- a timing baseline for "one operator pass + update + residual norm" per
 * iteration; the reference's own sweep is serial and dense).  work: 4*nv*d floats.  Returns iterations. */
int fem_ref_jacobi(int d, int64_t nv, int64_t ne, const int32_t* elems, const float* H, const float* mass, float dt,
                   const float* b, float* x, float* past_x, int iter_max, float tol, float* work, float* err_out) {
  const int64_t n = nv * d;
  float *ax = work, *diag = work + n, *scr = work + 2 * n;
  fem_ref_diag(d, nv, ne, elems, H, mass, dt, diag);
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) x[i] = 0.5f * b[i];
  fem_ref_apply(d, nv, ne, elems, H, mass, dt, 0, x, ax, scr);
  double s = 0.0;
#pragma omp parallel for reduction(+ : s) schedule(static)
  for (int64_t i = 0; i < n; ++i) s += (double)(b[i] - ax[i]) * (b[i] - ax[i]);
  float err = (float)sqrt(s), p_err = err;
  int it = 0;
  while (err > tol && it < iter_max) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
      const float a = diag[i];
      x[i] = fabsf(a) < 1e-6f ? 0.f : 0.75f * (b[i] - ax[i] + a * x[i]) / a + 0.25f * past_x[i];
    }
    fem_ref_apply(d, nv, ne, elems, H, mass, dt, 0, x, ax, scr);
    s = 0.0;
#pragma omp parallel for reduction(+ : s) schedule(static)
    for (int64_t i = 0; i < n; ++i) s += (double)(b[i] - ax[i]) * (b[i] - ax[i]);
    err = (float)sqrt(s);
    ++it;
    if (tol > 0.f && err >= p_err) { memcpy(x, past_x, n * sizeof(float)); break; }
    p_err = err;
    memcpy(past_x, x, n * sizeof(float));
  }
  if (err_out) *err_out = err;
  return it;
}
