/*
 * gravity_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C, fp64, CPU restatement of the reference's direct all-pairs gravity hot path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this library; the product (libgravsim.so) never links, loads or calls it.
 *
 * Parity pin: the reference ships no tests, golden vectors or fixtures, so
 * this restatement is pinned against the reference ITSELF: oracle/Makefile compiles
 * /root/reference/mpi.c where it lies (outputs into oracle/_ref/) and tests/test_oracle.py
 * checks gso_* against the reference's own calculate_force() and its own main() loop,
 * and against fixtures under tests/golden/ that were generated from that compiled reference
 * (tests/golden/make_golden.py).
 *
 * Every function cites the reference file:line it follows.
 *
 * Body layout: array of 7 doubles {px,py,pz,vx,vy,vz,mass} = reference `Particle`
 * (mpi.c:11-21), 56 bytes, no padding.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define GSO_STRIDE 7

/* Pair force on body i from body j, restating calculate_force (mpi.c:59-73):
 *   r = x_j - x_i (mpi.c:61); distance = |r| (mpi.c:62, :55-57);
 *   distance < 1e-10 -> zero (mpi.c:64-66);
 *   force_magnitude = (G*m_i*m_j)/(distance*distance) (mpi.c:68);
 *   scale = force_magnitude/distance (mpi.c:69); force = r*scale (mpi.c:70).
 * Extension (not in the reference): Plummer softening eps. With eps == 0 the code path is
 * exactly the reference's; with eps > 0 distance := sqrt(|r|^2 + eps^2) and the cutoff
 * never triggers. */
static inline void pair_force(const double *pi, const double *pj, double G, double eps,
                              double f[3]) {
    double rx = pj[0] - pi[0], ry = pj[1] - pi[1], rz = pj[2] - pi[2];
    double distance;
    f[0] = f[1] = f[2] = 0.0;
    if (eps == 0.0) {
        distance = sqrt(rx * rx + ry * ry + rz * rz);
        if (distance < 1e-10) return;
    } else {
        distance = sqrt(rx * rx + ry * ry + rz * rz + eps * eps);
    }
    double force_magnitude = (G * pi[6] * pj[6]) / (distance * distance);
    double scale = force_magnitude / distance;
    f[0] = rx * scale; f[1] = ry * scale; f[2] = rz * scale;
}

/* Total force row for body i over all j != i, restating mpi.c:197-204 (same j order). */
static inline void force_row(const double *p, int n, int i, double G, double eps, double F[3]) {
    F[0] = F[1] = F[2] = 0.0;
    for (int j = 0; j < n; j++) {
        if (i != j) {
            double f[3];
            pair_force(p + (size_t)i * GSO_STRIDE, p + (size_t)j * GSO_STRIDE, G, eps, f);
            F[0] += f[0]; F[1] += f[1]; F[2] += f[2];
        }
    }
}

/* Frozen-snapshot accelerations a_i = F_i * (1/m_i)  (mpi.c:197-206), all i. acc is 3n doubles. */
void gso_accel(const double *p, int n, double G, double eps, double *acc) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < n; i++) {
        double F[3];
        force_row(p, n, i, G, eps, F);
        double inv_m = 1.0 / p[(size_t)i * GSO_STRIDE + 6];   /* mpi.c:206 */
        acc[3 * (size_t)i + 0] = F[0] * inv_m;
        acc[3 * (size_t)i + 1] = F[1] * inv_m;
        acc[3 * (size_t)i + 2] = F[2] * inv_m;
    }
}

/* Same, for a sampled subset of i-bodies (full-size checks). acc is 3*nidx doubles. */
void gso_accel_subset(const double *p, int n, double G, double eps,
                      const int *idx, int nidx, double *acc) {
#pragma omp parallel for schedule(dynamic, 1)
    for (int k = 0; k < nidx; k++) {
        double F[3];
        int i = idx[k];
        force_row(p, n, i, G, eps, F);
        double inv_m = 1.0 / p[(size_t)i * GSO_STRIDE + 6];
        acc[3 * (size_t)k + 0] = F[0] * inv_m;
        acc[3 * (size_t)k + 1] = F[1] * inv_m;
        acc[3 * (size_t)k + 2] = F[2] * inv_m;
    }
}

/* Mass-free variant: a_i = sum_j G*m_j*(x_j-x_i)/d^3 (the force law with m_i divided out
 * analytically; equals gso_accel up to fp64 rounding and is defined for m_i == 0 where the
 * reference divides by zero, mpi.c:206). Sum of |a_ij| is also returned (norm3n may be NULL):
 * it is the natural floor for relative errors where |a_i| -> 0. */
void gso_accel_massfree_subset(const double *p, int n, double G, double eps,
                               const int *idx, int nidx, double *acc, double *normsum) {
#pragma omp parallel for schedule(dynamic, 1)
    for (int k = 0; k < nidx; k++) {
        int i = idx ? idx[k] : k;
        const double *pi = p + (size_t)i * GSO_STRIDE;
        double ax = 0, ay = 0, az = 0, ns = 0;
        for (int j = 0; j < n; j++) {
            if (j == i) continue;
            const double *pj = p + (size_t)j * GSO_STRIDE;
            double rx = pj[0] - pi[0], ry = pj[1] - pi[1], rz = pj[2] - pi[2];
            double d2 = rx * rx + ry * ry + rz * rz;
            double d;
            if (eps == 0.0) { d = sqrt(d2); if (d < 1e-10) continue; }
            else d = sqrt(d2 + eps * eps);
            double s = G * pj[6] / (d * d * d);
            ax += s * rx; ay += s * ry; az += s * rz;
            ns += s * sqrt(d2);
        }
        acc[3 * (size_t)k + 0] = ax; acc[3 * (size_t)k + 1] = ay; acc[3 * (size_t)k + 2] = az;
        if (normsum) normsum[k] = ns;
    }
}

/* One SYNCHRONOUS step: every body advanced from the same snapshot, with the reference's
 * integrator lines (mpi.c:206-215): a = F/m; v += a*dt; x += v*dt (new v).
 * Equals mpi.c as written when run with -np N (one body per rank).  This is the parity
 * target of the fused GPU step. */
void gso_step_sync(double *p, int n, double G, double eps, double dt) {
    double *acc = (double *)malloc(sizeof(double) * 3 * (size_t)n);
    gso_accel(p, n, G, eps, acc);
    for (int i = 0; i < n; i++) {
        double *b = p + (size_t)i * GSO_STRIDE;
        for (int c = 0; c < 3; c++) {
            b[3 + c] = b[3 + c] + acc[3 * (size_t)i + c] * dt;   /* mpi.c:207-210 */
            b[c] = b[c] + b[3 + c] * dt;                         /* mpi.c:212-215 */
        }
    }
    free(acc);
}

/* Synchronous kick-drift-kick leapfrog step (extension; not in the reference).
 * v += a(x) dt/2; x += v dt; v += a(x_new) dt/2. */
void gso_step_kdk(double *p, int n, double G, double eps, double dt) {
    double *acc = (double *)malloc(sizeof(double) * 3 * (size_t)n);
    gso_accel(p, n, G, eps, acc);
    for (int i = 0; i < n; i++) {
        double *b = p + (size_t)i * GSO_STRIDE;
        for (int c = 0; c < 3; c++) {
            b[3 + c] += acc[3 * (size_t)i + c] * (0.5 * dt);
            b[c] += b[3 + c] * dt;
        }
    }
    gso_accel(p, n, G, eps, acc);
    for (int i = 0; i < n; i++) {
        double *b = p + (size_t)i * GSO_STRIDE;
        for (int c = 0; c < 3; c++) b[3 + c] += acc[3 * (size_t)i + c] * (0.5 * dt);
    }
    free(acc);
}

/* One step of mpi.c AS WRITTEN with `np` ranks (mpi.c:184-187 partition, :196-216 in-place
 * update inside each rank's private copy, :227-231 all-gather).  Each rank sees its own
 * earlier bodies already advanced and every other rank's bodies at the old snapshot. */
void gso_step_aswritten(double *p, int n, double G, double dt, int np) {
    size_t bytes = sizeof(double) * GSO_STRIDE * (size_t)n;
    double *snapshot = (double *)malloc(bytes);
    double *rankcopy = (double *)malloc(bytes);
    memcpy(snapshot, p, bytes);
    int per = n / np, rem = n % np;                                /* mpi.c:184-185 */
    for (int rank = 0; rank < np; rank++) {
        int start = rank * per + (rank < rem ? rank : rem);        /* mpi.c:186 */
        int local = per + (rank < rem ? 1 : 0);                    /* mpi.c:187 */
        memcpy(rankcopy, snapshot, bytes);
        for (int i = start; i < start + local; i++) {              /* mpi.c:196 */
            double F[3];
            force_row(rankcopy, n, i, G, 0.0, F);                  /* mpi.c:197-204 */
            double *b = rankcopy + (size_t)i * GSO_STRIDE;
            double inv_m = 1.0 / b[6];                             /* mpi.c:206 */
            for (int c = 0; c < 3; c++) {
                b[3 + c] = b[3 + c] + (F[c] * inv_m) * dt;         /* mpi.c:207-210 */
                b[c] = b[c] + b[3 + c] * dt;                       /* mpi.c:212-215 */
            }
        }
        memcpy(p + (size_t)start * GSO_STRIDE, rankcopy + (size_t)start * GSO_STRIDE,
               sizeof(double) * GSO_STRIDE * (size_t)local);       /* mpi.c:227-231 */
    }
    free(snapshot); free(rankcopy);
}

void gso_run_sync(double *p, int n, double G, double eps, double dt, int steps) {
    for (int s = 0; s < steps; s++) gso_step_sync(p, n, G, eps, dt);
}
void gso_run_kdk(double *p, int n, double G, double eps, double dt, int steps) {
    for (int s = 0; s < steps; s++) gso_step_kdk(p, n, G, eps, dt);
}
void gso_run_aswritten(double *p, int n, double G, double dt, int steps, int np) {
    for (int s = 0; s < steps; s++) gso_step_aswritten(p, n, G, dt, np);
}

/* Diagnostics (extension; the reference computes none): kinetic energy, potential energy
 * -sum_{i<j} G m_i m_j / sqrt(r^2+eps^2) (pairs with eps==0 and r<1e-10 skipped, matching the
 * force cutoff), and linear momentum. */
void gso_energy(const double *p, int n, double G, double eps, double *ke, double *pe, double mom[3]) {
    double k = 0, mx = 0, my = 0, mz = 0;
    for (int i = 0; i < n; i++) {
        const double *b = p + (size_t)i * GSO_STRIDE;
        k += 0.5 * b[6] * (b[3] * b[3] + b[4] * b[4] + b[5] * b[5]);
        mx += b[6] * b[3]; my += b[6] * b[4]; mz += b[6] * b[5];
    }
    double u = 0;
#pragma omp parallel for schedule(dynamic, 16) reduction(+ : u)
    for (int i = 0; i < n; i++) {
        const double *pi = p + (size_t)i * GSO_STRIDE;
        double ui = 0;
        for (int j = i + 1; j < n; j++) {
            const double *pj = p + (size_t)j * GSO_STRIDE;
            double rx = pj[0] - pi[0], ry = pj[1] - pi[1], rz = pj[2] - pi[2];
            double d2 = rx * rx + ry * ry + rz * rz;
            double d;
            if (eps == 0.0) { d = sqrt(d2); if (d < 1e-10) continue; }
            else d = sqrt(d2 + eps * eps);
            ui -= G * pi[6] * pj[6] / d;
        }
        u += ui;
    }
    *ke = k; *pe = u; mom[0] = mx; mom[1] = my; mom[2] = mz;
}

/* The reference CUDA path's arithmetic AS WRITTEN, in fp32, with the data race serialised
 * (cuda.cu:32-60): per unordered pair i<j, force_magnitude = (G*m_i*m_j)/(r*r) in float
 * (cuda.cu:41) scattered to forces[3i..] and forces[3j..] (cuda.cu:43-49).  body28 is the
 * 7-float `Particle` (cuda.cu:14-29).  Kept only to document that this path overflows to
 * inf/NaN for solar-system masses; it is never a numerical oracle. */
void gso_cuda_literal_forces_f32(const float *body28, int n, float *forces3n) {
    const float Gf = 6.67430e-11f;                                   /* cuda.cu:11 */
    for (int i = 0; i < 3 * n; i++) forces3n[i] = 0.0f;              /* cuda.cu:152 */
    for (int i = 0; i < n; i++) {
        for (int j = i + 1; j < n; j++) {                            /* cuda.cu:57 */
            const float *a = body28 + 7 * (size_t)i, *b = body28 + 7 * (size_t)j;
            volatile float dx = b[0] - a[0], dy = b[1] - a[1], dz = b[2] - a[2];   /* cuda.cu:33-35 */
            volatile float r = sqrtf(dx * dx + dy * dy + dz * dz);   /* cuda.cu:37 */
            if (r < 1e-10f) continue;                                /* cuda.cu:39 */
            volatile float gm = Gf * a[6];
            volatile float gmm = gm * b[6];
            volatile float fm = gmm / (r * r);                       /* cuda.cu:41 */
            forces3n[3 * i + 0] += fm * dx / r; forces3n[3 * i + 1] += fm * dy / r;
            forces3n[3 * i + 2] += fm * dz / r;                      /* cuda.cu:43-45 */
            forces3n[3 * j + 0] -= fm * dx / r; forces3n[3 * j + 1] -= fm * dy / r;
            forces3n[3 * j + 2] -= fm * dz / r;                      /* cuda.cu:47-49 */
        }
    }
}

/* Host-side update of cuda.cu:63-78 in fp32: a = F/m; v += a*dt; x += v*dt. */
void gso_cuda_literal_update_f32(float *body28, int n, const float *forces3n, float dt) {
    for (int i = 0; i < n; i++) {
        float *p = body28 + 7 * (size_t)i;
        for (int c = 0; c < 3; c++) {
            float a = forces3n[3 * i + c] / p[6];                    /* cuda.cu:70 */
            p[3 + c] += a * dt;                                      /* cuda.cu:74 */
            p[c] += p[3 + c] * dt;                                   /* cuda.cu:75 */
        }
    }
}
