/*
 * gs_host.c -- plain-C host utilities of libgravsim (no CUDA): seeded initial conditions,
 * binary state files (checkpoint/resume) and final-position dumps in the reference's two text
 * formats.  SURVEY.md section 8(f) rows 1-2.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/gravsim.h"

/* ---- counter-based generator: splitmix64 of (seed, counter) -------------------------------- */
static uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
typedef struct { uint64_t key, ctr; } gs_rng;
static void rng_init(gs_rng *r, uint64_t seed, uint64_t stream) {
    r->key = mix64(seed ^ mix64(stream * 0xD1B54A32D192ED03ull + 1));
    r->ctr = 0;
}
static double rng_u01(gs_rng *r) {   /* [0,1), 53 bits */
    uint64_t x = mix64(r->key + (r->ctr++) * 0x9E3779B97F4A7C15ull);
    return (double)(x >> 11) * (1.0 / 9007199254740992.0);
}

typedef struct { float p[3], v[3], m; } body28;    /* cuda.cu:14-29 */
typedef struct { double p[3], v[3], m; } body56;   /* mpi.c:11-21   */

/* Sun / Earth / Mars, the constants shared by cuda.cu:82-93, mpi.c:76-94, pyspark.py:126-140. */
static void solar3(body28 *b, int n) {
    static const float init[3][7] = {
        {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 1.989e30f},
        {1.496e11f, 0.0f, 0.0f, 0.0f, 29.78e3f, 0.0f, 5.972e24f},
        {2.279e11f, 0.0f, 0.0f, 0.0f, 24.077e3f, 0.0f, 6.39e23f}};
    for (int i = 0; i < n && i < 3; i++) memcpy(&b[i], init[i], sizeof(body28));
}

int gs_ic_reference_uniform_f32(void *particles28, int n, uint64_t seed) {
    if (!particles28 || n <= 0) return GS_EINVAL;
    body28 *b = (body28 *)particles28;
    solar3(b, n);
    for (int i = 3; i < n; i++) {     /* one independent stream per body: order-free, reproducible */
        gs_rng r;
        rng_init(&r, seed, (uint64_t)i);
        for (int c = 0; c < 3; c++) b[i].p[c] = (float)(-3e11 + 6e11 * rng_u01(&r));   /* cuda.cu:129 */
        for (int c = 0; c < 3; c++) b[i].v[c] = (float)(-30e3 + 60e3 * rng_u01(&r));   /* cuda.cu:130 */
        b[i].m = (float)(1e23 + (1e25 - 1e23) * rng_u01(&r));                          /* cuda.cu:131 */
    }
    return GS_OK;
}

int gs_ic_reference_uniform_f64(void *particles56, int n, uint64_t seed) {
    if (!particles56 || n <= 0) return GS_EINVAL;
    body28 *tmp = (body28 *)malloc(sizeof(body28) * (size_t)n);
    if (!tmp) return GS_ENOMEM;
    gs_ic_reference_uniform_f32(tmp, n, seed);
    body56 *b = (body56 *)particles56;
    for (int i = 0; i < n; i++) {
        for (int c = 0; c < 3; c++) { b[i].p[c] = tmp[i].p[c]; b[i].v[c] = tmp[i].v[c]; }
        b[i].m = tmp[i].m;
    }
    free(tmp);
    return GS_OK;
}

static void unit_vector(gs_rng *r, double out[3]) {
    double z = 2.0 * rng_u01(r) - 1.0, phi = 6.283185307179586 * rng_u01(r);
    double s = sqrt(1.0 - z * z);
    out[0] = s * cos(phi); out[1] = s * sin(phi); out[2] = z;
}

/* One Plummer body (G = 1, total mass M, scale radius 1): Aarseth, Henon & Wielen (1974). */
static void plummer_body(gs_rng *r, double M, double rmax, double p[3], double v[3]) {
    double rad;
    do {
        double x1 = rng_u01(r);
        if (x1 < 1e-12) x1 = 1e-12;
        rad = 1.0 / sqrt(pow(x1, -2.0 / 3.0) - 1.0);
    } while (rad > rmax);
    double d[3];
    unit_vector(r, d);
    for (int c = 0; c < 3; c++) p[c] = rad * d[c];
    double q, g;
    do {
        q = rng_u01(r);
        g = 0.1 * rng_u01(r);
    } while (g > q * q * pow(1.0 - q * q, 3.5));
    double vesc = sqrt(2.0 * M) * pow(1.0 + rad * rad, -0.25);
    unit_vector(r, d);
    for (int c = 0; c < 3; c++) v[c] = q * vesc * d[c];
}

static int plummer_fill(body28 *b, int n, uint64_t seed, double rmax, double M, double mass_each,
                        const double cpos[3], const double cvel[3], uint64_t stream0) {
    double *tmp = (double *)malloc(sizeof(double) * 6 * (size_t)n);
    if (!tmp) return GS_ENOMEM;
    double com[6] = {0, 0, 0, 0, 0, 0};
    for (int i = 0; i < n; i++) {
        gs_rng r;
        rng_init(&r, seed, stream0 + (uint64_t)i);
        plummer_body(&r, M, rmax, tmp + 6 * (size_t)i, tmp + 6 * (size_t)i + 3);
        for (int c = 0; c < 6; c++) com[c] += tmp[6 * (size_t)i + c];
    }
    for (int c = 0; c < 6; c++) com[c] /= n;     /* equal masses: centre of mass = mean */
    for (int i = 0; i < n; i++) {
        for (int c = 0; c < 3; c++) {
            b[i].p[c] = (float)(tmp[6 * (size_t)i + c] - com[c] + cpos[c]);
            b[i].v[c] = (float)(tmp[6 * (size_t)i + 3 + c] - com[3 + c] + cvel[c]);
        }
        b[i].m = (float)mass_each;
    }
    free(tmp);
    return GS_OK;
}

int gs_ic_plummer_f32(void *particles28, int n, uint64_t seed, double rmax) {
    if (!particles28 || n <= 0 || !(rmax > 0)) return GS_EINVAL;
    const double zero[3] = {0, 0, 0};
    return plummer_fill((body28 *)particles28, n, seed, rmax, 1.0, 1.0 / n, zero, zero, 0);
}

int gs_ic_two_galaxy_f32(void *particles28, int n, uint64_t seed, double sep, double impact, double vrel) {
    if (!particles28 || n < 2) return GS_EINVAL;
    body28 *b = (body28 *)particles28;
    const int n1 = n / 2, n2 = n - n1;
    const double p1[3] = {-0.5 * sep, -0.5 * impact, 0}, p2[3] = {0.5 * sep, 0.5 * impact, 0};
    const double v1[3] = {0.5 * vrel, 0, 0}, v2[3] = {-0.5 * vrel, 0, 0};
    int rc = plummer_fill(b, n1, seed, 10.0, (double)n1 / n, 1.0 / n, p1, v1, 0);
    if (rc != GS_OK) return rc;
    return plummer_fill(b + n1, n2, seed, 10.0, (double)n2 / n, 1.0 / n, p2, v2, (uint64_t)n1);
}

/* ---- state files ---------------------------------------------------------------------------- */
static const char kMagic[8] = {'G', 'S', 'S', 'T', 'A', 'T', 'E', '1'};

int gs_state_write(const char *path, const gs_state_header *hdr, const void *records) {
    if (!path || !hdr || !records) return GS_EINVAL;
    if (hdr->n <= 0 || (hdr->record_bytes != 28 && hdr->record_bytes != 56)) return GS_EINVAL;
    gs_state_header h = *hdr;
    memcpy(h.magic, kMagic, 8);
    FILE *f = fopen(path, "wb");
    if (!f) return GS_EIO;
    size_t bytes = (size_t)h.n * (size_t)h.record_bytes;
    int ok = fwrite(&h, sizeof(h), 1, f) == 1 && fwrite(records, 1, bytes, f) == bytes;
    ok = (fclose(f) == 0) && ok;
    return ok ? GS_OK : GS_EIO;
}

int gs_state_read_header(const char *path, gs_state_header *hdr) {
    if (!path || !hdr) return GS_EINVAL;
    FILE *f = fopen(path, "rb");
    if (!f) return GS_EIO;
    int ok = fread(hdr, sizeof(*hdr), 1, f) == 1;
    fclose(f);
    if (!ok || memcmp(hdr->magic, kMagic, 8) != 0) return GS_EIO;
    if (hdr->n <= 0 || (hdr->record_bytes != 28 && hdr->record_bytes != 56)) return GS_EIO;
    return GS_OK;
}

int gs_state_read(const char *path, gs_state_header *hdr, void *records, size_t capacity_bytes) {
    gs_state_header h;
    int rc = gs_state_read_header(path, &h);
    if (rc != GS_OK) return rc;
    size_t bytes = (size_t)h.n * (size_t)h.record_bytes;
    if (!records || capacity_bytes < bytes) return GS_EINVAL;
    FILE *f = fopen(path, "rb");
    if (!f) return GS_EIO;
    int ok = fseek(f, (long)sizeof(h), SEEK_SET) == 0 && fread(records, 1, bytes, f) == bytes;
    fclose(f);
    if (!ok) return GS_EIO;
    if (hdr) *hdr = h;
    return GS_OK;
}

/* ---- reference-format dumps ------------------------------------------------------------------ */
int gs_dump_positions_cuda_format(void *stream, const void *particles28, int n, int print_particles) {
    if (!stream || !particles28 || n <= 0) return GS_EINVAL;
    FILE *f = (FILE *)stream;
    const body28 *b = (const body28 *)particles28;
    if (print_particles > n) print_particles = n;   /* the reference always prints 10 (cuda.cu:101) */
    fprintf(f, "\nFinal positions of first %d out of %d particles:\n", print_particles, n);   /* cuda.cu:100 */
    for (int i = 0; i < print_particles; i++)
        fprintf(f, "Particle %d: (%.14f, %.14f, %.14f)\n", i, b[i].p[0], b[i].p[1], b[i].p[2]);   /* cuda.cu:103-108 */
    return GS_OK;
}

int gs_dump_positions_mpi_format(void *stream, const void *particles56, int n) {
    if (!stream || !particles56 || n <= 0) return GS_EINVAL;
    FILE *f = (FILE *)stream;
    const body56 *b = (const body56 *)particles56;
    fprintf(f, "\nFinal positions:\n");                                                     /* mpi.c:249 */
    for (int i = 0; i < n; i++)
        fprintf(f, "Particle %d: (%e, %e, %e)\n", i, b[i].p[0], b[i].p[1], b[i].p[2]);      /* mpi.c:251-256 */
    return GS_OK;
}
