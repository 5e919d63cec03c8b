/*
 * gravsim -- plain-C host driver over the libgravsim C-ABI.
 *
 * Takes the place of the two reference drivers (cuda.cu:120-178, mpi.c:140-269): what they
 * hardcode (N, steps, dt: cuda.cu:121-123,155; mpi.c:146-148) is the command line here, the
 * "Step k" progress lines and the final-position dump keep the reference's formats.
 *
 *   gravsim N STEPS DT [--ic uniform|plummer|galaxies] [--seed S] [--eps E] [--G g]
 *           [--integrator euler|leapfrog] [--gpus P] [--in state.gs] [--out state.gs]
 *           [--print K] [--format cuda|mpi] [--energy] [--variant V] [--quiet]
 *
 * --gpus P runs one host thread per GPU, each with its own gs_ctx (rank = thread index), wired
 * together with one NCCL id: the same layout as `mpirun -np P` of mpi.c.
 */
#define _POSIX_C_SOURCE 200809L
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../../include/gravsim.h"

typedef struct {
    int n, steps, gpus, print_k, want_energy, quiet, variant, integrator;
    double dt, eps, G;
    const char *ic, *in_path, *out_path, *format;
    uint64_t seed;
} options;

typedef struct {
    const options *opt;
    int rank;
    const void *nccl_id;
    float *bodies;          /* full array, shared read-only for upload */
    float *result;          /* rank 0 only */
    double e0[5], e1[5];    /* ke, pe, mom */
    double loop_seconds;
    int rc;
    pthread_barrier_t *bar;
} worker;

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

#define CHECK(call)                                                                         \
    do {                                                                                    \
        int rc__ = (call);                                                                  \
        if (rc__ != GS_OK) {                                                                \
            fprintf(stderr, "gravsim: %s failed: %s (rank %d)\n", #call, gs_strerror(rc__), w->rank); \
            w->rc = rc__;                                                                   \
            return NULL;                                                                    \
        }                                                                                   \
    } while (0)

static void *run_rank(void *arg) {
    worker *w = (worker *)arg;
    const options *o = w->opt;
    gs_config cfg;
    gs_config_init(&cfg);
    cfg.n = o->n; cfg.dt = o->dt; cfg.G = o->G; cfg.eps = o->eps;
    cfg.integrator = o->integrator; cfg.variant = o->variant;
    cfg.device = w->rank; cfg.rank = w->rank; cfg.world_size = o->gpus; cfg.nccl_id = w->nccl_id;
    gs_ctx *ctx = NULL;
    CHECK(gs_create(&ctx, &cfg));
    CHECK(gs_upload_aos_f32(ctx, w->bodies, o->n));
    if (o->want_energy) CHECK(gs_energy(ctx, &w->e0[0], &w->e0[1], &w->e0[2]));
    CHECK(gs_sync(ctx));
    if (w->bar) pthread_barrier_wait(w->bar);
    const double t0 = now_s();
    for (int done = 0; done < o->steps;) {          /* "Step k" every 100 steps: cuda.cu:164, mpi.c:192 */
        if (w->rank == 0 && !o->quiet) printf("Step %d/%d\n", done, o->steps);
        int chunk = o->steps - done < 100 ? o->steps - done : 100;
        CHECK(gs_step(ctx, chunk));
        done += chunk;
    }
    CHECK(gs_sync(ctx));
    if (w->bar) pthread_barrier_wait(w->bar);
    w->loop_seconds = now_s() - t0;
    if (o->want_energy) CHECK(gs_energy(ctx, &w->e1[0], &w->e1[1], &w->e1[2]));
    float *dst = w->rank == 0 ? w->result : (float *)malloc(sizeof(float) * 7 * (size_t)o->n);
    if (!dst) { w->rc = GS_ENOMEM; return NULL; }
    CHECK(gs_download_aos_f32(ctx, dst, o->n));      /* collective: every rank takes part */
    if (w->rank != 0) free(dst);
    gs_destroy(ctx);
    return NULL;
}

static int usage(void) {
    fprintf(stderr,
            "usage: gravsim N STEPS DT [--ic uniform|plummer|galaxies] [--seed S] [--eps E] [--G g]\n"
            "               [--integrator euler|leapfrog] [--gpus P] [--in F] [--out F] [--print K]\n"
            "               [--format cuda|mpi] [--energy] [--variant V] [--quiet]\n");
    return 2;
}

int main(int argc, char **argv) {
    if (argc < 4) return usage();
    options o;
    memset(&o, 0, sizeof(o));
    o.n = atoi(argv[1]); o.steps = atoi(argv[2]); o.dt = atof(argv[3]);
    o.gpus = 1; o.print_k = 10; o.G = -1.0; o.ic = "uniform"; o.format = "cuda"; o.seed = 1234;
    for (int a = 4; a < argc; a++) {
        const char *k = argv[a];
        const char *v = (a + 1 < argc) ? argv[a + 1] : NULL;
        if (!strcmp(k, "--energy")) { o.want_energy = 1; continue; }
        if (!strcmp(k, "--quiet")) { o.quiet = 1; continue; }
        if (!v) return usage();
        if (!strcmp(k, "--ic")) o.ic = v;
        else if (!strcmp(k, "--seed")) o.seed = strtoull(v, NULL, 10);
        else if (!strcmp(k, "--eps")) o.eps = atof(v);
        else if (!strcmp(k, "--G")) o.G = atof(v);
        else if (!strcmp(k, "--integrator")) o.integrator = !strcmp(v, "leapfrog") ? GS_INTEGRATOR_LEAPFROG : GS_INTEGRATOR_EULER;
        else if (!strcmp(k, "--gpus")) o.gpus = atoi(v);
        else if (!strcmp(k, "--in")) o.in_path = v;
        else if (!strcmp(k, "--out")) o.out_path = v;
        else if (!strcmp(k, "--print")) o.print_k = atoi(v);
        else if (!strcmp(k, "--format")) o.format = v;
        else if (!strcmp(k, "--variant")) o.variant = atoi(v);
        else return usage();
        a++;
    }
    if (o.n <= 0 || o.steps < 0 || o.gpus < 1) return usage();
    const int nbody_units = !strcmp(o.ic, "plummer") || !strcmp(o.ic, "galaxies");
    if (o.G < 0) o.G = nbody_units ? 1.0 : GS_G_SI;

    float *bodies = (float *)malloc(sizeof(float) * 7 * (size_t)o.n);
    float *result = (float *)malloc(sizeof(float) * 7 * (size_t)o.n);
    if (!bodies || !result) { fprintf(stderr, "gravsim: out of host memory\n"); return 1; }
    int rc = GS_OK;
    int64_t step0 = 0;          /* resume: the clock of the input state carries over into the output */
    double time0 = 0.0;
    if (o.in_path) {
        gs_state_header h;
        rc = gs_state_read_header(o.in_path, &h);
        if (rc == GS_OK && (h.n != o.n || h.record_bytes != 28)) rc = GS_EINVAL;
        if (rc == GS_OK) rc = gs_state_read(o.in_path, &h, bodies, sizeof(float) * 7 * (size_t)o.n);
        if (rc == GS_OK) { step0 = h.step; time0 = h.time; }
    } else if (!strcmp(o.ic, "plummer")) rc = gs_ic_plummer_f32(bodies, o.n, o.seed, 10.0);
    else if (!strcmp(o.ic, "galaxies")) rc = gs_ic_two_galaxy_f32(bodies, o.n, o.seed, 8.0, 2.0, 0.5);
    else rc = gs_ic_reference_uniform_f32(bodies, o.n, o.seed);
    if (rc != GS_OK) { fprintf(stderr, "gravsim: initial state: %s\n", gs_strerror(rc)); return 1; }

    if (!o.quiet) printf("Starting gravity simulation\n");                     /* cuda.cu:143 */
    unsigned char nccl_id[GS_NCCL_ID_BYTES];
    if (o.gpus > 1 && (rc = gs_nccl_unique_id(nccl_id)) != GS_OK) {
        fprintf(stderr, "gravsim: %s\n", gs_strerror(rc));
        return 1;
    }
    worker *ws = (worker *)calloc((size_t)o.gpus, sizeof(worker));
    pthread_t *th = (pthread_t *)calloc((size_t)o.gpus, sizeof(pthread_t));
    pthread_barrier_t bar;
    if (o.gpus > 1) pthread_barrier_init(&bar, NULL, (unsigned)o.gpus);
    for (int r = 0; r < o.gpus; r++) {
        ws[r].opt = &o; ws[r].rank = r; ws[r].nccl_id = o.gpus > 1 ? nccl_id : NULL;
        ws[r].bodies = bodies; ws[r].result = result; ws[r].bar = o.gpus > 1 ? &bar : NULL;
        if (r > 0) pthread_create(&th[r], NULL, run_rank, &ws[r]);
    }
    run_rank(&ws[0]);
    for (int r = 1; r < o.gpus; r++) pthread_join(th[r], NULL);
    for (int r = 0; r < o.gpus; r++)
        if (ws[r].rc != GS_OK) return 1;

    const double secs = ws[0].loop_seconds;
    if (!o.quiet) {
        printf("Simulation took %f seconds\n", secs);                          /* cuda.cu:171 */
        printf("Average time per step: %.6f seconds\n", o.steps ? secs / o.steps : 0.0);   /* mpi.c:247 */
        printf("Interactions per second: %.4e (N=%d, %d GPU%s)\n",
               secs > 0 ? (double)o.n * ((double)o.n - 1.0) * o.steps / secs : 0.0, o.n, o.gpus, o.gpus > 1 ? "s" : "");
    }
    if (o.want_energy) {
        const double E0 = ws[0].e0[0] + ws[0].e0[1], E1 = ws[0].e1[0] + ws[0].e1[1];
        printf("Energy: initial %.10e final %.10e relative drift %.3e\n", E0, E1, E0 != 0 ? (E1 - E0) / (E0 < 0 ? -E0 : E0) : 0.0);
    }
    if (!strcmp(o.format, "mpi")) {
        double *d = (double *)malloc(sizeof(double) * 7 * (size_t)o.n);
        for (size_t i = 0; i < (size_t)o.n * 7; i++) d[i] = result[i];
        int np = o.print_k < o.n ? o.print_k : o.n;
        gs_dump_positions_mpi_format(stdout, d, np);
        free(d);
    } else {
        gs_dump_positions_cuda_format(stdout, result, o.n, o.print_k);
    }
    if (o.out_path) {
        gs_state_header h;
        memset(&h, 0, sizeof(h));
        h.n = o.n; h.record_bytes = 28; h.step = step0 + o.steps; h.time = time0 + o.steps * o.dt; h.dt = o.dt; h.G = o.G; h.eps = o.eps;
        if ((rc = gs_state_write(o.out_path, &h, result)) != GS_OK) {
            fprintf(stderr, "gravsim: writing %s: %s\n", o.out_path, gs_strerror(rc));
            return 1;
        }
    }
    if (!o.quiet) printf("Simulation completed successfully\n");              /* cuda.cu:175, mpi.c:259 */
    free(bodies); free(result); free(ws); free(th);
    return 0;
}
