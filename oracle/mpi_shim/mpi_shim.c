/*
 * mpi_shim.c -- TEST INFRASTRUCTURE. fork()-based implementation of the MPI subset in mpi.h.
 *
 *   GS_MPI_NP=<P>        number of ranks (forked in MPI_Init; parent is rank 0).
 *   GS_REF_STATE=<file>  raw array of N x 7 doubles (the reference `Particle`, mpi.c:11-21).
 *                        When set, rank 0 overwrites the buffer of the reference's SECOND
 *                        broadcast (the particle array, mpi.c:182) with the file's contents
 *                        before sending, so every implementation sees the same bodies (the
 *                        reference seeds its own generator from the clock, mpi.c:96).
 *   GS_REF_OUT=<file>    rank 0 writes the last all-gathered particle array (raw doubles) when the
 *                        reference frees its datatype (mpi.c:264), i.e. after the step loop:
 *                        the log file only holds positions to 7 digits (mpi.c:251).
 *   The span between the reference's two MPI_Wtime calls (mpi.c:189,239) is printed by rank 0 as
 *   "GS_REF_WTIME_SPAN <seconds>".
 *
 * Collectives copy through a shared staging area guarded by a process-shared pthread barrier.
 */
#define _GNU_SOURCE
#include "mpi.h"
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

typedef struct {
    pthread_barrier_t barrier;
    size_t staging_bytes;
} shim_shared;

static shim_shared *g_sh = NULL;
static unsigned char *g_staging = NULL;
static int g_rank = 0, g_size = 1;
static pid_t *g_children = NULL;
static size_t g_type_extent[64] = {0, sizeof(int), sizeof(double)};
static int g_ntypes = 3;
static int g_struct_bcasts = 0;
static void *g_last_gather = NULL;
static size_t g_last_gather_bytes = 0;
static int g_wtime_calls = 0;
static double g_wtime_first = 0.0;

static int env_int(const char *k, int dflt) { const char *v = getenv(k); return v ? atoi(v) : dflt; }
int gs_ref_n(void) { return env_int("GS_REF_N", 8); }
int gs_ref_steps(void) { return env_int("GS_REF_STEPS", 500); }
double gs_ref_dt(void) { const char *v = getenv("GS_REF_DT"); return v ? atof(v) : 3600.0; }

static void sync_ranks(void) { if (g_size > 1) pthread_barrier_wait(&g_sh->barrier); }

int MPI_Init(int *argc, char ***argv) {
    (void)argc; (void)argv;
    g_size = env_int("GS_MPI_NP", 1);
    if (g_size < 1) g_size = 1;
    size_t staging = (size_t)gs_ref_n() * 64 + 4096;
    if (staging < (1u << 20)) staging = 1u << 20;
    size_t total = sizeof(shim_shared) + staging;
    void *m = mmap(NULL, total, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    if (m == MAP_FAILED) { perror("mpi_shim mmap"); exit(2); }
    g_sh = (shim_shared *)m;
    g_sh->staging_bytes = staging;
    g_staging = (unsigned char *)m + sizeof(shim_shared);
    pthread_barrierattr_t at;
    pthread_barrierattr_init(&at);
    pthread_barrierattr_setpshared(&at, PTHREAD_PROCESS_SHARED);
    pthread_barrier_init(&g_sh->barrier, &at, (unsigned)g_size);
    fflush(NULL);
    g_children = (pid_t *)calloc((size_t)g_size, sizeof(pid_t));
    for (int r = 1; r < g_size; r++) {
        pid_t pid = fork();
        if (pid < 0) { perror("mpi_shim fork"); exit(2); }
        if (pid == 0) { g_rank = r; free(g_children); g_children = NULL; break; }
        g_children[r] = pid;
    }
    return MPI_SUCCESS;
}

int MPI_Finalize(void) {
    fflush(NULL);
    if (g_rank != 0) _exit(0);
    for (int r = 1; r < g_size; r++) { int st; waitpid(g_children[r], &st, 0); }
    return MPI_SUCCESS;
}

int MPI_Comm_rank(MPI_Comm c, int *rank) { (void)c; *rank = g_rank; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm c, int *size) { (void)c; *size = g_size; return MPI_SUCCESS; }

static void need(size_t bytes) {
    if (bytes > g_sh->staging_bytes) {
        fprintf(stderr, "mpi_shim: message of %zu bytes exceeds staging area (%zu)\n", bytes, g_sh->staging_bytes);
        exit(2);
    }
}

int MPI_Bcast(void *buf, int count, MPI_Datatype type, int root, MPI_Comm comm) {
    (void)comm;
    size_t bytes = (size_t)count * g_type_extent[type];
    if (type >= 3) {
        g_struct_bcasts++;
        const char *state = getenv("GS_REF_STATE");
        if (g_rank == root && state && g_struct_bcasts == 1) {
            FILE *f = fopen(state, "rb");
            if (!f || fread(buf, 1, bytes, f) != bytes) {
                fprintf(stderr, "mpi_shim: cannot read %zu bytes of state from %s\n", bytes, state);
                exit(2);
            }
            fclose(f);
        }
    }
    if (g_size == 1) return MPI_SUCCESS;
    need(bytes);
    if (g_rank == root) memcpy(g_staging, buf, bytes);
    sync_ranks();
    if (g_rank != root) memcpy(buf, g_staging, bytes);
    sync_ranks();
    return MPI_SUCCESS;
}

int MPI_Get_address(const void *loc, MPI_Aint *a) { *a = (MPI_Aint)(size_t)loc; return MPI_SUCCESS; }
MPI_Aint MPI_Aint_diff(MPI_Aint a, MPI_Aint b) { return a - b; }

int MPI_Type_create_struct(int count, const int bl[], const MPI_Aint disp[], const MPI_Datatype types[],
                           MPI_Datatype *newtype) {
    size_t extent = 0;
    for (int i = 0; i < count; i++) {
        size_t end = (size_t)disp[i] + (size_t)bl[i] * g_type_extent[types[i]];
        if (end > extent) extent = end;
    }
    g_type_extent[g_ntypes] = extent;
    *newtype = g_ntypes++;
    return MPI_SUCCESS;
}
int MPI_Type_commit(MPI_Datatype *t) { (void)t; return MPI_SUCCESS; }

int MPI_Type_free(MPI_Datatype *t) {
    (void)t;
    const char *out = getenv("GS_REF_OUT");
    if (g_rank == 0 && out && g_last_gather) {
        FILE *f = fopen(out, "wb");
        if (!f || fwrite(g_last_gather, 1, g_last_gather_bytes, f) != g_last_gather_bytes) {
            fprintf(stderr, "mpi_shim: cannot write %s\n", out);
            exit(2);
        }
        fclose(f);
    }
    return MPI_SUCCESS;
}

int MPI_Allgatherv(const void *sendbuf, int sendcount, MPI_Datatype sendtype, void *recvbuf,
                   const int recvcounts[], const int displs[], MPI_Datatype recvtype, MPI_Comm comm) {
    (void)comm; (void)sendtype;
    size_t ext = g_type_extent[recvtype];
    size_t total = ((size_t)displs[g_size - 1] + (size_t)recvcounts[g_size - 1]) * ext;
    g_last_gather = recvbuf; g_last_gather_bytes = total;
    if (g_size == 1) {
        /* send buffer aliases the receive buffer in the reference (mpi.c:228-229) */
        memmove((unsigned char *)recvbuf + (size_t)displs[0] * ext, sendbuf, (size_t)sendcount * ext);
        return MPI_SUCCESS;
    }
    need(total);
    memcpy(g_staging + (size_t)displs[g_rank] * ext, sendbuf, (size_t)sendcount * ext);
    sync_ranks();
    memcpy(recvbuf, g_staging, total);
    sync_ranks();
    return MPI_SUCCESS;
}

int MPI_Barrier(MPI_Comm comm) { (void)comm; sync_ranks(); return MPI_SUCCESS; }

double MPI_Wtime(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    double t = (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
    g_wtime_calls++;
    if (g_wtime_calls == 1) g_wtime_first = t;
    if (g_wtime_calls == 2 && g_rank == 0) {
        printf("GS_REF_WTIME_SPAN %.9f\n", t - g_wtime_first);
        fflush(stdout);
    }
    return t;
}
