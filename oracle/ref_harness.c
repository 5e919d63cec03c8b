/*
 * ref_harness.c -- TEST INFRASTRUCTURE. Calls the REFERENCE'S OWN calculate_force()
 * (compiled from /root/reference/mpi.c into oracle/_ref/, see oracle/Makefile) on a frozen
 * snapshot, so the restatement in gravity_oracle.c and the CUDA path can be checked against
 * the reference's arithmetic itself.
 *
 * Only the type declarations below are restated (ABI of mpi.c:11-21); every flop executes
 * inside the reference's object code.
 */
#include <stdlib.h>
#include <string.h>

typedef struct { double x, y, z; } Vector3;                              /* ABI of mpi.c:11-15 */
typedef struct { Vector3 position; Vector3 velocity; double mass; } Particle;   /* mpi.c:17-21 */

Vector3 calculate_force(Particle *p1, Particle *p2);                     /* mpi.c:59 */
Vector3 vector3_add(Vector3 a, Vector3 b);                               /* mpi.c:46 */
Vector3 vector3_multiply(Vector3 v, double scalar);                      /* mpi.c:37 */
Vector3 vector3_zero(void);                                              /* mpi.c:23 */

/* a_i = (sum_{j != i} calculate_force(p_i, p_j)) * (1/m_i), the expression at mpi.c:197-206,
 * evaluated for every i on the unmodified snapshot. */
static Vector3 accel_of(Particle *all, int n, int i) {
    Vector3 total = vector3_zero();
    for (int j = 0; j < n; j++) {
        if (i != j) total = vector3_add(total, calculate_force(&all[i], &all[j]));
    }
    return vector3_multiply(total, 1.0 / all[i].mass);
}

void gsref_accel(const double *aos56, int n, double *acc3n) {
    Particle *all = (Particle *)aos56;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < n; i++) {
        Vector3 a = accel_of(all, n, i);
        acc3n[3 * (size_t)i] = a.x; acc3n[3 * (size_t)i + 1] = a.y; acc3n[3 * (size_t)i + 2] = a.z;
    }
}

void gsref_accel_subset(const double *aos56, int n, const int *idx, int nidx, double *acc) {
    Particle *all = (Particle *)aos56;
#pragma omp parallel for schedule(dynamic, 1)
    for (int k = 0; k < nidx; k++) {
        Vector3 a = accel_of(all, n, idx[k]);
        acc[3 * (size_t)k] = a.x; acc[3 * (size_t)k + 1] = a.y; acc[3 * (size_t)k + 2] = a.z;
    }
}

/* Synchronous step built from the reference's own functions: accelerations from the frozen
 * snapshot, then the integrator expressions of mpi.c:207-215 via vector3_add/multiply. */
void gsref_step_sync(double *aos56, int n, double dt) {
    Particle *all = (Particle *)aos56;
    Vector3 *acc = (Vector3 *)malloc(sizeof(Vector3) * (size_t)n);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < n; i++) acc[i] = accel_of(all, n, i);
    for (int i = 0; i < n; i++) {
        all[i].velocity = vector3_add(all[i].velocity, vector3_multiply(acc[i], dt));
        all[i].position = vector3_add(all[i].position, vector3_multiply(all[i].velocity, dt));
    }
    free(acc);
}
