/*
 * cuda_ref_preamble.h -- TEST INFRASTRUCTURE. Force-included (nvcc -include) in front of the
 * reference's cuda.cu when oracle/Makefile builds the timing comparator oracle/_ref/cuda_ref.
 * The build-time patch touches three places only: N (cuda.cu:121), the step count (cuda.cu:155)
 * and, after the reference has generated its own random bodies (cuda.cu:127-138), one call that
 * overwrites them with the shared state file so it times the same inputs as everything else.
 * The kernel and host update are untouched.  Its output is non-finite: it is
 * a timing comparator, never a numerical oracle.
 */
#pragma once
#include <cstdio>
#include <cstdlib>
#include <vector>

static inline int gs_ref_n() { const char *v = getenv("GS_REF_N"); return v ? atoi(v) : 50000; }
static inline int gs_ref_steps() { const char *v = getenv("GS_REF_STEPS"); return v ? atoi(v) : 500; }

/* State file: raw N x 7 floats = the reference's 28-byte `Particle` (cuda.cu:14-29). */
template <class P>
static inline void gs_ref_load_state(std::vector<P> &particles) {
    const char *path = getenv("GS_REF_STATE");
    if (!path) return;
    static_assert(sizeof(P) == 28, "reference Particle is 7 floats");
    FILE *f = fopen(path, "rb");
    size_t want = particles.size() * sizeof(P);
    if (!f || fread((void *)particles.data(), 1, want, f) != want) {
        fprintf(stderr, "cuda_ref: cannot read %zu bytes from %s\n", want, path);
        exit(2);
    }
    fclose(f);
}
