/* oracle_matgen.c -- TEST INFRASTRUCTURE.  Restates matgen/random.cc. */
#include "oracle.h"

/* matgen/random.cc:16-40 (full_mul) and :53-75 (philox_2x64): 10 rounds,
 * state = {lo(R*M), hi(R*M) ^ key ^ L}, key bumped by 0xD2B7... after round 0. */
static void philox_2x64(uint64_t c0, uint64_t c1, uint64_t key, uint64_t out[2])
{
    const uint64_t bump = UINT64_C(0xD2B74407B1CE6E93);
    const uint64_t mult = UINT64_C(0x9E3779B97F4A7C15);
    for (int r = 0; r < 10; ++r) {
        if (r != 0) key += bump;
        unsigned __int128 prod = (unsigned __int128)c1 * mult;
        uint64_t lo = (uint64_t)prod, hi = (uint64_t)(prod >> 64);
        uint64_t L = c0;
        c0 = lo;
        c1 = hi ^ key ^ L;
    }
    out[0] = c0;
    out[1] = c1;
}

/* matgen/random.cc:80-89: top 53 bits / 2^53. */
static double to_real(uint64_t bits) { return (double)(bits >> 11) / 9007199254740992.0; }

static void gen(int dist, int64_t seed, int64_t i, int64_t j, double* re, double* im)
{
    uint64_t b[2];
    philox_2x64((uint64_t)i, (uint64_t)j, (uint64_t)seed, b);
    double r1 = to_real(b[0]), r2 = to_real(b[1]);
    if (dist == 2) { r1 = 2 * r1 - 1; r2 = 2 * r2 - 1; }   /* matgen/random.cc:119-122 */
    *re = r1;
    *im = r2;
}

void oracle_matgen_d(int dist, int64_t seed, int64_t m, int64_t n, int64_t ioff, int64_t joff,
                     double* A, int64_t lda)
{
    #pragma omp parallel for schedule(static)
    for (int64_t j = 0; j < n; ++j)
        for (int64_t i = 0; i < m; ++i) {
            double re, im;
            gen(dist, seed, i + ioff, j + joff, &re, &im);
            A[i + j * lda] = re;
        }
}

void oracle_matgen_z(int dist, int64_t seed, int64_t m, int64_t n, int64_t ioff, int64_t joff,
                     double* A, int64_t lda)
{
    #pragma omp parallel for schedule(static)
    for (int64_t j = 0; j < n; ++j)
        for (int64_t i = 0; i < m; ++i) {
            double re, im;
            gen(dist, seed, i + ioff, j + joff, &re, &im);
            A[2 * (i + j * lda)] = re;
            A[2 * (i + j * lda) + 1] = im;
        }
}
