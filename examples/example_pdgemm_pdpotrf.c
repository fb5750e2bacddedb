/* example_pdgemm_pdpotrf.c -- an unmodified-ScaLAPACK-style caller on the B200 path.
 *
 * The program below is what a ScaLAPACK user already has: local 2D block-cyclic arrays,
 * 9-integer descriptors, Fortran-convention calls to pdgemm_ and pdpotrf_.  Linked against
 * libslate_b200_scalapack.so (in front of, or instead of, libscalapack) those two calls run on
 * the GPU; everything else would fall through to ScaLAPACK.  Counterpart in the reference:
 * scalapack_api/example_pdgetrf.c.
 *
 * One rank per GPU.  Without BLACS (this container has no MPI) the process grid is declared
 * with slate_scalapack_grid_set(); with BLACS linked that call is not needed.  Multi-rank runs
 * pass the NCCL id from rank 0 (te_comm_unique_id + MPI_Bcast) to sb_initialize.
 *
 * Build (tests/test_abi.py compiles and links exactly this, no GPU needed for that):
 *   gcc -Iinclude examples/example_pdgemm_pdpotrf.c -Lslate_d35_b200/lib \
 *       -lslate_b200_scalapack -lslate_b200 -ltile_engine -Wl,-rpath,$PWD/slate_d35_b200/lib -lm
 */
#include <stdio.h>
#include <stdlib.h>

#include "slate_b200_scalapack.h"

/* host library entry points used for start-up (slate_d35_b200/host/src/c_api.cc) */
int sb_initialize(int device, int rank, int size, const void* nccl_id);
int sb_finalize(void);
const char* sb_last_error(void);

/* the Fortran names, as a ScaLAPACK header would declare them */
void pdgemm_(const char* transa, const char* transb, int* m, int* n, int* k, double* alpha, double* a,
             int* ia, int* ja, int* desca, double* b, int* ib, int* jb, int* descb, double* beta,
             double* c, int* ic, int* jc, int* descc);
void pdpotrf_(const char* uplo, int* n, double* a, int* ia, int* ja, int* desca, int* info);

int main(void)
{
    const int n = 1024, nb = 256, ictxt = 0;
    int one = 1, nn = n, info = -1;
    int desc[9] = {1, ictxt, n, n, nb, nb, 0, 0, n};      /* DTYPE, CTXT, M, N, MB, NB, RSRC, CSRC, LLD */
    double alpha = 1.0, beta = 0.0;
    double *a = malloc(sizeof(double) * n * n), *c = calloc((size_t)n * n, sizeof(double));
    if (!a || !c) return 2;
    for (int j = 0; j < n; ++j)
        for (int i = 0; i < n; ++i)
            a[i + (size_t)j * n] = (i == j) ? 2.0 : 1.0 / (1.0 + abs(i - j));

    if (sb_initialize(0, 0, 1, NULL) != 0) {              /* device 0, rank 0 of 1 */
        fprintf(stderr, "no GPU path: %s\n", sb_last_error());
        return 1;
    }
    slate_scalapack_grid_set(ictxt, 'C', 1, 1);

    /* C = A * A^T (symmetric positive definite), then its Cholesky factor */
    pdgemm_("N", "T", &nn, &nn, &nn, &alpha, a, &one, &one, desc, a, &one, &one, desc, &beta, c, &one,
            &one, desc);
    pdpotrf_("L", &nn, c, &one, &one, desc, &info);
    printf("pdgemm + pdpotrf on the GPU path: info = %d, L(0,0) = %.6f, status = %d\n", info, c[0],
           slate_scalapack_status());
    sb_finalize();
    free(a);
    free(c);
    return info != 0;
}
