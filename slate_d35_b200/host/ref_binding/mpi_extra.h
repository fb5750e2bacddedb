/* mpi_extra.h -- force-included when the reference's sources are compiled single rank
 * (-DSLATE_NO_MPI) for the drop-in build of slate_d35_b200/host/ref_binding.
 *
 * The reference's own no-MPI header (include/slate/internal/mpi.hh) is stale at this commit:
 * it lacks the three names below, which include/slate/internal/util.hh:279,
 * include/slate/BaseMatrix.hh:820 and src/internal/internal_reduce_info.cc:31 use.
 */
#ifndef REF_BINDING_MPI_EXTRA_H
#define REF_BINDING_MPI_EXTRA_H
#ifdef __cplusplus
extern "C" {
#endif
double MPI_Wtime(void);
#ifdef __cplusplus
}
#endif
#define MPI_COMM_SELF 1      /* == MPI_COMM_WORLD of the stub enum: one rank either way */
#define MPI_MIN       1001   /* any value distinct from the stub enum                   */
#endif
