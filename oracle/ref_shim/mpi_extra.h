/* mpi_extra.h -- TEST INFRASTRUCTURE (oracle/_ref build only).
 *
 * Force-included (-include) when the reference's HostTask path is compiled
 * single-rank with -DSLATE_NO_MPI.  The reference's own no-MPI header
 * (include/slate/internal/mpi.hh) is stale at this commit: it lacks the three
 * names below, which include/slate/internal/util.hh:279, BaseMatrix.hh:820 and
 * src/internal/internal_reduce_info.cc:31 use.  Nothing here is product code.
 */
#ifndef ORACLE_MPI_EXTRA_H
#define ORACLE_MPI_EXTRA_H
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
