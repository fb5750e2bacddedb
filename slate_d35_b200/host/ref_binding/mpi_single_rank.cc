// mpi_single_rank.cc -- a one-rank MPI for building the reference's own sources (unmodified)
// against the tile engine without an MPI installation (drop-in build, see Makefile).
// Stands in for the reference's src/stubs/mpi_stubs.cc, whose MPI_Waitall asserts
// (mpi_stubs.cc:222-225) although listBcast always calls it
// (include/slate/BaseMatrix.hh:2099-2100) and whose MPI_Allreduce handles only MPI_MAX of
// one element.  With one rank every collective is a local copy and no point-to-point
// message can be posted.
#include "slate/internal/mpi.hh"
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <complex>

int* MPI_STATUS_IGNORE = nullptr;

static size_t type_size(MPI_Datatype t)
{
    switch (t) {
        case MPI_BYTE: case MPI_CHAR: return 1;
        case MPI_INT: case MPI_UNSIGNED: case MPI_FLOAT: return 4;
        case MPI_LONG: case MPI_DOUBLE: case MPI_INT64_T: case MPI_2INT:
        case MPI_FLOAT_INT: case MPI_C_COMPLEX: return 8;
        case MPI_DOUBLE_INT: case MPI_C_DOUBLE_COMPLEX: return 16;
        default: return 0;   // derived types are never used with one rank
    }
}

static int no_p2p() { std::abort(); return MPI_SUCCESS; }   // unreachable with one rank

extern "C" {

double MPI_Wtime(void)
{
    using clk = std::chrono::steady_clock;
    return std::chrono::duration<double>(clk::now().time_since_epoch()).count();
}

int MPI_Type_contiguous(int, MPI_Datatype, MPI_Datatype* n) { *n = 0; return MPI_SUCCESS; }
int MPI_Type_vector(int, int, int, MPI_Datatype, MPI_Datatype* n) { *n = 0; return MPI_SUCCESS; }
int MPI_Type_commit(MPI_Datatype*) { return MPI_SUCCESS; }
int MPI_Type_free(MPI_Datatype*) { return MPI_SUCCESS; }

int MPI_Allreduce(const void* s, void* r, int count, MPI_Datatype t, MPI_Op, MPI_Comm)
{
    if (s != r) std::memmove(r, s, type_size(t) * (size_t)count);
    return MPI_SUCCESS;
}
int MPI_Reduce(void* s, void* r, int count, MPI_Datatype t, MPI_Op, int, MPI_Comm)
{
    if (s != r) std::memmove(r, s, type_size(t) * (size_t)count);
    return MPI_SUCCESS;
}
int MPI_Barrier(MPI_Comm) { return MPI_SUCCESS; }
int MPI_Bcast(void*, int, MPI_Datatype, int, MPI_Comm) { return MPI_SUCCESS; }

int MPI_Comm_create_group(MPI_Comm c, MPI_Group, int, MPI_Comm* n) { *n = c; return MPI_SUCCESS; }
int MPI_Comm_free(MPI_Comm*) { return MPI_SUCCESS; }
int MPI_Comm_group(MPI_Comm, MPI_Group* g) { *g = 0; return MPI_SUCCESS; }
int MPI_Comm_rank(MPI_Comm, int* r) { *r = 0; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm, int* s) { *s = 1; return MPI_SUCCESS; }
MPI_Fint MPI_Comm_f2c(MPI_Comm c) { return c; }

int MPI_Group_free(MPI_Group*) { return MPI_SUCCESS; }
int MPI_Group_incl(MPI_Group, int, const int[], MPI_Group* n) { *n = 0; return MPI_SUCCESS; }
int MPI_Group_translate_ranks(MPI_Group, int n, const int r1[], MPI_Group, int r2[])
{
    for (int i = 0; i < n; ++i) r2[i] = r1[i];
    return MPI_SUCCESS;
}

int MPI_Init(int*, char***) { return MPI_SUCCESS; }
int MPI_Init_thread(int*, char***, int, int* p) { *p = MPI_THREAD_MULTIPLE; return MPI_SUCCESS; }
int MPI_Initialized(int* f) { *f = 1; return MPI_SUCCESS; }
int MPI_Finalize(void) { return MPI_SUCCESS; }

int MPI_Irecv(void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Request*) { return no_p2p(); }
int MPI_Isend(const void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Request*) { return no_p2p(); }
int MPI_Recv(void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Status*) { return no_p2p(); }
int MPI_Send(const void*, int, MPI_Datatype, int, int, MPI_Comm) { return no_p2p(); }
int MPI_Sendrecv(const void*, int, MPI_Datatype, int, int, void*, int, MPI_Datatype, int, int,
                 MPI_Comm, MPI_Status*) { return no_p2p(); }

int MPI_Op_create(MPI_User_function*, int, MPI_Op* op) { *op = 0; return MPI_SUCCESS; }
int MPI_Op_free(MPI_Op*) { return MPI_SUCCESS; }
int MPI_Request_free(MPI_Request*) { return MPI_SUCCESS; }
int MPI_Wait(MPI_Request*, MPI_Status*) { return MPI_SUCCESS; }
int MPI_Waitall(int, MPI_Request[], MPI_Status[]) { return MPI_SUCCESS; }

int MPI_Error_string(int, char* s, int* len)
{
    std::strcpy(s, "single-rank MPI shim");
    *len = (int)std::strlen(s);
    return MPI_SUCCESS;
}

}  // extern "C"
