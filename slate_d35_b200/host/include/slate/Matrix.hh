// slate/Matrix.hh -- Matrix, HermitianMatrix, SymmetricMatrix, TriangularMatrix over
// BaseMatrix.  Constructors and factory functions keep the reference's argument lists
// (include/slate/Matrix.hh:41-55, 355-520, 832-847; HermitianMatrix.hh;
//  BaseTrapezoidMatrix.hh; TriangularMatrix.hh).
#ifndef SLATE_B200_MATRIX_HH
#define SLATE_B200_MATRIX_HH

#include "slate/BaseMatrix.hh"

namespace slate {

//------------------------------------------------------------------------------
template <typename scalar_t>
class Matrix : public BaseMatrix<scalar_t> {
public:
    Matrix() : BaseMatrix<scalar_t>() {}

    /// include/slate/Matrix.hh:41-55: m x n, square tiles nb, p x q grid (column-major ranks)
    Matrix(int64_t m, int64_t n, int64_t nb, int p, int q, MPI_Comm comm)
        : BaseMatrix<scalar_t>(m, n, nb, nb, GridOrder::Col, p, q, comm) {}
    Matrix(int64_t m, int64_t n, int64_t mb, int64_t nb, GridOrder order, int p, int q,
           MPI_Comm comm)
        : BaseMatrix<scalar_t>(m, n, mb, nb, order, p, q, comm) {}

    /// include/slate/Matrix.hh:33-38: tile sizes, owners and devices from the caller's functions
    /// (the tester's --nonuniform-nb / --dev-order paths, test/matrix_utils.cc:15-88).  Tile
    /// sizes need not be uniform: every kernel takes per-tile m, n, k, ld.
    Matrix(int64_t m, int64_t n, std::function<int64_t(int64_t)>& inTileMb,
           std::function<int64_t(int64_t)>& inTileNb, std::function<int(ij_tuple)>& inTileRank,
           std::function<int(ij_tuple)>& inTileDevice, MPI_Comm comm)
        : BaseMatrix<scalar_t>(m, n, inTileMb, inTileNb, inTileRank, inTileDevice, comm) {}

    /// include/slate/Matrix.hh:102-112: tiles are views into the caller's DEVICE arrays, one
    /// array per device of this rank holding that device's share of the 2D block-cyclic local
    /// matrix in column-major order with leading dimension lda.  One GPU per process here, so
    /// num_devices must be 1 and Aarray[0] is the rank's whole local array (ScaLAPACK layout, on
    /// the device).  The tiles are UserOwned: never freed, results land in the caller's memory.
    static Matrix fromDevices(int64_t m, int64_t n, scalar_t** Aarray, int num_devices, int64_t lda,
                              int64_t mb, int64_t nb, int p, int q, MPI_Comm comm)
    {
        slate_assert(num_devices == 1 && Aarray != nullptr);
        Matrix M(m, n, mb, nb, GridOrder::Col, p, q, comm);
        M.origin_ = Target::Devices;
        for (int64_t j = 0; j < M.nt(); ++j)
            for (int64_t i = 0; i < M.mt(); ++i)
                if (M.tileIsLocal(i, j))
                    M.tileInsert(i, j, 0, &Aarray[0][(i / p) * mb + (j / q) * nb * lda], lda);
        return M;
    }
    static Matrix fromDevices(int64_t m, int64_t n, scalar_t** Aarray, int num_devices, int64_t lda,
                              int64_t nb, int p, int q, MPI_Comm comm)
    {
        return fromDevices(m, n, Aarray, num_devices, lda, nb, nb, p, q, comm);
    }

    /// include/slate/Matrix.hh:355-384: view of a column-major host array; every rank
    /// passes its own array and owns the tiles the grid assigns to it.
    static Matrix fromLAPACK(int64_t m, int64_t n, scalar_t* A, int64_t lda, int64_t nb,
                             int p, int q, MPI_Comm comm)
    {
        Matrix M(m, n, nb, p, q, comm);
        M.origin_ = Target::Host;
        for (int64_t j = 0; j < M.nt(); ++j)
            for (int64_t i = 0; i < M.mt(); ++i)
                if (M.tileIsLocal(i, j))
                    M.tileInsert(i, j, HostNum, &A[i * nb + j * nb * lda], lda);
        return M;
    }

    /// include/slate/Matrix.hh:486-520: 2D block-cyclic local array, ScaLAPACK layout.
    static Matrix fromScaLAPACK(int64_t m, int64_t n, scalar_t* A, int64_t lda, int64_t nb,
                                int p, int q, MPI_Comm comm)
    {
        Matrix M(m, n, nb, p, q, comm);
        M.origin_ = Target::Host;
        for (int64_t j = 0; j < M.nt(); ++j)
            for (int64_t i = 0; i < M.mt(); ++i)
                if (M.tileIsLocal(i, j))
                    M.tileInsert(i, j, HostNum, &A[(i / p) * nb + (j / q) * nb * lda], lda);
        return M;
    }

    /// include/slate/Matrix.hh:486-520, full argument list: mb x nb tiles, grid order as
    /// BLACS reports it (scalapack_api/scalapack_slate.hh:58-78).
    static Matrix fromScaLAPACK(int64_t m, int64_t n, scalar_t* A, int64_t lda, int64_t mb,
                                int64_t nb, GridOrder order, int p, int q, MPI_Comm comm)
    {
        Matrix M(m, n, mb, nb, order, p, q, comm);
        M.origin_ = Target::Host;
        for (int64_t j = 0; j < M.nt(); ++j)
            for (int64_t i = 0; i < M.mt(); ++i)
                if (M.tileIsLocal(i, j))
                    M.tileInsert(i, j, HostNum, &A[(i / p) * mb + (j / q) * nb * lda], lda);
        return M;
    }

    /// include/slate/Matrix.hh:832-847: SLATE-owned tiles on the host or born on the device.
    void insertLocalTiles(Target origin = Target::Host)
    {
        this->origin_ = origin;
        const bool on_dev = (origin == Target::Devices);
        if (on_dev) {
            int64_t count = this->getMaxDeviceTiles(0);
            int64_t have = (int64_t)this->storage_->memory_.available(0);
            if (count > have) this->storage_->memory_.addDeviceBlocks(0, count - have);
        }
        for (int64_t j = 0; j < this->nt(); ++j)
            for (int64_t i = 0; i < this->mt(); ++i)
                if (this->tileIsLocal(i, j))
                    this->tileInsert(i, j, on_dev ? 0 : HostNum);
    }

    Matrix sub(int64_t i1, int64_t i2, int64_t j1, int64_t j2)
    {
        return Matrix(*this, i1, i2, j1, j2);
    }

    Matrix(BaseMatrix<scalar_t> const& orig, int64_t i1, int64_t i2, int64_t j1, int64_t j2)
        : BaseMatrix<scalar_t>(orig, i1, i2, j1, j2)
    {
        this->uplo_ = Uplo::General;
    }
};

//------------------------------------------------------------------------------
/// Upper or lower trapezoid storage: only tiles of the `uplo` triangle exist.
template <typename scalar_t>
class BaseTrapezoidMatrix : public BaseMatrix<scalar_t> {
public:
    BaseTrapezoidMatrix() : BaseMatrix<scalar_t>() {}

    /// BaseTrapezoidMatrix.hh: off-diagonal block as a general matrix
    Matrix<scalar_t> sub(int64_t i1, int64_t i2, int64_t j1, int64_t j2)
    {
        return Matrix<scalar_t>(*this, i1, i2, j1, j2);
    }

    void insertLocalTiles(Target origin = Target::Host)
    {
        this->origin_ = origin;
        const bool on_dev = (origin == Target::Devices);
        if (on_dev) {
            int64_t count = this->getMaxDeviceTiles(0);
            int64_t have = (int64_t)this->storage_->memory_.available(0);
            if (count > have) this->storage_->memory_.addDeviceBlocks(0, count - have);
        }
        for (int64_t j = 0; j < this->nt(); ++j)
            for (int64_t i = 0; i < this->mt(); ++i)
                if (this->tileIsLocal(i, j) && this->inStoredTriangle(i, j))
                    this->tileInsert(i, j, on_dev ? 0 : HostNum);
    }

protected:
    BaseTrapezoidMatrix(Uplo uplo, int64_t m, int64_t n, int64_t nb, int p, int q, MPI_Comm comm)
        : BaseMatrix<scalar_t>(m, n, nb, nb, GridOrder::Col, p, q, comm)
    {
        slate_assert(uplo == Uplo::Lower || uplo == Uplo::Upper);
        this->uplo_ = uplo;
    }
    BaseTrapezoidMatrix(Uplo uplo, int64_t m, int64_t n, int64_t nb, GridOrder order, int p, int q,
                        MPI_Comm comm)
        : BaseMatrix<scalar_t>(m, n, nb, nb, order, p, q, comm)
    {
        slate_assert(uplo == Uplo::Lower || uplo == Uplo::Upper);
        this->uplo_ = uplo;
    }
    BaseTrapezoidMatrix(Uplo uplo, int64_t m, int64_t n, std::function<int64_t(int64_t)>& inTileMb,
                        std::function<int64_t(int64_t)>& inTileNb,
                        std::function<int(ij_tuple)>& inTileRank,
                        std::function<int(ij_tuple)>& inTileDevice, MPI_Comm comm)
        : BaseMatrix<scalar_t>(m, n, inTileMb, inTileNb, inTileRank, inTileDevice, comm)
    {
        slate_assert(uplo == Uplo::Lower || uplo == Uplo::Upper);
        this->uplo_ = uplo;
    }
    BaseTrapezoidMatrix(BaseMatrix<scalar_t> const& orig, int64_t i1, int64_t i2)
        : BaseMatrix<scalar_t>(orig, i1, i2, i1, i2) {}
    BaseTrapezoidMatrix(Uplo uplo, BaseMatrix<scalar_t> const& orig) : BaseMatrix<scalar_t>(orig)
    {
        this->uplo_ = uplo;
    }

    void insertUserTiles(scalar_t* A, int64_t lda, int64_t nb)
    {
        this->origin_ = Target::Host;
        for (int64_t j = 0; j < this->nt(); ++j)
            for (int64_t i = 0; i < this->mt(); ++i)
                if (this->tileIsLocal(i, j) && this->inStoredTriangle(i, j))
                    this->tileInsert(i, j, HostNum, &A[i * nb + j * nb * lda], lda);
    }

    /// tiles of the stored triangle over a 2D block-cyclic local array
    /// (BaseTrapezoidMatrix.hh fromScaLAPACK constructor)
    void insertUserTilesScaLAPACK(scalar_t* A, int64_t lda, int64_t nb, int p, int q)
    {
        this->origin_ = Target::Host;
        for (int64_t j = 0; j < this->nt(); ++j)
            for (int64_t i = 0; i < this->mt(); ++i)
                if (this->tileIsLocal(i, j) && this->inStoredTriangle(i, j))
                    this->tileInsert(i, j, HostNum, &A[(i / p) * nb + (j / q) * nb * lda], lda);
    }
};

//------------------------------------------------------------------------------
template <typename scalar_t>
class HermitianMatrix : public BaseTrapezoidMatrix<scalar_t> {
public:
    HermitianMatrix() {}
    HermitianMatrix(Uplo uplo, int64_t n, int64_t nb, int p, int q, MPI_Comm comm)
        : BaseTrapezoidMatrix<scalar_t>(uplo, n, n, nb, p, q, comm) {}
    HermitianMatrix(Uplo uplo, int64_t n, int64_t nb, GridOrder order, int p, int q, MPI_Comm comm)
        : BaseTrapezoidMatrix<scalar_t>(uplo, n, n, nb, order, p, q, comm) {}

    /// HermitianMatrix.hh (lambda constructor): square, tile sizes inTileNb in both dimensions
    HermitianMatrix(Uplo uplo, int64_t n, std::function<int64_t(int64_t)>& inTileNb,
                    std::function<int(ij_tuple)>& inTileRank, std::function<int(ij_tuple)>& inTileDevice,
                    MPI_Comm comm)
        : BaseTrapezoidMatrix<scalar_t>(uplo, n, n, inTileNb, inTileNb, inTileRank, inTileDevice, comm) {}

    /// HermitianMatrix.hh fromDevices: the stored triangle as views into the caller's device
    /// array (see Matrix::fromDevices)
    static HermitianMatrix fromDevices(Uplo uplo, int64_t n, scalar_t** Aarray, int num_devices,
                                       int64_t lda, int64_t nb, int p, int q, MPI_Comm comm)
    {
        slate_assert(num_devices == 1 && Aarray != nullptr);
        HermitianMatrix M(uplo, n, nb, p, q, comm);
        M.origin_ = Target::Devices;
        for (int64_t j = 0; j < M.nt(); ++j)
            for (int64_t i = 0; i < M.mt(); ++i)
                if (M.tileIsLocal(i, j) && M.inStoredTriangle(i, j))
                    M.tileInsert(i, j, 0, &Aarray[0][(i / p) * nb + (j / q) * nb * lda], lda);
        return M;
    }

    /// HermitianMatrix.hh fromScaLAPACK (used by scalapack_api/*.cc)
    static HermitianMatrix fromScaLAPACK(Uplo uplo, int64_t n, scalar_t* A, int64_t lda, int64_t nb,
            GridOrder order, int p, int q, MPI_Comm comm)
    {
        HermitianMatrix M(uplo, n, nb, order, p, q, comm);
        M.insertUserTilesScaLAPACK(A, lda, nb, p, q);
        return M;
    }
    static HermitianMatrix fromLAPACK(Uplo uplo, int64_t n, scalar_t* A, int64_t lda, int64_t nb,
                                      int p, int q, MPI_Comm comm)
    {
        HermitianMatrix M(uplo, n, nb, p, q, comm);
        M.insertUserTiles(A, lda, nb);
        return M;
    }

    /// diagonal block [i1..i2] x [i1..i2]
    HermitianMatrix sub(int64_t i1, int64_t i2) { return HermitianMatrix(*this, i1, i2); }
    Matrix<scalar_t> sub(int64_t i1, int64_t i2, int64_t j1, int64_t j2)
    {
        return BaseTrapezoidMatrix<scalar_t>::sub(i1, i2, j1, j2);
    }

    HermitianMatrix(BaseMatrix<scalar_t> const& orig, int64_t i1, int64_t i2)
        : BaseTrapezoidMatrix<scalar_t>(orig, i1, i2) {}
    HermitianMatrix(Uplo uplo, BaseMatrix<scalar_t> const& orig)
        : BaseTrapezoidMatrix<scalar_t>(uplo, orig) {}
};

template <typename scalar_t>
class SymmetricMatrix : public BaseTrapezoidMatrix<scalar_t> {
public:
    SymmetricMatrix() {}
    SymmetricMatrix(Uplo uplo, int64_t n, int64_t nb, int p, int q, MPI_Comm comm)
        : BaseTrapezoidMatrix<scalar_t>(uplo, n, n, nb, p, q, comm) {}
    SymmetricMatrix(Uplo uplo, int64_t n, int64_t nb, GridOrder order, int p, int q, MPI_Comm comm)
        : BaseTrapezoidMatrix<scalar_t>(uplo, n, n, nb, order, p, q, comm) {}
    /// SymmetricMatrix.hh fromScaLAPACK (used by scalapack_api/*.cc)
    static SymmetricMatrix fromScaLAPACK(Uplo uplo, int64_t n, scalar_t* A, int64_t lda, int64_t nb,
            GridOrder order, int p, int q, MPI_Comm comm)
    {
        SymmetricMatrix M(uplo, n, nb, order, p, q, comm);
        M.insertUserTilesScaLAPACK(A, lda, nb, p, q);
        return M;
    }
    static SymmetricMatrix fromLAPACK(Uplo uplo, int64_t n, scalar_t* A, int64_t lda, int64_t nb,
                                      int p, int q, MPI_Comm comm)
    {
        SymmetricMatrix M(uplo, n, nb, p, q, comm);
        M.insertUserTiles(A, lda, nb);
        return M;
    }
    SymmetricMatrix sub(int64_t i1, int64_t i2) { return SymmetricMatrix(*this, i1, i2); }
    Matrix<scalar_t> sub(int64_t i1, int64_t i2, int64_t j1, int64_t j2)
    {
        return BaseTrapezoidMatrix<scalar_t>::sub(i1, i2, j1, j2);
    }
    SymmetricMatrix(BaseMatrix<scalar_t> const& orig, int64_t i1, int64_t i2)
        : BaseTrapezoidMatrix<scalar_t>(orig, i1, i2) {}
};

template <typename scalar_t>
class TriangularMatrix : public BaseTrapezoidMatrix<scalar_t> {
public:
    TriangularMatrix() : diag_(Diag::NonUnit) {}
    TriangularMatrix(Uplo uplo, Diag diag, int64_t n, int64_t nb, int p, int q, MPI_Comm comm)
        : BaseTrapezoidMatrix<scalar_t>(uplo, n, n, nb, p, q, comm), diag_(diag) {}
    TriangularMatrix(Uplo uplo, Diag diag, int64_t n, int64_t nb, GridOrder order, int p, int q,
                     MPI_Comm comm)
        : BaseTrapezoidMatrix<scalar_t>(uplo, n, n, nb, order, p, q, comm), diag_(diag) {}

    /// TriangularMatrix.hh fromScaLAPACK (used by scalapack_api/*.cc)
    static TriangularMatrix fromScaLAPACK(Uplo uplo, Diag diag, int64_t n, scalar_t* A, int64_t lda, int64_t nb,
            GridOrder order, int p, int q, MPI_Comm comm)
    {
        TriangularMatrix M(uplo, diag, n, nb, order, p, q, comm);
        M.insertUserTilesScaLAPACK(A, lda, nb, p, q);
        return M;
    }
    static TriangularMatrix fromLAPACK(Uplo uplo, Diag diag, int64_t n, scalar_t* A, int64_t lda,
                                       int64_t nb, int p, int q, MPI_Comm comm)
    {
        TriangularMatrix M(uplo, diag, n, nb, p, q, comm);
        M.insertUserTiles(A, lda, nb);
        return M;
    }

    /// TriangularMatrix.hh: conversion sharing the tiles (potrf.cc:113-114)
    /// TriangularMatrix.hh: the uplo triangle of a square general matrix, sharing its tiles
    /// (getrf_nopiv.cc:104, 137: the L and U factors held in A(k, k))
    TriangularMatrix(Uplo uplo, Diag diag, Matrix<scalar_t> const& orig)
        : BaseTrapezoidMatrix<scalar_t>(uplo, orig), diag_(diag)
    {
        slate_assert(orig.mt() == orig.nt());
    }
    TriangularMatrix(Diag diag, BaseTrapezoidMatrix<scalar_t> const& orig)
        : BaseTrapezoidMatrix<scalar_t>(orig), diag_(diag) {}

    Diag diag() const { return diag_; }
    TriangularMatrix sub(int64_t i1, int64_t i2) { return TriangularMatrix(*this, i1, i2); }
    TriangularMatrix(TriangularMatrix const& orig, int64_t i1, int64_t i2)
        : BaseTrapezoidMatrix<scalar_t>(orig, i1, i2), diag_(orig.diag_) {}
    Matrix<scalar_t> sub(int64_t i1, int64_t i2, int64_t j1, int64_t j2)
    {
        return BaseTrapezoidMatrix<scalar_t>::sub(i1, i2, j1, j2);
    }

protected:
    Diag diag_;
};

//------------------------------------------------------------------------------
// include/slate/BaseMatrix.hh:3247-3290: shallow views with the op toggled.
namespace detail {
template <typename M>
struct ViewAccess : public M {
    static void set_op(M& A, Op op) { static_cast<ViewAccess&>(A).op_ = op; }
};
}  // namespace detail

template <typename MatrixType>
MatrixType transpose(MatrixType& A)
{
    using scalar_t = decltype(A(0, 0)(0, 0));
    MatrixType AT = A;
    Op op = A.op();
    if (op == Op::NoTrans) op = Op::Trans;
    else if (op == Op::Trans || !is_complex<scalar_t>::value) op = Op::NoTrans;
    else slate_error("unsupported operation, results in conjugate-no-transpose");
    detail::ViewAccess<MatrixType>::set_op(AT, op);
    return AT;
}
template <typename MatrixType>
MatrixType transpose(MatrixType&& A) { return transpose(A); }

template <typename MatrixType>
MatrixType conj_transpose(MatrixType& A)
{
    using scalar_t = decltype(A(0, 0)(0, 0));
    MatrixType AT = A;
    Op op = A.op();
    if (op == Op::NoTrans) op = Op::ConjTrans;
    else if (op == Op::ConjTrans || !is_complex<scalar_t>::value) op = Op::NoTrans;
    else slate_error("unsupported operation, results in conjugate-no-transpose");
    detail::ViewAccess<MatrixType>::set_op(AT, op);
    return AT;
}
template <typename MatrixType>
MatrixType conj_transpose(MatrixType&& A) { return conj_transpose(A); }

}  // namespace slate

#endif
