// slate/Tile.hh -- the tile data contract (reference: include/slate/Tile.hh:129-420).
// A Tile is a light handle: extents, stride, data pointer, the op/uplo view flags,
// where the data lives (HostNum or the device ordinal) and who owns it.
#ifndef SLATE_B200_TILE_HH
#define SLATE_B200_TILE_HH

#include "slate/enums.hh"

#include <utility>

namespace slate {

template <typename scalar_t>
class Tile {
public:
    Tile()
        : mb_(0), nb_(0), stride_(0), user_stride_(0), op_(Op::NoTrans), uplo_(Uplo::General),
          data_(nullptr), kind_(TileKind::UserOwned), layout_(Layout::ColMajor), device_(HostNum)
    {}

    /// include/slate/Tile.hh:430-470 (mb, nb, data, lda, device, kind, layout)
    Tile(int64_t mb, int64_t nb, scalar_t* A, int64_t lda, int device, TileKind kind,
         Layout layout = Layout::ColMajor)
        : mb_(mb), nb_(nb), stride_(lda), user_stride_(lda), op_(Op::NoTrans),
          uplo_(Uplo::General), data_(A), kind_(kind), layout_(layout), device_(device)
    {
        slate_assert(mb >= 0 && nb >= 0);
        slate_assert(A != nullptr || mb * nb == 0);
        slate_assert(lda >= (layout == Layout::ColMajor ? mb : nb));
    }

    /// rows / cols of op(tile)
    int64_t mb() const { return op_ == Op::NoTrans ? mb_ : nb_; }
    int64_t nb() const { return op_ == Op::NoTrans ? nb_ : mb_; }
    int64_t stride() const { return stride_; }
    scalar_t* data() { return data_; }
    scalar_t const* data() const { return data_; }
    Op op() const { return op_; }
    void op(Op o) { op_ = o; }
    Uplo uplo() const { return uploLogical(); }
    Uplo uploLogical() const
    {
        if (uplo_ == Uplo::General) return Uplo::General;
        if ((uplo_ == Uplo::Lower) == (op_ == Op::NoTrans)) return Uplo::Lower;
        return Uplo::Upper;
    }
    Uplo uploPhysical() const { return uplo_; }
    void uplo(Uplo u) { uplo_ = u; }          // physical
    int device() const { return device_; }
    TileKind kind() const { return kind_; }
    bool origin() const { return kind_ != TileKind::Workspace; }
    bool workspace() const { return kind_ == TileKind::Workspace; }
    Layout layout() const { return layout_; }
    /// After a transposition of the stored array (tileLayoutConvert): RowMajor data of an
    /// mb x nb tile is the column-major array of its transpose, nb x mb with stride >= nb.
    /// `data` replaces the buffer when the conversion was done out of place.
    void setLayout(Layout layout, int64_t stride, scalar_t* data = nullptr)
    {
        slate_assert(stride >= (layout == Layout::ColMajor ? mb_ : nb_));
        layout_ = layout;
        stride_ = stride;
        if (data) data_ = data;
    }
    size_t bytes() const { return sizeof(scalar_t) * (size_t)mb_ * (size_t)nb_; }
    bool isContiguous() const { return stride_ == (layout_ == Layout::ColMajor ? mb_ : nb_); }

    /// host element access of op(tile) (include/slate/Tile.hh:593-640); host tiles only
    scalar_t operator()(int64_t i, int64_t j) const
    {
        slate_assert(device_ == HostNum);
        if (op_ == Op::NoTrans) return data_[i + j * stride_];
        scalar_t v = data_[j + i * stride_];
        return op_ == Op::ConjTrans ? conj(v) : v;
    }
    scalar_t& at(int64_t i, int64_t j)
    {
        slate_assert(device_ == HostNum);
        return op_ == Op::NoTrans ? data_[i + j * stride_] : data_[j + i * stride_];
    }

    template <typename T> friend Tile<T> transpose(Tile<T> const& A);
    template <typename T> friend Tile<T> conj_transpose(Tile<T> const& A);

private:
    int64_t mb_, nb_, stride_, user_stride_;
    Op op_;
    Uplo uplo_;
    scalar_t* data_;
    TileKind kind_;
    Layout layout_;
    int device_;
};

/// include/slate/Tile.hh:60-85: toggles the op; Trans of ConjTrans is not representable.
template <typename T>
Tile<T> transpose(Tile<T> const& A)
{
    Tile<T> AT = A;
    if (AT.op_ == Op::NoTrans) AT.op_ = Op::Trans;
    else if (AT.op_ == Op::Trans || !is_complex<T>::value) AT.op_ = Op::NoTrans;
    else slate_error("unsupported operation, results in conjugate-no-transpose");
    return AT;
}

template <typename T>
Tile<T> conj_transpose(Tile<T> const& A)
{
    Tile<T> AT = A;
    if (AT.op_ == Op::NoTrans) AT.op_ = Op::ConjTrans;
    else if (AT.op_ == Op::ConjTrans || !is_complex<T>::value) AT.op_ = Op::NoTrans;
    else slate_error("unsupported operation, results in conjugate-no-transpose");
    return AT;
}

}  // namespace slate

#endif
