// slate/enums.hh -- the enums, Options and exceptions of SLATE's public surface, as the
// B200 engine mirrors them.  Names, values and meaning follow the reference
// (include/slate/enums.hh:34-144, include/slate/types.hh:22-58,
//  include/slate/Exception.hh:53-120, blaspp/include/blas/util.hh enums).
#ifndef SLATE_B200_ENUMS_HH
#define SLATE_B200_ENUMS_HH

#include <complex>
#include <cstdint>
#include <exception>
#include <map>
#include <string>

namespace slate {

// ---- blas:: enums (char values as in blaspp) -------------------------------------------
enum class Layout : char { ColMajor = 'C', RowMajor = 'R' };
enum class Op     : char { NoTrans  = 'N', Trans    = 'T', ConjTrans = 'C' };
enum class Uplo   : char { Upper    = 'U', Lower    = 'L', General   = 'G' };
enum class Diag   : char { NonUnit  = 'N', Unit     = 'U' };
enum class Side   : char { Left     = 'L', Right    = 'R' };

// ---- include/slate/enums.hh:34-40 --------------------------------------------------------
enum class Target : char {
    Host      = 'H',
    HostTask  = 'T',
    HostNest  = 'N',
    HostBatch = 'B',
    Devices   = 'D',
};

namespace internal {
template <Target> class TargetType {};       // include/slate/enums.hh:42-48
}

// ---- include/slate/enums.hh:63-99 (the keys gemm / potrf read, same numbering) -----------
enum class Option : char {
    ChunkSize = 0,
    Lookahead,
    BlockSize,
    InnerBlocking,
    MaxPanelThreads,
    Tolerance,
    Target,
    HoldLocalWorkspace,
    Depth,
    MaxIterations,
    UseFallbackSolver,
    PivotThreshold,
    MethodGemm = 63,
};

/// include/slate/method.hh:77-98
typedef int Method;
namespace MethodGemm {
const Method Auto = 0;
const Method GemmA = 1;      ///< A stationary (src/gemmA.cc)
const Method GemmC = 2;      ///< C stationary (src/gemmC.cc)
}  // namespace MethodGemm

enum class TileKind : char { Workspace = 'w', SlateOwned = 'o', UserOwned = 'u' };

// include/slate/enums.hh:138-144
enum MOSI : short { Invalid = 0x0001, Shared = 0x0010, Modified = 0x0100, OnHold = 0x1000 };

enum class LayoutConvert : char { ColMajor = 'C', RowMajor = 'R', None = 'N' };
enum class GridOrder : char { Col = 'C', Row = 'R', Unknown = 'U' };

const int HostNum = -1;                       // include/slate/enums.hh:133
const int AnyDevice = -3;

// ---- include/slate/types.hh:22-58 ---------------------------------------------------------
class OptionValue {
public:
    OptionValue() : i_(0), d_(0) {}
    OptionValue(int i) : i_(i), d_(i) {}
    OptionValue(int64_t i) : i_(i), d_((double)i) {}
    OptionValue(double d) : i_((int64_t)d), d_(d) {}
    OptionValue(bool b) : i_(b), d_(b) {}
    OptionValue(Target t) : i_(int(t)), d_(int(t)) {}
    int64_t i_;
    double d_;
};
using Options = std::map<Option, OptionValue>;

template <typename T>
T get_option(Options const& opts, Option key, T def)
{
    auto it = opts.find(key);
    if (it == opts.end()) return def;
    return T(it->second.i_);
}
inline Target get_target(Options const& opts, Target def = Target::HostTask)
{
    auto it = opts.find(Option::Target);
    return it == opts.end() ? def : Target(char(it->second.i_));
}

// ---- include/slate/Exception.hh -------------------------------------------------------------
class Exception : public std::exception {
public:
    Exception() {}
    Exception(std::string const& msg) : msg_(msg) {}
    Exception(std::string const& msg, const char* func, const char* file, int line)
        : msg_(msg + ", in function " + func + ", in file " + file + ", line " + std::to_string(line)) {}
    const char* what() const noexcept override { return msg_.c_str(); }
protected:
    std::string msg_;
};
class NotImplemented : public Exception {
public:
    NotImplemented(const char* what, const char* func, const char* file, int line)
        : Exception(std::string("SLATE ERROR: Not yet implemented: ") + what, func, file, line) {}
};
class FalseConditionException : public Exception {
public:
    FalseConditionException(const char* cond, const char* func, const char* file, int line)
        : Exception(std::string("SLATE ERROR: Error check '") + cond + "' failed", func, file, line) {}
};
/// A C-ABI call (include/tile_engine.h) returned a non-zero status.
class EngineException : public Exception {
public:
    EngineException(const char* call, int status, const char* detail, const char* func,
                    const char* file, int line)
        : Exception(std::string("SLATE B200 ENGINE ERROR: ") + call + " failed with status "
                    + std::to_string(status) + " (" + detail + ")", func, file, line) {}
};

#define slate_error(msg) \
    do { throw slate::Exception(msg, __func__, __FILE__, __LINE__); } while (0)
#define slate_assert(cond) \
    do { if (!(cond)) throw slate::FalseConditionException(#cond, __func__, __FILE__, __LINE__); } while (0)
#define slate_error_if(cond) \
    do { if (cond) throw slate::Exception(std::string("SLATE ERROR: Error condition '") + #cond \
                                          + "' occured", __func__, __FILE__, __LINE__); } while (0)
#define slate_not_implemented(msg) \
    do { throw slate::NotImplemented(msg, __func__, __FILE__, __LINE__); } while (0)

// ---- scalar traits -----------------------------------------------------------------------
template <typename T> struct is_complex : std::false_type {};
template <typename T> struct is_complex<std::complex<T>> : std::true_type {};
template <typename T> struct real_type_traits { using type = T; };
template <typename T> struct real_type_traits<std::complex<T>> { using type = T; };
template <typename T> using real_type = typename real_type_traits<T>::type;
inline double conj(double x) { return x; }
inline std::complex<double> conj(std::complex<double> x) { return std::conj(x); }

}  // namespace slate

#endif
