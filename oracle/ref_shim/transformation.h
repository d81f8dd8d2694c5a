// TEST INFRASTRUCTURE (oracle/): stand-in for the reference's
// utils/modules_to_build/transformations/include/transformation.h, found first on the include
// path when oracle/Makefile compiles the reference's OWN src/transformation.cpp where it lies
// under /root/reference.  The reference header pulls in pybind11/eigen.h and Eigen3; Eigen is not
// installed in this image, and map_events (transformation.cpp:35-85) uses it for storage only --
// rows(), cols(), data(), row(i).data(), IsRowMajor, Zero(r, c); no arithmetic.  This header
// provides exactly those container members, so the function body that is compiled and run is the
// reference's, byte for byte.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <tuple>
#include <unordered_map>
#include <utility>
#include <vector>

namespace pybind11 { namespace literals {} }

namespace Eigen {
enum { Dynamic = -1 };
enum { ColMajor = 0, RowMajor = 1 };

template <class S> struct RowView {
    S* p;
    S* data() const { return p; }
};

template <class S, int R, int C, int Opt = ColMajor> class Array {
public:
    typedef S Scalar;
    enum { IsRowMajor = (Opt & RowMajor) ? 1 : 0 };
    Array() : rows_(0), cols_(0) {}
    Array(long r, long c) : rows_(r), cols_(c), store_((size_t)r * (size_t)c, S(0)) {}
    static Array Zero(long r, long c) { return Array(r, c); }
    long rows() const { return rows_; }
    long cols() const { return cols_; }
    S* data() { return store_.data(); }
    const S* data() const { return store_.data(); }
    RowView<S> row(long i) { return RowView<S>{store_.data() + (size_t)i * (size_t)cols_}; }   // row-major storage
private:
    long rows_, cols_;
    std::vector<S> store_;
};
template <class S, int R, int C, int Opt = ColMajor> using Matrix = Array<S, R, C, Opt>;

// non-owning view of caller memory with the layout of T (what pybind11 hands to an Eigen::Ref argument)
template <class T> class Ref {
public:
    typedef typename T::Scalar S;
    enum { IsRowMajor = T::IsRowMajor };
    Ref(S* p, long r, long c) : p_(p), rows_(r), cols_(c) {}
    long rows() const { return rows_; }
    long cols() const { return cols_; }
    S* data() const { return p_; }
    RowView<S> row(long i) const { return RowView<S>{p_ + (size_t)i * (size_t)cols_}; }
private:
    S* p_;
    long rows_, cols_;
};
}  // namespace Eigen

typedef Eigen::Array<float_t, Eigen::Dynamic, Eigen::Dynamic, Eigen::RowMajor> Event_t;
typedef Eigen::Matrix<uint64_t, Eigen::Dynamic, 1> VectorXull;

Event_t map_events(Eigen::Ref<Event_t>& in_events, std::tuple<int, int> in_shape,
                   Eigen::Ref<VectorXull>& in_src, Eigen::Ref<VectorXull>& in_dst);
