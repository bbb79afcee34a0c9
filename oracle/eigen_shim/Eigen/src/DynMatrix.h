// TEST INFRASTRUCTURE ONLY. Dynamic column-major dense matrix for the Eigen shim (see ../Core).
#ifndef GRAVOMG_EIGEN_SHIM_DYNMATRIX
#define GRAVOMG_EIGEN_SHIM_DYNMATRIX

namespace Eigen {

template <typename T> class DynMatrix;

// rowwise() range with a random-access iterator (Eigen 3.4's is random-access too, which makes
// libstdc++'s std::transform_reduce take its four-way unrolled path at src/multigrid.cpp:129).
template <typename T> struct RowwiseRange {
    const DynMatrix<T>* m;
    struct iterator {
        using iterator_category = std::random_access_iterator_tag;
        using value_type = StridedView<const T>;
        using difference_type = Index;
        using pointer = void;
        using reference = StridedView<const T>;
        const DynMatrix<T>* m;
        Index i;
        reference operator*() const { return m->row(i); }
        reference operator[](Index k) const { return m->row(i + k); }
        iterator& operator++() { ++i; return *this; }
        iterator operator++(int) { iterator t = *this; ++i; return t; }
        iterator& operator--() { --i; return *this; }
        iterator& operator+=(Index k) { i += k; return *this; }
        iterator& operator-=(Index k) { i -= k; return *this; }
        iterator operator+(Index k) const { return {m, i + k}; }
        iterator operator-(Index k) const { return {m, i - k}; }
        Index operator-(const iterator& o) const { return i - o.i; }
        bool operator==(const iterator& o) const { return i == o.i; }
        bool operator!=(const iterator& o) const { return i != o.i; }
        bool operator<(const iterator& o) const { return i < o.i; }
    };
    iterator begin() const { return {m, 0}; }
    iterator end() const { return {m, m->rows()}; }
};

template <typename T> class DynMatrix {
    Index r_ = 0, c_ = 0;
    std::vector<T> d_;

public:
    using Scalar = T;
    DynMatrix() {}
    template <typename A, typename B> DynMatrix(const A& rows, const B& cols)
        : r_(Index(rows)), c_(Index(cols)), d_(size_t(Index(rows) * Index(cols))) {}
    // one-argument form: a column vector (VectorXd values{n})
    template <typename A, typename = std::enable_if_t<std::is_integral_v<A>>>
    explicit DynMatrix(const A& rows) : r_(Index(rows)), c_(1), d_(size_t(rows)) {}

    Index rows() const { return r_; }
    Index cols() const { return c_; }
    Index size() const { return r_ * c_; }
    T* data() { return d_.data(); }
    const T* data() const { return d_.data(); }
    void resize(Index r, Index c) { r_ = r; c_ = c; d_.assign(size_t(r * c), T()); }

    T& operator()(Index i, Index j) { return d_[size_t(i + j * r_)]; }
    const T& operator()(Index i, Index j) const { return d_[size_t(i + j * r_)]; }
    T& operator()(Index i) { return d_[size_t(i)]; }
    const T& operator()(Index i) const { return d_[size_t(i)]; }
    T& operator[](Index i) { return d_[size_t(i)]; }
    const T& operator[](Index i) const { return d_[size_t(i)]; }

    StridedView<T> row(Index i) { return {d_.data() + i, c_, r_}; }
    StridedView<const T> row(Index i) const { return {d_.data() + i, c_, r_}; }
    StridedView<T> col(Index j) { return {d_.data() + j * r_, r_, 1}; }
    StridedView<const T> col(Index j) const { return {d_.data() + j * r_, r_, 1}; }
    RowwiseRange<T> rowwise() const { return {this}; }

    void setConstant(const T& v) { std::fill(d_.begin(), d_.end(), v); }
    T minCoeff() const { return *std::min_element(d_.begin(), d_.end()); }
    T maxCoeff() const { return *std::max_element(d_.begin(), d_.end()); }

    DynMatrix operator-(const DynMatrix& o) const {
        assert(r_ == o.r_ && c_ == o.c_);
        DynMatrix r(r_, c_);
        for (size_t i = 0; i < d_.size(); ++i) r.d_[i] = d_[i] - o.d_[i];
        return r;
    }
    DynMatrix operator*(const T& s) const {
        DynMatrix r(r_, c_);
        for (size_t i = 0; i < d_.size(); ++i) r.d_[i] = d_[i] * s;
        return r;
    }
};

using MatrixXd = DynMatrix<double>;
using MatrixXi = DynMatrix<int>;
using VectorXd = DynMatrix<double>;
using VectorXi = DynMatrix<int>;

}  // namespace Eigen

#endif
