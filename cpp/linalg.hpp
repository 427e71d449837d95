// cpp/linalg.hpp -- header-only C++ mirror of the reference's public surface over the C ABI.
//
// The reference is compiled code (Rust); its toolchain is absent from this image, so the host
// side above the C ABI is mirrored here in C++ with the same names, argument meaning and error
// behaviour: linalg::Vector<T> (src/vector.rs), linalg::Matrix<T> (src/matrix.rs), dot(),
// mul_transpose_mat(), mul_column_row(), Parameter<T>/Real (src/traits.rs).  Shape mismatches
// throw linalg::Panic where the reference's assert_eq! panics; every operator blocks until the
// device is done (event.wait_for(), vector.rs:108); out-of-place operators allocate a fresh result.
// Rust's `&a + &b` is `a + b` here; `a += &b` is `a += b`.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../include/linalg_b200.h"

namespace linalg {

struct Panic : std::runtime_error { using std::runtime_error::runtime_error; };

inline void check(la_status s) {           // the reference unwrap()s every OpenCL call
    if (s != LA_OK) throw Panic("liblinalg_b200: status " + std::to_string(s) + ": " + la_last_error());
}
inline void assert_eq(size_t left, size_t right) {
    if (left != right) throw Panic("assertion failed: `(left == right)` left: `" + std::to_string(left) + "`, right: `" + std::to_string(right) + "`");
}

// ---- traits.rs:13-57: element type -> device type
template <typename T> struct Parameter;
#define LINALG_PARAM(T, CODE, NAME) \
    template <> struct Parameter<T> { static constexpr int DTYPE = CODE; static const char* type_to_str() { return NAME; } };
LINALG_PARAM(uint8_t, LA_U8, "uchar")
LINALG_PARAM(int8_t, LA_U8, "uchar")        // traits.rs:36: i8 -> "uchar" (kept for drop-in parity)
LINALG_PARAM(uint16_t, LA_U16, "ushort")
LINALG_PARAM(int16_t, LA_I16, "short")
LINALG_PARAM(uint32_t, LA_U32, "uint")
LINALG_PARAM(int32_t, LA_I32, "int")
LINALG_PARAM(uint64_t, LA_U64, "ulong")
LINALG_PARAM(int64_t, LA_I64, "long")
LINALG_PARAM(float, LA_F32, "float")
LINALG_PARAM(double, LA_F64, "double")
#undef LINALG_PARAM

template <typename T> class Matrix;
template <typename T> class Vector;

// ---- user kernels: create_kernel (lib.rs:156-171) and the named-argument `ocl::Kernel` it returns.  `source` is the
// text of the user's src/extra_kernels.cl ({T}, IS_<TYPE>, `i`: lib.rs:158-163); declare the arguments in
// signature order with arg_buf_named / arg_scl_named, as lib.rs:189-274 does for the built-in kernels.
template <typename T>
class Kernel {
  public:
    Kernel(const std::string& kernel_name, const std::string& source) {
        check(la_user_kernel_create(Parameter<T>::DTYPE, source.c_str(), kernel_name.c_str(), &k_));
    }
    Kernel(const Kernel&) = delete;
    Kernel& operator=(const Kernel&) = delete;
    Kernel(Kernel&& o) noexcept : k_(o.k_), names_(std::move(o.names_)), slots_(std::move(o.slots_)) { o.k_ = nullptr; }
    ~Kernel() { if (k_) la_user_kernel_destroy(k_); }

    Kernel& arg_buf_named(const std::string& name) { names_.push_back(name); slots_.emplace_back(sizeof(void*), 0); return *this; }
    template <typename S> Kernel& arg_scl_named(const std::string& name, S value) {
        names_.push_back(name); slots_.emplace_back(sizeof(S), 0); return set_arg_scl_named(name, value);
    }
    Kernel& set_arg_buf_named(const std::string& name, const void* dev_ptr) { return store(name, &dev_ptr, sizeof(dev_ptr)); }
    template <typename S> Kernel& set_arg_scl_named(const std::string& name, S value) { return store(name, &value, sizeof(S)); }
    void enq(size_t gws) {                                          // kernel.cmd().gws(n).enq() + event.wait_for()
        std::vector<void*> argv;
        for (auto& sl : slots_) argv.push_back(sl.data());
        check(la_user_kernel_launch(k_, gws, argv.data(), nullptr));
        check(la_stream_sync(nullptr));
    }

  private:
    Kernel& store(const std::string& name, const void* bytes, size_t n) {
        for (size_t j = 0; j < names_.size(); ++j)
            if (names_[j] == name) {
                if (slots_[j].size() != n) throw Panic("kernel argument '" + name + "' has another size");
                std::memcpy(slots_[j].data(), bytes, n);
                return *this;
            }
        throw Panic("kernel has no argument named '" + name + "'");
    }
    la_user_kernel* k_ = nullptr;
    std::vector<std::string> names_;
    std::vector<std::vector<unsigned char>> slots_;
};

inline std::string load_extra_src(const std::string& path = "src/extra_kernels.cl") {      // lib.rs:173-180; missing file = ""
    std::string text;
    if (std::FILE* f = std::fopen(path.c_str(), "rb")) {
        char buf[4096]; size_t n;
        while ((n = std::fread(buf, 1, sizeof(buf), f)) > 0) text.append(buf, n);
        std::fclose(f);
    }
    return text;
}
template <typename T> Kernel<T> create_kernel(const std::string& kernel_name) { return Kernel<T>(kernel_name, load_extra_src()); }   // lib.rs:165-171

template <typename T>
class Vector {                              // vector.rs:23-27
  public:
    Vector(const Vector&) = delete;
    Vector& operator=(const Vector&) = delete;
    Vector(Vector&& o) noexcept : ptr_(o.ptr_), len_(o.len_) { o.ptr_ = nullptr; o.len_ = 0; }
    Vector& operator=(Vector&& o) noexcept { std::swap(ptr_, o.ptr_); std::swap(len_, o.len_); return *this; }
    ~Vector() { if (ptr_) la_free(ptr_); }

    static Vector uninitialized(size_t size) {                      // vector.rs:36-51
        void* p = nullptr; check(la_malloc(&p, size * sizeof(T))); return Vector(p, size);
    }
    static Vector create(T default_val, size_t size) {              // Vector::new, vector.rs:30-34
        Vector v = uninitialized(size); check(la_fill(D, v.ptr_, &default_val, size, nullptr)); sync(); return v;
    }
    static Vector from_vec(const std::vector<T>& v) {               // vector.rs:53-69
        Vector r = uninitialized(v.size()); check(la_upload(r.ptr_, v.data(), v.size() * sizeof(T), nullptr)); sync(); return r;
    }
    std::vector<T> to_vec() const {                                 // vector.rs:71-78
        std::vector<T> out(len_); check(la_download(out.data(), ptr_, len_ * sizeof(T), nullptr)); return out;
    }
    size_t len() const { return len_; }
    void* ptr() const { return ptr_; }
    Vector clone() const {                                          // vector.rs:700-708
        Vector r = uninitialized(len_); check(la_copy(r.ptr_, ptr_, len_ * sizeof(T), nullptr)); sync(); return r;
    }

    // vector.rs:366-643
    Vector operator+(const Vector& o) const { return each2(o, LA_ADD); }
    Vector operator-(const Vector& o) const { return each2(o, LA_SUB); }
    Vector operator*(const Vector& o) const { return each2(o, LA_MUL); }
    Vector operator/(const Vector& o) const { return each2(o, LA_DIV); }
    Vector& operator+=(const Vector& o) { return each_mut(o, LA_ADD); }
    Vector& operator-=(const Vector& o) { return each_mut(o, LA_SUB); }
    Vector& operator*=(const Vector& o) { return each_mut(o, LA_MUL); }
    Vector& operator/=(const Vector& o) { return each_mut(o, LA_DIV); }
    Vector operator*(T s) const { return scl(s, LA_MUL); }
    Vector operator/(T s) const { return scl(s, LA_DIV); }
    Vector& operator*=(T s) { check(la_ew_assign_vec_scl(D, LA_MUL, ptr_, &s, len_, nullptr)); sync(); return *this; }
    Vector& operator/=(T s) { check(la_ew_assign_vec_scl(D, LA_DIV, ptr_, &s, len_, nullptr)); sync(); return *this; }
    Vector operator*(const Matrix<T>& m) const;                     // vector.rs:508-539

    bool operator==(const Vector& o) const {                        // vector.rs:645-675
        if (len_ != o.len_) return false;
        int eq = 0; check(la_eq_vec(D, ptr_, o.ptr_, len_, &eq, nullptr)); return eq != 0;
    }
    bool operator!=(const Vector& o) const { return !(*this == o); }

    T sum() const { T out{}; check(la_sum_vec(D, ptr_, len_, &out, nullptr)); return out; }      // vector.rs:256-293

    // persistence (vector.rs:183-245, util.rs:6-21): native-endian u64 element size, u64 count, raw data
    void write_to_file(std::FILE* f) const {
        const uint64_t hdr[2] = {sizeof(T), (uint64_t)len_};
        const std::vector<T> v = to_vec();
        if (std::fwrite(hdr, 8, 2, f) != 2 || std::fwrite(v.data(), sizeof(T), v.size(), f) != v.size()) throw Panic("write failed");
    }
    static Vector read_from_file(std::FILE* f) {
        uint64_t hdr[2];
        if (std::fread(hdr, 8, 2, f) != 2) throw Panic("Failed to read all data from file");
        if (hdr[0] != sizeof(T)) throw Panic("Elem size from buffer does not seem to match, what was expected!Missmatch in endiannes?");   // vector.rs:215-220
        std::vector<T> v(hdr[1]);
        if (std::fread(v.data(), sizeof(T), v.size(), f) != v.size()) throw Panic("Failed to read all data from file");
        return from_vec(v);
    }
    void save(const std::string& path) const { std::FILE* f = std::fopen(path.c_str(), "wb"); if (!f) throw Panic("cannot create " + path); write_to_file(f); std::fclose(f); }
    static Vector open(const std::string& path) {
        std::FILE* f = std::fopen(path.c_str(), "rb"); if (!f) throw Panic("Failed to open file");
        Vector v = read_from_file(f); std::fclose(f); return v;
    }
    // user kernels (vector.rs:80-151): slots named "C" (result / self), "A", "B"
    static Vector generate(Kernel<T>& k, size_t size) {             // vector.rs:80-90
        Vector r = uninitialized(size); k.set_arg_buf_named("C", r.ptr_).enq(size); return r;
    }
    static Vector from_for_each2(const Vector& a, const Vector& b, Kernel<T>& k) {              // vector.rs:96-112
        assert_eq(a.len_, b.len_);
        Vector r = uninitialized(a.len_);
        k.set_arg_buf_named("C", r.ptr_).set_arg_buf_named("A", a.ptr_).set_arg_buf_named("B", b.ptr_).enq(r.len_);
        return r;
    }
    void for_each_mut(const Vector& o, Kernel<T>& k) {              // vector.rs:114-125
        assert_eq(len_, o.len_);
        k.set_arg_buf_named("C", ptr_).set_arg_buf_named("B", o.ptr_).enq(len_);
    }
    void map_mut(Kernel<T>& k) { k.set_arg_buf_named("C", ptr_).enq(len_); }                   // vector.rs:127-136
    Vector map(Kernel<T>& k) const {                                // vector.rs:138-151
        Vector r = uninitialized(len_); k.set_arg_buf_named("C", r.ptr_).set_arg_buf_named("B", ptr_).enq(len_); return r;
    }

    T length() const;                                               // vector.rs:303-305
    Vector normalized() const { T inv = T(1) / length(); return *this * inv; }                  // vector.rs:320-323
    Vector normalize() && { T inv = T(1) / length(); *this *= inv; return std::move(*this); }   // vector.rs:313-317

  private:
    static constexpr int D = Parameter<T>::DTYPE;
    Vector(void* p, size_t n) : ptr_(p), len_(n) {}
    static void sync() { check(la_stream_sync(nullptr)); }
    Vector each2(const Vector& o, int op) const {                   // from_for_each2_lock_free, vector.rs:96-112
        assert_eq(len_, o.len_);
        Vector r = uninitialized(len_); check(la_ew_vec_vec(D, op, r.ptr_, ptr_, o.ptr_, len_, nullptr)); sync(); return r;
    }
    Vector& each_mut(const Vector& o, int op) {                     // for_each_mut, vector.rs:114-125
        assert_eq(len_, o.len_);
        check(la_ew_assign_vec_vec(D, op, ptr_, o.ptr_, len_, nullptr)); sync(); return *this;
    }
    Vector scl(T s, int op) const {
        Vector r = uninitialized(len_); check(la_ew_vec_scl(D, op, r.ptr_, ptr_, &s, len_, nullptr)); sync(); return r;
    }
    void* ptr_; size_t len_;
    template <typename U> friend class Matrix;
    template <typename U> friend U dot(const Vector<U>&, const Vector<U>&);
    template <typename U> friend Vector<U> mul_transpose_mat(const Vector<U>&, const Matrix<U>&);
    template <typename U> friend Matrix<U> mul_column_row(const Vector<U>&, const Vector<U>&);
};

template <typename T>
T dot(const Vector<T>& a, const Vector<T>& b) {                     // vector.rs:326-363
    assert_eq(a.len(), b.len());
    T out{}; check(la_dot_vec_vec(Parameter<T>::DTYPE, a.ptr_, b.ptr_, a.len_, &out, nullptr)); return out;
}
template <typename T> T Vector<T>::length() const { return std::sqrt(dot(*this, *this)); }

template <typename T>
class Matrix {                                                      // matrix.rs:9-13
  public:
    Matrix(Matrix&&) noexcept = default;
    Matrix& operator=(Matrix&&) noexcept = default;
    static Matrix create(T v, size_t rows, size_t cols) { return Matrix(Vector<T>::create(v, rows * cols), rows, cols); }
    static Matrix uninitialized(size_t rows, size_t cols) { return Matrix(Vector<T>::uninitialized(rows * cols), rows, cols); }
    static Matrix from_vec(const std::vector<T>& v, size_t rows, size_t cols) {     // matrix.rs:39-46
        assert_eq(v.size(), rows * cols); return Matrix(Vector<T>::from_vec(v), rows, cols);
    }
    std::pair<size_t, size_t> len() const { return {rows_, cols_}; }
    size_t get_row_count() const { return rows_; }
    size_t get_col_count() const { return cols_; }
    std::vector<T> to_vec() const { return data_.to_vec(); }
    const Vector<T>& data() const { return data_; }

    // matrix.rs:134-218, 268-277: shape asserts, then delegate to Vector
    Matrix operator+(const Matrix& o) const { same(o); return Matrix(data_ + o.data_, rows_, cols_); }
    Matrix operator-(const Matrix& o) const { same(o); return Matrix(data_ - o.data_, rows_, cols_); }
    Matrix& operator+=(const Matrix& o) { same(o); data_ += o.data_; return *this; }
    Matrix& operator-=(const Matrix& o) { same(o); data_ -= o.data_; return *this; }
    Matrix operator*(T s) const { return Matrix(data_ * s, rows_, cols_); }
    Matrix& operator*=(T s) { data_ *= s; return *this; }
    Matrix operator/(T s) const { return Matrix(data_ / s, rows_, cols_); }
    Matrix operator*(const Matrix& o) const {                       // matrix.rs:221-232
        assert_eq(cols_, o.rows_); return mul_helper(data_.ptr_, o.data_.ptr_, rows_, cols_, o.cols_);
    }
    bool operator==(const Matrix& o) const { return rows_ == o.rows_ && cols_ == o.cols_ && data_ == o.data_; }   // matrix.rs:279-287
    bool operator!=(const Matrix& o) const { return !(*this == o); }
    // persistence (matrix.rs:76-130): u64 element size, u64 rows, u64 cols, raw row-major data
    void save(const std::string& path) const {
        std::FILE* f = std::fopen(path.c_str(), "wb"); if (!f) throw Panic("cannot create " + path);
        const uint64_t hdr[3] = {sizeof(T), (uint64_t)rows_, (uint64_t)cols_};
        const std::vector<T> v = to_vec();
        const bool ok = std::fwrite(hdr, 8, 3, f) == 3 && std::fwrite(v.data(), sizeof(T), v.size(), f) == v.size();
        std::fclose(f);
        if (!ok) throw Panic("write failed");
    }
    static Matrix open(const std::string& path) {
        std::FILE* f = std::fopen(path.c_str(), "rb"); if (!f) throw Panic("Failed to open file");
        uint64_t hdr[3];
        if (std::fread(hdr, 8, 3, f) != 3) { std::fclose(f); throw Panic("Failed to read all data from file"); }
        if (hdr[0] != sizeof(T)) { std::fclose(f); throw Panic("Elem size from buffer does not seem to match, what was expected!Missmatch in endiannes?"); }   // matrix.rs:97-100
        std::vector<T> v(hdr[1] * hdr[2]);
        const bool ok = std::fread(v.data(), sizeof(T), v.size(), f) == v.size();
        std::fclose(f);
        if (!ok) throw Panic("Failed to read all data from file");
        return from_vec(v, hdr[1], hdr[2]);
    }
    // One B, many products (la_gemm_prepare_b; no counterpart in the reference): `a.mul_prepared(b.prepare())` == `a * b`, bit for
    // bit, with the part of the product that only depends on b done once.  The handle borrows b: b must outlive it unchanged.
    class Prepared {
      public:
        explicit Prepared(const Matrix& b) : rows_(b.rows_), cols_(b.cols_) { check(la_gemm_prepare_b(Parameter<T>::DTYPE, b.data_.ptr_, b.rows_, b.cols_, nullptr, &h_)); }
        Prepared(const Prepared&) = delete;
        Prepared& operator=(const Prepared&) = delete;
        Prepared(Prepared&& o) noexcept : h_(o.h_), rows_(o.rows_), cols_(o.cols_) { o.h_ = nullptr; }
        ~Prepared() { if (h_) la_gemm_prepared_free(h_); }
      private:
        la_gemm_prepared* h_ = nullptr; size_t rows_, cols_;
        friend class Matrix;
    };
    Prepared prepare() const { return Prepared(*this); }
    Matrix mul_prepared(const Prepared& b) const {
        assert_eq(cols_, b.rows_);
        Matrix c = uninitialized(rows_, b.cols_);
        check(la_mul_mat_mat_prepared(c.data_.ptr_, data_.ptr_, b.h_, rows_, nullptr));
        check(la_stream_sync(nullptr));
        return c;
    }
    Matrix transpose() const {                                      // not in the reference (north star)
        Matrix r = uninitialized(cols_, rows_);
        check(la_transpose(Parameter<T>::DTYPE, r.data_.ptr_, data_.ptr_, rows_, cols_, nullptr)); check(la_stream_sync(nullptr)); return r;
    }
    static Matrix mul_helper(const void* a, const void* b, size_t a_rows, size_t a_cols, size_t b_cols) {   // matrix.rs:240-265
        Matrix r = uninitialized(a_rows, b_cols);
        check(la_mul_mat_mat(Parameter<T>::DTYPE, r.data_.ptr_, a, b, a_rows, a_cols, b_cols, nullptr)); check(la_stream_sync(nullptr)); return r;
    }

  private:
    Matrix(Vector<T>&& d, size_t r, size_t c) : data_(std::move(d)), rows_(r), cols_(c) {}
    void same(const Matrix& o) const { assert_eq(rows_, o.rows_); assert_eq(cols_, o.cols_); }
    Vector<T> data_; size_t rows_, cols_;
    template <typename U> friend class Vector;
    template <typename U> friend Vector<U> mul_transpose_mat(const Vector<U>&, const Matrix<U>&);
};

template <typename T>
Vector<T> Vector<T>::operator*(const Matrix<T>& m) const {          // vector.rs:508-539
    assert_eq(len_, m.get_row_count());
    Vector r = uninitialized(m.get_col_count());
    check(la_mul_vec_mat(D, r.ptr_, ptr_, m.data_.ptr_, m.rows_, m.cols_, nullptr)); sync(); return r;
}
template <typename T>
Vector<T> mul_transpose_mat(const Vector<T>& v, const Matrix<T>& m) {   // vector.rs:542-569
    assert_eq(v.len(), m.get_col_count());
    Vector<T> r = Vector<T>::uninitialized(m.get_row_count());
    check(la_mul_vec_transpose_mat(Parameter<T>::DTYPE, r.ptr_, v.ptr_, m.data_.ptr_, m.rows_, m.cols_, nullptr));
    check(la_stream_sync(nullptr)); return r;
}
template <typename T>
Matrix<T> mul_column_row(const Vector<T>& column, const Vector<T>& row) {   // matrix.rs:236-238
    return Matrix<T>::mul_helper(column.ptr_, row.ptr_, column.len(), 1, row.len());
}

// C = A x B on host vectors in one streamed call (la_mul_mat_mat_host): from_vec, from_vec, *, to_vec with the
// copies overlapped with the multiply.
template <typename T>
std::vector<T> mul_host(const std::vector<T>& a, const std::vector<T>& b, size_t a_rows, size_t a_cols, size_t b_cols) {
    assert_eq(a.size(), a_rows * a_cols); assert_eq(b.size(), a_cols * b_cols);
    std::vector<T> c(a_rows * b_cols);
    check(la_mul_mat_mat_host(Parameter<T>::DTYPE, c.data(), a.data(), b.data(), a_rows, a_cols, b_cols));
    return c;
}

}  // namespace linalg
