// m4ri_b200.hpp — C++ mirror of m4ri-rust's friendly operator API on top of the
// C-ABI of libm4ri_b200.so (include/m4ri_b200.h). Header-only.
//
// No Rust toolchain exists in the build image, so the host side above the C-ABI
// is written in C++ with the reference's names, argument meaning and error
// behaviour (a Rust `panic!`/`assert!` becomes a thrown std::logic_error):
//   BinMatrix  — m4ri-rust/src/friendly/binary_matrix.rs:32-573
//   BinVector  — m4ri-rust/src/friendly/binary_vector.rs:12-215
// Every product goes through the same extern "C" call the Rust macro `mul_impl!`
// expands to (binary_matrix.rs:53-95): mzd_mul (default feature `strassen_mul`),
// mzd_mul_m4rm (`m4rm_mul`) or mzd_mul_naive (`naive_mul`), selected at compile
// time with -DM4RI_RUST_M4RM_MUL / -DM4RI_RUST_NAIVE_MUL like the cargo features.
#ifndef M4RI_B200_HPP
#define M4RI_B200_HPP

#include <cstdint>
#include <cstring>
#include <functional>
#include <random>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "m4ri_b200.h"

namespace m4ri_rust {

class BinMatrix;

// mul_impl! (binary_matrix.rs:53-95): one FFI call, extra int argument 0
inline mzd_t *mul_impl(mzd_t *C, const mzd_t *A, const mzd_t *B) {
#if defined(M4RI_RUST_M4RM_MUL)
  return mzd_mul_m4rm(C, A, B, 0);
#elif defined(M4RI_RUST_NAIVE_MUL)
  return mzd_mul_naive(C, A, B);
#else
  return mzd_mul(C, A, B, 0);
#endif
}

// ---- BinVector: a bit vector in 64-bit blocks, LSB first (Vob<usize>) -----------------
class BinVector {
 public:
  BinVector() : len_(0) {}                                    // binary_vector.rs:44
  static BinVector from_elem(size_t len, bool elem) {         // :56
    BinVector v;
    v.len_ = len;
    v.words_.assign((len + 63) / 64, elem ? ~uint64_t(0) : 0);
    v.mask_last_block();
    return v;
  }
  static BinVector from_bools(const std::vector<bool> &bools) {  // :62
    BinVector v = from_elem(bools.size(), false);
    for (size_t i = 0; i < bools.size(); i++)
      if (bools[i]) v.words_[i / 64] |= uint64_t(1) << (i % 64);
    return v;
  }
  static BinVector from_function(size_t len, const std::function<bool(size_t)> &f) {  // :77
    BinVector v = from_elem(len, false);
    for (size_t i = 0; i < len; i++)
      if (f(i)) v.words_[i / 64] |= uint64_t(1) << (i % 64);
    return v;
  }
  static BinVector random(size_t len) {                        // :87
    static thread_local std::mt19937_64 rng{std::random_device{}()};
    BinVector v = from_elem(len, false);
    for (auto &w : v.words_) w = rng();
    v.mask_last_block();
    return v;
  }
  // bytes are MSB-first within each byte, like Vob::from_bytes (:104)
  static BinVector from_bytes(const std::vector<uint8_t> &bytes) {
    BinVector v = from_elem(bytes.size() * 8, false);
    for (size_t i = 0; i < bytes.size() * 8; i++)
      if ((bytes[i / 8] >> (7 - i % 8)) & 1) v.words_[i / 64] |= uint64_t(1) << (i % 64);
    return v;
  }
  static BinVector from_storage(std::vector<uint64_t> words, size_t len) {
    BinVector v;
    v.len_ = len;
    v.words_ = std::move(words);
    v.words_.resize((len + 63) / 64);
    v.mask_last_block();
    return v;
  }

  size_t len() const { return len_; }
  bool is_empty() const { return len_ == 0; }
  bool get(size_t i) const {
    if (i >= len_) throw std::out_of_range("BinVector::get");
    return (words_[i / 64] >> (i % 64)) & 1;
  }
  void set(size_t i, bool b) {
    if (i >= len_) throw std::out_of_range("BinVector::set");
    words_[i / 64] = (words_[i / 64] & ~(uint64_t(1) << (i % 64))) | (uint64_t(b) << (i % 64));
  }
  void push(bool b) {
    if (len_ % 64 == 0) words_.push_back(0);
    len_++;
    set(len_ - 1, b);
  }
  const std::vector<uint64_t> &get_storage() const { return words_; }
  uint32_t count_ones() const {                                // :112
    uint32_t c = 0;
    for (uint64_t w : words_) c += (uint32_t)__builtin_popcountll(w);
    return c;
  }
  void extend_from_binvec(const BinVector &other) {            // :119
    for (size_t i = 0; i < other.len_; i++) push(other.get(i));
  }
  uint32_t as_u32() const {                                    // :140
    if (len_ > 32) throw std::logic_error("Can't convert this to a >32 bit number");
    return words_.empty() ? 0 : (uint32_t)words_[0];
  }
  uint64_t as_u64() const {                                    // :148
    if (len_ > 64) throw std::logic_error("Can't convert this to a >64 bit number");
    return words_.empty() ? 0 : words_[0];
  }
  inline BinMatrix as_matrix() const;                          // :130 (1 x n)
  inline BinMatrix as_column_matrix() const;                   // :135 (n x 1)

  bool operator==(const BinVector &o) const { return len_ == o.len_ && words_ == o.words_; }
  bool operator!=(const BinVector &o) const { return !(*this == o); }
  BinVector &operator+=(const BinVector &o) {                  // :178-192, XOR
    if (len_ != o.len_) throw std::logic_error("unequal length vectors");
    for (size_t i = 0; i < words_.size(); i++) words_[i] ^= o.words_[i];
    return *this;
  }
  BinVector operator+(const BinVector &o) const {              // :156-176
    BinVector r = *this;
    r += o;
    return r;
  }
  // inner product over GF(2) (:194-215)
  bool operator*(const BinVector &o) const {
    if (len_ != o.len_) throw std::logic_error("unequal length vectors");
    uint64_t acc = 0;
    for (size_t i = 0; i < words_.size(); i++) acc ^= words_[i] & o.words_[i];
    return __builtin_popcountll(acc) & 1;
  }

 private:
  void mask_last_block() {
    if (len_ % 64 && !words_.empty()) words_.back() &= (uint64_t(1) << (len_ % 64)) - 1;
  }
  std::vector<uint64_t> words_;
  size_t len_;
};

// ---- BinMatrix: owns one mzd_t (binary_matrix.rs:32-51) ---------------------------------
class BinMatrix {
 public:
  // zero matrix (binary_matrix.rs:99-106); "Can't create a 0 matrix"
  static BinMatrix zero(size_t rows, size_t cols) {
    if (rows == 0 || cols == 0) throw std::logic_error("Can't create a 0 matrix");
    return BinMatrix(nonnull(mzd_init((rci_t)rows, (rci_t)cols)));
  }
  // from row vectors (:108-121)
  static BinMatrix from_rows(const std::vector<BinVector> &rows) {
    if (rows.empty()) throw std::logic_error("Can't create a 0 matrix");
    std::vector<std::vector<uint64_t>> storage;
    for (const auto &r : rows) storage.push_back(r.get_storage());
    return from_slices(storage, rows[0].len());
  }
  // from slices of words, written straight through rows[] with the last block masked (:124-167)
  static BinMatrix from_slices(const std::vector<std::vector<uint64_t>> &rows, size_t rowlen) {
    if (rows.empty() || rowlen == 0) throw std::logic_error("Can't create a 0 matrix");
    mzd_t *p = nonnull(mzd_init((rci_t)rows.size(), (rci_t)rowlen));
    const size_t blocks_per_row = rowlen / 64 + (rowlen % 64 == 0 ? 0 : 1);
    for (size_t r = 0; r < rows.size(); r++) {
      if (rows[r].size() * 64 < rowlen) {
        mzd_free(p);
        throw std::logic_error("expected len " + std::to_string(rowlen) + " bits but got only " +
                               std::to_string(rows[r].size()) + " blocks");
      }
      for (size_t b = 0; b < blocks_per_row; b++) {
        uint64_t w = rows[r][b];
        if (b == rowlen / 64) w &= (uint64_t(1) << (rowlen % 64)) - 1;
        p->rows[r][b] = w;
      }
    }
    return BinMatrix(p);
  }
  static BinMatrix from_slice(const uint64_t *words, size_t nwords, size_t rowlen) {
    return from_slices({std::vector<uint64_t>(words, words + nwords)}, rowlen);
  }
  static BinMatrix random(size_t rows, size_t columns) {  // :192-199
    BinMatrix m = zero(rows, columns);
    mzd_randomize(m.mzd_);
    return m;
  }
  static BinMatrix from_mzd(mzd_t *p) { return BinMatrix(nonnull(p)); }  // :202-206 (takes ownership)
  static BinMatrix identity(size_t rows) {                               // :209-217
    BinMatrix m = zero(rows, rows);
    mzd_set_ui(m.mzd_, 1);
    return m;
  }

  BinMatrix(const BinMatrix &o) : mzd_(nonnull(mzd_copy(nullptr, o.mzd_))) {}  // Clone :452-457
  BinMatrix(BinMatrix &&o) noexcept : mzd_(o.mzd_) { o.mzd_ = nullptr; }
  BinMatrix &operator=(BinMatrix o) noexcept {
    std::swap(mzd_, o.mzd_);
    return *this;
  }
  ~BinMatrix() {  // Drop :41-45
    if (mzd_) mzd_free(mzd_);
  }

  size_t nrows() const { return (size_t)mzd_->nrows; }  // :285-288
  size_t ncols() const { return (size_t)mzd_->ncols; }  // :293-296
  mzd_t *mzd() const { return mzd_; }
  uint64_t get_word(size_t row, size_t column) const {  // :298-305
    if (row >= nrows() || column >= (size_t)mzd_->width) throw std::out_of_range("BinMatrix::get_word");
    return mzd_->rows[row][column];
  }
  uint64_t &get_word_mut(size_t row, size_t column) {   // :315-320
    if (row >= nrows() || column >= (size_t)mzd_->width) throw std::out_of_range("BinMatrix::get_word_mut");
    return mzd_->rows[row][column];
  }
  bool bit(size_t row, size_t col) const {              // :364-368, mzd_read_bit mzd.rs:263-269
    return (mzd_->rows[row][col / 64] >> (col % 64)) & 1;
  }
  uint32_t count_ones() const {                         // :171-189
    uint32_t c = 0;
    for (rci_t r = 0; r < mzd_->nrows; r++)
      for (wi_t w = 0; w < mzd_->width; w++) {
        uint64_t x = mzd_->rows[r][w];
        if (w == mzd_->width - 1) x &= mzd_->high_bitmask;
        c += (uint32_t)__builtin_popcountll(x);
      }
    return c;
  }
  BinMatrix transposed() const { return BinMatrix(nonnull(mzd_transpose(nullptr, mzd_))); }  // :272-279
  // [A] [B] => [A B] (:219-232) and A over B (:234-245)
  BinMatrix augmented(const BinMatrix &o) const {
    if (nrows() != o.nrows()) throw std::logic_error("The rows need to be equal");
    return BinMatrix(nonnull(mzd_concat(nullptr, mzd_, o.mzd_)));
  }
  BinMatrix stacked(const BinMatrix &o) const { return BinMatrix(nonnull(mzd_stack(nullptr, mzd_, o.mzd_))); }
  // echelonize in place, return the rank (:255-261); rank() does it on a clone (:247-253)
  size_t echelonize() { return (size_t)mzd_echelonize(mzd_, 0); }
  size_t rank() const { return clone().echelonize(); }
  // the inverse as a new matrix (:263-268); "null mzd_t" when the matrix is singular, as nonnull! panics
  BinMatrix inverted() const { return BinMatrix(nonnull(mzd_inv_m4ri(nullptr, mzd_, 0))); }
  BinMatrix clone() const { return BinMatrix(*this); }

  // 1 x n or n x 1 matrix -> vector; the column case transposes first (:332-361)
  BinVector as_vector() const {
    if (nrows() != 1) {
      if (ncols() != 1) throw std::logic_error("needs to have only one column or row");
      return transposed().as_vector();
    }
    std::vector<uint64_t> words(mzd_->rows[0], mzd_->rows[0] + mzd_->width);
    return BinVector::from_storage(std::move(words), ncols());
  }
  // copy of a window, bit by bit like the reference (:371-392)
  BinMatrix get_window(size_t start_row, size_t start_col, size_t high_row, size_t high_col) const {
    BinMatrix w = zero(high_row - start_row, high_col - start_col);
    for (size_t i = start_row; i < high_row; i++)
      for (size_t j = start_col; j < high_col; j++)
        if (bit(i, j)) w.mzd_->rows[i - start_row][(j - start_col) / 64] |= uint64_t(1) << ((j - start_col) % 64);
    return w;
  }
  void set_window(size_t start_row, size_t start_col, const BinMatrix &other) {  // :398-413
    const size_t highr = start_row + other.nrows(), highc = start_col + other.ncols();
    if (ncols() < highc) throw std::logic_error("This matrix is too small!");
    if (nrows() < highr) throw std::logic_error("This matrix has too few rows !");
    for (size_t r = start_row; r < highr; r++)
      for (size_t c = start_col; c < highc; c++) {
        uint64_t &w = mzd_->rows[r][c / 64];
        w = (w & ~(uint64_t(1) << (c % 64))) | (uint64_t(other.bit(r - start_row, c - start_col)) << (c % 64));
      }
  }
  // A * x^T for x given as words: 1 x n matrix, transposed, then mzd_mul_naive — the exact
  // call sequence of the reference (:416-431). Result is nrows x 1.
  BinMatrix mul_slice(const uint64_t *other, size_t nwords) const {
    if (ncols() > nwords * 64)
      throw std::logic_error("Mismatched sizes: (" + std::to_string(nrows()) + "x" + std::to_string(ncols()) +
                             ") * (" + std::to_string(nwords * 64) + "x1) (too big)");
    BinMatrix x = from_slice(other, nwords, ncols()).transposed();
    return from_mzd(mzd_mul_naive(nullptr, mzd_, x.mzd_));
  }
  BinMatrix mul_slice(const std::vector<uint64_t> &other) const { return mul_slice(other.data(), other.size()); }

  bool operator==(const BinMatrix &o) const { return mzd_equal(mzd_, o.mzd_) != 0; }  // :434-438
  bool operator!=(const BinMatrix &o) const { return !(*this == o); }

  // &A * &B: one FFI call, NULL result = "Multiplication failed" (:459-472)
  BinMatrix operator*(const BinMatrix &o) const {
    mzd_t *p = mul_impl(nullptr, mzd_, o.mzd_);
    if (!p) throw std::logic_error("Multiplication failed");
    return BinMatrix(p);
  }
  BinMatrix operator+(const BinMatrix &o) const {  // :474-489
    return BinMatrix(nonnull(mzd_add(nullptr, mzd_, o.mzd_)));
  }
  BinMatrix &operator+=(const BinMatrix &o) {      // :507-526: mzd_add with C aliasing A
    mzd_add(mzd_, mzd_, o.mzd_);
    return *this;
  }
  // &A * &v through mul_slice and as_vector (:528-550)
  BinVector operator*(const BinVector &v) const {
    return mul_slice(v.get_storage()).as_vector();
  }

 private:
  explicit BinMatrix(mzd_t *p) : mzd_(p) {}
  static mzd_t *nonnull(mzd_t *p) {
    if (!p) throw std::logic_error("null mzd_t from libm4ri_b200");
    return p;
  }
  mzd_t *mzd_;
};

inline BinMatrix BinVector::as_matrix() const { return BinMatrix::from_rows({*this}); }
inline BinMatrix BinVector::as_column_matrix() const { return as_matrix().transposed(); }

// Solve AX = B for X; B holds the solution afterwards; true on success (binary_matrix.rs:575-586)
inline bool solve_left(BinMatrix a, BinMatrix &b) { return mzd_solve_left(a.mzd(), b.mzd(), 0, 1) == 0; }

// &v * &A: a 1 x n matrix through mul_impl! (binary_matrix.rs:552-573)
inline BinVector operator*(const BinVector &v, const BinMatrix &m) {
  return (v.as_matrix() * m).as_vector();
}

}  // namespace m4ri_rust
#endif  // M4RI_B200_HPP
