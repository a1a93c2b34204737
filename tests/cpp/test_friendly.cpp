// Replay of m4ri-rust's own unit tests through the C++ mirror of its friendly API
// (include/m4ri_b200.hpp) — same names, same assertions, cited by reference line.
// Every product below is an FFI call into libm4ri_b200.so and runs on the GPU.
//   m4ri-rust/src/friendly/binary_matrix.rs:588-775, binary_vector.rs:217-289
#include <cstdio>
#include <cstdlib>
#include <random>

#include "m4ri_b200.hpp"

using namespace m4ri_rust;

static int g_failed = 0, g_run = 0;
#define CHECK(cond)                                                          \
  do {                                                                       \
    if (!(cond)) {                                                           \
      std::fprintf(stderr, "FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond);   \
      g_failed++;                                                            \
    }                                                                        \
  } while (0)
#define TEST(name) static void name(); static struct name##_reg { name##_reg() { tests().push_back({#name, name}); } } name##_inst; static void name()
struct T { const char *n; void (*f)(); };
static std::vector<T> &tests() { static std::vector<T> t; return t; }

// ---- binary_matrix.rs -----------------------------------------------------------
TEST(bm_new) {  // :594-600
  BinMatrix m = BinMatrix::from_rows({BinVector::from_bools({true, false, true}),
                                      BinVector::from_bools({true, true, true})});
  CHECK(m.nrows() == 2 && m.ncols() == 3);
}
TEST(bm_identity) {  // :602-659
  std::vector<BinVector> rows;
  for (size_t i = 0; i < 10; i++) rows.push_back(BinVector::from_function(10, [i](size_t j) { return i == j; }));
  BinMatrix id = BinMatrix::from_rows(rows), id_gen = BinMatrix::identity(10);
  CHECK(id.nrows() == id_gen.nrows());
  CHECK(id.ncols() == id_gen.ncols());
  for (size_t i = 0; i < 8; i++)
    for (size_t j = 0; j < 8; j++) CHECK(id.bit(i, j) == id_gen.bit(i, j));
  CHECK(mzd_equal(id.mzd(), id_gen.mzd()) != 0);
  CHECK(id == id_gen);
}
TEST(bm_mul) {  // :661-670
  BinMatrix m1 = BinMatrix::identity(8), m2 = BinMatrix::identity(8), m3 = BinMatrix::identity(8);
  BinMatrix prod = m1 * m2;
  CHECK(mzd_equal(prod.mzd(), m3.mzd()) != 0);
}
TEST(bm_vecmul) {  // :672-686
  BinMatrix m1 = BinMatrix::identity(10);
  BinVector binvec = BinVector::from_elem(10, true);
  BinVector result = m1 * binvec;
  CHECK(result == binvec);
  result = binvec * m1;
  CHECK(result == binvec);
  BinMatrix m2 = BinMatrix::random(10, 3);
  result = binvec * m2;
  CHECK(result.len() == 3);
}
TEST(bm_test_random) { BinMatrix::random(10, 1); }  // :688-691
TEST(bm_test_as_vector_column) {  // :701-709
  for (size_t i = 1; i < 25; i++) {
    BinMatrix m1 = BinMatrix::random(i, 1);
    BinVector vec = m1.as_vector();
    CHECK(vec.len() == i);
    CHECK(m1 == vec.as_column_matrix());
  }
}
TEST(bm_test_as_vector_row) {  // :711-719
  for (size_t i = 1; i < 25; i++) {
    BinMatrix m1 = BinMatrix::random(1, i);
    BinVector vec = m1.as_vector();
    CHECK(vec.len() == i);
    CHECK(m1 == vec.as_matrix());
  }
}
TEST(bm_zero) {  // :721-729
  BinMatrix m1 = BinMatrix::zero(10, 3);
  for (size_t i = 0; i < 10; i++)
    for (size_t j = 0; j < 3; j++) CHECK(m1.bit(i, j) == false);
  bool threw = false;
  try { BinMatrix::zero(0, 3); } catch (const std::logic_error &) { threw = true; }  // :100-102
  CHECK(threw);
}
TEST(bm_set_window) {  // :731-755
  BinMatrix m1 = BinMatrix::zero(10, 10);
  m1.set_window(5, 5, BinMatrix::identity(5));
  for (size_t i = 0; i < 5; i++)
    for (size_t j = 0; j < 5; j++) CHECK(m1.bit(i, j) == false);
  for (size_t i = 5; i < 10; i++)
    for (size_t j = 5; j < 10; j++) CHECK(m1.bit(i, j) == (i == j));
  BinMatrix m2 = BinMatrix::random(10, 10);
  m2.set_window(5, 5, BinMatrix::identity(5));
  for (size_t i = 5; i < 10; i++)
    for (size_t j = 5; j < 10; j++) CHECK(m2.bit(i, j) == (i == j));
  CHECK(m2.get_window(5, 5, 10, 10) == BinMatrix::identity(5));
}
TEST(bm_test_random_unequal) {  // :757-762
  CHECK(BinMatrix::random(100, 100) != BinMatrix::random(100, 100));
}
TEST(bm_test_count_ones) {  // :764-774
  std::mt19937 rng(7);
  for (int it = 0; it < 60; it++) {
    size_t size = 1 + rng() % 999;
    BinVector v = BinVector::random(size);
    CHECK(v.count_ones() == v.as_matrix().count_ones());
    CHECK(v.count_ones() == v.as_column_matrix().count_ones());
  }
}
// ---- binary_vector.rs -----------------------------------------------------------
TEST(bv_init) { CHECK(BinVector::from_elem(10, false).len() == 10); }  // :222-227
TEST(bv_from_bytes) {  // :229-237
  CHECK(BinVector::from_bytes({0xFF}).len() == 8);
  BinVector b = BinVector::from_bytes({0x80});
  CHECK(b.get(0) == true);
  CHECK(b.get(1) == false);
}
TEST(bv_add) {  // :239-249
  BinVector a = BinVector::from_elem(10, false), b = BinVector::from_elem(10, false);
  BinVector c = a + b;
  CHECK(c.len() == 10);
  CHECK(c == BinVector::from_elem(10, false));
}
TEST(bv_mul) {  // :251-260
  CHECK((BinVector::from_elem(10, true) * BinVector::from_elem(10, false)) == false);
}
TEST(bv_as_matrix) {  // :262-269
  BinVector a = BinVector::random(10);
  BinMatrix amat = a.as_matrix();
  CHECK(amat.ncols() == 10 && amat.nrows() == 1);
  CHECK(amat.as_vector() == a);
}
TEST(bv_as_column_matrix) {  // :271-278
  BinVector a = BinVector::random(10);
  BinMatrix amat = a.as_column_matrix();
  CHECK(amat.ncols() == 1 && amat.nrows() == 10);
  CHECK(amat.as_vector() == a);
}
TEST(bv_count_ones) {  // :280-287
  CHECK(BinVector::from_elem(10, true).count_ones() == 10);
  CHECK(BinVector::from_elem(10, false).count_ones() == 0);
  CHECK(BinVector::from_bytes({0xA8}).count_ones() == 3);
}
// ---- the operator API on the reference's bench shapes (benches/binary_matrix.rs:26-80):
// checked with identities that need no second implementation -----------------------
TEST(ops_bench_shapes) {
  const size_t shapes[][3] = {{10, 10, 10}, {100, 10, 10}, {100, 10, 100}, {1000, 64, 1000},
                              {1000, 10, 1000}, {1000, 64, 1}, {1, 64, 1000}, {10, 1000, 10},
                              {1024, 1024, 1024}, {300, 2000, 130}};
  for (auto &s : shapes) {
    BinMatrix a = BinMatrix::random(s[0], s[1]), b = BinMatrix::random(s[1], s[2]),
              c = BinMatrix::random(s[1], s[2]);
    BinMatrix ab = a * b;
    CHECK(ab.nrows() == s[0] && ab.ncols() == s[2]);
    CHECK((a * (b + c)) == (ab + a * c));                        // distributivity
    CHECK(ab.transposed() == (b.transposed() * a.transposed())); // (AB)^T = B^T A^T
    CHECK((BinMatrix::identity(s[0]) * a) == a);
    CHECK((a * BinMatrix::identity(s[1])) == a);
    BinVector v = BinVector::random(s[1]);                        // A*v == column of A*[v]
    CHECK((a * v) == (a * v.as_column_matrix()).as_vector());
    BinVector u = BinVector::random(s[0]);                        // u*A == row of [u]*A
    CHECK((u * a) == (a.transposed() * u));
    BinMatrix acc = ab.clone();
    acc += ab;
    CHECK(acc == BinMatrix::zero(s[0], s[2]));
  }
  BinMatrix m = BinMatrix::random(34, 128);                       // matrix_mul_slice bench (:45-56)
  BinVector v = BinVector::random(500);
  BinMatrix r = m.mul_slice(v.get_storage());
  CHECK(r.nrows() == 34 && r.ncols() == 1);
  BinVector v128 = BinVector::from_storage(v.get_storage(), 128);
  CHECK(r.as_vector() == (m * v128));
}

// ---- the consumers of the product (binary_matrix.rs:219-268, 575-586): the reference holds no
// unit test for them, so they are checked with identities that need no second implementation ----
TEST(consumers_of_the_product) {
  BinMatrix a = BinMatrix::random(40, 70), b = BinMatrix::random(40, 9), c = BinMatrix::random(13, 70);
  BinMatrix ab = a.augmented(b), ac = a.stacked(c);
  CHECK(ab.nrows() == 40 && ab.ncols() == 79 && ac.nrows() == 53 && ac.ncols() == 70);
  for (size_t i = 0; i < 40; i++) {
    for (size_t j = 0; j < 70; j++) CHECK(ab.bit(i, j) == a.bit(i, j) && ac.bit(i, j) == a.bit(i, j));
    for (size_t j = 0; j < 9; j++) CHECK(ab.bit(i, 70 + j) == b.bit(i, j));
  }
  for (size_t i = 0; i < 13; i++)
    for (size_t j = 0; j < 70; j++) CHECK(ac.bit(40 + i, j) == c.bit(i, j));
  CHECK(BinMatrix::identity(77).rank() == 77);
  CHECK(BinMatrix::zero(5, 300).rank() == 0);
  CHECK(a.stacked(a).rank() == a.rank());                       // repeated rows add nothing
  CHECK((a.transposed() * a).rank() <= a.rank());
  // a random square matrix is invertible with probability ~0.29: try until one is
  for (int tries = 0; tries < 200; tries++) {
    BinMatrix m = BinMatrix::random(96, 96);
    if (m.rank() < 96) continue;
    BinMatrix inv = m.inverted();
    CHECK((m * inv) == BinMatrix::identity(96));
    CHECK((inv * m) == BinMatrix::identity(96));
    BinMatrix rhs = BinMatrix::random(96, 17), x = rhs.clone();
    CHECK(solve_left(m.clone(), x));
    CHECK((m * x) == rhs);                                      // AX = B
    BinMatrix e = m.clone();
    CHECK(e.echelonize() == 96 && e == BinMatrix::identity(96)); // full rank: the reduced form is I
    return;
  }
  CHECK(false && "no invertible 96 x 96 matrix in 200 tries");
}

int main() {
  if (m4ri_b200_device_count() <= 0) {
    std::fprintf(stderr, "test_friendly: no CUDA device (libm4ri_b200 has no CPU fallback)\n");
    return 2;
  }
  unsigned long long l0 = m4ri_b200_launch_count();
  for (auto &t : tests()) {
    int before = g_failed;
    t.f();
    g_run++;
    std::printf("%s %s\n", g_failed == before ? "ok  " : "FAIL", t.n);
  }
  std::printf("%d tests, %d failed checks, %llu kernel launches\n", g_run, g_failed,
              m4ri_b200_launch_count() - l0);
  return g_failed ? 1 : 0;
}
