// tests/reference_suite.cpp -- /root/reference/src/tests.rs restated against cpp/linalg.hpp:
// same inputs, same expected values, same assertions (assert_eq!/assert_ne! -> CHECK).  Built and
// run by tests/test_gpu_cpp_mirror.py on a GPU box.
#include <cmath>
#include <cstdio>
#include <numeric>
#include <vector>

#include "../cpp/linalg.hpp"

using namespace linalg;
typedef uint32_t TestType;      // tests.rs:3
static int failures = 0;
#define CHECK(cond) do { if (!(cond)) { std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); ++failures; } } while (0)
template <typename F> static bool panics(F f) { try { f(); } catch (const Panic&) { return true; } return false; }

static void f32_vec_mat_mul() {                  // tests.rs:5-21
    auto a = Vector<float>::from_vec({1, 2, 3, 4}); auto b = Vector<float>::from_vec({5, 4, 9});
    auto m = Matrix<float>::from_vec({1, 0, 0, 0, 2, 0, 0, 0, 3, 1, 0, 0}, 4, 3);
    CHECK(a * m == b);
}
static void vec_eq() {                           // tests.rs:23-44
    auto a = Vector<TestType>::from_vec({1, 2, 3, 4}); auto b = Vector<TestType>::from_vec({1, 2, 3, 5});
    auto c = Vector<TestType>::from_vec({1, 2, 3}); auto d = Vector<TestType>::from_vec({1, 2, 3, 4});
    CHECK(a == a); CHECK(b == b); CHECK(a != c); CHECK(c != a); CHECK(a != b); CHECK(b != a); CHECK(a == d); CHECK(d == a);
}
static void vec_sum() {                          // tests.rs:46-57
    for (TestType size : {1u, 2u, 8u, 16u, 64u, 128u, 971u}) {
        std::vector<TestType> v(size); std::iota(v.begin(), v.end(), 0u);
        TestType a = std::accumulate(v.begin(), v.end(), 0u);
        CHECK(a == Vector<TestType>::from_vec(v).sum());
    }
}
static void vec_vec_dot() {                      // tests.rs:59-79
    std::vector<TestType> ra, rb; TestType a = 0;
    for (TestType i = 1; i < 971; ++i) ra.push_back(i);
    for (TestType i = 972; i >= 3; --i) rb.push_back(i);
    for (size_t i = 0; i < ra.size(); ++i) a += ra[i] * rb[i];
    CHECK(a == dot(Vector<TestType>::from_vec(ra), Vector<TestType>::from_vec(rb)));
    CHECK(a == 153524810u);
}
static void mat_eq() {                           // tests.rs:81-123
    auto a = Matrix<TestType>::from_vec({0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11}, 4, 3);
    auto b = Matrix<TestType>::from_vec({0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 12}, 4, 3);
    auto c = Matrix<TestType>::from_vec({0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11}, 3, 4);
    auto d = Matrix<TestType>::from_vec({0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11}, 4, 3);
    CHECK(a == a); CHECK(b == b); CHECK(a != c); CHECK(c != a); CHECK(a != b); CHECK(b != a); CHECK(a == d); CHECK(d == a);
}
static void vec_mat_mul() {                      // tests.rs:125-141
    auto a = Vector<TestType>::from_vec({1, 2, 3, 4}); auto b = Vector<TestType>::from_vec({5, 4, 9});
    auto m = Matrix<TestType>::from_vec({1, 0, 0, 0, 2, 0, 0, 0, 3, 1, 0, 0}, 4, 3);
    CHECK(a * m == b);
}
static void mat_mat_mul() {                      // tests.rs:143-164
    auto a = Matrix<TestType>::from_vec({1, 2, 3, 4, 5, 6}, 2, 3); auto b = Matrix<TestType>::from_vec({7, 8, 9, 10, 11, 12}, 3, 2);
    auto c = Matrix<TestType>::from_vec({58, 64, 139, 154}, 2, 2);
    CHECK(c == a * b);
}
static void vec_transpose_mat_mul() {            // tests.rs:166-181
    auto a = Vector<TestType>::from_vec({1, 2, 3, 4}); auto b = Vector<TestType>::from_vec({5, 4, 9});
    auto m = Matrix<TestType>::from_vec({1, 0, 0, 1, 0, 2, 0, 0, 0, 0, 3, 0}, 3, 4);
    CHECK(mul_transpose_mat(a, m) == b);
}
static void vec_add() {                          // tests.rs:183-193
    auto a = Vector<TestType>::from_vec({1, 2, 3, 4}); auto r = Vector<TestType>::from_vec({2, 4, 6, 8});
    CHECK(r == a + a);
}
static void vec_mul_scl() {                      // tests.rs:195-206
    auto a = Vector<TestType>::from_vec({1, 2, 3, 4}); auto r = Vector<TestType>::from_vec({2, 4, 6, 8});
    CHECK(r == a * 2u);
}
static void save_load_vec() {                    // tests.rs:208-221
    const char* path = "save_load_vec.tmp";
    auto a = Vector<TestType>::from_vec({0x01234567u, 0x89ABCDEFu, 0xFEDCBA98u, 0x76543210u});
    a.save(path);
    auto b = Vector<TestType>::open(path);
    std::remove(path);
    CHECK(a == b);
}
static void save_load_mat() {                    // tests.rs:223-238
    const char* path = "save_load_mat.tmp";
    auto a = Matrix<TestType>::from_vec({0x01234567u, 0x89ABCDEFu, 0x13579BDFu, 0xFEDCBA98u, 0x76543210u, 0x02468ACEu}, 2, 3);
    a.save(path);
    auto b = Matrix<TestType>::open(path);
    std::remove(path);
    CHECK(a == b);
    CHECK(panics([&] { a.save(path); auto w = Matrix<uint64_t>::open(path); }));   // matrix.rs:97-100
    std::remove(path);
}
static void normalize_vec() {                    // tests.rs:240-252
    for (int size : {1, 2, 8, 16, 64, 128}) {
        std::vector<float> v; for (int x = 1; x <= size; ++x) v.push_back((float)x);
        auto a = Vector<float>::from_vec(v).normalized();
        auto b = Vector<float>::from_vec(v).normalize();
        CHECK(std::fabs(a.length() - 1.0f) < 0.01f); CHECK(std::fabs(b.length() - 1.0f) < 0.01f);
    }
}
static void length_vec() {                       // tests.rs:254-271
    for (int size : {1, 2, 8, 16, 64, 128}) {
        std::vector<float> v; float b = 0; for (int x = 0; x < size; ++x) { v.push_back((float)x); b += (float)x * (float)x; }
        CHECK(std::fabs(Vector<float>::from_vec(v).length() - std::sqrt(b)) < 0.0001f);
    }
}
static void vec_clone() {                        // tests.rs:273-284
    auto a = Vector<TestType>::from_vec({1, 2, 3, 4}); auto b = Vector<TestType>::from_vec({1, 2, 3, 5});
    auto c = a.clone();
    CHECK(a == c); CHECK(a != b);
}
static void panics_like_the_reference() {
    auto a = Vector<TestType>::from_vec({1, 2, 3}); auto b = Vector<TestType>::from_vec({1, 2});
    auto m = Matrix<TestType>::from_vec({1, 2, 3, 4, 5, 6}, 2, 3);
    CHECK(panics([&] { a + b; }));                         // vector.rs:97
    CHECK(panics([&] { dot(a, b); }));                     // vector.rs:327
    CHECK(panics([&] { a * m; }));                         // vector.rs:513
    CHECK(panics([&] { mul_transpose_mat(b, m); }));       // vector.rs:545
    CHECK(panics([&] { m * m; }));                         // matrix.rs:229
    CHECK(panics([&] { Matrix<TestType>::from_vec({1, 2, 3}, 2, 2); }));   // matrix.rs:40
    auto t = m.transpose();
    CHECK(t == Matrix<TestType>::from_vec({1, 4, 2, 5, 3, 6}, 3, 2));
    auto hc = mul_host<TestType>({1, 2, 3, 4, 5, 6}, {7, 8, 9, 10, 11, 12}, 2, 3, 2);
    CHECK((hc == std::vector<TestType>{58, 64, 139, 154}));
    auto o = mul_column_row(Vector<TestType>::from_vec({1, 2}), Vector<TestType>::from_vec({3, 4, 5}));
    CHECK(o == Matrix<TestType>::from_vec({3, 4, 5, 6, 8, 10}, 2, 3));
}

static void user_kernels() {                     // lib.rs:156-171 + vector.rs:80-151 (the reference has no test of these)
    const std::string src =
        "kernel void {T}_twice(global {T}* C, global {T}* B) { C[i] = B[i] * ({T})(2); }\n"
        "kernel void {T}_blend(global {T}* C, global {T}* A, global {T}* B) { C[i] = A[i] * B[i] + A[i]; }\n"
        "kernel void {T}_ramp(global {T}* C, {T} first) { C[i] = first + ({T})i; }\n";
    const std::string ty = linalg::Parameter<TestType>::type_to_str();
    linalg::Kernel<TestType> twice(ty + "_twice", src), blend(ty + "_blend", src), ramp(ty + "_ramp", src);
    twice.arg_buf_named("C").arg_buf_named("B");
    blend.arg_buf_named("C").arg_buf_named("A").arg_buf_named("B");
    ramp.arg_buf_named("C").arg_scl_named("first", (TestType)10);
    auto a = Vector<TestType>::from_vec({1, 2, 3, 4, 5}); auto b = Vector<TestType>::from_vec({5, 4, 3, 2, 1});
    CHECK(a.map(twice) == Vector<TestType>::from_vec({2, 4, 6, 8, 10}));
    CHECK(Vector<TestType>::from_for_each2(a, b, blend) == Vector<TestType>::from_vec({6, 10, 12, 12, 10}));
    auto c = a.clone(); c.for_each_mut(a, twice);
    CHECK(c == Vector<TestType>::from_vec({2, 4, 6, 8, 10}));
    auto g = Vector<TestType>::generate(ramp, 300);
    std::vector<TestType> want; for (int x = 0; x < 300; ++x) want.push_back((TestType)(10 + x));
    CHECK(g == Vector<TestType>::from_vec(want));
    g.map_mut(ramp);
    CHECK(g == Vector<TestType>::from_vec(want));
    CHECK(panics([&] { linalg::Kernel<TestType>(ty + "_absent", src); }));       // lib.rs:169 panic
    CHECK(panics([&] { linalg::Kernel<TestType>(ty + "_twice", "kernel void {T}_twice(global {T}* C) { C[i] = oops; }"); }));
    CHECK(panics([&] { Vector<TestType>::from_for_each2(a, Vector<TestType>::from_vec({1}), blend); }));
}

static void prepared_b() {                       // no counterpart in tests.rs: la_gemm_prepare_b gives the same bits as `*`
    const size_t M = 512, K = 1024, N = 512;
    std::vector<float> a(M * K), b(K * N);
    for (size_t i = 0; i < a.size(); ++i) a[i] = (float)((int)(i * 2654435761u % 2001) - 1000) / 1000.0f;
    for (size_t i = 0; i < b.size(); ++i) b[i] = (float)((int)(i * 40503u % 1999) - 999) / 999.0f;
    auto A = Matrix<float>::from_vec(a, M, K); auto B = Matrix<float>::from_vec(b, K, N);
    auto P = B.prepare();
    CHECK(A.mul_prepared(P) == A * B);
    std::vector<TestType> ai(M * K), bi(K * N);
    for (size_t i = 0; i < ai.size(); ++i) ai[i] = (TestType)(i * 2654435761u);
    for (size_t i = 0; i < bi.size(); ++i) bi[i] = (TestType)(i * 40503u + 7u);
    auto Ai = Matrix<TestType>::from_vec(ai, M, K); auto Bi = Matrix<TestType>::from_vec(bi, K, N);
    auto Pi = Bi.prepare();
    CHECK(Ai.mul_prepared(Pi) == Ai * Bi);
}

int main() {
    try {
        check(la_init(0));
        f32_vec_mat_mul(); vec_eq(); vec_sum(); vec_vec_dot(); mat_eq(); vec_mat_mul(); mat_mat_mul(); vec_transpose_mat_mul();
        vec_add(); vec_mul_scl(); save_load_vec(); save_load_mat(); normalize_vec(); length_vec(); vec_clone(); panics_like_the_reference(); user_kernels(); prepared_b();
    } catch (const std::exception& e) { std::printf("EXCEPTION: %s\n", e.what()); return 2; }
    std::printf(failures ? "%d checks failed\n" : "all reference tests passed (%d failures)\n", failures);
    return failures ? 1 : 0;
}
