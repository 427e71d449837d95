// tests/comm_two_proc.cpp -- the multi-GPU entry points of include/linalg_b200.h driven from plain C++, one process per GPU,
// with NO torch / NCCL / MPI anywhere: the `exchange` callback la_comm_create asks for is an all-gather through files in a
// scratch directory.  tests/test_gpu_comm_cpp.py builds this with g++ and starts `world` copies.
//   usage: comm_two_proc <rank> <world> <scratch dir>
// Every rank builds the same global operands from a counter-based generator, keeps its row block, and checks the sharded
// results against plain host loops over the global operands: integers bit for bit (wrap-around arithmetic), f32 dot within
// 2^-20 of sum|a||b|.
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../include/linalg_b200.h"

#define CHECK(call) do { la_status s_ = (call); if (s_ != LA_OK) { fprintf(stderr, "rank %d: %s failed: %s\n", g_rank, #call, la_last_error()); exit(2); } } while (0)
#define EXPECT(cond, what) do { if (!(cond)) { fprintf(stderr, "rank %d: CHECK FAILED: %s (%s:%d)\n", g_rank, what, __FILE__, __LINE__); exit(3); } } while (0)

static int g_rank = 0, g_world = 1;
static std::string g_dir;
static unsigned long g_round = 0;

// all-gather through files: rank r writes <dir>/<round>.<r>.tmp, renames it to <round>.<r>, then reads everybody's.
static int file_exchange(void*, const void* send, void* recv, size_t bytes) {
    const unsigned long round = g_round++;
    const std::string mine = g_dir + "/" + std::to_string(round) + "." + std::to_string(g_rank);
    FILE* f = fopen((mine + ".tmp").c_str(), "wb");
    if (!f) return 1;
    if (bytes && fwrite(send, 1, bytes, f) != bytes) { fclose(f); return 2; }
    fclose(f);
    if (rename((mine + ".tmp").c_str(), mine.c_str()) != 0) return 3;
    for (int p = 0; p < g_world; ++p) {
        const std::string theirs = g_dir + "/" + std::to_string(round) + "." + std::to_string(p);
        const auto t0 = std::chrono::steady_clock::now();
        for (;;) {
            FILE* g = fopen(theirs.c_str(), "rb");
            if (g) {
                const size_t got = bytes ? fread((char*)recv + (size_t)p * bytes, 1, bytes, g) : 0;
                fclose(g);
                if (got == bytes) break;
            }
            if (std::chrono::steady_clock::now() - t0 > std::chrono::seconds(60)) return 4;
            std::this_thread::sleep_for(std::chrono::microseconds(200));
        }
    }
    return 0;
}

static uint64_t mix(uint64_t x) { x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull; return x ^ (x >> 31); }
static std::vector<int32_t> ints(size_t n, uint64_t seed) { std::vector<int32_t> v(n); for (size_t i = 0; i < n; ++i) v[i] = (int32_t)(mix(seed ^ i) >> 32); return v; }
static std::vector<float> floats(size_t n, uint64_t seed) { std::vector<float> v(n); for (size_t i = 0; i < n; ++i) v[i] = (float)((double)(mix(seed ^ i) >> 40) / 8388608.0 - 1.0); return v; }

template <typename T> static void* to_device(const T* host, size_t n) {
    void* d = nullptr; CHECK(la_malloc(&d, n * sizeof(T)));
    CHECK(la_upload(d, host, n * sizeof(T), nullptr)); CHECK(la_stream_sync(nullptr));
    return d;
}

int main(int argc, char** argv) {
    if (argc != 4) { fprintf(stderr, "usage: %s rank world dir\n", argv[0]); return 1; }
    g_rank = atoi(argv[1]); g_world = atoi(argv[2]); g_dir = argv[3];
    int ndev = 0; CHECK(la_device_count(&ndev));
    EXPECT(ndev >= g_world, "one GPU per rank");
    CHECK(la_init(g_rank));
    la_comm* comm = nullptr;
    CHECK(la_comm_create(g_world, g_rank, file_exchange, nullptr, &comm));
    int w = 0, r = 0, peer = 0; CHECK(la_comm_info(comm, &w, &r, &peer));
    EXPECT(w == g_world && r == g_rank, "la_comm_info");
    const bool no_peer = getenv("LA_COMM_NO_PEER_MEMORY") != nullptr;       // second run of the test: the fallbacks through the callback
    EXPECT(peer == (no_peer ? 0 : 1), "peer memory path as expected");

    // ---- sum / dot, i32 (bit-exact mod 2^32) and f32
    const size_t n = (1u << 20) + 77;
    size_t lo, hi; la_partition(n, g_world, g_rank, &lo, &hi);
    {
        const auto a = ints(n, 1), b = ints(n, 2);
        uint32_t sum = 0, dot = 0;
        for (size_t i = 0; i < n; ++i) { sum += (uint32_t)a[i]; dot += (uint32_t)a[i] * (uint32_t)b[i]; }
        void* A = to_device(a.data() + lo, hi - lo); void* B = to_device(b.data() + lo, hi - lo);
        for (int round = 0; round < 20; ++round) {            // many rounds: the two mailbox parities get reused
            int32_t got = 0;
            CHECK(la_sum_vec_sharded(comm, LA_I32, A, hi - lo, &got, nullptr)); EXPECT((uint32_t)got == sum, "sharded i32 sum");
            CHECK(la_dot_vec_vec_sharded(comm, LA_I32, A, B, hi - lo, &got, nullptr)); EXPECT((uint32_t)got == dot, "sharded i32 dot");
        }
        int eq = -1;
        CHECK(la_eq_vec_sharded(comm, LA_I32, A, A, hi - lo, &eq, nullptr)); EXPECT(eq == 1, "A == A");
        CHECK(la_eq_vec_sharded(comm, LA_I32, A, B, hi - lo, &eq, nullptr)); EXPECT(eq == 0, "A != B");
        CHECK(la_free(A)); CHECK(la_free(B));
        const auto x = floats(n, 3), y = floats(n, 4);
        double truth = 0, scale = 0;
        for (size_t i = 0; i < n; ++i) { truth += (double)x[i] * y[i]; scale += fabs((double)x[i] * y[i]); }
        void* X = to_device(x.data() + lo, hi - lo); void* Y = to_device(y.data() + lo, hi - lo);
        float gotf = 0, again = 0;
        CHECK(la_dot_vec_vec_sharded(comm, LA_F32, X, Y, hi - lo, &gotf, nullptr));
        EXPECT(fabs((double)gotf - truth) <= scale / 1048576.0, "sharded f32 dot within 2^-20 sum|ab|");
        CHECK(la_dot_vec_vec_sharded(comm, LA_F32, X, Y, hi - lo, &again, nullptr));
        EXPECT(memcmp(&gotf, &again, 4) == 0, "sharded f32 dot is reproducible");
        CHECK(la_free(X)); CHECK(la_free(Y));
    }
    // ---- v x M (rows sharded) and transpose, i32
    {
        const size_t rows = 1037, cols = 2052;
        const auto m = ints(rows * cols, 5), v = ints(rows, 6);
        size_t rlo, rhi; la_partition(rows, g_world, g_rank, &rlo, &rhi);
        std::vector<uint32_t> want(cols, 0);
        for (size_t rr = 0; rr < rows; ++rr) for (size_t c = 0; c < cols; ++c) want[c] += (uint32_t)v[rr] * (uint32_t)m[rr * cols + c];
        void* M = to_device(m.data() + rlo * cols, (rhi - rlo) * cols); void* V = to_device(v.data() + rlo, rhi - rlo);
        void* C = nullptr; CHECK(la_malloc(&C, cols * 4));
        std::vector<uint32_t> got(cols);
        for (int round = 0; round < 3; ++round) {
            CHECK(la_mul_vec_mat_sharded(comm, LA_I32, C, V, M, rhi - rlo, cols, nullptr));
            CHECK(la_download(got.data(), C, cols * 4, nullptr));
            EXPECT(memcmp(got.data(), want.data(), cols * 4) == 0, "sharded v x M, i32");
        }
        size_t clo, chi; la_partition(cols, g_world, g_rank, &clo, &chi);
        void* T = nullptr; CHECK(la_malloc(&T, (chi - clo) * rows * 4));
        std::vector<int32_t> tgot((chi - clo) * rows);
        if (no_peer) EXPECT(la_transpose_sharded(comm, LA_I32, T, M, rows, cols, nullptr) == LA_ERR_UNSUPPORTED, "transpose needs peer memory");
        for (int round = 0; round < 2 && !no_peer; ++round) {
            CHECK(la_transpose_sharded(comm, LA_I32, T, M, rows, cols, nullptr));
            CHECK(la_download(tgot.data(), T, tgot.size() * 4, nullptr));
            bool same = true;
            for (size_t c = clo; c < chi && same; ++c) for (size_t rr = 0; rr < rows; ++rr) if (tgot[(c - clo) * rows + rr] != m[rr * cols + c]) { same = false; break; }
            EXPECT(same, "sharded transpose, i32");
        }
        CHECK(la_free(M)); CHECK(la_free(V)); CHECK(la_free(C)); CHECK(la_free(T));
    }
    // ---- v x M wider than the exported vector buffers (2 x 4 MiB at creation): they are grown, collectively
    {
        const size_t rows = 16 * (size_t)g_world + 3, cols = (1u << 20) + 8;
        const auto m = ints(rows * cols, 9), v = ints(rows, 10);
        size_t rlo, rhi; la_partition(rows, g_world, g_rank, &rlo, &rhi);
        std::vector<uint32_t> want(cols, 0);
        for (size_t rr = 0; rr < rows; ++rr) for (size_t c = 0; c < cols; ++c) want[c] += (uint32_t)v[rr] * (uint32_t)m[rr * cols + c];
        void* M = to_device(m.data() + rlo * cols, (rhi - rlo) * cols); void* V = to_device(v.data() + rlo, rhi - rlo);
        void* C = nullptr; CHECK(la_malloc(&C, cols * 4));
        std::vector<uint32_t> got(cols);
        for (int round = 0; round < 2; ++round) {
            CHECK(la_mul_vec_mat_sharded(comm, LA_I32, C, V, M, rhi - rlo, cols, nullptr));
            CHECK(la_download(got.data(), C, cols * 4, nullptr));
            EXPECT(memcmp(got.data(), want.data(), cols * 4) == 0, "sharded v x M wider than the initial exchange buffers");
        }
        CHECK(la_free(M)); CHECK(la_free(V)); CHECK(la_free(C));
    }
    // ---- M x M on host buffers, rows of A and C sharded, i32
    {
        const size_t Mg = 700, K = 64 * (size_t)g_world, N = 260;
        const auto a = ints(Mg * K, 7), b = ints(K * N, 8);
        size_t mlo, mhi; la_partition(Mg, g_world, g_rank, &mlo, &mhi);
        std::vector<int32_t> c((mhi - mlo) * N);
        CHECK(la_mul_mat_mat_host_sharded(comm, LA_I32, c.data(), a.data() + mlo * K, b.data(), Mg, K, N));
        bool same = true;
        for (size_t i = mlo; i < mhi && same; ++i) for (size_t j = 0; j < N; ++j) {
            uint32_t acc = 0;
            for (size_t k = 0; k < K; ++k) acc += (uint32_t)a[i * K + k] * (uint32_t)b[k * N + j];
            if ((uint32_t)c[(i - mlo) * N + j] != acc) { same = false; break; }
        }
        EXPECT(same, "sharded host GEMM, i32");
    }
    CHECK(la_comm_barrier(comm));
    CHECK(la_comm_destroy(comm));
    printf("rank %d/%d ok\n", g_rank, g_world);
    return 0;
}
