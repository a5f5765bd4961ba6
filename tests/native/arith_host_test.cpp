// TEST INFRASTRUCTURE: host build (g++) of the product's __host__ __device__ arithmetic
// (rust-lp_b200/csrc/relp_arith.cuh) so that tests/test_arith_host.py can check it against
// Python's fractions.Fraction without a GPU.  Not part of the product library.
#include "relp_arith.cuh"
#include <vector>
#include <algorithm>
#include <cstring>
using namespace relp;

static std::vector<uint32_t> g_heap(1 << 24);
static unsigned long long g_top = 0;
static int g_err = 0;
static Arena arena() { Arena a; a.top = &g_top; a.limit = g_heap.size(); a.err = &g_err; return a; }

extern "C" {
void t_reset() { g_top = 0; g_err = 0; }
int t_err() { return g_err; }
int t_opl() { return RELP_OPL; }
void t_make(int sign, const uint32_t* num, int nn, const uint32_t* den, int nd, int64_t* out) {
    Val v = store_val(sign, num, nn, den, nd, g_heap.data(), arena());
    out[0] = v.n; out[1] = v.d;
}
void t_fma(const int64_t* x, const int64_t* a, const int64_t* b, int s, int64_t* out) {
    Val v = val_fma(Val{x[0], x[1]}, Val{a[0], a[1]}, Val{b[0], b[1]}, s, g_heap.data(), arena());
    out[0] = v.n; out[1] = v.d;
}
void t_div(const int64_t* x, const int64_t* y, int64_t* out) {
    Val v = val_div(Val{x[0], x[1]}, Val{y[0], y[1]}, g_heap.data(), arena());
    out[0] = v.n; out[1] = v.d;
}
int t_cmp(const int64_t* x, const int64_t* y) { return val_cmp(Val{x[0], x[1]}, Val{y[0], y[1]}, g_heap.data()); }
int t_sign(const int64_t* x) { return val_sign(Val{x[0], x[1]}, g_heap.data()); }
// big value -> header + limbs
int t_read(const int64_t* x, uint32_t* out) {
    Val v{x[0], x[1]};
    unsigned w = big_words(g_heap.data(), v);
    memcpy(out, g_heap.data() + v.n, 4 * w);
    return (int)w;
}
uint64_t t_gcd64(uint64_t a, uint64_t b) { return gcd64(a, b); }
// multi-limb gcd and remainder on their own (out: limbs of the result, returns its count)
int t_mag_gcd(const uint32_t* a, int na, const uint32_t* b, int nb, uint32_t* out) {
    std::vector<uint32_t> u(a, a + na), v(b, b + nb), tmp(std::max(na, nb) + 2);
    u.resize(std::max(na, nb) + 2); v.resize(std::max(na, nb) + 2);
    uint32_t* g; int ng = mag_gcd(u.data(), na, v.data(), nb, &g, tmp.data());
    for (int i = 0; i < ng; i++) out[i] = g[i];
    return ng;
}
int t_mag_mod(const uint32_t* a, int na, const uint32_t* b, int nb, uint32_t* out) {
    std::vector<uint32_t> u(a, a + na), tmp(nb + 2); u.resize(na + 2);
    int nr = mag_mod(u.data(), na, b, nb, tmp.data());
    for (int i = 0; i < nr; i++) out[i] = u[i];
    return nr;
}
}
