// TEST INFRASTRUCTURE: host build (g++) of the product's __host__ __device__ arithmetic
// (rust-lp_b200/csrc/relp_arith.cuh) so that tests/test_arith_host.py can check it against
// Python's fractions.Fraction without a GPU.  Not part of the product library.
#define RELP_WCB_TEST_SLAB 1
#include "relp_arith.cuh"
#include <vector>
#include <algorithm>
#include <cstring>
using namespace relp;
namespace relp { uint32_t* relp_test_slab = nullptr; unsigned relp_test_slab_words = 0; }
static std::vector<uint32_t> g_slab;

static std::vector<uint32_t> g_heap(1 << 24);
static unsigned long long g_top = 16;
static ScratchExt g_ext;
static int g_errs[2] = {0, 0};      // err word + the "long values exist" word behind it (Arena::err[1])
#define g_err g_errs[0]
static Arena arena() {
    Arena a; a.top = &g_top; a.limit = g_heap.size(); a.err = g_errs; a.flag = g_errs + 1;
    g_ext.scr_top = &g_top; g_ext.scr_limit = g_heap.size(); g_ext.err = g_errs;     // scratch behind the values, same region
    const uint64_t addr = (uint64_t)(uintptr_t)&g_ext;
    g_heap[2] = (uint32_t)addr; g_heap[3] = (uint32_t)(addr >> 32);                  // how big_cmp finds it (heap_scratch)
    return a;
}

extern "C" {
// scratch slab of the cooperative operations (0 words = none: everything from the arena); a small one makes them spill midway
void t_set_slab(int words) { g_slab.assign((size_t)(words > 0 ? words : 1), 0xdeadbeefu); relp::relp_test_slab = g_slab.data(); relp::relp_test_slab_words = (unsigned)(words > 0 ? words : 0); }
void t_reset() { g_top = 16; g_err = 0; }
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

// ---- cooperative multi-limb tier (relp_wcb.cuh) on S host threads behind a barrier ------------------
// Every "lane" is a host thread running the same template code; shfl / ballot go through a slot array
// between two barriers.  This is how the carry-lookahead across lanes and the chunking are checked on the
// CPU for group sizes 1..32 before the code runs in a kernel.
#include <thread>
#include <atomic>
#include <mutex>
#include <condition_variable>
#include <functional>
namespace {
struct Barrier {       // sense-reversing spin barrier (a condition variable costs tens of microseconds per collective)
    std::atomic<int> count{0}; std::atomic<unsigned> gen{0}; int n = 1;
    void wait() {
        const unsigned g = gen.load(std::memory_order_acquire);
        if (count.fetch_add(1, std::memory_order_acq_rel) + 1 == n) { count.store(0, std::memory_order_relaxed); gen.store(g + 1, std::memory_order_release); }
        else { int spins = 0; while (gen.load(std::memory_order_acquire) == g) if (++spins > 200) { std::this_thread::yield(); spins = 0; } }
    }
};
struct Shared { Barrier bar; uint64_t slot[32]; int size = 1; };
struct ThreadGroup {
    Shared* sh; int rk;
    int rank() const { return rk; }
    int size() const { return sh->size; }
    void sync() const { sh->bar.wait(); }
    uint32_t ballot(bool p) const {
        sh->slot[rk] = p ? 1 : 0; sh->bar.wait();
        uint32_t m = 0; for (int i = 0; i < sh->size; i++) if (sh->slot[i]) m |= 1u << i;
        sh->bar.wait();
        return m;
    }
    template <class T> T shfl(T v, int src) const {
        static_assert(sizeof(T) <= 8, "slot size");
        uint64_t w = 0; memcpy(&w, &v, sizeof(T));
        sh->slot[rk] = w; sh->bar.wait();
        uint64_t o = sh->slot[src]; sh->bar.wait();
        T r; memcpy(&r, &o, sizeof(T)); return r;
    }
    template <class T> T shfl_up(T v, int d) const {
        static_assert(sizeof(T) <= 8, "slot size");
        uint64_t w = 0; memcpy(&w, &v, sizeof(T));
        sh->slot[rk] = w; sh->bar.wait();
        uint64_t o = sh->slot[rk >= d ? rk - d : rk]; sh->bar.wait();
        T r; memcpy(&r, &o, sizeof(T)); return r;
    }
};
template <class F> void run_group(int S, F f) {
    Shared sh; sh.size = S; sh.bar.n = S;
    std::vector<std::thread> th;
    for (int r = 0; r < S; r++) th.emplace_back([&sh, r, &f] { ThreadGroup g{&sh, r}; f(g); });
    for (auto& t : th) t.join();
}
}  // namespace

extern "C" {
// op: 0 add, 1 sub (a >= b), 2 mul, 3 mod, 4 gcd, 5 divexact (b | a), 6 mul1 (b[0]), 7 shl (b[0] bits), 8 shr (b[0] bits, exact or not), 9 cmp
int t_w_op(int S, int op, const uint32_t* a, int na, const uint32_t* b, int nb, uint32_t* out) {
    const int cap = 2 * (na + nb) + 8 + (op == 7 ? (int)b[0] / 32 + 2 : 0);
    std::vector<uint32_t> A(cap, 0), B(cap, 0), R(cap, 0), T1(cap, 0), T2(cap, 0), T3(cap, 0);
    std::copy(a, a + na, A.begin()); std::copy(b, b + nb, B.begin());
    int result = 0; uint32_t* rp = R.data();
    run_group(S, [&](const ThreadGroup& g) {
        int n = 0; uint32_t* p = R.data();
        switch (op) {
        case 0: n = w_add(g, R.data(), A.data(), na, B.data(), nb); break;
        case 1: n = w_sub(g, R.data(), A.data(), na, B.data(), nb); break;
        case 2: n = w_mul(g, R.data(), A.data(), na, B.data(), nb); break;
        case 3: n = w_mod(g, A.data(), na, B.data(), nb, T1.data(), T2.data()); p = A.data(); break;
        case 4: n = w_gcd(g, A.data(), na, B.data(), nb, T1.data(), T2.data(), T3.data(), &p); break;
        case 5: n = w_divexact(g, R.data(), A.data(), na, B.data(), nb, T1.data(), T2.data()); break;
        case 6: n = w_mul1(g, R.data(), A.data(), na, b[0]); break;
        case 7: n = w_shl(g, A.data(), na, (int)b[0]); p = A.data(); break;
        case 8: n = w_shr(g, A.data(), na, (int)b[0]); p = A.data(); break;
        case 9: n = w_cmp(g, A.data(), na, B.data(), nb); break;
        }
        if (g.rank() == 0) { result = n; rp = p; }
    });
    if (op != 9) for (int i = 0; i < result; i++) out[i] = rp[i];
    return result;
}
// the cooperative rational operations on a group of S host threads (values live in g_heap)
void t_w_fma(int S, const int64_t* x, const int64_t* a, const int64_t* b, int s, int64_t* out) {
    Val r = val_zero();
    const Arena ar = arena();
    run_group(S, [&](const ThreadGroup& g) {
        Val v = wcb_fma(g, Val{x[0], x[1]}, Val{a[0], a[1]}, Val{b[0], b[1]}, s, g_heap.data(), ar);
        if (g.rank() == 0) r = v;
    });
    out[0] = r.n; out[1] = r.d;
}
void t_w_div(int S, const int64_t* x, const int64_t* y, int64_t* out) {
    Val r = val_zero();
    const Arena ar = arena();
    run_group(S, [&](const ThreadGroup& g) {
        Val v = wcb_div(g, Val{x[0], x[1]}, Val{y[0], y[1]}, g_heap.data(), ar);
        if (g.rank() == 0) r = v;
    });
    out[0] = r.n; out[1] = r.d;
}
int t_w_cmp(int S, const int64_t* x, const int64_t* y) {
    int r = 0;
    const Arena ar = arena();
    run_group(S, [&](const ThreadGroup& g) {
        int c = wcb_cmp(g, Val{x[0], x[1]}, Val{y[0], y[1]}, g_heap.data(), ar);
        if (g.rank() == 0) r = c;
    });
    return r;
}
}
