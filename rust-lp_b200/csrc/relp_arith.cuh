// Exact rational arithmetic for the device-resident Carry matrix.
//
// Replaces, value for value, the reference's `RationalBig` = newtype over `num::BigRational`
// (/root/reference/src/data/number_types/rational/big/mod.rs:17-29; operators forwarded in
// big/wrapping.rs:47-311): every stored value is canonical -- gcd-reduced, denominator > 0,
// zero = 0/1 -- so results are bit-identical to the reference's by uniqueness of that form,
// whatever algorithm computes them.
//
// Two tiers, chosen per element:
//   small  : Val{n, d} with int64 numerator and int64 denominator d >= 1; all intermediates in
//            128 bits with explicit overflow detection ("does the reduced result fit?").
//   big    : Val{handle, 0}; `handle` is a word offset into the limb heap where
//            [header | numerator limbs | denominator limbs] live (32-bit limbs, little endian).
//            header = nn | nd << 14 | sign << 31.   A value is stored big only if it does NOT fit
//            small, so equal values have equal representations up to the handle.
//   absent : Val{*, -1} marks "no candidate" in reductions.
//
// Everything here is __host__ __device__ so that the same code is unit-tested on the CPU
// (tests/test_arith_host.py builds csrc/arith_host_test.cpp with g++) before it runs in kernels.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define RELP_HD __host__ __device__ __forceinline__
#define RELP_HDN __host__ __device__ __noinline__
#else
#define RELP_HD inline
#define RELP_HDN
#endif

// The headers can be compiled twice into one library: RELP_NS names the namespace of this copy.  relp_fastloop.cu
// builds the persistent loop kernel with RELP_NO_WCB: the same code WITHOUT the arbitrary-length tier behind it (that
// tier's call graph costs the hot int64 path 20 % -- measured -- through register allocation and code layout), for as
// long as no stored value is longer than RELP_LONG_TRIGGER limbs; then the copy with the full tier takes over.
#ifndef RELP_NS
#define RELP_NS relp
#endif
#ifndef RELP_LONG_TRIGGER
#define RELP_LONG_TRIGGER 12     // a stored numerator / denominator longer than this raises DState::long_seen (Arena::flag)
#endif
#ifndef RELP_OPL
#define RELP_OPL 64              // operands up to this many limbs may be handled by ONE thread (fixed-size stack buffers);
                                 // longer ones go to the cooperative routines of relp_wcb.cuh, which have no fixed size
#endif
#define RELP_TL (3 * RELP_OPL + 2)   // max limbs of an intermediate
#define RELP_MAX_LIMBS 0x3fff        // what the header's 14-bit length fields can say (512 Kbit); the limb heap is the real bound

namespace RELP_NS {

typedef unsigned __int128 u128;
typedef __int128 i128;

struct Val { int64_t n; int64_t d; };

enum { ERR_NONE = 0, ERR_LIMB_CAPACITY = 1, ERR_HEAP_EXHAUSTED = 2, ERR_ZERO_PIVOT = 4, ERR_PEER = 8, ERR_PEER_BIG = 16 };

// Bump allocator over a region [*top, limit) of the limb heap.
// Scratch for the temporaries of the cooperative multi-limb operations (relp_wcb.cuh): one slab per resident warp
// (taken with a lock word for the duration of an operation) and, behind it, the transient region of the heap.
// Lives in memory (one per context); only the slow paths ever read it.
struct ScratchExt {
    uint32_t* wslab = nullptr; unsigned wslab_words = 0; int* wlock = nullptr; int wslots = 0;
    unsigned long long* scr_top = nullptr; unsigned long long scr_limit = 0; int* err = nullptr; int* flag = nullptr;
    int coop_from = RELP_OPL;    // operands longer than this many limbs leave the per-thread routines for the cooperative ones
    unsigned long long heap_words = 0;   // size of the limb heap (sanity checks of handles in the cooperative entry points)
    int dbg = 0;                 // first failed sanity check (0 = none): see w_check
    // one slab per THREAD (indexed by the global thread id, no lock) for long operations done by a single thread: without
    // it their temporaries come from the transient region, which is only recycled at the end of an iteration -- 674 000
    // queued elements of 70 limbs exhausted a 4 GiB region (netlib 25FV47, phase two).  Allocated when a solve turns long.
    uint32_t* tslab = nullptr; unsigned tslab_words = 0; unsigned tslots = 0;
};
// The limb heap's first words are never a value (offset 0 is not a handle): words 2..3 hold the address of the
// ScratchExt of the heap's owner (null = none).  The slow paths find it there, so that the Arena every stage carries
// in registers stays three words, and comparisons -- which have no arena argument -- find room for long cross products.
RELP_HD const ScratchExt* heap_scratch(const uint32_t* heap) {
    const uint64_t a = (uint64_t)heap[2] | ((uint64_t)heap[3] << 32);
    return (const ScratchExt*)(uintptr_t)a;
}
struct Arena {
    unsigned long long* top = nullptr;
    unsigned long long limit = 0;
    int* err = nullptr;          // sticky error bits (ERR_*)
    int* flag = nullptr;         // set to 1 when a value longer than RELP_LONG_TRIGGER limbs is stored (DState::long_seen)
};

RELP_HD void set_err(int* err, int bit) {
    if (!err) return;
#if defined(__CUDA_ARCH__)
    atomicOr(err, bit);
#else
    *err |= bit;
#endif
}

RELP_HD long long arena_alloc(const Arena& a, unsigned words) {
#if defined(__CUDA_ARCH__)
    unsigned long long off = atomicAdd(a.top, (unsigned long long)words);
#else
    unsigned long long off = *a.top; *a.top += words;
#endif
    if (off + words > a.limit) { set_err(a.err, ERR_HEAP_EXHAUSTED); return -1; }
    return (long long)off;
}

RELP_HD bool is_small(const Val& v) { return v.d > 0; }
RELP_HD bool is_big(const Val& v) { return v.d == 0; }
RELP_HD bool is_absent(const Val& v) { return v.d < 0; }
RELP_HD Val make_small(int64_t n, int64_t d) { Val v; v.n = n; v.d = d; return v; }
RELP_HD Val val_zero() { return make_small(0, 1); }
RELP_HD Val val_absent() { return make_small(0, -1); }
RELP_HD bool val_is_zero(const Val& v) { return v.d == 1 && v.n == 0; }

RELP_HD int ctz64(uint64_t x) {
#if defined(__CUDA_ARCH__)
    return __ffsll((long long)x) - 1;
#else
    return __builtin_ctzll(x);
#endif
}
RELP_HD int ctz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __ffs((int)x) - 1;
#else
    return __builtin_ctz(x);
#endif
}

// Binary GCD on 64-bit magnitudes (Stein, with count-trailing-zero shifts).
RELP_HD uint64_t gcd64(uint64_t a, uint64_t b) {
    if (a == 0) return b;
    if (b == 0) return a;
    if (a == 1 || b == 1) return 1;
    int sh = ctz64(a | b);
    a >>= ctz64(a);
    do {
        b >>= ctz64(b);
        if (a > b) { uint64_t t = a; a = b; b = t; }
        b -= a;
    } while (b);
    return a << sh;
}

RELP_HD uint64_t abs64(int64_t x) { return x < 0 ? (uint64_t)0 - (uint64_t)x : (uint64_t)x; }

// ------------------------------------------------------------------------------------------
// small tier: int64/int64 with 128-bit intermediates.  Each returns false when the canonical
// result (or an intermediate product) does not fit, in which case the caller escalates.
// ------------------------------------------------------------------------------------------

#define RELP_I64_MAX ((i128)0x7fffffffffffffffLL)

// out = x + s * a * b, s = +1 or -1      (sparse.rs:213-248 add_multiple_of_row element,
//                                          carry/mod.rs:283-313 update_b, :326-333 cost row,
//                                          dense.rs:81-92 / sparse.rs:82-106 inner products)
// The general (rational) case is a separate, non-inlined function: it is long (three gcds, 128-bit
// divisions) and the persistent kernel calls fma from dozens of places -- inlined everywhere it made
// the kernel a megabyte of code that no instruction cache holds.  The integer case stays inline.
// 32-bit binary GCD (same algorithm as gcd64 on half-width registers)
RELP_HD uint32_t gcd32(uint32_t a, uint32_t b) {
    if (a == 0) return b;
    if (b == 0) return a;
    if (a == 1 || b == 1) return 1;
    int sh = ctz32(a | b);
    a >>= ctz32(a);
    do {
        b >>= ctz32(b);
        if (a > b) { uint32_t t = a; a = b; b = t; }
        b -= a;
    } while (b);
    return a << sh;
}

RELP_HDN bool small_fma_rational(const Val& x, const Val& a, const Val& b, int s, Val* out) {
    // canonical product tn/td of a*b (cross-reduce, then multiply)
    uint64_t an = abs64(a.n), bn = abs64(b.n);
    // All six operands below 2^31 and the cross-reduced product too (the usual case while values are small): the
    // same computation in 32-bit gcds and 64-bit products -- no 128-bit arithmetic, and nothing can overflow.
    if (((an | bn | abs64(x.n) | (uint64_t)a.d | (uint64_t)b.d | (uint64_t)x.d) >> 31) == 0) {
        const uint32_t g1 = b.d == 1 ? 1u : gcd32((uint32_t)an, (uint32_t)b.d);
        const uint32_t g2 = a.d == 1 ? 1u : gcd32((uint32_t)bn, (uint32_t)a.d);
        // (a division by 1 is the common case and costs a reciprocal on the quarter-rate XU pipe: skip it)
        const uint32_t an1 = g1 == 1 ? (uint32_t)an : (uint32_t)an / g1, bd1 = g1 == 1 ? (uint32_t)b.d : (uint32_t)b.d / g1;
        const uint32_t bn1 = g2 == 1 ? (uint32_t)bn : (uint32_t)bn / g2, ad1 = g2 == 1 ? (uint32_t)a.d : (uint32_t)a.d / g2;
        const uint64_t tn = (uint64_t)an1 * bn1;
        const uint64_t td = (uint64_t)ad1 * bd1;
        if (((tn | td) >> 31) == 0) {
            const bool tneg = ((a.n < 0) != (b.n < 0)) != (s < 0);
            if (x.n == 0) { *out = make_small(tneg ? -(int64_t)tn : (int64_t)tn, (int64_t)td); return true; }
            const uint32_t xd = (uint32_t)x.d, tdd = (uint32_t)td;
            const uint32_t g = (xd == 1 || tdd == 1) ? 1u : gcd32(xd, tdd);
            const uint32_t xdg = g == 1 ? xd : xd / g, tdg = g == 1 ? tdd : tdd / g;
            const int64_t p = x.n * (int64_t)tdg;                  // |.| < 2^62
            const int64_t q = (int64_t)tn * (int64_t)xdg;          // < 2^62
            int64_t num = tneg ? p - q : p + q;                    // |.| < 2^63
            if (num == 0) { *out = val_zero(); return true; }
            uint64_t den = (uint64_t)xdg * tdd;                    // < 2^62
            if (g > 1) {                                           // gcd(num, den) = gcd(num, g)  (Knuth 4.5.1)
                const uint64_t na = num < 0 ? (uint64_t)0 - (uint64_t)num : (uint64_t)num;
                const uint32_t g3 = gcd32((na >> 32) == 0 ? (uint32_t)na % g : (uint32_t)(na % g), g);
                if (g3 > 1) {
                    if ((na >> 32) == 0 && (den >> 32) == 0) {     // 32-bit quotients when they fit
                        const uint32_t qn = (uint32_t)na / g3;
                        num = num < 0 ? -(int64_t)qn : (int64_t)qn; den = (uint32_t)den / g3;
                    } else { num /= (int64_t)g3; den /= g3; }
                }
            }
            *out = make_small(num, (int64_t)den);
            return true;
        }
    }
    uint64_t g1 = b.d == 1 ? 1 : gcd64(an, (uint64_t)b.d);
    uint64_t g2 = a.d == 1 ? 1 : gcd64(bn, (uint64_t)a.d);
    u128 tn = (u128)(an / g1) * (bn / g2);
    u128 td = (u128)((uint64_t)a.d / g2) * ((uint64_t)b.d / g1);
    if (tn > (u128)RELP_I64_MAX || td > (u128)RELP_I64_MAX) return false;
    bool tneg = ((a.n < 0) != (b.n < 0)) != (s < 0);
    uint64_t tnn = (uint64_t)tn, tdd = (uint64_t)td;
    if (x.n == 0) { *out = make_small(tneg ? -(int64_t)tnn : (int64_t)tnn, (int64_t)tdd); return true; }
    // x.n/x.d + (+-tn)/td
    uint64_t xd = (uint64_t)x.d;
    uint64_t g = (xd == 1 || tdd == 1) ? 1 : gcd64(xd, tdd);
    uint64_t xdg = xd / g, tdg = tdd / g;
    i128 p = (i128)x.n * (i128)tdg;                // |.| < 2^126
    i128 q = (i128)tnn * (i128)xdg;                // < 2^126
    i128 num = tneg ? p - q : p + q;               // |.| < 2^127
    if (num == 0) { *out = val_zero(); return true; }
    u128 den = (u128)xdg * tdd;                    // < 2^126
    if (g > 1) {                                   // gcd(num, den) = gcd(num, g)  (Knuth 4.5.1)
        u128 na = num < 0 ? (u128)0 - (u128)num : (u128)num;
        uint64_t g3 = gcd64((uint64_t)(na % g), g);
        if (g3 > 1) { num /= (i128)g3; den /= g3; }
    }
    if (num > RELP_I64_MAX || num < -RELP_I64_MAX || den > (u128)RELP_I64_MAX) return false;
    *out = make_small((int64_t)num, (int64_t)den);
    return true;
}

RELP_HD bool small_fma(const Val& x, const Val& a, const Val& b, int s, Val* out) {
    if (a.n == 0 || b.n == 0) { *out = x; return true; }
    if (x.d == 1 && a.d == 1 && b.d == 1) {                 // integer fast path
#ifndef RELP_NO_FMA32
        if (a.n == (int64_t)(int32_t)a.n && b.n == (int64_t)(int32_t)b.n) {
            // both factors fit 32 bits: one 32 x 32 -> 64 multiply (IMAD.WIDE) instead of an emulated
            // 64 x 64 -> 128 product, and a sign test instead of a 128-bit range check
            const int64_t t = (int64_t)(int32_t)a.n * (int64_t)(int32_t)b.n;
            int64_t r; bool ovf;
            if (s > 0) { r = (int64_t)((uint64_t)x.n + (uint64_t)t); ovf = ((x.n ^ r) & (t ^ r)) < 0; }
            else       { r = (int64_t)((uint64_t)x.n - (uint64_t)t); ovf = ((x.n ^ t) & (x.n ^ r)) < 0; }
            if (ovf || r == INT64_MIN) return false;
            *out = make_small(r, 1);
            return true;
        }
#endif
        i128 t = (i128)a.n * b.n;
        i128 r = s > 0 ? (i128)x.n + t : (i128)x.n - t;
        if (r > RELP_I64_MAX || r < -RELP_I64_MAX) return false;
        *out = make_small((int64_t)r, 1);
        return true;
    }
    return small_fma_rational(x, a, b, s, out);
}

// out = x / y  (y != 0)      (sparse.rs:291-301 element_wise_divide, tableau/mod.rs:229 ratio)
RELP_HDN bool small_div_general(const Val& x, const Val& y, Val* out) {
    uint64_t xn = abs64(x.n), yn = abs64(y.n);
    uint64_t g1 = gcd64(xn, yn);
    uint64_t g2 = (x.d == 1 || y.d == 1) ? 1 : gcd64((uint64_t)x.d, (uint64_t)y.d);
    u128 n = (u128)(xn / g1) * ((uint64_t)y.d / g2);
    u128 d = (u128)((uint64_t)x.d / g2) * (yn / g1);
    if (n > (u128)RELP_I64_MAX || d > (u128)RELP_I64_MAX) return false;
    bool neg = (x.n < 0) != (y.n < 0);
    *out = make_small(neg ? -(int64_t)(uint64_t)n : (int64_t)(uint64_t)n, (int64_t)(uint64_t)d);
    return true;
}

RELP_HD bool small_div(const Val& x, const Val& y, Val* out) {
    if (x.n == 0) { *out = val_zero(); return true; }
    if (y.d == 1 && (y.n == 1 || y.n == -1)) {          // unit pivots (totally unimodular bases)
        *out = make_small(y.n > 0 ? x.n : -x.n, x.d); return true;
    }
    return small_div_general(x, y, out);
}

RELP_HD int small_cmp(const Val& x, const Val& y) {
    if (x.d == y.d) return x.n < y.n ? -1 : (x.n > y.n ? 1 : 0);
    i128 l = (i128)x.n * y.d, r = (i128)y.n * x.d;
    return l < r ? -1 : (l > r ? 1 : 0);
}

// ------------------------------------------------------------------------------------------
// big tier: magnitudes as little-endian uint32 arrays (n limbs, no leading zero limb; 0 <=> n=0)
// ------------------------------------------------------------------------------------------

RELP_HD int mag_trim(const uint32_t* a, int n) { while (n > 0 && a[n - 1] == 0) n--; return n; }

RELP_HD int mag_cmp(const uint32_t* a, int na, const uint32_t* b, int nb) {
    if (na != nb) return na < nb ? -1 : 1;
    for (int i = na - 1; i >= 0; i--) if (a[i] != b[i]) return a[i] < b[i] ? -1 : 1;
    return 0;
}

// r = a + b (r may alias a or b); returns limbs of r (capacity max(na,nb)+1)
RELP_HD int mag_add(uint32_t* r, const uint32_t* a, int na, const uint32_t* b, int nb) {
    if (na < nb) { const uint32_t* t = a; a = b; b = t; int ti = na; na = nb; nb = ti; }
    uint64_t c = 0;
    int i = 0;
    for (; i < nb; i++) { c += (uint64_t)a[i] + b[i]; r[i] = (uint32_t)c; c >>= 32; }
    for (; i < na; i++) { c += a[i]; r[i] = (uint32_t)c; c >>= 32; }
    if (c) { r[i++] = (uint32_t)c; }
    return i;
}

// r = a - b, requires a >= b (r may alias a); returns trimmed limbs
RELP_HD int mag_sub(uint32_t* r, const uint32_t* a, int na, const uint32_t* b, int nb) {
    int64_t c = 0;
    int i = 0;
    for (; i < nb; i++) { c += (int64_t)a[i] - (int64_t)b[i]; r[i] = (uint32_t)c; c >>= 32; }
    for (; i < na; i++) { c += (int64_t)a[i]; r[i] = (uint32_t)c; c >>= 32; }
    return mag_trim(r, na);
}

// r = a * b (r must not alias); returns limbs (capacity na+nb)
RELP_HD int mag_mul(uint32_t* r, const uint32_t* a, int na, const uint32_t* b, int nb) {
    if (na == 0 || nb == 0) return 0;
    for (int i = 0; i < na + nb; i++) r[i] = 0;
    for (int i = 0; i < na; i++) {
        uint64_t c = 0, ai = a[i];
        for (int j = 0; j < nb; j++) {
            c += ai * b[j] + r[i + j];
            r[i + j] = (uint32_t)c; c >>= 32;
        }
        r[i + nb] = (uint32_t)c;
    }
    return mag_trim(r, na + nb);
}

RELP_HD int mag_ctz(const uint32_t* a, int n) {     // n > 0
    int i = 0;
    while (a[i] == 0) i++;
    return 32 * i + ctz32(a[i]);
}

// a >>= s in place; returns limbs
RELP_HD int mag_shr(uint32_t* a, int n, int s) {
    int w = s >> 5, b = s & 31;
    if (w >= n) return 0;
    int m = n - w;
    if (b == 0) { for (int i = 0; i < m; i++) a[i] = a[i + w]; }
    else {
        for (int i = 0; i < m - 1; i++) a[i] = (a[i + w] >> b) | (a[i + w + 1] << (32 - b));
        a[m - 1] = a[n - 1] >> b;
    }
    return mag_trim(a, m);
}

// a <<= s in place (capacity n + s/32 + 1); returns limbs
RELP_HD int mag_shl(uint32_t* a, int n, int s) {
    if (n == 0) return 0;
    int w = s >> 5, b = s & 31;
    if (b == 0) { for (int i = n - 1; i >= 0; i--) a[i + w] = a[i]; n += w; }
    else {
        a[n + w] = a[n - 1] >> (32 - b);
        for (int i = n - 1; i > 0; i--) a[i + w] = (a[i] << b) | (a[i - 1] >> (32 - b));
        a[w] = a[0] << b;
        n += w + 1;
    }
    for (int i = 0; i < w; i++) a[i] = 0;
    return mag_trim(a, n);
}

RELP_HD uint64_t mag_lo64(const uint32_t* a, int n) { return n == 0 ? 0 : (n == 1 ? (uint64_t)a[0] : ((uint64_t)a[1] << 32) | a[0]); }
// u mod v for a 64-bit v != 0
RELP_HD uint64_t mag_mod64(const uint32_t* u, int nu, uint64_t v) {
    if (v >> 32 == 0) {                       // 32-bit divisor: 64-bit steps
        uint64_t r = 0;
        for (int i = nu - 1; i >= 0; i--) r = ((r << 32) | u[i]) % v;
        return r;
    }
    u128 r = 0;
    for (int i = nu - 1; i >= 0; i--) r = ((r << 32) | u[i]) % v;
    return (uint64_t)r;
}
RELP_HD int mag_set64(uint32_t* a, uint64_t x) { a[0] = (uint32_t)x; a[1] = (uint32_t)(x >> 32); return a[1] ? 2 : (a[0] ? 1 : 0); }

RELP_HD int clz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __clz((int)x);
#else
    return __builtin_clz(x);
#endif
}

// u = u mod v in place (Knuth 4.3.1 D), nv >= 2, v[nv-1] != 0; u needs one spare limb above nu;
// vn: scratch of nv limbs.  Returns the trimmed limb count of the remainder.
RELP_HD int mag_mod(uint32_t* u, int nu, const uint32_t* v, int nv, uint32_t* vn) {
    if (nu < nv) return nu;
    const int sft = clz32(v[nv - 1]);
    if (sft) {
        for (int i = nv - 1; i > 0; i--) vn[i] = (v[i] << sft) | (v[i - 1] >> (32 - sft));
        vn[0] = v[0] << sft;
        u[nu] = u[nu - 1] >> (32 - sft);
        for (int i = nu - 1; i > 0; i--) u[i] = (u[i] << sft) | (u[i - 1] >> (32 - sft));
        u[0] <<= sft;
    } else {
        for (int i = 0; i < nv; i++) vn[i] = v[i];
        u[nu] = 0;
    }
    const uint64_t vtop = vn[nv - 1], vsec = vn[nv - 2];
    for (int j = nu - nv; j >= 0; j--) {
        uint64_t num = ((uint64_t)u[j + nv] << 32) | u[j + nv - 1];
        uint64_t qhat = num / vtop, rhat = num - qhat * vtop;
        while (qhat >> 32 || qhat * vsec > ((rhat << 32) | u[j + nv - 2])) {
            qhat--; rhat += vtop;
            if (rhat >> 32) break;
        }
        int64_t borrow = 0; uint64_t carry = 0;
        for (int i = 0; i < nv; i++) {
            uint64_t pr = qhat * vn[i] + carry;
            carry = pr >> 32;
            int64_t t = (int64_t)u[i + j] - (int64_t)(uint32_t)pr + borrow;
            u[i + j] = (uint32_t)t;
            borrow = t >> 32;                       // 0 or -1
        }
        int64_t t = (int64_t)u[j + nv] - (int64_t)carry + borrow;
        u[j + nv] = (uint32_t)t;
        if (t < 0) {                                // qhat was one too large: add v back
            uint64_t c = 0;
            for (int i = 0; i < nv; i++) { c += (uint64_t)u[i + j] + vn[i]; u[i + j] = (uint32_t)c; c >>= 32; }
            u[j + nv] += (uint32_t)c;
        }
    }
    if (sft) {
        for (int i = 0; i < nv - 1; i++) u[i] = (u[i] >> sft) | (u[i + 1] << (32 - sft));
        u[nv - 1] >>= sft;
    }
    return mag_trim(u, nv);
}

// Inner loop of Lehmer's gcd (Knuth 4.5.2 L, steps L2-L4) on the 62-bit leading parts uh >= vh: cofactors with
// |.| < 2^30 such that (A u + B v, C u + D v) continues the Euclidean sequence of the full numbers.  Knuth's test is
// floor((uh + A) / (vh + C)) == floor((uh + B) / (vh + D)).  A 64-bit division is ~120 instructions on the GPU and the
// loop used to do two per quotient, which made it 90 % of a 50-limb gcd: quotients 0..3 (two thirds of all, by the
// Gauss-Kuzmin law) are now found by comparison, larger ones by ONE division, and the second quotient is never formed
// -- q d2 <= n2 < (q + 1) d2 is checked by a multiplication.
RELP_HD void lehmer_cofactors(int64_t uh, int64_t vh, int64_t& A, int64_t& B, int64_t& C, int64_t& D) {
    for (;;) {
        if (vh + C <= 0 || vh + D <= 0 || uh + A < 0 || uh + B < 0) break;
        const uint64_t n1 = (uint64_t)(uh + A), d1 = (uint64_t)(vh + C), n2 = (uint64_t)(uh + B), d2 = (uint64_t)(vh + D);   // all < 2^62 + 2^30
        uint64_t q, lo;
        if (n1 < d1) { q = 0; lo = 0; }
        else if (n1 - d1 < d1) { q = 1; lo = d2; }
        else if (n1 - 2 * d1 < d1) { q = 2; lo = 2 * d2; }
        else if (n1 - 3 * d1 < d1) { q = 3; lo = 3 * d2; }
        else {
            q = n1 / d1;
            if (q > 0x7fffffffULL) break;
#if defined(__CUDA_ARCH__)
            if (__umul64hi(q, d2)) break;
#else
            if ((unsigned __int128)q * d2 >> 64) break;
#endif
            lo = q * d2;
        }
        if (lo > n2 || n2 - lo >= d2) break;                                   // the two quotients differ
        const int64_t sq = (int64_t)q;
        const int64_t nA = A - sq * C, nB = B - sq * D;
        if (nA > 0x3fffffffLL || nA < -0x3fffffffLL || nB > 0x3fffffffLL || nB < -0x3fffffffLL) break;
        const int64_t nh = uh - sq * vh;
        A = C; C = nA; B = D; D = nB; uh = vh; vh = nh;
    }
}

// GCD of two nonzero magnitudes (Lehmer, Knuth 4.5.2 L, with 62-bit leading parts and 31-bit cofactors);
// destroys both buffers; *res points at the buffer (u or v) holding the result; returns its limb
// count.  Each buffer needs one spare limb beyond max(nu, nv).  `tmp`: min(nu, nv) limbs.
// One outer step removes about 30 bits from both operands for two passes over the limbs; as soon as the
// smaller operand fits 64 bits the rest is one modular reduction and a 64-bit gcd.
RELP_HD int mag_gcd(uint32_t* u, int nu, uint32_t* v, int nv, uint32_t** res, uint32_t* tmp) {
    if (mag_cmp(u, nu, v, nv) < 0) { uint32_t* t = u; u = v; v = t; int ti = nu; nu = nv; nv = ti; }
    while (nv > 2) {                                 // invariant: u >= v > 0
        // leading 62 bits of u and the bits of v in the same position
        const int top = nu - 1;
        const int sft = clz32(u[top]);
        uint64_t uw = ((uint64_t)u[top] << 32) | u[top - 1], vw = 0;
        uint32_t u3 = top >= 2 ? u[top - 2] : 0, v3 = 0;
        if (nv == nu) { vw = ((uint64_t)v[top] << 32) | v[top - 1]; v3 = top >= 2 ? v[top - 2] : 0; }
        else if (nv == nu - 1) { vw = v[top - 1]; v3 = top >= 2 ? v[top - 2] : 0; }
        if (sft) { uw = (uw << sft) | (u3 >> (32 - sft)); vw = (vw << sft) | (v3 >> (32 - sft)); }
        int64_t A = 1, B = 0, C = 0, D = 1;
        if (nv >= nu - 1) lehmer_cofactors((int64_t)(uw >> 2), (int64_t)(vw >> 2), A, B, C, D);
        if (B == 0) {                                // no usable cofactors: one full remainder step
            int nr = mag_mod(u, nu, v, nv, tmp);
            uint32_t* t = u; u = v; v = t; nu = nv; nv = nr;
            if (nv == 0) { *res = u; return nu; }
            continue;
        }
        // (u, v) <- (A u + B v, C u + D v); both are non-negative and smaller than before
        int64_t cu = 0, cv = 0;
        for (int i = 0; i < nu; i++) {
            int64_t ui = u[i], vi = i < nv ? v[i] : 0;
            int64_t tu = A * ui + B * vi + cu, tv = C * ui + D * vi + cv;
            u[i] = (uint32_t)tu; cu = tu >> 32;
            v[i] = (uint32_t)tv; cv = tv >> 32;
        }
        int n2 = mag_trim(v, nu);
        nu = mag_trim(u, nu); nv = n2;
        if (nv == 0) { *res = u; return nu; }
        if (mag_cmp(u, nu, v, nv) < 0) { uint32_t* t = u; u = v; v = t; int ti = nu; nu = nv; nv = ti; }
    }
    uint64_t bq = mag_lo64(v, nv);
    uint64_t aq = nu <= 2 ? mag_lo64(u, nu) : mag_mod64(u, nu, bq);
    *res = u;
    return mag_set64(u, gcd64(aq, bq));
}

// inverse of an odd 32-bit number modulo 2^32 (Newton)
RELP_HD uint32_t inv32(uint32_t d) {
    uint32_t x = d;                 // correct to 3 bits
    x *= 2u - d * x; x *= 2u - d * x; x *= 2u - d * x; x *= 2u - d * x;
    return x;
}

// q = a / d where d divides a exactly (Jebelean: multiply by d^-1 modulo 2^(32 nq)).
// `a` is destroyed (used as the running remainder).  q must not alias a.  Returns limbs of q.
RELP_HD int mag_divexact(uint32_t* q, uint32_t* a, int na, const uint32_t* d, int nd, uint32_t* dtmp) {
    if (na == 0) return 0;
    if (nd == 1 && d[0] == 1) { for (int i = 0; i < na; i++) q[i] = a[i]; return na; }
    int z = mag_ctz(d, nd);
    const uint32_t* dd = d;
    if (z) {                                     // strip the common power of two
        for (int i = 0; i < nd; i++) dtmp[i] = d[i];
        nd = mag_shr(dtmp, nd, z);
        na = mag_shr(a, na, z);
        dd = dtmp;
    }
    int nq = na - nd + 1;                        // quotient has nq or nq-1 limbs
    uint32_t inv = inv32(dd[0]);
    for (int i = 0; i < nq; i++) {
        uint32_t qi = a[i] * inv;
        q[i] = qi;
        if (qi) {                                // a[i..nq) -= qi * dd   (truncated at limb nq)
            uint64_t borrow = 0;
            int lim = nq - i < nd ? nq - i : nd;
            int j = 0;
            for (; j < lim; j++) {
                uint64_t p = (uint64_t)qi * dd[j] + borrow;
                uint32_t lo = (uint32_t)p;
                borrow = (p >> 32) + (a[i + j] < lo);
                a[i + j] -= lo;
            }
            for (int k = i + j; borrow && k < nq; k++) {
                uint32_t lo = (uint32_t)borrow;
                borrow = (borrow >> 32) + (a[k] < lo);
                a[k] -= lo;
            }
        }
    }
    return mag_trim(q, nq);
}

// ------------------------------------------------------------------------------------------
// Views of stored values
// ------------------------------------------------------------------------------------------
struct Ref {
    const uint32_t* num; int nn;
    const uint32_t* den; int nd;
    int sign;                    // -1, 0, +1
    uint32_t tmp[4];             // backing store for small values
};

RELP_HD void ref_load(Ref& r, const Val& v, const uint32_t* heap) {
    if (v.d > 0) {
        uint64_t a = abs64(v.n);
        r.tmp[0] = (uint32_t)a; r.tmp[1] = (uint32_t)(a >> 32);
        r.tmp[2] = (uint32_t)v.d; r.tmp[3] = (uint32_t)((uint64_t)v.d >> 32);
        r.num = r.tmp; r.nn = r.tmp[1] ? 2 : (r.tmp[0] ? 1 : 0);
        r.den = r.tmp + 2; r.nd = r.tmp[3] ? 2 : 1;
        r.sign = v.n > 0 ? 1 : (v.n < 0 ? -1 : 0);
    } else {
        const uint32_t* p = heap + v.n;
        uint32_t h = p[0];
        r.nn = (int)(h & 0x3fff); r.nd = (int)((h >> 14) & 0x3fff);
        r.sign = (h >> 31) ? -1 : 1;
        r.num = p + 1; r.den = p + 1 + r.nn;
    }
}

RELP_HD unsigned big_words(const uint32_t* heap, const Val& v) {
    uint32_t h = heap[v.n];
    return 1u + (h & 0x3fff) + ((h >> 14) & 0x3fff);
}

RELP_HD int val_sign(const Val& v, const uint32_t* heap) {
    if (v.d > 0) return v.n > 0 ? 1 : (v.n < 0 ? -1 : 0);
    return (heap[v.n] >> 31) ? -1 : 1;
}

// Store a canonical (already reduced, den > 0) value: small if it fits, else in the arena.
RELP_HD Val store_val(int sign, const uint32_t* num, int nn, const uint32_t* den, int nd,
                      uint32_t* heap, const Arena& ar) {
    if (nn == 0) return val_zero();
    if (nn <= 2 && nd <= 2) {
        uint64_t a = num[0] | (nn > 1 ? (uint64_t)num[1] << 32 : 0);
        uint64_t b = den[0] | (nd > 1 ? (uint64_t)den[1] << 32 : 0);
        if (a <= 0x7fffffffffffffffULL && b <= 0x7fffffffffffffffULL)
            return make_small(sign < 0 ? -(int64_t)a : (int64_t)a, (int64_t)b);
    }
    if (nn > RELP_MAX_LIMBS || nd > RELP_MAX_LIMBS) { set_err(ar.err, ERR_LIMB_CAPACITY); return val_zero(); }
    if ((nn > RELP_LONG_TRIGGER || nd > RELP_LONG_TRIGGER) && ar.flag) *ar.flag = 1;      // "long values exist"
    long long off = arena_alloc(ar, 1u + nn + nd);
    if (off < 0) return val_zero();
    uint32_t* p = heap + off;
    p[0] = (uint32_t)nn | ((uint32_t)nd << 14) | (sign < 0 ? 0x80000000u : 0u);
    for (int i = 0; i < nn; i++) p[1 + i] = num[i];
    for (int i = 0; i < nd; i++) p[1 + nn + i] = den[i];
    Val v; v.n = off; v.d = 0;
    return v;
}

// Copy a value into another arena (persistent <-> transient, compaction).
RELP_HD Val copy_val(const Val& v, uint32_t* heap, const Arena& ar) {
    if (v.d != 0) return v;
    unsigned w = big_words(heap, v);
    long long off = arena_alloc(ar, w);
    if (off < 0) return val_zero();
    for (unsigned i = 0; i < w; i++) heap[off + i] = heap[v.n + i];
    Val o; o.n = off; o.d = 0;
    return o;
}

// gcd of two magnitudes that must stay intact: copies into scratch (each max(na, nb) + 2 limbs) first.
// g == 1 is recognised without any work when either operand is 1.
RELP_HD int gcd_keep(const uint32_t* a, int na, const uint32_t* b, int nb, uint32_t* s1, uint32_t* s2, uint32_t** g, uint32_t* tmp) {
    if ((na == 1 && a[0] == 1) || (nb == 1 && b[0] == 1)) { s1[0] = 1; *g = s1; return 1; }
    for (int i = 0; i < na; i++) s1[i] = a[i];
    for (int i = 0; i < nb; i++) s2[i] = b[i];
    return mag_gcd(s1, na, s2, nb, g, tmp);
}
// q = a / g exactly, a stays intact (tmp: na limbs for the running remainder, dtmp: ng limbs)
RELP_HD int div_keep(uint32_t* q, const uint32_t* a, int na, const uint32_t* g, int ng, uint32_t* tmp, uint32_t* dtmp) {
    if (ng == 1 && g[0] == 1) { for (int i = 0; i < na; i++) q[i] = a[i]; return na; }
    for (int i = 0; i < na; i++) tmp[i] = a[i];
    return mag_divexact(q, tmp, na, g, ng, dtmp);
}

// out = x + s * a * b for arbitrary tiers.                      (slow path of K1/K2/K3/K5)
// Knuth 4.5.1 / GMP mpq: the product is formed from cross-reduced factors (two gcds of operand size),
// the sum needs gcd(x.d, t.d) and -- only when that is not 1 -- one gcd against it.  No gcd ever sees
// the full-size numerator and denominator of the unreduced result, which is what made the plain
// "compute, then reduce" version spend 98 % of its time in one 3x-length binary GCD.
RELP_HDN Val big_fma_thread(const Val& x, const Val& a, const Val& b, int s, uint32_t* heap, const Arena& ar) {
    Ref rx, ra, rb;
    ref_load(rx, x, heap); ref_load(ra, a, heap); ref_load(rb, b, heap);
    if (ra.sign == 0 || rb.sign == 0) return copy_val(x, heap, ar);
    if (rx.nn > RELP_OPL || rx.nd > RELP_OPL || ra.nn > RELP_OPL || ra.nd > RELP_OPL ||
        rb.nn > RELP_OPL || rb.nd > RELP_OPL) { set_err(ar.err, ERR_LIMB_CAPACITY); return val_zero(); }
    uint32_t s1[RELP_TL], s2[RELP_TL], s3[RELP_TL];          // scratch
    uint32_t f1[RELP_OPL + 2], f2[RELP_OPL + 2];             // reduced factors
    uint32_t tn[2 * RELP_OPL + 2], td[2 * RELP_OPL + 2];     // t = a * b in lowest terms
    uint32_t* g;
    // t.n = (a.n / g1)(b.n / g2), t.d = (a.d / g2)(b.d / g1), g1 = gcd(a.n, b.d), g2 = gcd(b.n, a.d)
    int ng = gcd_keep(ra.num, ra.nn, rb.den, rb.nd, s1, s2, &g, s3);
    int n1 = div_keep(f1, ra.num, ra.nn, g, ng, s3, g == s1 ? s2 : s1);
    int n2 = div_keep(f2, rb.den, rb.nd, g, ng, s3, g == s1 ? s2 : s1);
    // (f1 = a.n / g1, f2 = b.d / g1)  -- second pair goes through s-buffers, products are formed at once
    uint32_t h1[RELP_OPL + 2], h2[RELP_OPL + 2];
    ng = gcd_keep(rb.num, rb.nn, ra.den, ra.nd, s1, s2, &g, s3);
    int m1 = div_keep(h1, rb.num, rb.nn, g, ng, s3, g == s1 ? s2 : s1);
    int m2 = div_keep(h2, ra.den, ra.nd, g, ng, s3, g == s1 ? s2 : s1);
    int ntn = mag_mul(tn, f1, n1, h1, m1);
    int ntd = mag_mul(td, h2, m2, f2, n2);
    const int st = ra.sign * rb.sign * s;
    if (rx.sign == 0) return store_val(st, tn, ntn, td, ntd, heap, ar);
    // x + t
    ng = gcd_keep(rx.den, rx.nd, td, ntd, s1, s2, &g, s3);
    if (ng == 1 && g[0] == 1) {
        // coprime denominators: (x.n t.d + t.n x.d) / (x.d t.d) is already in lowest terms
        int np = mag_mul(s1, rx.num, rx.nn, td, ntd);
        int nq = mag_mul(s2, tn, ntn, rx.den, rx.nd);
        int sign;
        if (rx.sign == st) { np = mag_add(s1, s1, np, s2, nq); sign = st; }
        else {
            int c = mag_cmp(s1, np, s2, nq);
            if (c == 0) return val_zero();
            if (c > 0) { np = mag_sub(s1, s1, np, s2, nq); sign = rx.sign; }
            else { np = mag_sub(s1, s2, nq, s1, np); sign = st; }
        }
        int nden = mag_mul(s3, rx.den, rx.nd, td, ntd);
        return store_val(sign, s1, np, s3, nden, heap, ar);
    }
    // d1 = g: num = x.n (t.d / d1) + t.n (x.d / d1); d2 = gcd(num, d1); result (num / d2) / ((x.d / d1)(t.d / d2))
    uint32_t d1[RELP_OPL + 2];
    const int nd1 = ng;
    for (int i = 0; i < ng; i++) d1[i] = g[i];
    uint32_t xq[RELP_OPL + 2], tq[2 * RELP_OPL + 2];
    int nxq = div_keep(xq, rx.den, rx.nd, d1, nd1, s1, s2);
    int ntq = div_keep(tq, td, ntd, d1, nd1, s1, s2);
    int np = mag_mul(s1, rx.num, rx.nn, tq, ntq);
    int nq = mag_mul(s2, tn, ntn, xq, nxq);
    int sign;
    if (rx.sign == st) { np = mag_add(s1, s1, np, s2, nq); sign = st; }
    else {
        int c = mag_cmp(s1, np, s2, nq);
        if (c == 0) return val_zero();
        if (c > 0) { np = mag_sub(s1, s1, np, s2, nq); sign = rx.sign; }
        else { np = mag_sub(s1, s2, nq, s1, np); sign = st; }
    }
    // s1 = num (np limbs).  d2 = gcd(num, d1)
    uint32_t g3[RELP_OPL + 2];
    ng = gcd_keep(s1, np, d1, nd1, s2, s3, &g, tn);            // tn is free now; the remainder scratch needs min(np, nd1) limbs
    const int ng3 = ng;
    for (int i = 0; i < ng; i++) g3[i] = g[i];
    int nnum = div_keep(s2, s1, np, g3, ng3, s3, tn);             // tn is free now (used as divisor scratch)
    int ntd2 = div_keep(s1, td, ntd, g3, ng3, s3, tn);            // s1 = t.d / d2
    int nden = mag_mul(s3, xq, nxq, s1, ntd2);
    return store_val(sign, s2, nnum, s3, nden, heap, ar);
}

// out = x / y (y != 0) for arbitrary tiers: (x.n / g1)(y.d / g2) / ((x.d / g2)(y.n / g1)), g1 = gcd(x.n, y.n),
// g2 = gcd(y.d, x.d) -- in lowest terms without a gcd of the products.
RELP_HDN Val big_div_thread(const Val& x, const Val& y, uint32_t* heap, const Arena& ar) {
    Ref rx, ry;
    ref_load(rx, x, heap); ref_load(ry, y, heap);
    if (rx.sign == 0) return val_zero();
    if (rx.nn > RELP_OPL || rx.nd > RELP_OPL || ry.nn > RELP_OPL || ry.nd > RELP_OPL) {
        set_err(ar.err, ERR_LIMB_CAPACITY); return val_zero();
    }
    uint32_t s1[RELP_OPL + 2], s2[RELP_OPL + 2], s3[RELP_OPL + 2];
    uint32_t f1[RELP_OPL + 2], f2[RELP_OPL + 2], h1[RELP_OPL + 2], h2[RELP_OPL + 2];
    uint32_t num[2 * RELP_OPL + 2], den[2 * RELP_OPL + 2];
    uint32_t* g;
    int ng = gcd_keep(rx.num, rx.nn, ry.num, ry.nn, s1, s2, &g, s3);
    int n1 = div_keep(f1, rx.num, rx.nn, g, ng, s3, g == s1 ? s2 : s1);
    int n2 = div_keep(f2, ry.num, ry.nn, g, ng, s3, g == s1 ? s2 : s1);
    ng = gcd_keep(ry.den, ry.nd, rx.den, rx.nd, s1, s2, &g, s3);
    int m1 = div_keep(h1, ry.den, ry.nd, g, ng, s3, g == s1 ? s2 : s1);
    int m2 = div_keep(h2, rx.den, rx.nd, g, ng, s3, g == s1 ? s2 : s1);
    int nn = mag_mul(num, f1, n1, h1, m1);
    int nd = mag_mul(den, h2, m2, f2, n2);
    return store_val(rx.sign * ry.sign, num, nn, den, nd, heap, ar);
}

// sign(x - y) for arbitrary tiers.
RELP_HDN int big_cmp_thread(const Val& x, const Val& y, const uint32_t* heap) {
    Ref rx, ry;
    ref_load(rx, x, heap); ref_load(ry, y, heap);
    if (rx.sign != ry.sign) return rx.sign < ry.sign ? -1 : 1;
    if (rx.sign == 0) return 0;
    uint32_t l[2 * RELP_OPL], r[2 * RELP_OPL];
    int nl = mag_mul(l, rx.num, rx.nn, ry.den, ry.nd);
    int nr = mag_mul(r, ry.num, ry.nn, rx.den, rx.nd);
    int c = mag_cmp(l, nl, r, nr);
    return rx.sign > 0 ? c : -c;
}

#ifndef RELP_NO_WCB
}  // namespace RELP_NS
#include "relp_wcb.cuh"
namespace RELP_NS {

// ------------------------------------------------------------------------------------------
// Cooperative tier: the same three operations on operands of any length, by a group of lanes
// ------------------------------------------------------------------------------------------
#if defined(__CUDACC__) && !defined(RELP_EMUL)
struct WarpGroup {      // all 32 lanes of a warp, converged by the caller (one-dimensional blocks, a multiple of 32 threads)
    __device__ int rank() const { return (int)(threadIdx.x & 31u); }
    __device__ int size() const { return 32; }
    __device__ void sync() const { __syncwarp(); }
    __device__ uint32_t ballot(bool p) const { return __ballot_sync(0xffffffffu, p); }
    template <class T> __device__ T shfl(T v, int src) const { return __shfl_sync(0xffffffffu, v, src); }
    template <class T> __device__ T shfl_up(T v, int d) const { return __shfl_up_sync(0xffffffffu, v, d); }
};
#endif



#if defined(RELP_WCB_TEST_SLAB) && !defined(__CUDA_ARCH__)
extern uint32_t* relp_test_slab; extern unsigned relp_test_slab_words;
#endif
template <class G> struct WOp {          // scratch of one cooperative operation
    WScratch<G> sc; int slot; bool locked; const ScratchExt* ext;
};
template <class G> RELP_HD void wop_begin(const G& g, WOp<G>& op, uint32_t* heap, const Arena& ar) {
    op.sc.slab = nullptr; op.sc.cap = 0; op.sc.used = 0; op.sc.heap = heap; op.slot = -1; op.locked = false; op.sc.spills = 0;
    op.sc.fallback = ar;
    op.sc.fallback.flag = nullptr;
    const ScratchExt* x = heap_scratch(heap);
    op.ext = x;
    if (x && x->scr_top) { op.sc.fallback.top = x->scr_top; op.sc.fallback.limit = x->scr_limit; }
#if defined(__CUDA_ARCH__) && !defined(RELP_EMUL)
    if (g.size() == 32 && x && x->wslab && x->wslots > 0) {      // a whole warp: its slab (%warpid is unique within a CTA only, hence the lock)
        unsigned smid, warpid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        asm volatile("mov.u32 %0, %%warpid;" : "=r"(warpid));
        const int slot = (int)((smid * 64u + warpid) % (unsigned)x->wslots);
        int got = 0;
        if (g.rank() == 0) got = atomicCAS(&x->wlock[slot], 0, 1) == 0;
        got = g.shfl(got, 0);
        if (got) { op.sc.slab = x->wslab + (size_t)slot * x->wslab_words; op.sc.cap = x->wslab_words; op.slot = slot; op.locked = true; }
    }
    if (g.size() == 1 && x && x->tslab) {                         // one thread: its own slab
        const unsigned long long gt = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
        if (gt < x->tslots) { op.sc.slab = x->tslab + (size_t)gt * x->tslab_words; op.sc.cap = x->tslab_words; }
    }
#endif
#if defined(RELP_WCB_TEST_SLAB) && !defined(__CUDA_ARCH__)
    if (relp_test_slab_words) { op.sc.slab = relp_test_slab; op.sc.cap = relp_test_slab_words; }     // host tests of the slab's stack discipline
#endif
    (void)g;
}
template <class G> RELP_HD void wop_end(const G& g, WOp<G>& op, const Arena& ar) {
#if defined(__CUDA_ARCH__) && !defined(RELP_EMUL)
    if (op.locked) { g.sync(); if (g.rank() == 0) atomicExch(&op.ext->wlock[op.slot], 0); }
#endif
    (void)g; (void)op; (void)ar;
}

// store a canonical value (every lane holds the same lengths; the limbs are in memory)
RELP_W Val w_store(const G& g, int sign, const uint32_t* num, int nn, const uint32_t* den, int nd, uint32_t* heap, const Arena& ar) {
    if (nn == 0) return val_zero();
    if (nn <= 2 && nd <= 2) {
        const uint64_t a = num[0] | (nn > 1 ? (uint64_t)num[1] << 32 : 0), b = den[0] | (nd > 1 ? (uint64_t)den[1] << 32 : 0);
        if (a <= 0x7fffffffffffffffULL && b <= 0x7fffffffffffffffULL) return make_small(sign < 0 ? -(int64_t)a : (int64_t)a, (int64_t)b);
    }
    if (nn > RELP_MAX_LIMBS || nd > RELP_MAX_LIMBS) { set_err(ar.err, ERR_LIMB_CAPACITY); return val_zero(); }
    long long off = 0;
    if (g.rank() == 0) off = arena_alloc(ar, 1u + nn + nd);
    off = g.shfl(off, 0);
    if (off < 0) return val_zero();
    uint32_t* p = heap + off;
    if (g.rank() == 0) {
        p[0] = (uint32_t)nn | ((uint32_t)nd << 14) | (sign < 0 ? 0x80000000u : 0u);
        if ((nn > RELP_LONG_TRIGGER || nd > RELP_LONG_TRIGGER) && ar.flag) *ar.flag = 1;
    }
    for (int i = g.rank(); i < nn; i += g.size()) p[1 + i] = num[i];
    for (int i = g.rank(); i < nd; i += g.size()) p[1 + nn + i] = den[i];
    g.sync();
    Val v; v.n = off; v.d = 0;
    return v;
}
RELP_W Val w_copy_val(const G& g, const Val& v, uint32_t* heap, const Arena& ar) {
    if (v.d != 0) return v;
    const unsigned w = big_words(heap, v);
    long long off = 0;
    if (g.rank() == 0) off = arena_alloc(ar, w);
    off = g.shfl(off, 0);
    if (off < 0) return val_zero();
    for (unsigned i = g.rank(); i < w; i += g.size()) heap[off + i] = heap[v.n + i];
    g.sync();
    Val o; o.n = off; o.d = 0;
    return o;
}

// gcd(p, q) of two magnitudes that must stay intact, into scratch; *gp points at it, returns its limbs.
// gcd = 1 is recognised without any work when either operand is 1 (*gp = nullptr then, and 1 is returned).
RELP_W int w_gcd_of(const G& g, WScratch<G>& sc, const uint32_t* p, int np, const uint32_t* q, int nq, uint32_t** gp) {
    *gp = nullptr;
    if ((np == 1 && p[0] == 1) || (nq == 1 && q[0] == 1)) return 1;
    const int cap = (np > nq ? np : nq) + 3;
    const unsigned mk = sc.mark(), sp = sc.spills;
    uint32_t* u = sc.take(g, cap); uint32_t* v = sc.take(g, cap);
    uint32_t* t1 = sc.take(g, cap); uint32_t* t2 = sc.take(g, cap); uint32_t* t3 = sc.take(g, cap);
    if (!u || !v || !t1 || !t2 || !t3) return -1;
    w_copy(g, u, p, np); w_copy(g, v, q, nq);
    const int ng = w_gcd(g, u, np, v, nq, t1, t2, t3, gp);
    if (ng == 1 && (*gp)[0] == 1) *gp = nullptr;
    if (sc.releasable(sp)) {          // keep the result at the bottom of what was taken, give the other four buffers back
        g.sync();                     // (every lane is done reading them: what is given back is overwritten by the next take)
        if (*gp && *gp != u) { w_copy(g, u, *gp, ng); *gp = u; }
        sc.release_to(mk + (*gp ? (unsigned)ng : 0u));
    }
    return ng;
}
// p / d exactly into fresh scratch (p stays intact); d == nullptr means 1: the operand itself is returned
RELP_W const uint32_t* w_quot(const G& g, WScratch<G>& sc, const uint32_t* p, int np, const uint32_t* d, int nd, int* nout) {
    if (!d) { *nout = np; return p; }
    const unsigned mk = sc.mark(), sp = sc.spills;
    uint32_t* out = sc.take(g, np + 1); uint32_t* cp = sc.take(g, np + 1);
    uint32_t* dt = sc.take(g, nd + 1); uint32_t* pt = sc.take(g, nd + 2);
    if (!out || !cp || !dt || !pt) { *nout = -1; return nullptr; }
    w_copy(g, cp, p, np);
    *nout = w_divexact(g, out, cp, np, d, nd, dt, pt);
    if (sc.releasable(sp)) { g.sync(); sc.release_to(mk + (unsigned)(*nout > 0 ? *nout : 1)); }      // the quotient stays, its temporaries go (once every lane is done with them)
    return out;
}

// Sanity of an operand before a cooperative operation walks it: a handle must lie inside the heap and its header must
// describe limbs that do too.  A violation is reported (ERR_PEER bit + a location code in ext->dbg), not followed.
RELP_HD bool w_check(const Val& v, const uint32_t* heap, const Arena& ar, int where) {
    if (v.d != 0) return v.d > 0;
    const ScratchExt* x = heap_scratch(heap);
    if (!x || !x->heap_words) return true;
    bool ok = v.n >= 16 && (unsigned long long)v.n < x->heap_words;
    if (ok) { const uint32_t h = heap[v.n]; const unsigned nn = h & 0x3fff, nd = (h >> 14) & 0x3fff; ok = nn >= 1 && nd >= 1 && (unsigned long long)v.n + 1 + nn + nd <= x->heap_words; }
    if (!ok) { set_err(ar.err, ERR_PEER); const_cast<ScratchExt*>(x)->dbg = where; }
    return ok;
}

// out = x + s * a * b, cooperative (same steps as big_fma_thread: Knuth 4.5.1 / mpq)
RELP_W Val wcb_fma(const G& g, const Val& x, const Val& a, const Val& b, int s, uint32_t* heap, const Arena& ar) {
    if (!w_check(x, heap, ar, 1) || !w_check(a, heap, ar, 2) || !w_check(b, heap, ar, 3)) return val_zero();
    Ref rx, ra, rb;
    ref_load(rx, x, heap); ref_load(ra, a, heap); ref_load(rb, b, heap);
    if (ra.sign == 0 || rb.sign == 0) return w_copy_val(g, x, heap, ar);
    WOp<G> op; wop_begin(g, op, heap, ar);
    WScratch<G>& sc = op.sc;
    Val out = val_zero();
    bool fail = false;
    do {
        uint32_t* gp; int ng, n1, n2, m1, m2;
        // t = a * b in lowest terms: (a.n / g1)(b.n / g2) / ((a.d / g2)(b.d / g1))
        ng = w_gcd_of(g, sc, ra.num, ra.nn, rb.den, rb.nd, &gp); if (ng < 0) { fail = true; break; }
        const uint32_t* f1 = w_quot(g, sc, ra.num, ra.nn, gp, ng, &n1);
        const uint32_t* f2 = w_quot(g, sc, rb.den, rb.nd, gp, ng, &n2);
        ng = w_gcd_of(g, sc, rb.num, rb.nn, ra.den, ra.nd, &gp); if (ng < 0) { fail = true; break; }
        const uint32_t* h1 = w_quot(g, sc, rb.num, rb.nn, gp, ng, &m1);
        const uint32_t* h2 = w_quot(g, sc, ra.den, ra.nd, gp, ng, &m2);
        if (n1 < 0 || n2 < 0 || m1 < 0 || m2 < 0) { fail = true; break; }
        uint32_t* tn = sc.take(g, n1 + m1 + 1); uint32_t* td = sc.take(g, n2 + m2 + 1);
        if (!tn || !td) { fail = true; break; }
        const int ntn = w_mul(g, tn, f1, n1, h1, m1), ntd = w_mul(g, td, h2, m2, f2, n2);
        const int st = ra.sign * rb.sign * s;
        if (rx.sign == 0) { out = w_store(g, st, tn, ntn, td, ntd, heap, ar); break; }
        // x + t: d1 = gcd(x.d, t.d)
        uint32_t* d1; const int nd1 = w_gcd_of(g, sc, rx.den, rx.nd, td, ntd, &d1); if (nd1 < 0) { fail = true; break; }
        int nxq, ntq;
        const uint32_t* xq = w_quot(g, sc, rx.den, rx.nd, d1, nd1, &nxq);
        const uint32_t* tq = w_quot(g, sc, td, ntd, d1, nd1, &ntq);
        if (nxq < 0 || ntq < 0) { fail = true; break; }
        const int npq = (rx.nn + ntq > ntn + nxq ? rx.nn + ntq : ntn + nxq) + 2;     // P also receives P + Q or |P - Q|
        uint32_t* P = sc.take(g, npq); uint32_t* Q = sc.take(g, npq);
        if (!P || !Q) { fail = true; break; }
        int np = w_mul(g, P, rx.num, rx.nn, tq, ntq), nq = w_mul(g, Q, tn, ntn, xq, nxq);
        int sign;
        if (rx.sign == st) { np = w_add(g, P, P, np, Q, nq); sign = st; }
        else {
            const int c = w_cmp(g, P, np, Q, nq);
            if (c == 0) break;                                        // exact cancellation: zero
            if (c > 0) { np = w_sub(g, P, P, np, Q, nq); sign = rx.sign; }
            else { np = w_sub(g, P, Q, nq, P, np); sign = st; }
        }
        if (!d1) {      // coprime denominators: (x.n t.d + t.n x.d) / (x.d t.d) is in lowest terms
            uint32_t* den = sc.take(g, rx.nd + ntd + 1); if (!den) { fail = true; break; }
            const int nden = w_mul(g, den, rx.den, rx.nd, td, ntd);
            out = w_store(g, sign, P, np, den, nden, heap, ar);
            break;
        }
        // d2 = gcd(num, d1); result (num / d2) / ((x.d / d1)(t.d / d2))
        uint32_t* d2; const int nd2 = w_gcd_of(g, sc, P, np, d1, nd1, &d2); if (nd2 < 0) { fail = true; break; }
        int nnum, ntd2;
        const uint32_t* num = w_quot(g, sc, P, np, d2, nd2, &nnum);
        const uint32_t* td2 = w_quot(g, sc, td, ntd, d2, nd2, &ntd2);
        if (nnum < 0 || ntd2 < 0) { fail = true; break; }
        uint32_t* den = sc.take(g, nxq + ntd2 + 1); if (!den) { fail = true; break; }
        const int nden = w_mul(g, den, xq, nxq, td2, ntd2);
        out = w_store(g, sign, num, nnum, den, nden, heap, ar);
    } while (false);
    if (fail) set_err(ar.err, ERR_HEAP_EXHAUSTED);
    wop_end(g, op, ar);
    return out;
}

// out = x / y (y != 0), cooperative
RELP_W Val wcb_div(const G& g, const Val& x, const Val& y, uint32_t* heap, const Arena& ar) {
    if (!w_check(x, heap, ar, 4) || !w_check(y, heap, ar, 5)) return val_zero();
    Ref rx, ry;
    ref_load(rx, x, heap); ref_load(ry, y, heap);
    if (rx.sign == 0) return val_zero();
    WOp<G> op; wop_begin(g, op, heap, ar);
    WScratch<G>& sc = op.sc;
    Val out = val_zero();
    bool fail = false;
    do {
        uint32_t* gp; int ng, n1, n2, m1, m2;
        ng = w_gcd_of(g, sc, rx.num, rx.nn, ry.num, ry.nn, &gp); if (ng < 0) { fail = true; break; }
        const uint32_t* f1 = w_quot(g, sc, rx.num, rx.nn, gp, ng, &n1);
        const uint32_t* f2 = w_quot(g, sc, ry.num, ry.nn, gp, ng, &n2);
        ng = w_gcd_of(g, sc, ry.den, ry.nd, rx.den, rx.nd, &gp); if (ng < 0) { fail = true; break; }
        const uint32_t* h1 = w_quot(g, sc, ry.den, ry.nd, gp, ng, &m1);
        const uint32_t* h2 = w_quot(g, sc, rx.den, rx.nd, gp, ng, &m2);
        if (n1 < 0 || n2 < 0 || m1 < 0 || m2 < 0) { fail = true; break; }
        uint32_t* num = sc.take(g, n1 + m1 + 1); uint32_t* den = sc.take(g, n2 + m2 + 1);
        if (!num || !den) { fail = true; break; }
        const int nn = w_mul(g, num, f1, n1, h1, m1), nd = w_mul(g, den, h2, m2, f2, n2);
        out = w_store(g, rx.sign * ry.sign, num, nn, den, nd, heap, ar);
    } while (false);
    if (fail) set_err(ar.err, ERR_HEAP_EXHAUSTED);
    wop_end(g, op, ar);
    return out;
}

// sign(x - y), cooperative; `scr` says where the two cross products may live
RELP_W int wcb_cmp(const G& g, const Val& x, const Val& y, const uint32_t* heap, const Arena& scr) {
    Ref rx, ry;
    ref_load(rx, x, heap); ref_load(ry, y, heap);
    if (rx.sign != ry.sign) return rx.sign < ry.sign ? -1 : 1;
    if (rx.sign == 0) return 0;
    // lengths decide most comparisons: |x.n y.d| has nn + nd or nn + nd - 1 limbs
    const int ll = rx.nn + ry.nd, lr = ry.nn + rx.nd;
    if (ll > lr + 1) return rx.sign > 0 ? 1 : -1;
    if (lr > ll + 1) return rx.sign > 0 ? -1 : 1;
    WOp<G> op; wop_begin(g, op, const_cast<uint32_t*>(heap), scr);
    int c = 0;
    uint32_t* l = op.sc.take(g, ll + 1); uint32_t* r = op.sc.take(g, lr + 1);
    if (!l || !r) set_err(scr.err, ERR_HEAP_EXHAUSTED);
    else {
        const int nl = w_mul(g, l, rx.num, rx.nn, ry.den, ry.nd), nr = w_mul(g, r, ry.num, ry.nn, rx.den, rx.nd);
        c = w_cmp(g, l, nl, r, nr);
    }
    wop_end(g, op, scr);
    return rx.sign > 0 ? c : -c;
}

// ---- the multi-limb entry points: per-thread routines for short operands, the cooperative ones beyond ----
RELP_HD int ref_max_limbs(const Val& v, const uint32_t* heap) {
    if (v.d != 0) return 2;
    const uint32_t h = heap[v.n];
    const int nn = (int)(h & 0x3fff), nd = (int)((h >> 14) & 0x3fff);
    return nn > nd ? nn : nd;
}
RELP_HD int imax3(int a, int b, int c) { return a > b ? (a > c ? a : c) : (b > c ? b : c); }

#ifdef RELP_WCB_TRACE
#include <cstdio>
static void relp_wcb_trace(const char* op, const Val& x, const Val& a, const Val& b, int s, const uint32_t* heap) {
    static FILE* f = fopen("/root/repo/gpurun_out/wcb_trace.txt", "w");
    fprintf(f, "%s s=%d", op, s);
    const Val* vs[3] = {&x, &a, &b};
    for (int q = 0; q < 3; q++) {
        Ref r; ref_load(r, *vs[q], heap);
        fprintf(f, " | sign %d num", r.sign);
        for (int i = 0; i < r.nn; i++) fprintf(f, " %08x", r.num[i]);
        fprintf(f, " den");
        for (int i = 0; i < r.nd; i++) fprintf(f, " %08x", r.den[i]);
    }
    fprintf(f, "\n"); fflush(f);
}
#endif
#else   // RELP_NO_WCB: the copy of this header without the arbitrary-length tier (relp_fastloop.cu)
RELP_HD int ref_max_limbs(const Val& v, const uint32_t* heap) {
    if (v.d != 0) return 2;
    const uint32_t h = heap[v.n];
    const int nn = (int)(h & 0x3fff), nd = (int)((h >> 14) & 0x3fff);
    return nn > nd ? nn : nd;
}
RELP_HD int imax3(int a, int b, int c) { return a > b ? (a > c ? a : c) : (b > c ? b : c); }
// operands beyond RELP_OPL limbs cannot occur while DState::long_seen is clear (see RELP_LONG_TRIGGER); if one does,
// the per-thread routines report ERR_LIMB_CAPACITY and the host repeats the solve with the full tier
RELP_HD Val big_fma(const Val& x, const Val& a, const Val& b, int s, uint32_t* heap, const Arena& ar) { return big_fma_thread(x, a, b, s, heap, ar); }
RELP_HD Val big_div(const Val& x, const Val& y, uint32_t* heap, const Arena& ar) { return big_div_thread(x, y, heap, ar); }
RELP_HD int big_cmp(const Val& x, const Val& y, const uint32_t* heap) { return big_cmp_thread(x, y, heap); }
#endif
#ifndef RELP_NO_WCB
// From arbitrary (divergent) code a long operation is done by the calling lane alone, with the cooperative routines
// on a group of one: correct everywhere, no fixed size.  The stages that have many long operations call wcb_fma /
// wcb_div themselves with a converged WarpGroup (stage_update_big, relp_kernels.cuh).
// (An earlier version let whatever lanes had arrived together -- cooperative_groups::coalesced_threads() -- share each
// operation: right in a plain kernel, but inside the stages, where lanes arrive from loops with different trip counts,
// it hung or faulted on the device, while the same code on one lane per operation passed every test.)
RELP_HDN Val big_fma(const Val& x, const Val& a, const Val& b, int s, uint32_t* heap, const Arena& ar) {
    const ScratchExt* ext = heap_scratch(heap);
    const int thr = (ext && ext->coop_from < RELP_OPL) ? ext->coop_from : RELP_OPL;
    if (imax3(ref_max_limbs(x, heap), ref_max_limbs(a, heap), ref_max_limbs(b, heap)) <= thr) return big_fma_thread(x, a, b, s, heap, ar);
#ifdef RELP_WCB_TRACE
    relp_wcb_trace("fma", x, a, b, s, heap);
#endif
    return wcb_fma(SoloGroup(), x, a, b, s, heap, ar);
}

RELP_HDN Val big_div(const Val& x, const Val& y, uint32_t* heap, const Arena& ar) {
    const ScratchExt* ext = heap_scratch(heap);
    const int thr = (ext && ext->coop_from < RELP_OPL) ? ext->coop_from : RELP_OPL;
    const int mx = ref_max_limbs(x, heap), my = ref_max_limbs(y, heap);
    if ((mx > my ? mx : my) <= thr) return big_div_thread(x, y, heap, ar);
#ifdef RELP_WCB_TRACE
    relp_wcb_trace("div", x, y, y, 0, heap);
#endif
    return wcb_div(SoloGroup(), x, y, heap, ar);
}

RELP_HDN int big_cmp(const Val& x, const Val& y, const uint32_t* heap) {
    const int mx = ref_max_limbs(x, heap), my = ref_max_limbs(y, heap);
    if ((mx > my ? mx : my) <= RELP_OPL) return big_cmp_thread(x, y, heap);
    const ScratchExt* ext = heap_scratch(heap);
    Arena scr;                            // no result is stored: only the scratch provider matters
    if (ext) { scr.top = ext->scr_top; scr.limit = ext->scr_limit; scr.err = ext->err; }
    return wcb_cmp(SoloGroup(), x, y, heap, scr);
}
#endif   // RELP_NO_WCB

// ------------------------------------------------------------------------------------------
// Tier-dispatching front ends
// ------------------------------------------------------------------------------------------
RELP_HDN Val val_fma_slow(const Val& x, const Val& a, const Val& b, int s, uint32_t* heap, const Arena& ar) {
    Val out;
    if (x.d > 0 && a.d > 0 && b.d > 0) {
        if (a.n == 0 || b.n == 0) return x;
        if (small_fma_rational(x, a, b, s, &out)) return out;
    }
    return big_fma(x, a, b, s, heap, ar);
}
RELP_HD Val val_fma(const Val& x, const Val& a, const Val& b, int s, uint32_t* heap, const Arena& ar) {
    if (x.d == 1 && a.d == 1 && b.d == 1) {                 // integers: one 128-bit multiply-add
        i128 t = (i128)a.n * b.n;
        i128 r = s > 0 ? (i128)x.n + t : (i128)x.n - t;
        if (r <= RELP_I64_MAX && r >= -RELP_I64_MAX) return make_small((int64_t)r, 1);
    }
    return val_fma_slow(x, a, b, s, heap, ar);
}
RELP_HDN Val val_div_slow(const Val& x, const Val& y, uint32_t* heap, const Arena& ar) {
    Val out;
    if (x.d > 0 && y.d > 0 && small_div_general(x, y, &out)) return out;
    return big_div(x, y, heap, ar);
}
RELP_HD Val val_div(const Val& x, const Val& y, uint32_t* heap, const Arena& ar) {
    if (x.d > 0 && y.d == 1 && (y.n == 1 || y.n == -1)) return make_small(y.n > 0 ? x.n : -x.n, x.d);
    if (x.d == 1 && x.n == 0) return x;
    return val_div_slow(x, y, heap, ar);
}
RELP_HD int val_cmp(const Val& x, const Val& y, const uint32_t* heap) {
    if (x.d > 0 && y.d > 0) return small_cmp(x, y);
    return big_cmp(x, y, heap);
}
// b / d (d > 0) for COMPARISON only (ratio test, tableau/mod.rs:229): the unreduced fraction
// (b.n d.d) / (b.d d.n) when it fits -- val_cmp cross-multiplies, it does not need lowest terms, and
// reducing costs two gcds and four 64-bit divisions per row.  Falls back to the exact division.
RELP_HD Val val_ratio(const Val& b, const Val& d, uint32_t* heap, const Arena& ar) {
    if (b.d > 0 && d.d > 0) {
        i128 n = (i128)b.n * d.d, q = (i128)b.d * d.n;
        if (n <= RELP_I64_MAX && n >= -RELP_I64_MAX && q <= RELP_I64_MAX) return make_small((int64_t)n, (int64_t)q);
    }
    return val_div(b, d, heap, ar);
}

}  // namespace RELP_NS
