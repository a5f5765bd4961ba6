// Warp-cooperative multi-limb arithmetic ("WCB"): the arbitrary-length tier of the exact rationals
// (north star: "multi-limb big-integer kernels, GCD normalisation done
// warp-cooperatively").  Included by relp_arith.cuh.
//
// A number is a little-endian array of 32-bit limbs in memory (the limb heap or scratch).  A GROUP of
// lanes works on ONE operation at a time: limb i belongs to lane i mod S (S = lanes in the group), so a
// pass over a number is a loop over "chunks" of S consecutive limbs -- coalesced loads and stores --
// and carries cross lanes through one ballot per chunk (carry-lookahead: generate / propagate masks,
// carries-in = ((G << 1 | cin) + P) ^ P).  Lengths are plain ints, identical in every lane.  Nothing
// is sized at compile time: a value may have up to 16383 limbs (the header's field), bounded in
// practice by the limb heap -- the reference's RationalBig is unbounded too
// (/root/reference/src/data/number_types/rational/big/mod.rs:17-29).
//
// The group type G provides rank(), size(), sync(), ballot(pred) (bit r = lane of rank r),
// shfl(v, src_rank), shfl_up(v, delta).  On the device it is WarpGroup -- the 32 converged lanes of a warp,
// in the stages that queue many long operations (stage_update_big) -- or SoloGroup (one lane: host code, the
// CPU emulation build, and a device thread that meets a long value in divergent code).
// tests/native/arith_host_test.cpp runs the same templates on 1..32 host threads behind a barrier and
// checks them against Python integers before they ever run in a kernel.
#pragma once

namespace RELP_NS {

struct SoloGroup {
    RELP_HD int rank() const { return 0; }
    RELP_HD int size() const { return 1; }
    RELP_HD void sync() const {}
    RELP_HD uint32_t ballot(bool p) const { return p ? 1u : 0u; }
    template <class T> RELP_HD T shfl(T v, int) const { return v; }
    template <class T> RELP_HD T shfl_up(T v, int) const { return v; }
};

#if defined(__CUDACC__)
#define RELP_W template <class G> __host__ __device__ __noinline__
#else
#define RELP_W template <class G>
#endif

// Bump allocator for the temporaries of one operation: a slab that belongs to the group for the duration
// of the operation and, past its end, the transient region of the limb heap (so no operation fails for
// lack of scratch before the heap itself is full).  Every lane carries an identical copy.
template <class G> struct WScratch {
    uint32_t* slab; unsigned cap, used;
    uint32_t* heap; Arena fallback;
    unsigned spills = 0;          // takes served by the fallback arena (those cannot be given back before the iteration ends)
    // Stack discipline for the slab: a routine notes mark() on entry, takes its temporaries, and gives everything behind
    // its result back with release_to() -- possible only if nothing was spilled in between.
    RELP_HD unsigned mark() const { return used; }
    RELP_HD bool releasable(unsigned spills_at_mark) const { return slab != nullptr && spills == spills_at_mark; }
    RELP_HD void release_to(unsigned new_used) { used = new_used; }
    RELP_HD uint32_t* take(const G& g, int words) {
        unsigned w = (unsigned)(words < 1 ? 1 : words);
#if defined(RELP_WCB_TEST_MALLOC) && !defined(__CUDA_ARCH__)
        // host test build under AddressSanitizer: every buffer on its own, exactly as large as asked for
        uint64_t q = 0;
        if (g.rank() == 0) q = (uint64_t)(uintptr_t)(new uint32_t[w]);
        q = g.shfl(q, 0);
        return (uint32_t*)(uintptr_t)q;
#endif
        if (slab && used + w <= cap) { uint32_t* p = slab + used; used += w; return p; }
        spills++;
        if (!fallback.top) return nullptr;
        long long off = 0;
        if (g.rank() == 0) off = arena_alloc(fallback, w);
        off = g.shfl(off, 0);
        return off < 0 ? nullptr : heap + off;
    }
};

// ---- passes over one or two numbers ---------------------------------------------------------------
RELP_W int w_trim(const G& g, const uint32_t* a, int n) {
    const int S = g.size(), rk = g.rank();
    while (n > 0) {
        const int base = ((n - 1) / S) * S, i = base + rk;
        const uint32_t m = g.ballot(i < n && a[i] != 0);
        if (m) return base + 32 - clz32(m);
        n = base;
    }
    return 0;
}

RELP_W int w_cmp(const G& g, const uint32_t* a, int na, const uint32_t* b, int nb) {
    if (na != nb) return na < nb ? -1 : 1;
    const int S = g.size(), rk = g.rank();
    for (int n = na; n > 0;) {
        const int base = ((n - 1) / S) * S, i = base + rk;
        uint32_t x = 0, y = 0;
        if (i < n) { x = a[i]; y = b[i]; }
        const uint32_t m = g.ballot(x != y);
        if (m) {
            const int t = 31 - clz32(m);
            const uint32_t xt = g.shfl(x, t), yt = g.shfl(y, t);
            return xt < yt ? -1 : 1;
        }
        n = base;
    }
    return 0;
}

RELP_W void w_copy(const G& g, uint32_t* d, const uint32_t* s, int n) {
    for (int i = g.rank(); i < n; i += g.size()) d[i] = s[i];
    g.sync();
}

// number of trailing zero bits (n > 0, a != 0)
RELP_W int w_ctz(const G& g, const uint32_t* a, int n) {
    const int S = g.size(), rk = g.rank();
    for (int base = 0; base < n; base += S) {
        const int i = base + rk;
        const uint32_t x = i < n ? a[i] : 0u;
        const uint32_t m = g.ballot(x != 0);
        if (m) { const int t = ctz32(m); return 32 * (base + t) + ctz32(g.shfl(x, t)); }
    }
    return 32 * n;
}

// a[0..n) += b[0..nb) (nb <= n); returns the carry out.
RELP_W uint32_t w_add_n(const G& g, uint32_t* a, const uint32_t* b, int nb, int n) {
    const int S = g.size(), rk = g.rank();
    uint32_t cin = 0;
    for (int base = 0; base < n; base += S) {
        const int i = base + rk;
        const uint32_t x = i < n ? a[i] : 0u, y = i < nb ? b[i] : 0u;
        uint32_t s = x + y;
        const uint64_t Gm = ((uint64_t)g.ballot(s < x) << 1) | cin, Pm = g.ballot(i < n && s == 0xffffffffu);
        const uint64_t C = (Gm + Pm) ^ Pm;
        s += (uint32_t)(C >> rk) & 1u;
        if (i < n) a[i] = s;
        const int last = n - base < S ? n - base : S;           // the carry out of this chunk leaves at its last LIVE lane
        cin = (uint32_t)(C >> last) & 1u;
    }
    g.sync();
    return cin;
}

// a[0..n) -= b[0..nb) (nb <= n); returns the borrow out.
RELP_W uint32_t w_sub_n(const G& g, uint32_t* a, const uint32_t* b, int nb, int n) {
    const int S = g.size(), rk = g.rank();
    uint32_t bin = 0;
    for (int base = 0; base < n; base += S) {
        const int i = base + rk;
        const uint32_t x = i < n ? a[i] : 0u, y = i < nb ? b[i] : 0u;
        uint32_t s = x - y;
        const uint64_t Gm = ((uint64_t)g.ballot(x < y) << 1) | bin, Pm = g.ballot(i < n && s == 0u);
        const uint64_t C = (Gm + Pm) ^ Pm;
        s -= (uint32_t)(C >> rk) & 1u;
        if (i < n) a[i] = s;
        const int last = n - base < S ? n - base : S;
        bin = (uint32_t)(C >> last) & 1u;
    }
    g.sync();
    return bin;
}

// r = a + b (r may alias a or b; capacity max(na, nb) + 1); returns the limbs of r
RELP_W int w_add(const G& g, uint32_t* r, const uint32_t* a, int na, const uint32_t* b, int nb) {
    if (na < nb) { const uint32_t* t = a; a = b; b = t; const int ti = na; na = nb; nb = ti; }
    const int S = g.size(), rk = g.rank();
    uint32_t cin = 0;
    for (int base = 0; base <= na; base += S) {          // one position beyond the longer operand takes the last carry
        const int i = base + rk;
        const uint32_t x = i < na ? a[i] : 0u, y = i < nb ? b[i] : 0u;
        uint32_t s = x + y;
        const uint64_t Gm = ((uint64_t)g.ballot(s < x) << 1) | cin, Pm = g.ballot(s == 0xffffffffu);
        const uint64_t C = (Gm + Pm) ^ Pm;
        s += (uint32_t)(C >> rk) & 1u;
        if (i <= na) r[i] = s;
        cin = (uint32_t)(C >> S) & 1u;
    }
    g.sync();
    return r[na] ? na + 1 : na;
}

// r = a - b, a >= b (r may alias a or b); returns the trimmed limbs of r
RELP_W int w_sub(const G& g, uint32_t* r, const uint32_t* a, int na, const uint32_t* b, int nb) {
    const int S = g.size(), rk = g.rank();
    uint32_t bin = 0;
    for (int base = 0; base < na; base += S) {
        const int i = base + rk;
        const uint32_t x = i < na ? a[i] : 0u, y = i < nb ? b[i] : 0u;
        uint32_t s = x - y;
        const uint64_t Gm = ((uint64_t)g.ballot(x < y) << 1) | bin, Pm = g.ballot(i < na && s == 0u);
        const uint64_t C = (Gm + Pm) ^ Pm;
        s -= (uint32_t)(C >> rk) & 1u;
        if (i < na) r[i] = s;
        bin = (uint32_t)(C >> S) & 1u;
    }
    g.sync();
    return w_trim(g, r, na);
}

// r[0..na] = a * m (r may alias a; na + 1 limbs are always written); returns the limbs of r
RELP_W int w_mul1(const G& g, uint32_t* r, const uint32_t* a, int na, uint32_t m) {
    const int S = g.size(), rk = g.rank();
    uint32_t cin = 0, hprev = 0;
    for (int base = 0; base <= na; base += S) {
        const int i = base + rk;
        const uint64_t p = i < na ? (uint64_t)a[i] * m : 0ull;
        const uint32_t lo = (uint32_t)p, hi = (uint32_t)(p >> 32);
        uint32_t up = g.shfl_up(hi, 1);
        if (rk == 0) up = hprev;
        uint32_t s = lo + up;
        const uint64_t Gm = ((uint64_t)g.ballot(s < lo) << 1) | cin, Pm = g.ballot(s == 0xffffffffu);
        const uint64_t C = (Gm + Pm) ^ Pm;
        s += (uint32_t)(C >> rk) & 1u;
        if (i <= na) r[i] = s;
        cin = (uint32_t)(C >> S) & 1u;
        hprev = g.shfl(hi, S - 1);
    }
    g.sync();
    return w_trim(g, r, na + 1);
}

// r = a * b (r must not alias; capacity na + nb): column sums.  Lane k of a chunk owns column k of the
// product, sum of a[k - j] * b[j] in a 96-bit accumulator (b[j]: one address for the group, a[k - j]:
// consecutive lanes read consecutive limbs); the three words of every column sum are then added at limb
// positions k, k + 1, k + 2 with one carry-lookahead per chunk.  No step depends on the previous column.
RELP_W int w_mul(const G& g, uint32_t* r, const uint32_t* a, int na, const uint32_t* b, int nb) {
    if (na == 0 || nb == 0) return 0;
    if (na < nb) { const uint32_t* t = a; a = b; b = t; const int ti = na; na = nb; nb = ti; }
    const int S = g.size(), rk = g.rank(), nr = na + nb;
    uint32_t pw1 = 0, pw2a = 0, pw2b = 0, pc = 0;          // w1 of column base-1, w2 of columns base-2 / base-1, small carry
    for (int base = 0; base < nr; base += S) {
        const int k = base + rk;
        uint64_t lo = 0; uint32_t hi = 0;
        int jlo = base - (na - 1); if (jlo < 0) jlo = 0;
        int jhi = base + S - 1; if (jhi > nb - 1) jhi = nb - 1;
        for (int j = jlo; j <= jhi; j++) {
            const int i = k - j;
            if (i >= 0 && i < na) {
                const uint64_t p = (uint64_t)a[i] * b[j];
                lo += p; hi += (lo < p);
            }
        }
        const uint32_t w0 = (uint32_t)lo, w1 = (uint32_t)(lo >> 32), w2 = hi;
        uint32_t y = g.shfl_up(w1, 1), z = g.shfl_up(w2, 2);
        if (rk == 0) { y = pw1; z = pw2a; }
        else if (rk == 1) z = pw2b;
        const uint64_t t = (uint64_t)w0 + y + z + (rk == 0 ? pc : 0u);
        const uint32_t tlo = (uint32_t)t, sm = (uint32_t)(t >> 32);
        uint32_t e = g.shfl_up(sm, 1);
        if (rk == 0) e = 0;
        uint32_t u = tlo + e;
        const uint64_t Gm = (uint64_t)g.ballot(u < tlo) << 1, Pm = g.ballot(u == 0xffffffffu);
        const uint64_t C = (Gm + Pm) ^ Pm;
        u += (uint32_t)(C >> rk) & 1u;
        if (k < nr) r[k] = u;
        const uint32_t last_w1 = g.shfl(w1, S - 1), last_w2 = g.shfl(w2, S - 1);
        const uint32_t prev_w2 = S >= 2 ? g.shfl(w2, S - 2) : pw2b;
        pw1 = last_w1; pw2a = prev_w2; pw2b = last_w2;
        pc = g.shfl(sm, S - 1) + ((uint32_t)(C >> S) & 1u);
    }
    g.sync();
    return w_trim(g, r, nr);
}

// a >>= s in place (s < 32 n); returns the limbs
RELP_W int w_shr(const G& g, uint32_t* a, int n, int s) {
    const int S = g.size(), rk = g.rank(), w = s >> 5, b = s & 31;
    if (w >= n) return 0;
    const int m = n - w;
    for (int base = 0; base < m; base += S) {
        const int i = base + rk;
        uint32_t v = 0;
        if (i < m) {
            const uint32_t lo = a[i + w], hi = i + w + 1 < n ? a[i + w + 1] : 0u;
            v = b ? (lo >> b) | (hi << (32 - b)) : lo;
        }
        g.sync();                    // every lane has read its limbs before any lane overwrites one
        if (i < m) a[i] = v;
    }
    g.sync();
    return w_trim(g, a, m);
}

// a <<= s in place (capacity n + s / 32 + 1); returns the limbs
RELP_W int w_shl(const G& g, uint32_t* a, int n, int s) {
    if (n == 0) return 0;
    const int S = g.size(), rk = g.rank(), w = s >> 5, b = s & 31;
    const int m = n + w + 1;
    for (int base = ((m - 1) / S) * S; base >= 0; base -= S) {
        const int i = base + rk;
        uint32_t v = 0;
        if (i < m) {
            const int q = i - w;
            const uint32_t cur = (q >= 0 && q < n) ? a[q] : 0u, low = (q - 1 >= 0 && q - 1 < n) ? a[q - 1] : 0u;
            v = b ? (cur << b) | (low >> (32 - b)) : cur;
        }
        g.sync();
        if (i < m) a[i] = v;
    }
    g.sync();
    return w_trim(g, a, m);
}

// u mod v for a single-limb v (every lane runs the same Horner chain on the same addresses)
RELP_W uint32_t w_mod1(const G& g, const uint32_t* u, int nu, uint32_t v) {
    (void)g;
    uint64_t r = 0;
    for (int i = nu - 1; i >= 0; i--) r = ((r << 32) | u[i]) % v;
    return (uint32_t)r;
}

// u <- u mod v (Knuth 4.3.1 D; one quotient digit per step: the digit is estimated by every lane from the
// same three limbs, the multiply-subtract is one pass of the group).  u: capacity nu + 2; vn: nv + 1 limbs
// of scratch (normalised divisor); pt: nv + 1 limbs of scratch.  Returns the trimmed limbs of the remainder.
RELP_W int w_mod(const G& g, uint32_t* u, int nu, const uint32_t* v, int nv, uint32_t* vn, uint32_t* pt) {
    if (nu < nv) return nu;
    const int rk = g.rank();
    if (nv == 1) {
        const uint32_t r = w_mod1(g, u, nu, v[0]);
        g.sync();
        if (rk == 0) u[0] = r;
        g.sync();
        return r ? 1 : 0;
    }
    const int sft = clz32(v[nv - 1]);
    w_copy(g, vn, v, nv);
    if (rk == 0) { u[nu] = 0; u[nu + 1] = 0; }
    g.sync();
    if (sft) { w_shl(g, vn, nv, sft); w_shl(g, u, nu, sft); }
    const uint64_t vtop = vn[nv - 1], vsec = vn[nv - 2];
    for (int j = nu - nv; j >= 0; j--) {
        const uint64_t num = ((uint64_t)u[j + nv] << 32) | u[j + nv - 1];
        const uint32_t u2 = u[j + nv - 2];
        uint64_t qhat = num / vtop, rhat = num - qhat * vtop;
        while ((qhat >> 32) || qhat * vsec > ((rhat << 32) | u2)) {
            qhat--; rhat += vtop;
            if (rhat >> 32) break;
        }
        if (qhat) {
            w_mul1(g, pt, vn, nv, (uint32_t)qhat);                       // nv + 1 limbs
            if (w_sub_n(g, u + j, pt, nv + 1, nv + 1)) w_add_n(g, u + j, vn, nv, nv + 1);   // qhat was one too large: add v back
        }
    }
    int nr = w_trim(g, u, nv);
    if (sft && nr) nr = w_shr(g, u, nr, sft);
    return nr;
}

// GCD of two nonzero magnitudes (Lehmer, Knuth 4.5.2 L; 62-bit leading parts, 31-bit cofactors: every lane
// derives the same cofactors from the same leading limbs, the two linear combinations are passes of the
// group).  u, v: the operands, destroyed; t1, t2, t3: scratch; all five need max(nu, nv) + 3 limbs.
// *res points at the buffer that holds the result; returns its limb count.
RELP_W int w_gcd(const G& g, uint32_t* u, int nu, uint32_t* v, int nv, uint32_t* t1, uint32_t* t2, uint32_t* t3, uint32_t** res) {
    const int rk = g.rank();
    if (w_cmp(g, u, nu, v, nv) < 0) { uint32_t* t = u; u = v; v = t; const int ti = nu; nu = nv; nv = ti; }
    while (nv > 2) {                                 // invariant: u >= v > 0
        const int top = nu - 1;
        const int sft = clz32(u[top]);
        uint64_t uw = ((uint64_t)u[top] << 32) | u[top - 1], vw = 0;
        const uint32_t u3 = top >= 2 ? u[top - 2] : 0u;
        uint32_t v3 = 0;
        if (nv == nu) { vw = ((uint64_t)v[top] << 32) | v[top - 1]; v3 = top >= 2 ? v[top - 2] : 0u; }
        else if (nv == nu - 1) { vw = v[top - 1]; v3 = top >= 2 ? v[top - 2] : 0u; }
        if (sft) { uw = (uw << sft) | (u3 >> (32 - sft)); vw = (vw << sft) | (v3 >> (32 - sft)); }
        int64_t A = 1, B = 0, C = 0, D = 1;
        if (nv >= nu - 1) lehmer_cofactors((int64_t)(uw >> 2), (int64_t)(vw >> 2), A, B, C, D);     // every lane alike
        if (B == 0) {                                // no usable cofactors: one full remainder step
            const int nr = w_mod(g, u, nu, v, nv, t1, t2);
            uint32_t* t = u; u = v; v = t; nu = nv; nv = nr;
            if (nv == 0) { *res = u; return nu; }
            continue;
        }
        // (u, v) <- (A u + B v, C u + D v): both non-negative, the cofactors of a pair have opposite signs
        const uint32_t aA = (uint32_t)(A < 0 ? -A : A), aB = (uint32_t)(B < 0 ? -B : B);
        const uint32_t aC = (uint32_t)(C < 0 ? -C : C), aD = (uint32_t)(D < 0 ? -D : D);
        int nx = w_mul1(g, t1, u, nu, aA), ny = w_mul1(g, t2, v, nv, aB);
        const int nun = B < 0 ? w_sub(g, t1, t1, nx, t2, ny) : w_sub(g, t1, t2, ny, t1, nx);
        nx = w_mul1(g, t2, u, nu, aC); ny = w_mul1(g, t3, v, nv, aD);
        const int nvn = D < 0 ? w_sub(g, t2, t2, nx, t3, ny) : w_sub(g, t2, t3, ny, t2, nx);
        uint32_t* ou = u; uint32_t* ov = v;
        u = t1; v = t2; t1 = ou; t2 = ov;
        nu = nun; nv = nvn;
        if (nv == 0) { *res = u; return nu; }
        if (w_cmp(g, u, nu, v, nv) < 0) { uint32_t* t = u; u = v; v = t; const int ti = nu; nu = nv; nv = ti; }
    }
    // v fits 64 bits: one remainder and a 64-bit gcd, by every lane alike
    if (nu > 2) { nu = w_mod(g, u, nu, v, nv, t1, t2); if (nu == 0) { *res = v; return nv; } }
    const uint64_t aq = mag_lo64(u, nu), bq = mag_lo64(v, nv);
    const uint64_t gg = gcd64(aq, bq);
    g.sync();
    if (rk == 0) { u[0] = (uint32_t)gg; u[1] = (uint32_t)(gg >> 32); }
    g.sync();
    *res = u;
    return (gg >> 32) ? 2 : 1;
}

// q = a / d where d divides a exactly (Jebelean: q_i = a_i d^-1 mod 2^32, then a -= q_i d B^i, truncated at
// the quotient's length).  a is destroyed (capacity na); q: na - nd + 1 limbs; dt: nd limbs, pt: nd + 1 limbs of scratch.
RELP_W int w_divexact(const G& g, uint32_t* q, uint32_t* a, int na, const uint32_t* d, int nd, uint32_t* dt, uint32_t* pt) {
    if (na == 0) return 0;
    if (nd == 1 && d[0] == 1) { w_copy(g, q, a, na); return na; }
    const int rk = g.rank();
    const int z = w_ctz(g, d, nd);
    const uint32_t* dd = d;
    if (z) {                                     // strip the common power of two
        w_copy(g, dt, d, nd);
        nd = w_shr(g, dt, nd, z);
        na = w_shr(g, a, na, z);
        dd = dt;
    }
    const int nq = na - nd + 1;
    const uint32_t inv = inv32(dd[0]);
    for (int i = 0; i < nq; i++) {
        const uint32_t qi = a[i] * inv;
        if (rk == 0) q[i] = qi;
        if (qi) {
            const int win = nq - i;
            const int L = win < nd ? win : nd;
            w_mul1(g, pt, dd, L, qi);                              // L + 1 limbs
            w_sub_n(g, a + i, pt, L + 1 < win ? L + 1 : win, win);   // modulo B^nq: the borrow out is dropped
        }
    }
    g.sync();
    return w_trim(g, q, nq);
}

}  // namespace RELP_NS
