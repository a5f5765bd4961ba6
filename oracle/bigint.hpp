// ORACLE (test infrastructure, not product): arbitrary-precision signed integer.
//
// Stands in for num-bigint 0.4 (crate `num = "0.4.0"`, /root/reference/Cargo.toml:16; not
// vendored under /root/reference, no Cargo.lock => exact version unpinned).  Only the VALUE
// semantics matter for parity (every RationalBig the reference stores is canonical,
// src/data/number_types/rational/big/mod.rs:17-29), so the algorithms here are plain
// schoolbook / Knuth D and need not match num-bigint's internals.
//
// Representation: small mode (limbs empty, value in `s`) or big mode (sign + magnitude in
// 32-bit limbs, little endian, no leading zero limb, magnitude does not fit in int64).
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include <algorithm>
#include <cstdlib>

namespace oracle {

typedef unsigned __int128 u128;
typedef __int128 i128;

class BigInt {
public:
    int64_t s = 0;                 // value when mag is empty
    bool neg = false;              // sign when mag is non-empty
    std::vector<uint32_t> mag;     // magnitude when not representable as int64

    BigInt() {}
    BigInt(int64_t v) : s(v) {}
    static BigInt from_i128(i128 v) {
        BigInt r;
        if (v >= INT64_MIN && v <= INT64_MAX) { r.s = (int64_t)v; return r; }
        r.neg = v < 0;
        u128 a = r.neg ? (u128)0 - (u128)v : (u128)v;
        while (a) { r.mag.push_back((uint32_t)a); a >>= 32; }
        return r;
    }
    bool is_small() const { return mag.empty(); }
    bool is_zero() const { return mag.empty() && s == 0; }
    int sign() const { return mag.empty() ? (s > 0) - (s < 0) : (neg ? -1 : 1); }
    bool is_one() const { return mag.empty() && s == 1; }

    // ---- magnitude helpers -------------------------------------------------------------
    static void trim(std::vector<uint32_t>& a) { while (!a.empty() && a.back() == 0) a.pop_back(); }
    static int cmp_mag(const std::vector<uint32_t>& a, const std::vector<uint32_t>& b) {
        if (a.size() != b.size()) return a.size() < b.size() ? -1 : 1;
        for (size_t i = a.size(); i-- > 0;) if (a[i] != b[i]) return a[i] < b[i] ? -1 : 1;
        return 0;
    }
    static std::vector<uint32_t> add_mag(const std::vector<uint32_t>& a, const std::vector<uint32_t>& b) {
        const std::vector<uint32_t>& x = a.size() >= b.size() ? a : b;
        const std::vector<uint32_t>& y = a.size() >= b.size() ? b : a;
        std::vector<uint32_t> r(x.size() + 1);
        uint64_t c = 0;
        for (size_t i = 0; i < x.size(); i++) {
            c += (uint64_t)x[i] + (i < y.size() ? y[i] : 0);
            r[i] = (uint32_t)c; c >>= 32;
        }
        r[x.size()] = (uint32_t)c; trim(r); return r;
    }
    // a >= b required
    static std::vector<uint32_t> sub_mag(const std::vector<uint32_t>& a, const std::vector<uint32_t>& b) {
        std::vector<uint32_t> r(a.size());
        int64_t c = 0;
        for (size_t i = 0; i < a.size(); i++) {
            c += (int64_t)a[i] - (i < b.size() ? (int64_t)b[i] : 0);
            r[i] = (uint32_t)c; c >>= 32;   // arithmetic shift keeps the borrow
        }
        trim(r); return r;
    }
    static std::vector<uint32_t> mul_mag(const std::vector<uint32_t>& a, const std::vector<uint32_t>& b) {
        std::vector<uint32_t> r(a.size() + b.size(), 0);
        for (size_t i = 0; i < a.size(); i++) {
            uint64_t c = 0, ai = a[i];
            for (size_t j = 0; j < b.size(); j++) {
                c += ai * b[j] + r[i + j];
                r[i + j] = (uint32_t)c; c >>= 32;
            }
            r[i + b.size()] = (uint32_t)c;
        }
        trim(r); return r;
    }
    // Knuth algorithm D; q = a / b, rem = a % b (magnitudes); b non-empty.
    static void divmod_mag(const std::vector<uint32_t>& a, const std::vector<uint32_t>& b,
                           std::vector<uint32_t>& q, std::vector<uint32_t>& rem) {
        if (cmp_mag(a, b) < 0) { q.clear(); rem = a; return; }
        if (b.size() == 1) {
            q.assign(a.size(), 0); uint64_t r = 0;
            for (size_t i = a.size(); i-- > 0;) { r = (r << 32) | a[i]; q[i] = (uint32_t)(r / b[0]); r %= b[0]; }
            trim(q); rem.clear(); if (r) rem.push_back((uint32_t)r); return;
        }
        int sh = __builtin_clz(b.back());
        size_t n = b.size(), m = a.size() - n;
        std::vector<uint32_t> u(a.size() + 1), v(n);
        for (size_t i = n; i-- > 0;) v[i] = (b[i] << sh) | (sh && i ? (b[i - 1] >> (32 - sh)) : 0);
        u[a.size()] = sh ? (a.back() >> (32 - sh)) : 0;
        for (size_t i = a.size(); i-- > 0;) u[i] = (a[i] << sh) | (sh && i ? (a[i - 1] >> (32 - sh)) : 0);
        q.assign(m + 1, 0);
        for (size_t j = m + 1; j-- > 0;) {
            uint64_t num = ((uint64_t)u[j + n] << 32) | u[j + n - 1];
            uint64_t qh = num / v[n - 1], rh = num % v[n - 1];
            while (qh >= (1ull << 32) || qh * v[n - 2] > ((rh << 32) | u[j + n - 2])) {
                qh--; rh += v[n - 1]; if (rh >= (1ull << 32)) break;
            }
            int64_t borrow = 0; uint64_t carry = 0;
            for (size_t i = 0; i < n; i++) {
                uint64_t p = qh * v[i] + carry; carry = p >> 32;
                int64_t t = (int64_t)u[i + j] - (int64_t)(uint32_t)p + borrow;
                u[i + j] = (uint32_t)t; borrow = t >> 32;
            }
            int64_t t = (int64_t)u[j + n] - (int64_t)carry + borrow;
            u[j + n] = (uint32_t)t;
            if (t < 0) {   // add back
                qh--; uint64_t c = 0;
                for (size_t i = 0; i < n; i++) { c += (uint64_t)u[i + j] + v[i]; u[i + j] = (uint32_t)c; c >>= 32; }
                u[j + n] += (uint32_t)c;
            }
            q[j] = (uint32_t)qh;
        }
        trim(q);
        rem.assign(n, 0);
        for (size_t i = 0; i < n; i++) rem[i] = (u[i] >> sh) | (sh ? ((uint64_t)u[i + 1] << (32 - sh)) : 0);
        trim(rem);
    }

    // ---- conversions ---------------------------------------------------------------------
    void to_mag(std::vector<uint32_t>& m, bool& n) const {
        if (!mag.empty()) { m = mag; n = neg; return; }
        n = s < 0; uint64_t a = n ? 0ull - (uint64_t)s : (uint64_t)s;
        m.clear(); while (a) { m.push_back((uint32_t)a); a >>= 32; }
    }
    static BigInt from_mag(std::vector<uint32_t>& m, bool n) {
        trim(m);
        BigInt r;
        if (m.size() <= 2) {
            uint64_t a = 0; for (size_t i = m.size(); i-- > 0;) a = (a << 32) | m[i];
            if (!n && a <= (uint64_t)INT64_MAX) { r.s = (int64_t)a; return r; }
            if (n && a <= (1ull << 63)) { r.s = (int64_t)(0ull - a); return r; }
        }
        r.neg = n; r.mag.swap(m); return r;
    }
    size_t bits() const {
        if (mag.empty()) { uint64_t a = s < 0 ? 0ull - (uint64_t)s : (uint64_t)s; return a ? 64 - __builtin_clzll(a) : 0; }
        return 32 * (mag.size() - 1) + (32 - __builtin_clz(mag.back()));
    }

    // ---- arithmetic ------------------------------------------------------------------------
    BigInt operator-() const {
        if (mag.empty()) { if (s != INT64_MIN) return BigInt(-s); return from_i128(-(i128)s); }
        BigInt r(*this); r.neg = !neg;
        // -(2^63) in big mode becomes small INT64_MIN
        if (r.neg && r.mag.size() == 2 && r.mag[0] == 0 && r.mag[1] == 0x80000000u) { BigInt t; t.s = INT64_MIN; return t; }
        return r;
    }
    static BigInt addsub(const BigInt& a, const BigInt& b, bool sub) {
        if (a.mag.empty() && b.mag.empty()) {
            int64_t r;
            if (!(sub ? __builtin_sub_overflow(a.s, b.s, &r) : __builtin_add_overflow(a.s, b.s, &r))) return BigInt(r);
            return from_i128(sub ? (i128)a.s - b.s : (i128)a.s + b.s);
        }
        std::vector<uint32_t> ma, mb; bool na, nb;
        a.to_mag(ma, na); b.to_mag(mb, nb); if (sub) nb = !nb;
        std::vector<uint32_t> r; bool nr;
        if (na == nb) { r = add_mag(ma, mb); nr = na; }
        else {
            int c = cmp_mag(ma, mb);
            if (c == 0) return BigInt(0);
            if (c > 0) { r = sub_mag(ma, mb); nr = na; } else { r = sub_mag(mb, ma); nr = nb; }
        }
        return from_mag(r, nr);
    }
    BigInt operator+(const BigInt& o) const { return addsub(*this, o, false); }
    BigInt operator-(const BigInt& o) const { return addsub(*this, o, true); }
    BigInt operator*(const BigInt& o) const {
        if (mag.empty() && o.mag.empty()) {
            int64_t r;
            if (!__builtin_mul_overflow(s, o.s, &r)) return BigInt(r);
            return from_i128((i128)s * o.s);
        }
        if (is_zero() || o.is_zero()) return BigInt(0);
        std::vector<uint32_t> ma, mb; bool na, nb;
        to_mag(ma, na); o.to_mag(mb, nb);
        std::vector<uint32_t> r = mul_mag(ma, mb);
        return from_mag(r, na != nb);
    }
    // truncated division (C semantics); divisor non-zero
    static void divmod(const BigInt& a, const BigInt& b, BigInt& q, BigInt& r) {
        if (a.mag.empty() && b.mag.empty() && !(a.s == INT64_MIN && b.s == -1)) { q = BigInt(a.s / b.s); r = BigInt(a.s % b.s); return; }
        std::vector<uint32_t> ma, mb, mq, mr; bool na, nb;
        a.to_mag(ma, na); b.to_mag(mb, nb);
        divmod_mag(ma, mb, mq, mr);
        q = from_mag(mq, na != nb); r = from_mag(mr, na);
    }
    BigInt operator/(const BigInt& o) const { BigInt q, r; divmod(*this, o, q, r); return q; }
    BigInt operator%(const BigInt& o) const { BigInt q, r; divmod(*this, o, q, r); return r; }
    BigInt abs() const { return sign() < 0 ? -*this : *this; }

    static int cmp(const BigInt& a, const BigInt& b) {
        if (a.mag.empty() && b.mag.empty()) return (a.s > b.s) - (a.s < b.s);
        int sa = a.sign(), sb = b.sign();
        if (sa != sb) return sa < sb ? -1 : 1;
        if (a.mag.empty()) return sa > 0 ? -1 : 1;   // |a| < |b| since b is big
        if (b.mag.empty()) return sa > 0 ? 1 : -1;
        int c = cmp_mag(a.mag, b.mag);
        return sa > 0 ? c : -c;
    }
    bool operator==(const BigInt& o) const { return cmp(*this, o) == 0; }
    bool operator!=(const BigInt& o) const { return cmp(*this, o) != 0; }
    bool operator<(const BigInt& o) const { return cmp(*this, o) < 0; }

    static uint64_t gcd_u64(uint64_t a, uint64_t b) {
        if (!a) return b; if (!b) return a;
        int sh = __builtin_ctzll(a | b); a >>= __builtin_ctzll(a);
        do { b >>= __builtin_ctzll(b); if (a > b) std::swap(a, b); b -= a; } while (b);
        return a << sh;
    }
    // non-negative gcd; gcd(0,0) = 0
    static BigInt gcd(const BigInt& a, const BigInt& b) {
        if (a.mag.empty() && b.mag.empty()) {
            uint64_t x = a.s < 0 ? 0ull - (uint64_t)a.s : (uint64_t)a.s;
            uint64_t y = b.s < 0 ? 0ull - (uint64_t)b.s : (uint64_t)b.s;
            uint64_t g = gcd_u64(x, y);
            return from_i128((i128)g);
        }
        BigInt x = a.abs(), y = b.abs();
        while (!y.is_zero()) {
            if (x.mag.empty() && y.mag.empty()) return gcd(x, y);
            BigInt r = x % y; x = y; y = r;
        }
        return x;
    }

    std::string to_string() const {
        if (mag.empty()) return std::to_string(s);
        std::vector<uint32_t> m = mag; std::string out;
        while (!m.empty()) {   // divide by 10^9 repeatedly
            uint64_t r = 0;
            for (size_t i = m.size(); i-- > 0;) { r = (r << 32) | m[i]; m[i] = (uint32_t)(r / 1000000000u); r %= 1000000000u; }
            trim(m);
            for (int k = 0; k < 9; k++) { out.push_back('0' + r % 10); r /= 10; if (m.empty() && r == 0) break; }
        }
        if (neg) out.push_back('-');
        std::reverse(out.begin(), out.end());
        return out;
    }
    static BigInt from_string(const char* p) {
        bool n = false; if (*p == '-') { n = true; p++; } else if (*p == '+') p++;
        BigInt r(0), ten9(1000000000);
        while (*p >= '0' && *p <= '9') {
            int64_t chunk = 0, mul = 1; int k = 0;
            while (k < 9 && *p >= '0' && *p <= '9') { chunk = chunk * 10 + (*p - '0'); mul *= 10; p++; k++; }
            r = r * BigInt(mul) + BigInt(chunk);
        }
        return n ? -r : r;
    }
};

}  // namespace oracle
