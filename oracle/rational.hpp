// ORACLE (test infrastructure, not product): canonical arbitrary-precision rational.
//
// Restates the VALUE semantics of `RationalBig` = newtype over `num::BigRational`
// (/root/reference/src/data/number_types/rational/big/mod.rs:17-29, operators forwarded in
// big/wrapping.rs:47-311): every stored value is gcd-reduced, denominator > 0, zero = 0/1.
// The KATs that pin this boundary in the reference (rational/mod.rs:14-104: 3/2 == 6/4,
// 0/2 == 0/5, 3/2 + 6/4 = 3, 3/2 * 6/4 = 9/4, (3/2)/(6/4) = 1) are replayed in
// tests/test_oracle_rational.py.
#pragma once
#include "bigint.hpp"

namespace oracle {

class Rat {
public:
    BigInt n, d;   // d > 0, gcd(|n|, d) = 1
    Rat() : n(0), d(1) {}
    Rat(int64_t v) : n(v), d(1) {}
    Rat(const BigInt& nn, const BigInt& dd) : n(nn), d(dd) { normalize(); }
    static Rat raw(const BigInt& nn, const BigInt& dd) { Rat r; r.n = nn; r.d = dd; return r; }
    void normalize() {
        if (d.sign() < 0) { n = -n; d = -d; }
        if (n.is_zero()) { d = BigInt(1); return; }
        if (d.is_one()) return;
        BigInt g = BigInt::gcd(n, d);
        if (!g.is_one()) { n = n / g; d = d / g; }
    }
    bool is_zero() const { return n.is_zero(); }
    int sign() const { return n.sign(); }
    bool is_int() const { return d.is_one(); }
    Rat operator-() const { return raw(-n, d); }

    Rat operator*(const Rat& o) const {
        if (is_zero() || o.is_zero()) return Rat();
        if (d.is_one() && o.d.is_one()) return raw(n * o.n, BigInt(1));
        BigInt g1 = BigInt::gcd(n, o.d), g2 = BigInt::gcd(o.n, d);
        BigInt a = g1.is_one() ? n : n / g1, b = g2.is_one() ? o.n : o.n / g2;
        BigInt c = g2.is_one() ? d : d / g2, e = g1.is_one() ? o.d : o.d / g1;
        return raw(a * b, c * e);
    }
    Rat inv() const { return sign() < 0 ? raw(-d, -n) : raw(d, n); }
    Rat operator/(const Rat& o) const { return *this * o.inv(); }
    static Rat addsub(const Rat& x, const Rat& y, bool sub) {
        if (y.is_zero()) return x;
        if (x.is_zero()) return sub ? -y : y;
        if (x.d.is_one() && y.d.is_one()) return raw(sub ? x.n - y.n : x.n + y.n, BigInt(1));
        BigInt g = BigInt::gcd(x.d, y.d);
        if (g.is_one()) {
            BigInt t = sub ? x.n * y.d - y.n * x.d : x.n * y.d + y.n * x.d;
            if (t.is_zero()) return Rat();
            return raw(t, x.d * y.d);
        }
        BigInt xd = x.d / g, yd = y.d / g;
        BigInt t = sub ? x.n * yd - y.n * xd : x.n * yd + y.n * xd;
        if (t.is_zero()) return Rat();
        BigInt g2 = BigInt::gcd(t, g);
        if (g2.is_one()) return raw(t, xd * y.d);
        return raw(t / g2, xd * (y.d / g2));
    }
    Rat operator+(const Rat& o) const { return addsub(*this, o, false); }
    Rat operator-(const Rat& o) const { return addsub(*this, o, true); }
    Rat& operator+=(const Rat& o) { *this = *this + o; return *this; }
    Rat& operator-=(const Rat& o) { *this = *this - o; return *this; }

    static int cmp(const Rat& a, const Rat& b) {
        int sa = a.sign(), sb = b.sign();
        if (sa != sb) return sa < sb ? -1 : 1;
        if (a.d == b.d) return BigInt::cmp(a.n, b.n);
        return BigInt::cmp(a.n * b.d, b.n * a.d);
    }
    bool operator==(const Rat& o) const { return n == o.n && d == o.d; }
    bool operator!=(const Rat& o) const { return !(*this == o); }
    bool operator<(const Rat& o) const { return cmp(*this, o) < 0; }

    std::string to_string() const { return d.is_one() ? n.to_string() : n.to_string() + "/" + d.to_string(); }
    static Rat from_string(const char* p) {
        BigInt a = BigInt::from_string(p);
        const char* s = p; while (*s && *s != '/') s++;
        if (*s == '/') return Rat(a, BigInt::from_string(s + 1));
        return Rat(a, BigInt(1));
    }
};

}  // namespace oracle
