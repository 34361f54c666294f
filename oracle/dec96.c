/*
 * dec96.c -- TEST INFRASTRUCTURE ONLY (CPU oracle); see dec96.h for what is restated
 * and why parity is "unpinned".  Every operator is "compute the exact result in a wide
 * integer, then fit it back into 96 bits / scale <= 28 with one half-even rounding".
 */
#include "dec96.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ---- wide unsigned integers: 10 x 32 bits = 320 bits (96-bit mantissa * 10^57 < 2^286) */
#define WN 10
typedef struct {
    uint32_t w[WN];
} wide;

static void w_zero(wide *a) { memset(a->w, 0, sizeof a->w); }

static int w_len(const wide *a) {
    int n = WN;
    while (n > 0 && a->w[n - 1] == 0) n--;
    return n;
}

static int w_is_zero(const wide *a) { return w_len(a) == 0; }

static int w_bits(const wide *a) {
    int n = w_len(a);
    if (n == 0) return 0;
    return 32 * (n - 1) + (32 - __builtin_clz(a->w[n - 1]));
}

static int w_cmp(const wide *a, const wide *b) {
    for (int i = WN - 1; i >= 0; i--) {
        if (a->w[i] != b->w[i]) return a->w[i] < b->w[i] ? -1 : 1;
    }
    return 0;
}

static void w_add(wide *a, const wide *b) {
    uint64_t c = 0;
    for (int i = 0; i < WN; i++) {
        c += (uint64_t)a->w[i] + b->w[i];
        a->w[i] = (uint32_t)c;
        c >>= 32;
    }
}

/* a -= b, requires a >= b */
static void w_sub(wide *a, const wide *b) {
    int64_t borrow = 0;
    for (int i = 0; i < WN; i++) {
        int64_t d = (int64_t)a->w[i] - b->w[i] + borrow;
        a->w[i] = (uint32_t)d;
        borrow = d < 0 ? -1 : 0;
    }
}

static void w_add_small(wide *a, uint32_t x) {
    uint64_t c = x;
    for (int i = 0; i < WN && c; i++) {
        c += a->w[i];
        a->w[i] = (uint32_t)c;
        c >>= 32;
    }
}

static void w_mul_small(wide *a, uint32_t f) {
    uint64_t c = 0;
    for (int i = 0; i < WN; i++) {
        c += (uint64_t)a->w[i] * f;
        a->w[i] = (uint32_t)c;
        c >>= 32;
    }
}

/* in-place quotient, returns the remainder */
static uint32_t w_divmod_small(wide *a, uint32_t d) {
    uint64_t r = 0;
    for (int i = WN - 1; i >= 0; i--) {
        uint64_t cur = (r << 32) | a->w[i];
        a->w[i] = (uint32_t)(cur / d);
        r = cur % d;
    }
    return (uint32_t)r;
}

static void w_mul_pow10(wide *a, int e) {
    while (e >= 9) {
        w_mul_small(a, 1000000000u);
        e -= 9;
    }
    static const uint32_t p10[9] = {1, 10, 100, 1000, 10000, 100000, 1000000, 10000000, 100000000};
    if (e > 0) w_mul_small(a, p10[e]);
}

static void w_pow10(wide *out, int e) {
    w_zero(out);
    out->w[0] = 1;
    w_mul_pow10(out, e);
}

static void w_mul(const wide *a, const wide *b, wide *out) {
    uint32_t tmp[2 * WN];
    memset(tmp, 0, sizeof tmp);
    int la = w_len(a), lb = w_len(b);
    for (int i = 0; i < la; i++) {
        uint64_t c = 0;
        for (int j = 0; j < lb; j++) {
            c += (uint64_t)a->w[i] * b->w[j] + tmp[i + j];
            tmp[i + j] = (uint32_t)c;
            c >>= 32;
        }
        tmp[i + lb] = (uint32_t)c;
    }
    memcpy(out->w, tmp, sizeof out->w);
}

/* Schoolbook long division on 32-bit limbs (Knuth vol.2 algorithm D). */
static void w_divmod(const wide *num, const wide *den, wide *q, wide *r) {
    int n = w_len(den);
    int m = w_len(num);
    w_zero(q);
    if (n == 0) { /* caller guarantees den != 0 */
        w_zero(r);
        return;
    }
    if (w_cmp(num, den) < 0) {
        *r = *num;
        return;
    }
    if (n == 1) {
        *q = *num;
        uint32_t rem = w_divmod_small(q, den->w[0]);
        w_zero(r);
        r->w[0] = rem;
        return;
    }
    int s = __builtin_clz(den->w[n - 1]);
    uint32_t vn[WN], un[WN + 1];
    for (int i = n - 1; i > 0; i--)
        vn[i] = (den->w[i] << s) | (s ? (uint32_t)((uint64_t)den->w[i - 1] >> (32 - s)) : 0);
    vn[0] = den->w[0] << s;
    un[m] = s ? (uint32_t)((uint64_t)num->w[m - 1] >> (32 - s)) : 0;
    for (int i = m - 1; i > 0; i--)
        un[i] = (num->w[i] << s) | (s ? (uint32_t)((uint64_t)num->w[i - 1] >> (32 - s)) : 0);
    un[0] = num->w[0] << s;

    for (int j = m - n; j >= 0; j--) {
        uint64_t top = ((uint64_t)un[j + n] << 32) | un[j + n - 1];
        uint64_t qhat = top / vn[n - 1];
        uint64_t rhat = top % vn[n - 1];
        while (qhat >= (1ull << 32) || qhat * vn[n - 2] > ((rhat << 32) | un[j + n - 2])) {
            qhat--;
            rhat += vn[n - 1];
            if (rhat >= (1ull << 32)) break;
        }
        int64_t borrow = 0;
        uint64_t carry = 0;
        for (int i = 0; i < n; i++) {
            uint64_t p = qhat * vn[i] + carry;
            carry = p >> 32;
            int64_t t = (int64_t)un[i + j] - (int64_t)(uint32_t)p + borrow;
            un[i + j] = (uint32_t)t;
            borrow = t < 0 ? -1 : 0;
        }
        int64_t t = (int64_t)un[j + n] - (int64_t)carry + borrow;
        un[j + n] = (uint32_t)t;
        if (t < 0) { /* qhat was one too large: add the divisor back */
            qhat--;
            uint64_t c = 0;
            for (int i = 0; i < n; i++) {
                c += (uint64_t)un[i + j] + vn[i];
                un[i + j] = (uint32_t)c;
                c >>= 32;
            }
            un[j + n] += (uint32_t)c;
        }
        q->w[j] = (uint32_t)qhat;
    }
    w_zero(r);
    for (int i = 0; i < n; i++)
        r->w[i] = (un[i] >> s) | (s ? (uint32_t)((uint64_t)un[i + 1] << (32 - s)) : 0);
}

/* ---- fit: value = W / 10^sc (+ a nonzero tail below W's last digit if sticky) -> dec96 */
static int fit(const wide *W, int sc, int sticky, int neg, dec96 *out) {
    int k = sc > 28 ? sc - 28 : 0;
    int bits = w_bits(W);
    if (bits > 97) { /* need 10^k > W / 2^96 >= 2^(bits-97) */
        int kest = (int)((bits - 97) * 0.30102999);
        if (kest > k) k = kest;
    }
    for (;; k++) {
        if (sc - k < 0) return DEC96_OVERFLOW;
        wide q;
        if (k == 0) {
            q = *W; /* sticky is never set with k == 0 (division always carries >= 1 guard digit) */
        } else {
            wide P, r, half;
            w_pow10(&P, k);
            w_divmod(W, &P, &q, &r);
            half = P;
            w_divmod_small(&half, 2);
            int c = w_cmp(&r, &half);
            if (c > 0 || (c == 0 && (sticky || (q.w[0] & 1)))) w_add_small(&q, 1);
        }
        if (w_bits(&q) <= 96) {
            out->m[0] = q.w[0];
            out->m[1] = q.w[1];
            out->m[2] = q.w[2];
            out->scale = (uint8_t)(sc - k);
            out->neg = (uint8_t)((neg && !w_is_zero(&q)) ? 1 : 0);
            return DEC96_OK;
        }
    }
}

static void to_wide(const dec96 *a, wide *w) {
    w_zero(w);
    w->w[0] = a->m[0];
    w->w[1] = a->m[1];
    w->w[2] = a->m[2];
}

dec96 dec96_zero(void) {
    dec96 z;
    memset(&z, 0, sizeof z);
    return z;
}

dec96 dec96_one(void) {
    dec96 z = dec96_zero();
    z.m[0] = 1;
    return z;
}

int dec96_is_zero(const dec96 *a) { return (a->m[0] | a->m[1] | a->m[2]) == 0; }

int dec96_from_str(const char *s, dec96 *out) {
    while (*s == ' ') s++;
    int neg = 0;
    if (*s == '+' || *s == '-') neg = (*s++ == '-');
    wide W;
    w_zero(&W);
    int sc = 0, seen_point = 0, ndig = 0, sticky = 0;
    for (; *s; s++) {
        if (*s >= '0' && *s <= '9') {
            if (w_bits(&W) > 280) { /* absurdly long literal: fold the tail into sticky */
                if (*s != '0') sticky = 1;
                if (!seen_point) return DEC96_OVERFLOW;
                continue;
            }
            w_mul_small(&W, 10);
            w_add_small(&W, (uint32_t)(*s - '0'));
            if (seen_point) sc++;
            ndig++;
        } else if (*s == '.' && !seen_point) {
            seen_point = 1;
        } else if (*s == '_') {
            continue;
        } else {
            break;
        }
    }
    if (ndig == 0) return DEC96_PARSE;
    if (*s == 'e' || *s == 'E') {
        char *end;
        long ex = strtol(s + 1, &end, 10);
        if (end == s + 1) return DEC96_PARSE;
        s = end;
        if (ex > 60 || ex < -400) return ex > 0 ? DEC96_OVERFLOW : DEC96_PARSE;
        sc -= (int)ex;
    }
    while (*s == ' ') s++;
    if (*s) return DEC96_PARSE;
    if (sc < 0) {
        if (w_bits(&W) + (-sc) * 4 > 300) return DEC96_OVERFLOW;
        w_mul_pow10(&W, -sc);
        sc = 0;
    }
    if (sc > 57 + 28) { /* far below the smallest representable step: rounds to zero */
        *out = dec96_zero();
        return DEC96_OK;
    }
    if (sticky && sc <= 28) return DEC96_OVERFLOW;
    return fit(&W, sc, sticky, neg, out);
}

static size_t emit(const dec96 *a, char *buf, size_t buflen, int strip) {
    char digits[40];
    int nd = 0;
    wide W;
    to_wide(a, &W);
    while (!w_is_zero(&W)) digits[nd++] = (char)('0' + w_divmod_small(&W, 10));
    while (nd <= a->scale) digits[nd++] = '0'; /* at least one integer digit */
    char tmp[48];
    size_t p = 0;
    if (a->neg && !dec96_is_zero(a)) tmp[p++] = '-';
    for (int i = nd - 1; i >= 0; i--) {
        tmp[p++] = digits[i];
        if (i == a->scale && i != 0) tmp[p++] = '.';
    }
    if (strip && a->scale > 0) {
        while (tmp[p - 1] == '0') p--;
        if (tmp[p - 1] == '.') p--;
    }
    tmp[p] = 0;
    if (buflen) {
        size_t n = p < buflen - 1 ? p : buflen - 1;
        memcpy(buf, tmp, n);
        buf[n] = 0;
    }
    return p;
}

size_t dec96_to_str(const dec96 *a, char *buf, size_t buflen) { return emit(a, buf, buflen, 0); }
size_t dec96_to_str_normalized(const dec96 *a, char *buf, size_t buflen) { return emit(a, buf, buflen, 1); }

double dec96_to_double(const dec96 *a) {
    char buf[48];
    dec96_to_str(a, buf, sizeof buf);
    return strtod(buf, NULL);
}

int dec96_from_f32(float x, dec96 *out) {
    char buf[64];
    if (!(x == x) || x - x != 0.0f) return DEC96_PARSE; /* NaN / inf */
    for (int p = 0; p < 9; p++) {
        snprintf(buf, sizeof buf, "%.*e", p, (double)x);
        if (strtof(buf, NULL) == x) break;
    }
    return dec96_from_str(buf, out);
}

int dec96_from_f64(double x, dec96 *out) {
    char buf[64];
    if (!(x == x) || x - x != 0.0) return DEC96_PARSE;
    for (int p = 0; p < 17; p++) {
        snprintf(buf, sizeof buf, "%.*e", p, x);
        if (strtod(buf, NULL) == x) break;
    }
    return dec96_from_str(buf, out);
}

static int add_signed(const dec96 *a, int aneg, const dec96 *b, int bneg, dec96 *out) {
    wide A, B;
    to_wide(a, &A);
    to_wide(b, &B);
    int sc = a->scale > b->scale ? a->scale : b->scale;
    w_mul_pow10(&A, sc - a->scale);
    w_mul_pow10(&B, sc - b->scale);
    int neg;
    if (aneg == bneg) {
        w_add(&A, &B);
        neg = aneg;
    } else if (w_cmp(&A, &B) >= 0) {
        w_sub(&A, &B);
        neg = aneg;
    } else {
        w_sub(&B, &A);
        A = B;
        neg = bneg;
    }
    return fit(&A, sc, 0, neg, out);
}

int dec96_add(const dec96 *a, const dec96 *b, dec96 *out) { return add_signed(a, a->neg, b, b->neg, out); }
int dec96_sub(const dec96 *a, const dec96 *b, dec96 *out) { return add_signed(a, a->neg, b, !b->neg, out); }

int dec96_mul(const dec96 *a, const dec96 *b, dec96 *out) {
    wide A, B, P;
    to_wide(a, &A);
    to_wide(b, &B);
    w_mul(&A, &B, &P);
    return fit(&P, a->scale + b->scale, 0, a->neg != b->neg, out);
}

int dec96_div(const dec96 *a, const dec96 *b, dec96 *out) {
    if (dec96_is_zero(b)) return DEC96_DIV_BY_ZERO;
    if (dec96_is_zero(a)) {
        *out = dec96_zero();
        return DEC96_OK;
    }
    /* a/b * 10^29 = a.m * 10^(29 + b.scale - a.scale) / b.m : one guard digit past scale 28 */
    wide N, D, Q, R;
    to_wide(a, &N);
    to_wide(b, &D);
    w_mul_pow10(&N, 29 + b->scale - a->scale);
    w_divmod(&N, &D, &Q, &R);
    return fit(&Q, 29, !w_is_zero(&R), a->neg != b->neg, out);
}

int dec96_cmp(const dec96 *a, const dec96 *b) {
    int az = dec96_is_zero(a), bz = dec96_is_zero(b);
    int an = a->neg && !az, bn = b->neg && !bz;
    if (an != bn) return an ? -1 : 1;
    wide A, B;
    to_wide(a, &A);
    to_wide(b, &B);
    int sc = a->scale > b->scale ? a->scale : b->scale;
    w_mul_pow10(&A, sc - a->scale);
    w_mul_pow10(&B, sc - b->scale);
    int c = w_cmp(&A, &B);
    return an ? -c : c;
}
