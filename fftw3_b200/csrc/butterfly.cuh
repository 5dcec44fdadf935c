// csrc/butterfly.cuh -- register-resident radix-R DFT butterflies (forward sign, natural
// order in, natural order out, unnormalised).  These hand-written routines take the place
// of the reference's generated no-twiddle / twiddle codelets n1_* and t1_*
// (fftw/dft_scalar/codelets/n1_4.c:94-132, t1_4.c:125-187, fftw/dft_simd/common/*.c): a
// GPU thread keeps the R points in registers; the twiddle multiply is done by the caller.
// The backward transform is obtained, as in the reference (fftw_extract_reim,
// fftw/fftw_api.c:14555-14564), by swapping re/im on the way in and out.
#pragma once
#include "b2_common.h"

template <int R, typename T> struct Bf;

template <typename T> struct Bf<1, T> {
    static B2_HD void run(cpx<T> *v) { (void)v; }
};

template <typename T> struct Bf<2, T> {
    static B2_HD void run(cpx<T> *v) {
        cpx<T> a = v[0];
        v[0] = cadd(a, v[1]);
        v[1] = csub(a, v[1]);
    }
};

template <typename T> struct Bf<4, T> {
    static B2_HD void run(cpx<T> *v) {
        cpx<T> s0 = cadd(v[0], v[2]), d0 = csub(v[0], v[2]);
        cpx<T> s1 = cadd(v[1], v[3]), d1 = cmul_mi(csub(v[1], v[3]));
        v[0] = cadd(s0, s1);
        v[2] = csub(s0, s1);
        v[1] = cadd(d0, d1);
        v[3] = csub(d0, d1);
    }
};

template <typename T> struct Bf<8, T> {
    static B2_HD void run(cpx<T> *v) {
        const T h = (T)0.70710678118654752440;
        cpx<T> e[4] = {v[0], v[2], v[4], v[6]};
        cpx<T> o[4] = {v[1], v[3], v[5], v[7]};
        Bf<4, T>::run(e);
        Bf<4, T>::run(o);
        cpx<T> o1 = cpx<T>{(o[1].x + o[1].y) * h, (o[1].y - o[1].x) * h};
        cpx<T> o2 = cmul_mi(o[2]);
        cpx<T> o3 = cpx<T>{(o[3].y - o[3].x) * h, -(o[3].x + o[3].y) * h};
        v[0] = cadd(e[0], o[0]);
        v[4] = csub(e[0], o[0]);
        v[1] = cadd(e[1], o1);
        v[5] = csub(e[1], o1);
        v[2] = cadd(e[2], o2);
        v[6] = csub(e[2], o2);
        v[3] = cadd(e[3], o3);
        v[7] = csub(e[3], o3);
    }
};

// 16 = 4 x 4: four radix-4 butterflies over the stride-4 sub-sequences, the w16^(j2*k1)
// twiddles, then four radix-4 butterflies; X[k1 + 4 k2].
template <typename T> struct Bf<16, T> {
    static B2_HD void run(cpx<T> *v) {
        const T h = (T)0.70710678118654752440;
        const T c1 = (T)0.92387953251128675613, s1 = (T)0.38268343236508977173;
        cpx<T> b[4][4];  // b[j2][k1]
#pragma unroll
        for (int j2 = 0; j2 < 4; ++j2) {
            b[j2][0] = v[j2];
            b[j2][1] = v[4 + j2];
            b[j2][2] = v[8 + j2];
            b[j2][3] = v[12 + j2];
            Bf<4, T>::run(b[j2]);
        }
        // twiddles w16^(j2*k1)
        b[1][1] = cmul(b[1][1], cpx<T>{c1, -s1});
        b[1][2] = cpx<T>{(b[1][2].x + b[1][2].y) * h, (b[1][2].y - b[1][2].x) * h};  // w^2
        b[1][3] = cmul(b[1][3], cpx<T>{s1, -c1});                                      // w^3
        b[2][1] = cpx<T>{(b[2][1].x + b[2][1].y) * h, (b[2][1].y - b[2][1].x) * h};  // w^2
        b[2][2] = cmul_mi(b[2][2]);                                                    // w^4
        b[2][3] = cpx<T>{(b[2][3].y - b[2][3].x) * h, -(b[2][3].x + b[2][3].y) * h}; // w^6
        b[3][1] = cmul(b[3][1], cpx<T>{s1, -c1});                                      // w^3
        b[3][2] = cpx<T>{(b[3][2].y - b[3][2].x) * h, -(b[3][2].x + b[3][2].y) * h}; // w^6
        b[3][3] = cmul(b[3][3], cpx<T>{-c1, s1});                                      // w^9
#pragma unroll
        for (int k1 = 0; k1 < 4; ++k1) {
            cpx<T> r[4] = {b[0][k1], b[1][k1], b[2][k1], b[3][k1]};
            Bf<4, T>::run(r);
            v[k1] = r[0];
            v[k1 + 4] = r[1];
            v[k1 + 8] = r[2];
            v[k1 + 12] = r[3];
        }
    }
};

// Odd primes: X[k] = a0 + sum_{j=1..(P-1)/2} (a_j + a_{P-j}) cos(2 pi jk/P)
//                                          - i (a_j - a_{P-j}) sin(2 pi jk/P)
template <int P> struct PrimeTab;
template <> struct PrimeTab<3> {
    static B2_HD double c(int m) { const double t[3] = {1, -0.5, -0.5}; return t[m]; }
    static B2_HD double s(int m) { const double t[3] = {0, 0.86602540378443864676, -0.86602540378443864676}; return t[m]; }
};
template <> struct PrimeTab<5> {
    static B2_HD double c(int m) {
        const double t[5] = {1, 0.30901699437494742410, -0.80901699437494742410, -0.80901699437494742410, 0.30901699437494742410};
        return t[m];
    }
    static B2_HD double s(int m) {
        const double t[5] = {0, 0.95105651629515357212, 0.58778525229247312917, -0.58778525229247312917, -0.95105651629515357212};
        return t[m];
    }
};
template <> struct PrimeTab<7> {
    static B2_HD double c(int m) {
        const double t[7] = {1, 0.62348980185873353053, -0.22252093395631440429, -0.90096886790241912624,
                             -0.90096886790241912624, -0.22252093395631440429, 0.62348980185873353053};
        return t[m];
    }
    static B2_HD double s(int m) {
        const double t[7] = {0, 0.78183148246802980871, 0.97492791218182360702, 0.43388373911755812048,
                             -0.43388373911755812048, -0.97492791218182360702, -0.78183148246802980871};
        return t[m];
    }
};
template <> struct PrimeTab<11> {
    static B2_HD double c(int m) {
        const double t[11] = {1, 0.84125353283118116886, 0.41541501300188642553, -0.14231483827328514044,
                              -0.65486073394528506406, -0.95949297361449738989, -0.95949297361449738989,
                              -0.65486073394528506406, -0.14231483827328514044, 0.41541501300188642553,
                              0.84125353283118116886};
        return t[m];
    }
    static B2_HD double s(int m) {
        const double t[11] = {0, 0.54064081745559758211, 0.90963199535451837141, 0.98982144188093273238,
                              0.75574957435425828377, 0.28173255684142969771, -0.28173255684142969771,
                              -0.75574957435425828377, -0.98982144188093273238, -0.90963199535451837141,
                              -0.54064081745559758211};
        return t[m];
    }
};
template <> struct PrimeTab<13> {
    static B2_HD double c(int m) {
        const double t[13] = {1, 0.88545602565320989590, 0.56806474673115580251, 0.12053668025532305335,
                              -0.35460488704253562597, -0.74851074817110109863, -0.97094181742605202716,
                              -0.97094181742605202716, -0.74851074817110109863, -0.35460488704253562597,
                              0.12053668025532305335, 0.56806474673115580251, 0.88545602565320989590};
        return t[m];
    }
    static B2_HD double s(int m) {
        const double t[13] = {0, 0.46472317204376854566, 0.82298386589365639458, 0.99270887409805399280,
                              0.93501624268541482344, 0.66312265824079520238, 0.23931566428755776715,
                              -0.23931566428755776715, -0.66312265824079520238, -0.93501624268541482344,
                              -0.99270887409805399280, -0.82298386589365639458, -0.46472317204376854566};
        return t[m];
    }
};

template <int P, typename T> struct BfPrime {
    static B2_HD void run(cpx<T> *v) {
        constexpr int H = (P - 1) / 2;
        cpx<T> sp[H], sm[H];
#pragma unroll
        for (int j = 1; j <= H; ++j) {
            sp[j - 1] = cadd(v[j], v[P - j]);
            sm[j - 1] = csub(v[j], v[P - j]);
        }
        cpx<T> a0 = v[0];
        cpx<T> dc = a0;
#pragma unroll
        for (int j = 0; j < H; ++j) dc = cadd(dc, sp[j]);
        v[0] = dc;
#pragma unroll
        for (int k = 1; k <= H; ++k) {
            cpx<T> re = a0;            // sum of cos terms
            cpx<T> im = cpx<T>{0, 0};  // sum of sin terms (to be multiplied by -i)
#pragma unroll
            for (int j = 1; j <= H; ++j) {
                const int m = (j * k) % P;
                const T c = (T)PrimeTab<P>::c(m), s = (T)PrimeTab<P>::s(m);
                re.x += sp[j - 1].x * c;
                re.y += sp[j - 1].y * c;
                im.x += sm[j - 1].x * s;
                im.y += sm[j - 1].y * s;
            }
            cpx<T> t = cmul_mi(im);
            v[k] = cadd(re, t);
            v[P - k] = csub(re, t);
        }
    }
};
template <typename T> struct Bf<3, T> : BfPrime<3, T> {};
template <typename T> struct Bf<5, T> : BfPrime<5, T> {};
template <typename T> struct Bf<7, T> : BfPrime<7, T> {};
template <typename T> struct Bf<11, T> : BfPrime<11, T> {};
template <typename T> struct Bf<13, T> : BfPrime<13, T> {};

// Composite radices with coprime factors R = R1 * R2 (Good-Thomas / prime-factor form: no twiddles between
// the two stages).  With a = R2^-1 mod R1, b = R1^-1 mod R2:
//   input index  j = (R2*j1 + R1*j2) mod R,   output index  k = (R2*a*k1 + R1*b*k2) mod R,
//   X[k] = sum_{j1,j2} x[j] w_R1^(j1 k1) w_R2^(j2 k2).
// Every index is a compile-time constant after unrolling.  These stand where the reference has the generated
// codelets n1_6, n1_10, n1_12, n1_15, n1_20 (fftw/dft_scalar/codelets/codlist.c:57-109), which the generator
// derives the same way.
constexpr int b2_inv_mod(int a, int m) {
    for (int x = 1; x < m; ++x)
        if ((a * x) % m == 1) return x;
    return 0;
}
template <int R1, int R2, typename T> struct BfPFA {
    static constexpr int R = R1 * R2;
    static constexpr int A = R2 * b2_inv_mod(R2 % R1, R1), B = R1 * b2_inv_mod(R1 % R2, R2);
    static_assert(A % R1 == 1 % R1 && B % R2 == 1 % R2, "the factors must be coprime");
    static B2_HD void run(cpx<T> *v) {
        cpx<T> b[R2][R1];  // b[j2][k1]
#pragma unroll
        for (int j2 = 0; j2 < R2; ++j2) {
#pragma unroll
            for (int j1 = 0; j1 < R1; ++j1) b[j2][j1] = v[(R2 * j1 + R1 * j2) % R];
            Bf<R1, T>::run(b[j2]);
        }
#pragma unroll
        for (int k1 = 0; k1 < R1; ++k1) {
            cpx<T> r[R2];
#pragma unroll
            for (int j2 = 0; j2 < R2; ++j2) r[j2] = b[j2][k1];
            Bf<R2, T>::run(r);
#pragma unroll
            for (int k2 = 0; k2 < R2; ++k2) v[(A * k1 + B * k2) % R] = r[k2];
        }
    }
};
template <typename T> struct Bf<6, T> : BfPFA<3, 2, T> {};
template <typename T> struct Bf<10, T> : BfPFA<5, 2, T> {};
template <typename T> struct Bf<12, T> : BfPFA<4, 3, T> {};
template <typename T> struct Bf<15, T> : BfPFA<5, 3, T> {};
template <typename T> struct Bf<20, T> : BfPFA<4, 5, T> {};
template <typename T> struct Bf<24, T> : BfPFA<8, 3, T> {};
template <typename T> struct Bf<30, T> : BfPFA<5, 6, T> {};
template <typename T> struct Bf<14, T> : BfPFA<7, 2, T> {};
template <typename T> struct Bf<21, T> : BfPFA<7, 3, T> {};
template <typename T> struct Bf<28, T> : BfPFA<7, 4, T> {};

// 9 = 3 x 3 (Cooley-Tukey: x[3 j1 + j2] -> three radix-3 butterflies over j1, twiddles w9^(j2 k1), three over
// j2 -> X[k1 + 3 k2]).  Reference codelet: n1_9.
template <typename T> struct Bf<9, T> {
    static B2_HD void run(cpx<T> *v) {
        const T c1 = (T)0.76604444311897803520, s1 = (T)0.64278760968653932632;   // 2 pi / 9
        const T c2 = (T)0.17364817766693034885, s2 = (T)0.98480775301220805937;   // 4 pi / 9
        const T c4 = (T)-0.93969262078590838405, s4 = (T)0.34202014332566873304;  // 8 pi / 9
        cpx<T> b[3][3];  // b[j2][k1]
#pragma unroll
        for (int j2 = 0; j2 < 3; ++j2) {
            b[j2][0] = v[j2];
            b[j2][1] = v[3 + j2];
            b[j2][2] = v[6 + j2];
            Bf<3, T>::run(b[j2]);
        }
        b[1][1] = cmul(b[1][1], cpx<T>{c1, -s1});
        b[1][2] = cmul(b[1][2], cpx<T>{c2, -s2});
        b[2][1] = cmul(b[2][1], cpx<T>{c2, -s2});
        b[2][2] = cmul(b[2][2], cpx<T>{c4, -s4});
#pragma unroll
        for (int k1 = 0; k1 < 3; ++k1) {
            cpx<T> r[3] = {b[0][k1], b[1][k1], b[2][k1]};
            Bf<3, T>::run(r);
            v[k1] = r[0];
            v[k1 + 3] = r[1];
            v[k1 + 6] = r[2];
        }
    }
};
