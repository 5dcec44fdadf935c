// csrc/blue_tables.h -- host-side tables for the Bluestein path (blue_kernel.cuh): the chirp
// c[j] = exp(-i pi j^2 / n) with j^2 reduced mod 2n in exact integer arithmetic (the reference
// does the same reduction, dft_bluestein_sequence fftw/fftw_api.c:1598-1611), and
// bhat = FFT_M(conj(c) wrapped to length M) / M, computed once per (n, M) in long double by a
// plain radix-2 FFT on the host and rounded once to the working precision.
#pragma once
#include <complex>

#include "twiddle.h"

namespace b2 {

inline void host_fft_pow2(std::vector<std::complex<long double>> &a) {
    const size_t n = a.size();
    for (size_t i = 1, j = 0; i < n; ++i) {  // bit reversal
        size_t bit = n >> 1;
        for (; j & bit; bit >>= 1) j ^= bit;
        j ^= bit;
        if (i < j) std::swap(a[i], a[j]);
    }
    std::vector<std::complex<long double>> w(n / 2 ? n / 2 : 1);
    for (size_t k = 0; k < n / 2; ++k) {
        long double c, s;
        unit_root(k, n, &c, &s);
        w[k] = std::complex<long double>(c, s);
    }
    for (size_t len = 2; len <= n; len <<= 1) {
        const size_t half = len / 2, step = n / len;
        for (size_t i = 0; i < n; i += len)
            for (size_t k = 0; k < half; ++k) {
                const std::complex<long double> u = a[i + k], v = a[i + k + half] * w[k * step];
                a[i + k] = u + v;
                a[i + k + half] = u - v;
            }
    }
}

struct BlueTables {
    std::mutex mu;
    // keyed by device as well (see TwiddleCache); a failed upload is NOT remembered: the next plan tries again
    std::map<std::pair<int, unsigned long long>, cpx_t *> chirps;
    std::map<std::tuple<int, unsigned long long, unsigned long long>, cpx_t *> bhats;

    static cpx_t *upload(const std::vector<cpx_t> &h) {
        cpx_t *d = (cpx_t *)dev_malloc(h.size() * sizeof(cpx_t));
        if (!d) return nullptr;
        if (dev_upload(d, h.data(), h.size() * sizeof(cpx_t)) != 0) {
            dev_free(d);
            return nullptr;
        }
        return d;
    }
    cpx_t *chirp(unsigned long long n) {
        std::lock_guard<std::mutex> lk(mu);
        const auto ckey = std::make_pair(current_device(), n);
        auto it = chirps.find(ckey);
        if (it != chirps.end()) return it->second;
        std::vector<cpx_t> h(n);
        for (unsigned long long j = 0; j < n; ++j) {
            long double c, s;
            unit_root((j * j) % (2 * n), 2 * n, &c, &s);  // exp(-2 pi i (j^2 mod 2n) / 2n)
            h[j].x = (real_t)c;
            h[j].y = (real_t)s;
        }
        cpx_t *d = upload(h);
        if (d) chirps[ckey] = d;
        return d;
    }
    cpx_t *bhat(unsigned long long n, unsigned long long M) {
        std::lock_guard<std::mutex> lk(mu);
        auto key = std::make_tuple(current_device(), n, M);
        auto it = bhats.find(key);
        if (it != bhats.end()) return it->second;
        std::vector<std::complex<long double>> b(M, std::complex<long double>(0, 0));
        for (unsigned long long d = 0; d < n; ++d) {
            long double c, s;
            unit_root((d * d) % (2 * n), 2 * n, &c, &s);
            const std::complex<long double> v(c, -s);  // conj of the chirp
            b[d] = v;
            if (d) b[M - d] = v;
        }
        host_fft_pow2(b);
        std::vector<cpx_t> h(M);
        for (unsigned long long k = 0; k < M; ++k) {
            h[k].x = (real_t)(b[k].real() / (long double)M);
            h[k].y = (real_t)(b[k].imag() / (long double)M);
        }
        cpx_t *d = upload(h);
        if (d) bhats[key] = d;
        return d;
    }
    void clear() {
        std::lock_guard<std::mutex> lk(mu);
        for (auto &kv : chirps) dev_free(kv.second);
        for (auto &kv : bhats) dev_free(kv.second);
        chirps.clear();
        bhats.clear();
    }
};
inline BlueTables &blue_tables() {
    static BlueTables t;
    return t;
}

}  // namespace b2
