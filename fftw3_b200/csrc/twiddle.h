// csrc/twiddle.h -- twiddle tables, generated on the host in extended precision and rounded
// ONCE to the working precision, then cached on the device per (kind, n).  Same accuracy
// contract as the reference's trig generator, which evaluates sin/cos in `trigreal` with an
// octant reduction and rounds once (real_cexp, fftw/fftw_api.c:18850-18892; twiddle cache
// fftw_twiddle_awake, fftw_api.c:19253): no recurrences, no fast-math intrinsics.
#pragma once
#include <math.h>

#include <map>
#include <tuple>
#include <mutex>
#include <utility>
#include <vector>

#include "device_rt.h"

namespace b2 {

// exp(-2 pi i num/den).  num/den = q/4 + f with |f| <= 1/8: the quarter-turn part is applied
// exactly by symmetry, and f's numerator (4 num - q den) is formed in exact integer
// arithmetic, so the only rounding is in cosl/sinl of a small argument (long double).
inline void unit_root(unsigned long long num, unsigned long long den, long double *c, long double *s) {
    num %= den;  // den < 2^60 assumed
    const long double twopi = 6.283185307179586476925286766559005768L;
    const long long q = (long long)((8ull * num + den) / (2ull * den));  // round(4 num / den)
    const long long fnum = (long long)(4ull * num) - q * (long long)den;   // exact
    const long double f = (long double)fnum / (4.0L * (long double)den);
    const long double cr = cosl(twopi * f), sr = sinl(twopi * f);
    long double cc, ss;
    switch (q & 3) {
    case 0: cc = cr; ss = sr; break;
    case 1: cc = -sr; ss = cr; break;
    case 2: cc = -cr; ss = -sr; break;
    default: cc = sr; ss = -cr; break;
    }
    *c = cc;
    *s = -ss;  // exp(-i theta)
}

struct TwiddleCache {
    std::mutex mu;
    // keyed by the device the table lives on as well: a table cudaMalloc'ed on one GPU is not addressable from another
    std::map<std::tuple<int, unsigned long long, unsigned long long, unsigned long long>, cpx_t *> tabs;

    // count entries w_den^(j*mult), j < count
    cpx_t *get(unsigned long long den, unsigned long long count, unsigned long long mult) {
        std::lock_guard<std::mutex> lk(mu);
        auto key = std::make_tuple(current_device(), den, count, mult);
        auto it = tabs.find(key);
        if (it != tabs.end()) return it->second;
        std::vector<cpx_t> h(count ? count : 1);
        for (unsigned long long j = 0; j < count; ++j) {
            long double c, s;
            unit_root((j * mult) % den, den, &c, &s);
            h[j].x = (real_t)c;
            h[j].y = (real_t)s;
        }
        cpx_t *d = (cpx_t *)dev_malloc(h.size() * sizeof(cpx_t));
        if (!d) return nullptr;
        if (dev_upload(d, h.data(), h.size() * sizeof(cpx_t)) != 0) {
            dev_free(d);
            return nullptr;
        }
        tabs[key] = d;
        return d;
    }
    void clear() {
        std::lock_guard<std::mutex> lk(mu);
        for (auto &kv : tabs) dev_free(kv.second);
        tabs.clear();
    }
};
inline TwiddleCache &twiddles() {
    static TwiddleCache c;
    return c;
}

}  // namespace b2
