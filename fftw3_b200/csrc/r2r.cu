// csrc/r2r.cu -- real-to-real transforms (fftw_plan_r2r*, fftw_plan_many_r2r, fftw_plan_guru_r2r,
// fftw_execute_r2r): R2HC, HC2R, DHT and the eight DCT / DST kinds, any rank, any kind per dimension.
//
// The reference has a family of dedicated solvers (rdft-dht, reodft010e-r2hc, reodft11e-radix2,
// redft00e-r2hc-pad, ... fftw/fftw_api.c:11700-14160).  This engine keeps ONE fast real transform (the
// r2c / c2r tile kernels) and reduces every kind to it by the symmetric-extension identities the
// reference's manual defines the kinds with (fftw/doc/reference.texi "1d Real-even DFTs (DCTs)",
// "1d Real-odd DFTs (DSTs)"), one dimension at a time on a dense work array:
//
//   embed   x (n reals) -> y (N reals): the even / odd extension of the kind (N = n, 2(n-1), 2(n+1));
//           types II / III (x10, x01) use the length-2n extension and a quarter-sample twiddle
//           exp(-+ i pi k / 2n) after the r2c (II) or before the c2r (III); type IV the 8n zero-stuffed one
//   r2c_N   batched over all lines of the array                    [the hot kernels]
//   extract X[k] = Re or -Im of F at the kind's index map
//
// (HC2R: build the Hermitian half from the halfcomplex line, c2r.)  Correct for every kind and size
// the r2c path accepts; the 4n / 8n extensions cost bandwidth, which is acceptable for a family that
// SURVEY.md section 8(f) ranks last -- a dedicated n-point pre/post-twiddle kernel is the obvious
// follow-up where a kind matters.
#include <math.h>

#include <algorithm>
#include <vector>

#include "b200fft.h"
#include "device_rt.h"
#include "plan_struct.h"

using namespace b2;

enum { K_R2HC = 0, K_HC2R = 1, K_DHT = 2, K_REDFT00 = 3, K_REDFT01 = 4, K_REDFT10 = 5, K_REDFT11 = 6,
       K_RODFT00 = 7, K_RODFT01 = 8, K_RODFT10 = 9, K_RODFT11 = 10 };

static long long r2r_ext_len(int kind, long long n) {
    switch (kind) {
    case K_R2HC: case K_HC2R: case K_DHT: return n;
    case K_REDFT00: return 2 * (n - 1);
    case K_RODFT00: return 2 * (n + 1);
    case K_REDFT10: case K_REDFT01: case K_RODFT10: case K_RODFT01: return 2 * n;  // + a quarter-sample twiddle
    case K_REDFT11: case K_RODFT11: return 8 * n;
    }
    return -1;
}

#define B2_R2R_MAXD (2 * B2FFT_MAXRANK)
struct R2RCopyParams {
    int nd;
    long long n[B2_R2R_MAXD], st[B2_R2R_MAXD];
    long long total;
};
struct R2RLineParams {
    int kind;
    long long n, N, sd, lines;  // sd: dense stride of the transformed dim; lines = total / n
};

B2_HD long long b2_r2r_offset(const R2RCopyParams &p, long long i) {
    long long off = 0;
    for (int d = p.nd - 1; d >= 0; --d) {
        const long long q = i / p.n[d];
        off += (i - q * p.n[d]) * p.st[d];
        i = q;
    }
    return off;
}
B2_HD long long b2_r2r_line_base(const R2RLineParams &p, long long l) {
    const long long o = l / p.sd;
    return o * p.n * p.sd + (l - o * p.sd);
}

// y[q] of the extension of kind for the line x (stride sd)
B2_HD real_t b2_r2r_embed(const R2RLineParams &p, const real_t *x, long long q) {
    const long long n = p.n, N = p.N, sd = p.sd;
    switch (p.kind) {
    case K_R2HC: case K_DHT: return x[q * sd];
    case K_REDFT00: return x[(q <= N / 2 ? q : N - q) * sd];
    case K_REDFT10: return q < n ? x[q * sd] : x[(2 * n - 1 - q) * sd];
    case K_REDFT11:
        if (!(q & 1)) return 0;
        if (q < 2 * n) return x[((q - 1) / 2) * sd];
        if (q > 6 * n) return x[((8 * n - q - 1) / 2) * sd];
        return 0;
    case K_RODFT00:
        if (q == 0 || q == n + 1) return 0;
        return q <= n ? x[(q - 1) * sd] : -x[(N - q - 1) * sd];
    case K_RODFT10: return q < n ? x[q * sd] : -x[(2 * n - 1 - q) * sd];
    case K_RODFT11:
        if (!(q & 1)) return 0;
        if (q < 2 * n) return x[((q - 1) / 2) * sd];
        if (q > 6 * n) return -x[((8 * n - q - 1) / 2) * sd];
        return 0;
    }
    return 0;
}
// X[k] from the half spectrum F (N/2+1 complex) of the extension
B2_HD real_t b2_r2r_extract(const R2RLineParams &p, const cpx_t *F, long long k) {
    const long long n = p.n;
    switch (p.kind) {
    case K_R2HC: return 2 * k <= n ? F[k].x : F[n - k].y;
    case K_DHT: return 2 * k <= n ? F[k].x - F[k].y : F[n - k].x + F[n - k].y;
    case K_REDFT00: return F[k].x;
    case K_REDFT11: return F[2 * k + 1].x;
    case K_RODFT00: return -F[k + 1].y;
    case K_RODFT11: return -F[2 * k + 1].y;
    case K_REDFT10: {  // F_k = exp(i pi k / 2n) X_k
        const double a = 3.14159265358979323846 * (double)k / (double)(2 * n);
        return (real_t)((double)F[k].x * cos(a) + (double)F[k].y * sin(a));
    }
    case K_RODFT10: {  // F_m = -i exp(i pi m / 2n) X_k, m = k + 1
        const double a = 3.14159265358979323846 * (double)(k + 1) / (double)(2 * n);
        return (real_t)((double)F[k + 1].x * sin(a) - (double)F[k + 1].y * cos(a));
    }
    }
    return 0;
}
B2_HD bool b2_r2r_is_c2r(int kind) { return kind == K_HC2R || kind == K_REDFT01 || kind == K_RODFT01; }
// the kinds computed by a c2r: Hermitian half F[k], k <= N/2, built from the line h
//   HC2R     the halfcomplex line itself
//   REDFT01  F_j = x_j exp(i pi j / 2n), F_n = 0:          c2r_2n(F)[k] = x_0 + 2 sum x_j cos(pi j (k+1/2) / n)
//   RODFT01  F_j = -i x_(j-1) exp(i pi j / 2n), F_0 = 0, F_n = x_(n-1)
B2_HD cpx_t b2_r2r_hc(const R2RLineParams &p, const real_t *h, long long k) {
    const long long n = p.n, sd = p.sd;
    if (p.kind == K_HC2R) {
        if (k == 0 || 2 * k == n) return cpx_t{h[k * sd], 0};
        return cpx_t{h[k * sd], h[(n - k) * sd]};
    }
    const double a = 3.14159265358979323846 * (double)k / (double)(2 * n);
    if (p.kind == K_REDFT01) {
        if (k == n) return cpx_t{0, 0};
        const double v = (double)h[k * sd];
        return cpx_t{(real_t)(v * cos(a)), (real_t)(v * sin(a))};
    }
    if (k == 0) return cpx_t{0, 0};
    const double v = (double)h[(k - 1) * sd];
    if (k == n) return cpx_t{(real_t)v, 0};
    return cpx_t{(real_t)(v * sin(a)), (real_t)(-v * cos(a))};  // -i v exp(i a)
}

#if B2_CUDA
#define B2_GRID_LOOP(i, total) \
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (total); i += (long long)gridDim.x * blockDim.x)
__global__ void __launch_bounds__(256) b2_r2r_copy_kernel(const R2RCopyParams p, const real_t *B2_RESTRICT src,
                                                          real_t *B2_RESTRICT dst, int gather) {
    B2_GRID_LOOP(i, p.total) {
        const long long off = b2_r2r_offset(p, i);
        if (gather)
            dst[i] = src[off];
        else
            dst[off] = src[i];
    }
}
__global__ void __launch_bounds__(256) b2_r2r_embed_kernel(const R2RLineParams p, const real_t *B2_RESTRICT D,
                                                           real_t *B2_RESTRICT W1, cpx_t *B2_RESTRICT W2) {
    if (b2_r2r_is_c2r(p.kind)) {
        const long long hc = p.N / 2 + 1;
        B2_GRID_LOOP(i, p.lines * hc) {
            const long long l = i / hc, k = i - l * hc;
            W2[i] = b2_r2r_hc(p, D + b2_r2r_line_base(p, l), k);
        }
    } else {
        B2_GRID_LOOP(i, p.lines * p.N) {
            const long long l = i / p.N, q = i - l * p.N;
            W1[i] = b2_r2r_embed(p, D + b2_r2r_line_base(p, l), q);
        }
    }
}
__global__ void __launch_bounds__(256) b2_r2r_extract_kernel(const R2RLineParams p, const real_t *B2_RESTRICT W1,
                                                             const cpx_t *B2_RESTRICT W2, real_t *B2_RESTRICT D) {
    B2_GRID_LOOP(i, p.lines * p.n) {
        // consecutive threads walk the lines (the dense dim), not the points of one line
        const long long k = i / p.lines, l = i - k * p.lines;
        real_t *x = D + b2_r2r_line_base(p, l);
        x[k * p.sd] = b2_r2r_is_c2r(p.kind) ? W1[l * p.N + k] : b2_r2r_extract(p, W2 + l * (p.N / 2 + 1), k);
    }
}
#endif

struct R2RDim {
    R2RLineParams lp;
    b2fft_plan *plan = nullptr;
};
struct b2fft_r2r_s {
    std::mutex exec_mutex;  // executes of one plan share its work arrays: serialised
    R2RCopyParams in_cp, out_cp;
    long long in_extent = 0, out_extent = 0, total = 0;
    std::vector<R2RDim> dims;
    real_t *D = nullptr, *W1 = nullptr;
    cpx_t *W2 = nullptr;
    real_t *stage[2] = {nullptr, nullptr};
    double flops = 0;
    std::string desc;
};

static unsigned r2r_grid(long long total) {
    return (unsigned)std::max<long long>(1, std::min<long long>((total + 255) / 256, (long long)sm_count() * 32));
}

extern "C" void b2fft_r2r_plan_destroy(b2fft_r2r *r) {
    if (!r) return;
    for (R2RDim &d : r->dims) b2fft_plan_destroy(d.plan);
    dev_free(r->D);
    dev_free(r->W1);
    dev_free(r->W2);
    for (real_t *p : r->stage) dev_free(p);
    delete r;
}

// dims / vdims: strides in reals; kinds[i]: fftw_r2r_kind of transform dim i
extern "C" b2fft_r2r *b2fft_r2r_plan_create(int rank, const b2fft_dim *dims, int vrank, const b2fft_dim *vdims,
                                            const int *kinds) {
    if (rank < 0 || rank > B2FFT_MAXRANK || vrank < 0 || vrank > B2FFT_MAXRANK || (rank > 0 && (!dims || !kinds))) return nullptr;
    DeviceScope caller;
    if (ensure_device() != 0) return nullptr;
    b2fft_r2r *r = new b2fft_r2r_s();
    // dense order: batch dims outermost, then the transform dims
    std::vector<b2fft_dim> all;
    for (int i = 0; i < vrank; ++i) all.push_back(vdims[i]);
    for (int i = 0; i < rank; ++i) all.push_back(dims[i]);
    r->in_cp.nd = r->out_cp.nd = (int)all.size();
    r->total = 1;
    r->in_extent = r->out_extent = 1;
    bool ok = true;
    for (size_t i = 0; i < all.size(); ++i) {
        if (all[i].n < (i < (size_t)vrank ? 0 : 1) || all[i].is < 0 || all[i].os < 0) ok = false;
        r->in_cp.n[i] = r->out_cp.n[i] = all[i].n;
        r->in_cp.st[i] = all[i].is;
        r->out_cp.st[i] = all[i].os;
        r->total *= all[i].n;
        r->in_extent += (all[i].n - 1) * all[i].is;
        r->out_extent += (all[i].n - 1) * all[i].os;
    }
    if (!ok) {
        set_error("r2r: bad dimensions or negative strides");
        b2fft_r2r_plan_destroy(r);
        return nullptr;
    }
    r->in_cp.total = r->out_cp.total = r->total;
    long long w1 = 1, w2 = 1;
    long long sd = 1;
    std::string kinds_s;
    static const char *kn[] = {"R2HC", "HC2R", "DHT", "REDFT00", "REDFT01", "REDFT10", "REDFT11", "RODFT00", "RODFT01", "RODFT10", "RODFT11"};
    for (int i = rank - 1; i >= 0 && ok && r->total > 0; --i) {
        R2RDim d;
        d.lp.kind = kinds[i];
        d.lp.n = dims[i].n;
        d.lp.N = r2r_ext_len(kinds[i], dims[i].n);
        d.lp.sd = sd;
        d.lp.lines = r->total / dims[i].n;
        sd *= dims[i].n;
        if (kinds[i] < 0 || kinds[i] > 10 || d.lp.N < 1) {
            // REDFT00 needs n >= 2, like the reference (fftw_api.c: redft00e requires n > 1)
            set_error("r2r: unsupported kind %d for size %td", kinds[i], dims[i].n);
            ok = false;
            break;
        }
        kinds_s = std::string(kn[kinds[i]]) + (kinds_s.empty() ? "" : " x ") + kinds_s;
        if (dims[i].n == 1 && kinds[i] <= K_DHT) continue;  // length-1 R2HC / HC2R / DHT: the identity
        b2fft_geom g;
        memset(&g, 0, sizeof g);
        g.kind = b2_r2r_is_c2r(kinds[i]) ? B2FFT_C2R : B2FFT_R2C;
        g.sign = b2_r2r_is_c2r(kinds[i]) ? +1 : -1;
        g.rank = 1;
        g.vrank = 1;
        const long long hc = d.lp.N / 2 + 1;
        if (b2_r2r_is_c2r(kinds[i])) {
            g.dims[0] = b2fft_dim{(ptrdiff_t)d.lp.N, 1, 1};
            g.vdims[0] = b2fft_dim{(ptrdiff_t)d.lp.lines, (ptrdiff_t)hc, (ptrdiff_t)d.lp.N};
        } else {
            g.dims[0] = b2fft_dim{(ptrdiff_t)d.lp.N, 1, 1};
            g.vdims[0] = b2fft_dim{(ptrdiff_t)d.lp.lines, (ptrdiff_t)d.lp.N, (ptrdiff_t)hc};
        }
        d.plan = b2fft_plan_create(&g);
        if (!d.plan) {
            ok = false;
            break;
        }
        w1 = std::max(w1, d.lp.lines * d.lp.N);
        w2 = std::max(w2, d.lp.lines * hc);
        r->flops += 2.5 * (double)d.lp.n * (d.lp.n > 1 ? log2((double)d.lp.n) : 0) * (double)d.lp.lines;
        r->dims.push_back(d);
    }
    if (ok) {
        r->D = (real_t *)dev_malloc((size_t)std::max<long long>(r->total, 1) * sizeof(real_t));
        r->W1 = (real_t *)dev_malloc((size_t)w1 * sizeof(real_t));
        r->W2 = (cpx_t *)dev_malloc((size_t)w2 * sizeof(cpx_t));
        ok = r->D && r->W1 && r->W2;
    }
    if (!ok) {
        b2fft_r2r_plan_destroy(r);
        return nullptr;
    }
    char buf[256];
    snprintf(buf, sizeof buf, "r2r %s on %lld reals: per dimension embed -> batched r2c of the extension -> extract\n",
             kinds_s.empty() ? "(copy)" : kinds_s.c_str(), r->total);
    r->desc = buf;
    return r;
}

extern "C" int b2fft_r2r_execute(b2fft_r2r *r, void *in_, void *out_) {
    if (!r) return -1;
    std::lock_guard<std::mutex> guard(r->exec_mutex);
    DeviceScope caller;
    if (ensure_device() != 0) return -1;
    if (r->total <= 0) return 0;
    real_t *in = (real_t *)in_, *out = (real_t *)out_;
    real_t *din = in, *dout = out;
#if B2_CUDA
    const bool hin = !is_device_ptr(in), hout = !is_device_ptr(out);
    if (hin) {
        if (!r->stage[0]) r->stage[0] = (real_t *)dev_malloc((size_t)r->in_extent * sizeof(real_t));
        if (!r->stage[0]) return -3;
        din = r->stage[0];
        B2_CK(cudaMemcpy(din, in, (size_t)r->in_extent * sizeof(real_t), cudaMemcpyHostToDevice));
    }
    if (hout) {
        if (!r->stage[1]) r->stage[1] = (real_t *)dev_malloc((size_t)r->out_extent * sizeof(real_t));
        if (!r->stage[1]) return -3;
        dout = r->stage[1];
        if (r->out_extent != r->total)  // keep whatever lives in the gaps of a strided output
            B2_CK(cudaMemcpy(dout, out, (size_t)r->out_extent * sizeof(real_t), cudaMemcpyHostToDevice));
    }
    b2_r2r_copy_kernel<<<r2r_grid(r->total), 256>>>(r->in_cp, din, r->D, 1);
    ++launch_counter();
    B2_CK(cudaGetLastError());
#else
    for (long long i = 0; i < r->total; ++i) r->D[i] = din[b2_r2r_offset(r->in_cp, i)];
#endif
    for (R2RDim &d : r->dims) {
        const R2RLineParams &p = d.lp;
        const long long hc = p.N / 2 + 1;
#if B2_CUDA
        b2_r2r_embed_kernel<<<r2r_grid(p.lines * (b2_r2r_is_c2r(p.kind) ? hc : p.N)), 256>>>(p, r->D, r->W1, r->W2);
        ++launch_counter();
        B2_CK(cudaGetLastError());
#else
        if (b2_r2r_is_c2r(p.kind)) {
            for (long long i = 0; i < p.lines * hc; ++i) r->W2[i] = b2_r2r_hc(p, r->D + b2_r2r_line_base(p, i / hc), i % hc);
        } else {
            for (long long i = 0; i < p.lines * p.N; ++i) r->W1[i] = b2_r2r_embed(p, r->D + b2_r2r_line_base(p, i / p.N), i % p.N);
        }
#endif
        int rc = b2_r2r_is_c2r(p.kind) ? b2fft_execute_async(d.plan, r->W2, r->W1, nullptr)
                                  : b2fft_execute_async(d.plan, r->W1, r->W2, nullptr);
        if (rc) return rc;
#if B2_CUDA
        b2_r2r_extract_kernel<<<r2r_grid(p.lines * p.n), 256>>>(p, r->W1, r->W2, r->D);
        ++launch_counter();
        B2_CK(cudaGetLastError());
#else
        for (long long i = 0; i < p.lines * p.n; ++i) {
            const long long k = i / p.lines, l = i % p.lines;
            real_t *x = r->D + b2_r2r_line_base(p, l);
            x[k * p.sd] = b2_r2r_is_c2r(p.kind) ? r->W1[l * p.N + k] : b2_r2r_extract(p, r->W2 + l * hc, k);
        }
#endif
    }
#if B2_CUDA
    b2_r2r_copy_kernel<<<r2r_grid(r->total), 256>>>(r->out_cp, r->D, dout, 0);
    ++launch_counter();
    B2_CK(cudaGetLastError());
    if (hout) B2_CK(cudaMemcpy(out, dout, (size_t)r->out_extent * sizeof(real_t), cudaMemcpyDeviceToHost));
    B2_CK(cudaDeviceSynchronize());
#else
    for (long long i = 0; i < r->total; ++i) dout[b2_r2r_offset(r->out_cp, i)] = r->D[i];
#endif
    return 0;
}

extern "C" int b2fft_r2r_describe(const b2fft_r2r *r, char *buf, size_t n) {
    return r ? snprintf(buf, n, "%s", r->desc.c_str()) : 0;
}
extern "C" double b2fft_r2r_flops(const b2fft_r2r *r) { return r ? r->flops : 0.0; }
