// csrc/engine.cu -- the CUDA engine behind include/b200fft.h: builds plans with the selector,
// launches the sm_100a kernels, stages host arrays over PCIe when the caller passes host
// pointers.  Compiled by nvcc for the product; compiled by g++ -DB2_EMULATE (-x c++) into the
// test-only emulation library.  There is deliberately no CPU execution path in the product
// build: without a CUDA device b2fft_plan_create returns NULL.
#include <chrono>
#include <mutex>

#include "b200fft.h"
#include "generic_kernel.cuh"
#include "plan_build_real.h"
#include "plan_struct.h"
#include "realpp_kernel.cuh"
#include "blue_kernel.cuh"
#include "chain_build.h"

using namespace b2;

// ------------------------------------------------------------------------- launching
// bfirst / bcount: a sub-range of the step's blocks (grouped execution); bcount < 0: all of them
static int launch_step(const StepPlan &s, const real_t *src, real_t *dst, stream_t stream, long long bfirst = 0, long long bcount = -1) {
    // host loops over the line dims that did not fit in the kernel descriptor (almost never any)
    long long idx_store[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    std::vector<long long> idx_heap;
    long long *idx = idx_store;
    if (s.outer.size() > 8) {
        idx_heap.assign(s.outer.size(), 0);
        idx = idx_heap.data();
    }
    for (;;) {
        long long so = 0, dof = 0;
        for (size_t i = 0; i < s.outer.size(); ++i) {
            so += idx[i] * s.outer[i].s;
            dof += idx[i] * s.outer[i].d;
        }
        const real_t *sp = src + so;
        real_t *dp = dst + dof;
        // a chain member was re-tiled for the chained launch: on its own it runs the generic kernel
        bool fast = s.kind == SK_PASS && s.fast_id >= 0 && s.chain_id < 0;
        bool al_s = ((uintptr_t)sp % sizeof(cpx_t)) == 0, al_d = ((uintptr_t)dp % sizeof(cpx_t)) == 0;
        if (!al_s || !al_d) fast = false;  // vector access needs 2-real alignment of the base addresses
        PassParams pp = fast ? s.pp : s.gen_pp;
        if (!al_s) pp.vec_ld = 0;
        if (!al_d) pp.vec_st = 0;
        long long nblocks = (s.kind == SK_PASS && !fast) ? s.gen_nblocks : s.nblocks;
        if (bcount >= 0) nblocks = std::min(nblocks, bfirst + bcount);
        const long long maxgrid = 1ll << 30;
        for (long long b0 = bfirst; b0 < nblocks; b0 += maxgrid) {
            const long long nb = std::min(maxgrid, nblocks - b0);
            if (s.kind == SK_PASS && fast) {
                int rc = b2_launch_tile_kernel(s.fast_id, pp, sp, dp, b0, nb, s.nthreads, s.smem, stream);
                if (rc) return rc;
            } else if (s.kind == SK_PASS) {
                const size_t smem = b2_generic_smem_bytes(pp.C, pp.lstride);
#if B2_CUDA
                b2_generic_pass_kernel<<<(unsigned)nb, s.gen_threads, smem, stream>>>(pp, sp, dp, b0);
#else
                std::vector<char> sm(smem + 64);
                for (long long b = 0; b < nb; ++b) b2_generic_pass_body(pp, sp, dp, b0 + b, sm.data(), s.gen_threads);
#endif
            } else if (s.kind == SK_BLUE) {
#if B2_CUDA
                b2_blue_kernel<<<(unsigned)nb, s.nthreads, 0, stream>>>(s.bp, sp, dp, b0);
#else
                for (long long b = 0; b < nb; ++b) b2_blue_body(s.bp, sp, dp, b0 + b, s.nthreads);
#endif
            } else {
#if B2_CUDA
                b2_realpp_kernel<<<(unsigned)nb, s.nthreads, 0, stream>>>(s.rp, s.kind == SK_RPOST, sp, dp, b0);
#else
                for (long long b = 0; b < nb; ++b) b2_realpp_body(s.rp, s.kind == SK_RPOST, sp, dp, b0 + b, s.nthreads);
#endif
            }
            ++launch_counter();  // (launch errors are collected once per execute: run_steps / b2_plan_execute_scatter)
            (void)stream;
        }
        size_t i = 0;
        for (; i < s.outer.size(); ++i) {
            if (++idx[i] < s.outer[i].n) break;
            idx[i] = 0;
        }
        if (i == s.outer.size()) break;
    }
    return 0;
}

static int run_chain(b2fft_plan *p, const Chain &ch, real_t *const *buf, stream_t stream) {
    ChainParams cp;
    memset(&cp, 0, sizeof cp);
    cp.nsteps = ch.count;
    cp.lag = ch.lag;
    cp.G = ch.G;
    cp.bpg = ch.bpg;
    cp.counters = ch.counters;
    cp.errflag = ch.counters + (size_t)ch.count * ch.G;
    for (int k = 0; k < ch.count; ++k) {
        const StepPlan &s = p->pc.steps[ch.first + k];
        ChainStep &c = cp.st[k];
        c.kind = s.kind;
        c.tpg = ch.tpg[k];
        c.nthreads = s.kind == SK_PASS ? s.nthreads : ch.threads;
        c.src = buf[s.src];
        c.dst = buf[s.dst];
        if (s.kind == SK_PASS) {
            const TileEntry &e = b2_tile_table()[s.fast_id];
            c.mode = e.mode;
            c.epi = e.epi;
            c.pp = s.pp;
            b2_pp_prepare(c.pp);
            c.pp.pf_ahead = 0;
        } else
            c.rp = s.rp;
    }
    return b2_launch_chain(cp, ch.threads, ch.smem, stream);
}

// Grouped execution (selector.h Step::grp_L).  Worth it only when the array between the two steps is much larger
// than the L2; group size: FFTW_B200_GROUP_MIB (default 32 MiB of that array per group).
// (the environment is read per call, but only for plans that have a groupable pair: tests lower the thresholds)
static long long group_slabs(const b2fft_plan *p, const StepPlan &s) {
    const char *e = getenv("FFTW_B200_GROUP_KIB");
    const double group_bytes = e ? 1024.0 * std::max(1ll, atoll(e)) : 32.0 * (1 << 20);
    const double total = (double)(s.dst == B_SCR ? p->pc.scr_reals : p->pc.out_reals) * sizeof(real_t);
    const double slab = total / (double)s.grp_L;
    return std::max<long long>(1, std::min<long long>(s.grp_L, (long long)(group_bytes / slab + 0.5)));
}
static bool group_pair_ok(const b2fft_plan *p, const StepPlan &s, const StepPlan &t) {
    // Opt-in (FFTW_B200_GROUPS=1): measured on B200 (profiles/r2_group_probe.md) every group size loses -- 512^3 1.39-2.02 ms
    // against 1.19, r2c 2^20 x 256 1.56-1.93 against 1.21: ~6 us of ramp-up and tail per small launch outweigh the HBM
    // traffic that the L2-resident intermediate saves.
    if (s.grp_L <= 1 || s.grp_L != t.grp_L || !getenv("FFTW_B200_GROUPS") || getenv("FFTW_B200_NO_GROUPS")) return false;
    if (s.kind != SK_PASS || t.kind != SK_PASS || s.fast_id < 0 || t.fast_id < 0 || s.chain_id >= 0 || t.chain_id >= 0) return false;
    if (!s.outer.empty() || !t.outer.empty() || s.dst != t.src) return false;
    if (s.nblocks % s.grp_L || t.nblocks % t.grp_L) return false;  // a tile must not straddle two slabs
    for (const StepPlan *q : {&s, &t})                             // (slab dim merged into the tile dim: whole tiles per slab)
        if (q->pp.nld == 1 && (q->pp.ln[0] % q->grp_L || (q->pp.ln[0] / q->grp_L) % q->pp.C)) return false;
    const double total = (double)(s.dst == B_SCR ? p->pc.scr_reals : p->pc.out_reals) * sizeof(real_t);
    const char *e = getenv("FFTW_B200_GROUP_MIN_KIB");
    if (total < (e ? 1024.0 * atoll(e) : 192.0 * (1 << 20))) return false;  // fits the L2 (or nearly): nothing to gain
    return group_slabs(p, s) < s.grp_L;
}

static int run_steps(b2fft_plan *p, real_t *in, real_t *out, stream_t stream) {
    const long long woff = p->pc.neg ? p->pc.out_off : 0;  // work arrays in the output layout: element 0 sits woff reals in
    real_t *buf[B_NBUF] = {in, out, p->scr ? p->scr + woff : nullptr, p->wrk ? p->wrk + woff : nullptr};
    const bool aligned = ((uintptr_t)in % sizeof(cpx_t)) == 0 && ((uintptr_t)out % sizeof(cpx_t)) == 0;
    if (!aligned && p->needs_aligned) {
        // arrays that are not aligned to a complex element: the same problem planned without the
        // steps that need vector access (built on first use)
        if (!p->alt) {
            no_two_pass_real() = true;
            p->alt = b2fft_plan_create(&p->pc.g);
            no_two_pass_real() = false;
            if (!p->alt) return -4;
        }
        return run_steps(p->alt, in, out, stream);
    }
    const size_t ns = p->pc.steps.size();
    for (size_t i = 0; i < ns; ++i) {
        const StepPlan &s = p->pc.steps[i];
        if (i + 1 < ns && aligned && group_pair_ok(p, s, p->pc.steps[i + 1])) {
            // ---- grouped execution: steps i and i+1 slab group by slab group, the array between them stays in L2
            const StepPlan &t = p->pc.steps[i + 1];
            const long long L = s.grp_L, bs = s.nblocks / L, bt = t.nblocks / L;
            const long long spg = group_slabs(p, s);
            for (long long a = 0; a < L; a += spg) {
                const long long cnt = std::min(spg, L - a);
                // scratch between the two steps: every group sits at the start of the scratch (never written back)
                const long long shift = (s.grp_slab > 0 && s.dst == B_SCR && t.src == B_SCR) ? a * s.grp_slab : 0;
                int rc = launch_step(s, buf[s.src], buf[s.dst] - shift, stream, a * bs, cnt * bs);
                if (!rc) rc = launch_step(t, buf[t.src] - shift, buf[t.dst], stream, a * bt, cnt * bt);
                if (rc) return rc;
            }
            ++i;
            continue;
        }
        if (s.chain_id >= 0 && s.chain_pos == 0 && aligned && p->chains) {
            const Chain &ch = (*p->chains)[s.chain_id];
            int rc = run_chain(p, ch, buf, stream);
            if (rc) return rc;
            i += ch.count - 1;
            continue;
        }
        int rc = launch_step(s, buf[s.src], buf[s.dst], stream);
        if (rc) return rc;
    }
#if B2_CUDA
    B2_CK(cudaGetLastError());
#endif
    return 0;
}

// ------------------------------------------------------------------------- scattered last step
int b2_plan_enable_scatter(b2fft_plan_s *p, int nowners, long long blk, long long plane_stride, const long long *pstride,
                           int self, const long long *rstride, int local) {
    if (!p || p->pc.steps.empty() || nowners < 1 || nowners > B2_MAXSCAT || blk < 1) return -1;
    if (p->chains && !p->chains->empty()) return -1;
    const StepPlan &last = p->pc.steps.back();
    if (last.kind != SK_PASS || last.fast_id < 0 || !last.outer.empty() || last.dst != B_OUT) return -1;
    const TileEntry &e = b2_tile_table()[last.fast_id];
    if (e.mode != MODE_COL || e.epi != 0) return -1;
    StepPlan *st = new StepPlan(last);
    PassParams &q = st->pp;
    q = last.gen_pp;  // as built by make_pass_step, before the tile selection
    q.scat_dim = -1;
    for (int d = 1; d < q.nld; ++d)
        if (q.ld[d] == plane_stride && q.scat_dim < 0) q.scat_dim = d;
    if (q.nld == 1 && q.ln[0] > 0) q.scat_dim = -1;  // a single plane: coordinate 0
    else if (q.scat_dim < 0) {
        delete st;
        return -1;
    }
    if (q.scat_dim >= 0) q.ld[q.scat_dim] = 0;  // the plane offset is per owner
    q.scat_n = nowners;
    q.scat_self1 = self >= 0 ? self + 1 : 0;
    q.scat_lplane = plane_stride;
    q.scat_blk = (int)blk;
    b2_fastdiv_make((unsigned long long)blk, q.scat_m, q.scat_s);
    for (int o = 0; o < nowners; ++o) {
        q.scat_pstride[o] = pstride[o];
        q.scat_rstride[o] = rstride ? rstride[o] : q.d_axis;
    }
    q.scat_local = local;
    b2_select_tile_kernel(*st);
    if (st->fast_id < 0 || b2_tile_table()[st->fast_id].epi != 8) {
        delete st;
        return -1;
    }
    delete p->scat;
    p->scat = st;
    return 0;
}

int b2_plan_execute_scatter(b2fft_plan_s *p, void *in, void *out, real_t *const *bases, long long coord0,
                            stream_t stream) {
    if (!p || !p->scat) return -1;
    real_t *buf[B_NBUF] = {(real_t *)in, (real_t *)out, p->scr, p->wrk};
    const size_t ns = p->pc.steps.size();
    for (size_t i = 0; i + 1 < ns; ++i) {
        const StepPlan &s = p->pc.steps[i];
        int rc = launch_step(s, buf[s.src], buf[s.dst], stream);
        if (rc) return rc;
    }
    StepPlan s = *p->scat;
    for (int o = 0; o < s.pp.scat_n; ++o) s.pp.scat_base[o] = bases[o];
    s.pp.scat_coord0 = coord0;
    int rc = launch_step(s, buf[s.src], buf[s.dst], stream);
#if B2_CUDA
    if (!rc) B2_CK(cudaGetLastError());
#endif
    return rc;
}

#if B2_CUDA
static void pipe_destroy(b2fft_plan *p);
#endif

// ------------------------------------------------------------------------- plan create
#if B2_CUDA
// function attributes (opt-in dynamic shared memory) are per DEVICE: once per device a plan is created on
static void init_kernel_attrs() {
    static std::mutex mu;
    static bool done[64] = {false};
    const int d = current_device();
    std::lock_guard<std::mutex> lk(mu);
    if (d >= 0 && d < 64 && done[d]) return;
    cudaFuncSetAttribute(b2_generic_pass_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem_optin());
    b2_init_tile_kernels();
    cudaGetLastError();  // an attribute a device refuses shows up as that kernel's launch error, not as the next execute's
    if (d >= 0 && d < 64) done[d] = true;
}
#endif

extern "C" b2fft_plan *b2fft_plan_create(const b2fft_geom *g) {
    if (!g) return nullptr;
    DeviceScope caller;
    if (ensure_device() != 0) return nullptr;
#if B2_CUDA
    init_kernel_attrs();
#endif
    if (g->rank < 0 || g->rank > B2FFT_MAXRANK || g->vrank < 0 || g->vrank > B2FFT_MAXRANK) {
        set_error("rank out of range");
        return nullptr;
    }
    b2fft_plan *p = new b2fft_plan_s();
    p->device = current_device();
    p->pc.g = *g;
    p->pc.in_place = g->in_place != 0;
    auto build = [&]() {
        p->pc.steps.clear();
        bool ok = (g->kind == B2FFT_C2C) ? build_c2c(p->pc) : build_real(p->pc);
        if (!ok) return false;
        for (StepPlan &s : p->pc.steps) {
            s.gen_threads = s.nthreads;
            s.gen_pp = s.pp;
            s.gen_nblocks = s.nblocks;
            if (s.kind == SK_PASS) b2_select_tile_kernel(s);
        }
        return true;
    };
    bool ok = build();
    if (ok) {
        // steps of the two-pass real path exist only as tile kernels: if the selector turned one down
        // after all, take the three-pass path
        bool missing = false;
        for (const StepPlan &s : p->pc.steps)
            if (s.needs_tile_kernel && s.fast_id < 0) missing = true;
        if (missing) {
            no_two_pass_real() = true;
            ok = build();
            no_two_pass_real() = false;
        }
        for (const StepPlan &s : p->pc.steps)
            if (s.needs_tile_kernel) p->needs_aligned = true;
    }
    if (ok) {
        if (p->pc.scr_reals > 0) {
            p->scr = (real_t *)dev_malloc((size_t)p->pc.scr_reals * sizeof(real_t));
            if (!p->scr) ok = false;
        }
        if (ok && p->pc.wrk_reals > 0) {
            p->wrk = (real_t *)dev_malloc((size_t)p->pc.wrk_reals * sizeof(real_t));
            if (!p->wrk) ok = false;
        }
        p->chains = new std::vector<Chain>();
        build_chains(p->pc.steps, *p->chains);
        for (Chain &ch : *p->chains) {
            const size_t nb = sizeof(unsigned) * ((size_t)ch.count * ch.G + 1);
            ch.counters = (unsigned *)dev_malloc(nb);
            if (!ch.counters) ok = false;
#if B2_CUDA
            else if (cudaMemset(ch.counters, 0, nb) != cudaSuccess) ok = false;
#endif
        }
    }
    if (!ok) {
        delete p;
        return nullptr;
    }
    std::string d;
    static const char *bn[] = {"in", "out", "scratch", "work"};
    for (size_t i = 0; i < p->pc.steps.size(); ++i) {
        const StepPlan &s = p->pc.steps[i];
        d += "  pass " + std::to_string(i) + ": " + s.desc + " [" + bn[s.src] + "->" + bn[s.dst] + "]";
        if (s.chain_id >= 0) {
            const Chain &ch = (*p->chains)[s.chain_id];
            d += " {chained launch " + std::to_string(s.chain_id) + ", step " + std::to_string(s.chain_pos) + ": " +
                 std::to_string(ch.G) + " groups x " + std::to_string(ch.tpg[s.chain_pos]) + " tiles, " +
                 std::to_string(ch.threads) + " threads, lag " + std::to_string(ch.lag) + "}";
        }
        d += "\n";
    }
    p->pc.desc = d;
    return p;
}

// ------------------------------------------------------------------------- FFTW_MEASURE
static const char *cfg_name(int bit) {
    switch (bit) {
    case B2_CFG_NO_TMA: return "round-1 column kernels instead of TMA";
    case B2_CFG_NO_PIPE: return "no async-copy pipelined column kernels";
    case B2_CFG_NO_PREFER3: return "two passes instead of three for a long axis";
    case B2_CFG_THREE_PASS_REAL: return "three-pass real transform";
    case B2_CFG_NO_TWO_PASS_REAL: return "no two-pass real transform";
    case B2_CFG_TWO_PASS_C2R: return "two-pass c2r";
    case B2_CFG_NO_MIX: return "generic kernel instead of the mixed-radix tile kernels";
    }
    return "?";
}
static b2fft_plan *create_with_cfg(const b2fft_geom *g, int cfg) {
    const int saved = cfg_override();
    cfg_override() = cfg;
    b2fft_plan *p = b2fft_plan_create(g);
    cfg_override() = saved;
    if (p) p->cfg = cfg;
    return p;
}
// average time of one execute on device arrays, in ms; < 0: could not be measured
static double time_plan(b2fft_plan *p, real_t *in, real_t *out) {
    real_t *ri = in + p->pc.in_off, *ro = (p->pc.in_place ? in : out) + p->pc.out_off;
    const int reps = 3;
#if B2_CUDA
    cudaEvent_t e0, e1;
    if (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess) return -1;
    float ms = -1;
    if (run_steps(p, ri, ro, 0) == 0 && cudaEventRecord(e0, 0) == cudaSuccess) {
        bool ok = true;
        for (int i = 0; i < reps && ok; ++i) ok = run_steps(p, ri, ro, 0) == 0;
        ok = ok && cudaEventRecord(e1, 0) == cudaSuccess && cudaEventSynchronize(e1) == cudaSuccess &&
             cudaEventElapsedTime(&ms, e0, e1) == cudaSuccess;
        if (!ok) ms = -1;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaGetLastError();
    return ms < 0 ? -1 : ms / reps;
#else
    if (run_steps(p, ri, ro, nullptr) != 0) return -1;
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < reps; ++i)
        if (run_steps(p, ri, ro, nullptr) != 0) return -1;
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count() / reps;
#endif
}

extern "C" b2fft_plan *b2fft_plan_create_tuned(const b2fft_geom *g, int cfg) {
    if (!g) return nullptr;
    if (cfg == B2FFT_CFG_DEFAULT) return b2fft_plan_create(g);
    DeviceScope caller;
    if (cfg >= 0) return create_with_cfg(g, cfg & B2_CFG_ALL);  // replay (wisdom): that configuration, unmeasured
    // ---- measure: the default plus every single switch that changes the plan for this geometry
    b2fft_plan *best = create_with_cfg(g, 0);
    if (!best || best->pc.empty) return best;
    std::vector<b2fft_plan *> cand{best};
    for (int bit = 1; bit <= B2_CFG_ALL; bit <<= 1) {
        b2fft_plan *q = create_with_cfg(g, bit);
        if (q && q->pc.desc != best->pc.desc)
            cand.push_back(q);
        else
            b2fft_plan_destroy(q);
    }
    std::string note = "  measured (FFTW_MEASURE):";
    if (cand.size() > 1) {
        // zero-filled plan-owned arrays: the cost of an FFT does not depend on the data (the reference's own speed
        // test transforms zeros, libbench2/speed.c:42-53) and the caller's arrays stay untouched
        const size_t ib = (size_t)best->pc.in_reals * sizeof(real_t), ob = (size_t)best->pc.out_reals * sizeof(real_t);
        real_t *din = (real_t *)dev_malloc(std::max(ib, (size_t)16)), *dout = best->pc.in_place ? din : (real_t *)dev_malloc(std::max(ob, (size_t)16));
        bool ok = din && dout;
#if B2_CUDA
        ok = ok && cudaMemset(din, 0, ib) == cudaSuccess && (dout == din || cudaMemset(dout, 0, ob) == cudaSuccess);
#else
        if (ok) {
            memset(din, 0, ib);
            if (dout != din) memset(dout, 0, ob);
        }
#endif
        double tbest = 1e300;
        size_t ibest = 0;
        for (size_t i = 0; ok && i < cand.size(); ++i) {
            const double t = time_plan(cand[i], din, dout);
            char buf[160];
            snprintf(buf, sizeof buf, " [%s: %.4f ms]", cand[i]->cfg ? cfg_name(cand[i]->cfg) : "default", t);
            note += buf;
            if (t >= 0 && t < tbest) {
                tbest = t;
                ibest = i;
            }
        }
        if (dout != din) dev_free(dout);
        dev_free(din);
        if (!ok) {
            set_error("");
            note += " work arrays for the measurement could not be allocated: default configuration";
        }
        for (size_t i = 0; i < cand.size(); ++i)
            if (i != ibest) b2fft_plan_destroy(cand[i]);
        best = cand[ibest];
        best->measured = ok;
    } else {
        note += " no alternative configuration for this geometry";
        best->measured = true;
    }
    best->pc.desc += note + "\n";
    return best;
}
extern "C" int b2fft_plan_config(const b2fft_plan *p) { return p ? p->cfg : 0; }
extern "C" int b2fft_plan_measured(const b2fft_plan *p) { return p && p->measured ? 1 : 0; }

extern "C" void b2fft_plan_destroy(b2fft_plan *p) {
    if (!p) return;
    DeviceScope caller;
    ensure_device_of(p->device);
    if (p->slab) b2_slab_destroy(p);
#if B2_CUDA
    pipe_destroy(p);
#endif
    if (p->chains) {
        for (Chain &ch : *p->chains) dev_free(ch.counters);
        delete p->chains;
    }
    dev_free(p->scr);
    dev_free(p->wrk);
    if (p->alt) b2fft_plan_destroy(p->alt);
    delete p->scat;
    dev_free(p->dev_in);
    dev_free(p->dev_out);
#if B2_CUDA
    if (p->own_stream && p->stream) cudaStreamDestroy(p->stream);
#endif
    delete p;
}

// ------------------------------------------------------------------------- host pipeline
#if B2_CUDA
static void pipe_destroy(b2fft_plan *p) {
    HostPipe *hp = p->pipe;
    if (!hp) return;
    for (auto &sl : hp->slot) {
        b2fft_plan_destroy(sl.plan);
        b2fft_plan_destroy(sl.plan_tail);
        if (sl.din != sl.dout) dev_free(sl.din);  // in place: one buffer
        dev_free(sl.dout);
        if (sl.stream) cudaStreamDestroy(sl.stream);
    }
    delete hp;
    p->pipe = nullptr;
}

// Decide once whether host executes of this plan can be chunked along a batch dimension.
static void pipe_setup(b2fft_plan *p) {
    HostPipe *hp = new HostPipe();
    p->pipe = hp;
    hp->state = -1;
    const PlanCore &pc = p->pc;
    const b2fft_geom &g = pc.g;
    if (getenv("FFTW_B200_NO_PIPELINE") || pc.neg) return;
    // dense arrays only: a chunk of the batch must be one contiguous range on both sides
    if (pc.in_dense != pc.in_reals || pc.out_dense != pc.out_reals) return;
    int v = -1;
    for (int i = 0; i < g.vrank; ++i)
        if (g.vdims[i].n > 1 && (v < 0 || g.vdims[i].os > g.vdims[v].os)) v = i;
    if (v < 0 || g.vdims[v].n < 4) return;
    const long long iu = (g.kind == B2FFT_R2C) ? 1 : 2, ou = (g.kind == B2FFT_C2R) ? 1 : 2;
    hp->in_step = (long long)g.vdims[v].is * iu;
    hp->out_step = (long long)g.vdims[v].os * ou;
    // the batch dim must be the outermost one on both sides
    if (hp->in_step * g.vdims[v].n != pc.in_reals || hp->out_step * g.vdims[v].n != pc.out_reals) return;
    const double bytes_per = (double)(hp->in_step + hp->out_step) * sizeof(real_t);
    long long chunk = (long long)std::max(1.0, (96.0 * (1 << 20)) / bytes_per);
    chunk = std::min<long long>(chunk, (g.vdims[v].n + 3) / 4);
    hp->chunk = chunk;
    hp->nchunks = (g.vdims[v].n + chunk - 1) / chunk;
    hp->tail = g.vdims[v].n % chunk;
    hp->vdim = v;
    for (auto &sl : hp->slot) {
        b2fft_geom g2 = g;
        g2.vdims[v].n = chunk;
        sl.plan = b2fft_plan_create(&g2);
        if (!sl.plan) return;
        const size_t ib = (size_t)hp->in_step * chunk * sizeof(real_t), ob = (size_t)hp->out_step * chunk * sizeof(real_t);
        sl.dout = (real_t *)dev_malloc(pc.in_place ? std::max(ib, ob) : ob);
        sl.din = pc.in_place ? sl.dout : (real_t *)dev_malloc(ib);
        if (!sl.din || !sl.dout) return;
        if (cudaStreamCreateWithFlags(&sl.stream, cudaStreamNonBlocking) != cudaSuccess) return;
    }
    if (hp->tail) {
        b2fft_geom g2 = g;
        g2.vdims[v].n = hp->tail;
        HostPipe::Slot &sl = hp->slot[(hp->nchunks - 1) % 3];
        sl.plan_tail = b2fft_plan_create(&g2);
        if (!sl.plan_tail) return;
    }
    hp->state = 1;
}

static int pipe_execute(b2fft_plan *p, void *in, void *out) {
    HostPipe *hp = p->pipe;
    const bool same = p->pc.in_place;
    for (long long c = 0; c < hp->nchunks; ++c) {
        HostPipe::Slot &sl = hp->slot[c % 3];
        const long long cnt = (c == hp->nchunks - 1 && hp->tail) ? hp->tail : hp->chunk;
        b2fft_plan *sp = (cnt == hp->chunk) ? sl.plan : sl.plan_tail;
        const size_t ib = (size_t)hp->in_step * cnt * sizeof(real_t), ob = (size_t)hp->out_step * cnt * sizeof(real_t);
        const real_t *hin = (const real_t *)in + c * hp->chunk * hp->in_step;
        real_t *hout = (real_t *)out + c * hp->chunk * hp->out_step;
        B2_CK(cudaMemcpyAsync(sl.din, hin, ib, cudaMemcpyHostToDevice, sl.stream));
        int rc = run_steps(sp, sl.din, same ? sl.din : sl.dout, sl.stream);
        if (rc) return rc;
        B2_CK(cudaMemcpyAsync(hout, same ? sl.din : sl.dout, ob, cudaMemcpyDeviceToHost, sl.stream));
    }
    for (auto &sl : hp->slot) B2_CK(cudaStreamSynchronize(sl.stream));
    return 0;
}
#endif

// ------------------------------------------------------------------------- execute
extern "C" int b2fft_execute_async(b2fft_plan *p, void *in, void *out, void *stream) {
    if (!p) return -1;
    if (p->pc.empty) return 0;
    DeviceScope caller;
    if (ensure_device_of(p->device) != 0) return -1;
    return run_steps(p, (real_t *)in, (real_t *)out, (stream_t)stream);
}

extern "C" int b2fft_execute(b2fft_plan *p, void *in, void *out) {
    if (!p) return -1;
    std::lock_guard<std::mutex> guard(p->exec_mutex);
    DeviceScope caller;
    if (p->slab) return b2_slab_execute(p, in, out);
    if (p->pc.empty) return 0;
    if (ensure_device_of(p->device) != 0) return -1;
    const bool same = (in == out);
    if (same != p->pc.in_place) {
        set_error("execute: in-place-ness differs from the plan (reference contract, fftw_api.c:434-440)");
        return -2;
    }
#if B2_CUDA
    if (!p->cls_valid || in != p->cls_in || out != p->cls_out) {
        p->cls_din = is_device_ptr(in);
        p->cls_dout = (out == in) ? p->cls_din : (bool)is_device_ptr(out);
        p->cls_in = in;
        p->cls_out = out;
        p->cls_valid = true;
    }
    const bool din = p->cls_din, dout = p->cls_dout;
    if (!din && !dout) {
        if (!p->pipe) pipe_setup(p);
        if (p->pipe->state == 1) return pipe_execute(p, in, out);
    }
    real_t *ri = (real_t *)in, *ro = (real_t *)out;
    const size_t ib = (size_t)p->pc.in_reals * sizeof(real_t), ob = (size_t)p->pc.out_reals * sizeof(real_t);
    // (negative strides: the host spans start in_off / out_off reals BELOW the pointers; in place both are equal)
    const long long ioff = p->pc.in_off, ooff = p->pc.out_off;
    const real_t *hin = (const real_t *)in - ioff;
    real_t *hout = (real_t *)out - ooff;
    if (!dout) {
        if (!p->dev_out) p->dev_out = (real_t *)dev_malloc(std::max(ob, same ? ib : (size_t)0));
        if (!p->dev_out) return -3;
        ro = p->dev_out + ooff;
        // keep whatever lives in the gaps of a strided / embedded output
        if (!same && p->pc.out_dense != p->pc.out_reals) B2_CK(cudaMemcpyAsync(p->dev_out, hout, ob, cudaMemcpyHostToDevice, 0));
    }
    if (!din) {
        if (same)
            ri = ro;
        else {
            if (!p->dev_in) p->dev_in = (real_t *)dev_malloc(ib);
            if (!p->dev_in) return -3;
            ri = p->dev_in + ioff;
        }
        B2_CK(cudaMemcpyAsync(ri - ioff, hin, ib, cudaMemcpyHostToDevice, 0));
    } else if (same)
        ri = ro;
    int rc = run_steps(p, ri, ro, 0);
    if (rc) return rc;
    if (!dout) B2_CK(cudaMemcpyAsync(hout, ro - ooff, std::max(ob, same ? ib : (size_t)0), cudaMemcpyDeviceToHost, 0));
    B2_CK(cudaStreamSynchronize(0));
    return 0;
#else
    return run_steps(p, (real_t *)in, (real_t *)out, nullptr);
#endif
}

// ------------------------------------------------------------------------- introspection
extern "C" int b2fft_plan_num_passes(const b2fft_plan *p) { return p ? (int)p->pc.steps.size() : 0; }
extern "C" int b2fft_plan_num_hbm_sweeps(const b2fft_plan *p) { return p ? (int)p->pc.steps.size() : 0; }
extern "C" size_t b2fft_plan_scratch_bytes(const b2fft_plan *p) {
    return p ? (size_t)(p->pc.scr_reals + p->pc.wrk_reals) * sizeof(real_t) : 0;
}
extern "C" int b2fft_plan_describe(const b2fft_plan *p, char *buf, size_t n) {
    if (!p) return 0;
    return snprintf(buf, n, "b200fft %s plan, %zu kernel pass(es), scratch %zu bytes\n%s", B2_PREC_NAME,
                    p->pc.steps.size(), (size_t)(p->pc.scr_reals + p->pc.wrk_reals) * sizeof(real_t), p->pc.desc.c_str());
}
extern "C" double b2fft_plan_flops(const b2fft_plan *p) { return p ? p->pc.flops : 0.0; }
extern "C" unsigned long long b2fft_launch_count(void) { return launch_counter(); }

extern "C" void *b2fft_malloc(size_t n) {
    DeviceScope caller;
    return host_malloc(n);
}
extern "C" void b2fft_free(void *p) { host_free(p); }
extern "C" void *b2fft_device_malloc(size_t n) {
    DeviceScope caller;
    if (ensure_device() != 0) return nullptr;
    return dev_malloc(n);
}
extern "C" void b2fft_device_free(void *p) { dev_free(p); }
extern "C" int b2fft_device_count(void) {
#if B2_CUDA
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
#else
    return 1;
#endif
}
extern "C" int b2fft_set_device(int d) {
#if B2_CUDA
    // selects the device of plans created from now on (existing plans keep theirs; the table caches are per
    // device); the caller's own current device is left alone
    DeviceScope caller;
    if (cudaSetDevice(d) != cudaSuccess) {
        cudaGetLastError();
        return -1;
    }
    device_ordinal() = d;
    return 0;
#else
    (void)d;
    return 0;
#endif
}
extern "C" const char *b2fft_last_error(void) { return last_error().c_str(); }
extern "C" void b2fft_cleanup(void) {
    twiddles().clear();
    blue_tables().clear();
}
extern "C" const char *b2fft_precision(void) { return B2_PREC_NAME; }
