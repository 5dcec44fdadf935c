// csrc/chain_build.h -- decides which consecutive plan steps are fused into one chained launch
// (chain_kernel.cuh) and proves, per plan, that the fusion is legal: every access of group g
// to a buffer that the chain writes must stay inside an address interval owned by group g.
// The proof is a brute-force walk over all tiles at plan time -- cheap, and independent of how
// the geometry was produced.
#pragma once
#include <map>

#include "chain_kernel.cuh"
#include "plan_struct.h"

namespace b2 {

struct Chain {
    int first = 0, count = 0;
    int threads = 0, lag = 1;
    size_t smem = 0;
    long long G = 0, bpg = 0;
    long long tpg[B2_CHAIN_MAX] = {0, 0, 0};
    unsigned *counters = nullptr;  // [count][G] + 1 error flag
};

struct Interval {
    long long lo = (1ll << 62), hi = -(1ll << 62);
    void add(long long a, long long b) {
        lo = std::min(lo, a);
        hi = std::max(hi, b);
    }
};

// [lo, hi) in reals touched by tile `bid` of a step, on its source and destination buffers
inline void tile_footprint(const StepPlan &s, long long bid, Interval &src, Interval &dst) {
    if (s.kind == SK_PASS) {
        const PassParams &p = s.pp;
        const long long t0 = bid % p.tiles0;
        long long rest = bid / p.tiles0;
        const long long l0 = t0 * p.C;
        const long long nl = std::min<long long>(p.C, p.ln[0] - l0);
        long long sb = 0, db = 0;
        for (int d = 1; d < p.nld; ++d) {
            const long long c = rest % p.ln[d];
            rest /= p.ln[d];
            sb += c * p.ls[d];
            db += c * p.ld[d];
        }
        src.add(sb + l0 * p.ls[0], sb + (l0 + nl - 1) * p.ls[0] + (p.nload - 1) * p.s_axis + p.s_pair + 1);
        dst.add(db + l0 * p.ld[0], db + (l0 + nl - 1) * p.ld[0] + (p.nstore - 1) * p.d_axis + p.d_pair + 1);
    } else {
        const RealPPParams &r = s.rp;
        const long long chunks = (r.m / 2 + 1 + r.pts_per_block - 1) / r.pts_per_block;
        long long rest = bid / chunks;
        long long sb = 0, db = 0;
        for (int d = 0; d < r.nld; ++d) {
            const long long c = rest % r.ln[d];
            rest /= r.ln[d];
            sb += c * r.ls[d];
            db += c * r.ld[d];
        }
        src.add(sb, sb + (long long)r.m * r.s_axis + 2);  // pairs (k, m-k): the whole row
        dst.add(db, db + (long long)r.m * r.d_axis + 2);
    }
}

inline bool chain_groups_ok(const std::vector<StepPlan> &steps, int first, int count, long long G) {
    std::map<int, std::vector<Interval>> hull;  // buffer id -> per-group interval
    std::map<int, bool> written;
    for (int k = 0; k < count; ++k) written[steps[first + k].dst] = true;
    for (int k = 0; k < count; ++k) {
        const StepPlan &s = steps[first + k];
        if (s.nblocks % G) return false;
        const long long tpg = s.nblocks / G;
        for (long long bid = 0; bid < s.nblocks; ++bid) {
            const long long g = bid / tpg;
            Interval a, b;
            tile_footprint(s, bid, a, b);
            if (written.count(s.src)) {
                auto &v = hull[s.src];
                if (v.empty()) v.resize((size_t)G);
                v[(size_t)g].add(a.lo, a.hi);
            }
            auto &w = hull[s.dst];
            if (w.empty()) w.resize((size_t)G);
            w[(size_t)g].add(b.lo, b.hi);
        }
    }
    for (auto &kv : hull) {
        const std::vector<Interval> &v = kv.second;
        for (size_t g = 0; g + 1 < v.size(); ++g)
            if (v[g].hi > v[g + 1].lo) return false;  // group regions must be disjoint and ordered
    }
    return true;
}

// candidate group counts of a step: products of its slowest-varying dims, and splits of the tile dim
inline void chain_candidates(const StepPlan &s, std::vector<long long> &out) {
    if (s.kind == SK_PASS) {
        const PassParams &p = s.pp;
        long long prod = 1;
        for (int d = p.nld - 1; d >= 1; --d) {
            prod *= p.ln[d];
            out.push_back(prod);
        }
        for (long long q = 1; q <= p.tiles0; q *= 2)
            if (p.tiles0 % q == 0) out.push_back(prod * (p.tiles0 / q));
    } else {
        long long prod = 1;
        for (int d = s.rp.nld - 1; d >= 0; --d) {
            prod *= s.rp.ln[d];
            out.push_back(prod);
        }
    }
}

inline bool chainable_step(const StepPlan &s) {
    if (!s.outer.empty()) return false;
    if (s.kind == SK_PASS)  // (the chain kernel dispatches over the power-of-two list only: no mixed-radix rows)
        return s.fast_id >= 0 && !s.needs_tile_kernel && b2_tile_table()[s.fast_id].epi <= 1 && b2_tile_table()[s.fast_id].pipe != 3;
    return s.kind == SK_RPOST || s.kind == SK_RPRE;
}

// Re-tile a row-mapped step so that its CTA has `threads` threads.
inline void retile_rows(StepPlan &s, int threads) {
    if (s.kind != SK_PASS) return;
    const TileEntry &e = b2_tile_table()[s.fast_id];
    if (e.mode == MODE_COL) return;
    const int T = e.threads / e.c;
    int C = std::max(1, threads / T);
    while (C > 1 && e.smem_for(C) > (size_t)max_smem_optin() - 1024) C /= 2;
    if (e.mode == MODE_ROW) {
        // whole tiles only: a ragged tile dim is fine for one launch, but groups need C | lines
        while (C > 1 && s.pp.ln[0] % C) C /= 2;
    }
    PassParams &p = s.pp;
    p.C = C;
    p.tiles0 = (p.ln[0] + C - 1) / C;
    long long nb = p.tiles0;
    for (int d = 1; d < p.nld; ++d) nb *= p.ln[d];
    s.nblocks = nb;
    s.nthreads = C * T;
    s.smem = e.smem_for(C);
}

inline void build_chains(std::vector<StepPlan> &steps, std::vector<Chain> &chains) {
    chains.clear();
    // Opt-in (FFTW_B200_CHAIN=1).  Measured on B200 (profiles/r1_chain_512cube.md): the chained
    // z+y passes of 512^3 move half the DRAM bytes of the two separate launches (1.08 + 1.03 GB
    // instead of 4.3 GB) but run at 2 CTAs x 512 threads per SM, phase-locked on their barriers,
    // and take 1.08 ms against 0.83 ms for the two launches: the tile kernels are issue-bound
    // before they are DRAM-bound at that occupancy.  Kept for the TMA-fed variant.
    const char *on = getenv("FFTW_B200_CHAIN");
    if (!on || atoi(on) == 0 || getenv("FFTW_B200_NO_CHAIN")) return;
    const int ns = (int)steps.size();
    for (int i = 0; i < ns;) {
        int best = 0;
        Chain bc;
        std::vector<StepPlan> best_steps;
        for (int cnt = std::min(B2_CHAIN_MAX, ns - i); cnt >= 2 && !best; --cnt) {
            bool ok = true;
            for (int k = 0; k < cnt && ok; ++k) {
                ok = chainable_step(steps[i + k]);
                if (k > 0) ok = ok && steps[i + k].src == steps[i + k - 1].dst;
            }
            if (!ok) continue;
            std::vector<StepPlan> trial(steps.begin() + i, steps.begin() + i + cnt);
            int threads = 0;
            for (const StepPlan &s : trial)
                if (s.kind == SK_PASS) threads = std::max(threads, s.nthreads);
            if (threads == 0) continue;
            for (StepPlan &s : trial) retile_rows(s, threads);
            size_t smem = 0;
            for (const StepPlan &s : trial) {
                threads = std::max(threads, s.kind == SK_PASS ? s.nthreads : 0);
                smem = std::max(smem, s.smem);
            }
            std::vector<long long> cand;
            for (const StepPlan &s : trial) chain_candidates(s, cand);
            std::sort(cand.begin(), cand.end());
            cand.erase(std::unique(cand.begin(), cand.end()), cand.end());
            long long total = 0;
            for (const StepPlan &s : trial) total += s.nblocks;
            // prefer the finest grouping that still gives a slot a few dozen CTAs
            long long chosen = 0;
            for (auto it = cand.rbegin(); it != cand.rend(); ++it) {
                const long long G = *it;
                if (G < 2 || total / G < 8) continue;
                if (chain_groups_ok(trial, 0, cnt, G)) {
                    chosen = G;
                    break;
                }
            }
            if (!chosen) continue;
            best = cnt;
            bc.first = i;
            bc.count = cnt;
            bc.threads = threads;
            bc.smem = smem;
            bc.G = chosen;
            bc.bpg = 0;
            for (int k = 0; k < cnt; ++k) {
                bc.tpg[k] = trial[k].nblocks / chosen;
                bc.bpg += bc.tpg[k];
            }
            // producers should have left the SMs before their consumers are dispatched
            const int occ = std::max(1, std::min((int)(2048 / threads), (int)((size_t)max_smem_optin() / std::max<size_t>(smem, 1))));
            bc.lag = (int)std::max<long long>(1, (2ll * sm_count() * occ + bc.bpg - 1) / bc.bpg);
            bc.lag = std::min(bc.lag, 16);
            if (const char *e = getenv("FFTW_B200_CHAIN_LAG")) bc.lag = std::max(1, atoi(e));
            best_steps = trial;
        }
        if (best) {
            for (int k = 0; k < best; ++k) {
                steps[i + k] = best_steps[k];
                steps[i + k].chain_id = (int)chains.size();
                steps[i + k].chain_pos = k;
            }
            chains.push_back(bc);
            i += best;
        } else
            ++i;
    }
}

}  // namespace b2
