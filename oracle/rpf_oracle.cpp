// oracle/rpf_oracle.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see rpf_oracle.h).
//
// Plain C++ restatement of the reference's RPF sample filter. Every function cites the
// reference lines it follows (paths relative to /root/reference/src/custom/).
// Arithmetic is FP64 in the reference's operation order; compile WITHOUT -ffast-math and
// without FMA contraction (-ffp-contract=off) so sums and products round like the
// reference's x86-64 -O3 build.
#include "rpf_oracle.h"
#include "../include/rpf_subsample.h"

#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

namespace {

// A sample record is position (2) | colour (3) | NR random parameters | NF features. The reference's SampleData is NR = 2
// (pLens.xy), NF = 12 (n0, p0, n1, p1): sd.h:40-46,62-94. Every loop below runs over the layout's counts, so a wider
// record (time and light samples as random parameters, a texture colour as features) is filtered by the
// reference's formulas with SD_N_RANDOM / SD_N_FEATURES lifted -- there is no reference output for those, the default
// layout is the one pinned against oracle/_ref.
constexpr int NC = 3, NP = 2;
constexpr int CH_POS = 0, CH_COL = 2, CH_RND = 5;
constexpr int MAX_CH = 32, MAX_F = 27;
struct Layout {
    int NR = 2, NF = 12;
    int NCH = 19, CH_FEAT = 7;
    Layout() = default;
    Layout(int nr, int nf) : NR(nr > 0 ? nr : 2), NF(nf > 0 ? nf : 12) {
        NCH = NP + NC + NR + NF;
        CH_FEAT = CH_RND + NR;
    }
    int pairs() const { return NF * (NR + NP) + NC * (NR + NP + NF); }
};

// mi.cpp:53-58: bins default to max(1,(int)sqrt(N))
inline int bins_for(int n) {
    int b = static_cast<int>(std::sqrt(static_cast<double>(n)));
    return b < 1 ? 1 : b;
}

// mi.cpp:47-50 (std::min_element / max_element semantics: strict '<', first extreme wins)
inline void min_max(const double *v, int n, double &lo, double &hi) {
    lo = v[0];
    hi = v[0];
    for (int i = 1; i < n; ++i) {
        if (v[i] < lo) lo = v[i];
        if (hi < v[i]) hi = v[i];
    }
}

// mi.cpp:13-17 and mi.cpp:27-38: bin index of one value
inline int bin_of(double v, double lo, double hi, int bins) {
    if (hi == lo) return 0;
    int b = static_cast<int>((v - lo) / (hi - lo) * bins);
    if (b > bins - 1) b = bins - 1;
    if (b < 0) b = 0;
    return b;
}

// mi.cpp:65-87 given precomputed bin indices and marginal counts.
// joint is scratch of bins*bins ints.
double mi_from_bins(const int *bx, const int *by, const int *hx, const int *hy, int n, int bins,
                    std::vector<int> &joint) {
    joint.assign(static_cast<size_t>(bins) * bins, 0);
    for (int i = 0; i < n; ++i) joint[static_cast<size_t>(bx[i]) * bins + by[i]]++;
    const double total = static_cast<double>(n);
    double mi = 0.0;
    for (int i = 0; i < bins; ++i) {
        const double px = hx[i] / total;
        for (int j = 0; j < bins; ++j) {
            const double pxy = joint[static_cast<size_t>(i) * bins + j] / total;
            const double pp = px * (hy[j] / total);
            if (pxy > 0 && pp != 0) mi += pxy * std::log(pxy / pp);
        }
    }
    return mi;
}

struct ChannelBins {
    std::vector<int> bin;  // [19][n]
    std::vector<int> hist; // [19][bins]
};

// mi.cpp:5-21 per channel, once per neighbourhood (the reference redoes this inside each of the
// 96 MutualInformation calls with identical results).
void bin_channels(const double *z /*[n][NCH]*/, int n, int bins, ChannelBins &cb, const Layout &ly) {
    const int NCH = ly.NCH;
    cb.bin.assign(static_cast<size_t>(NCH) * n, 0);
    cb.hist.assign(static_cast<size_t>(NCH) * bins, 0);
    std::vector<double> col(n);
    for (int c = 0; c < NCH; ++c) {
        for (int i = 0; i < n; ++i) col[i] = z[static_cast<size_t>(i) * NCH + c];
        double lo, hi;
        min_max(col.data(), n, lo, hi);
        int *b = &cb.bin[static_cast<size_t>(c) * n];
        int *h = &cb.hist[static_cast<size_t>(c) * bins];
        for (int i = 0; i < n; ++i) {
            b[i] = bin_of(col[i], lo, hi, bins);
            h[b[i]]++;
        }
    }
}

// rpf.cpp:356-488. mi_log (optional, 96 doubles): every MutualInformation value in call order, which is
// feature i against r0, r1, p0, p1 at 4*i + {0..3} (rpf.cpp:416-427), then colour i against r0, r1, p0, p1, f0..f11 at
// 48 + 16*i + {0..15} (rpf.cpp:429-442).
void cf_weights(const double *z, int n, int flags, double alpha[NC], double *beta /*[NF]*/, double &wrc,
                ChannelBins &cb, std::vector<int> &joint, double eps = 0.0, double *mi_log = nullptr, const Layout &ly = Layout()) {
    const int NR = ly.NR, NF = ly.NF, CH_FEAT = ly.CH_FEAT;
    const int bins = bins_for(n);
    bin_channels(z, n, bins, cb, ly);
    int calls = 0;
    auto MI = [&](int ca, int cbn) {
        const double v = mi_from_bins(&cb.bin[static_cast<size_t>(ca) * n], &cb.bin[static_cast<size_t>(cbn) * n],
                                      &cb.hist[static_cast<size_t>(ca) * bins], &cb.hist[static_cast<size_t>(cbn) * bins],
                                      n, bins, joint);
        if (mi_log) mi_log[calls] = v;
        ++calls;
        return v;
    };
    double D_r_f[MAX_F], D_p_f[MAX_F], D_c_f[MAX_F];
    double D_r_c[NC], D_p_c[NC], D_f_c[NC];
    for (int k = 0; k < NF; ++k) D_r_f[k] = D_p_f[k] = D_c_f[k] = 0;
    for (int k = 0; k < NC; ++k) D_r_c[k] = D_p_c[k] = D_f_c[k] = 0;
    // rpf.cpp:416-427
    for (int i = 0; i < NF; ++i) {
        for (int j = 0; j < NR; ++j) D_r_f[i] += MI(CH_FEAT + i, CH_RND + j);
        for (int j = 0; j < NP; ++j) D_p_f[i] += MI(CH_FEAT + i, CH_POS + j);
    }
    // rpf.cpp:429-442
    for (int i = 0; i < NC; ++i) {
        for (int j = 0; j < NR; ++j) D_r_c[i] += MI(CH_COL + i, CH_RND + j);
        for (int j = 0; j < NP; ++j) D_p_c[i] += MI(CH_COL + i, CH_POS + j);
        for (int j = 0; j < NF; ++j) {
            const double m = MI(CH_COL + i, CH_FEAT + j);
            D_f_c[i] += m;
            D_c_f[j] += m;  // P1 accumulation axis
        }
    }
    // rpf.cpp:449-456
    double Dfc = 0, Drc = 0, Dpc = 0;
    for (int i = 0; i < NC; ++i) {
        Dfc += D_f_c[i];
        Drc += D_r_c[i];
        Dpc += D_p_c[i];
    }
    const bool p1 = flags & RPF_ORACLE_P1, p2 = flags & RPF_ORACLE_P2;
    auto ratio = [&](double num, double den) {
        if (eps != 0) den = den + eps;  // paper mode: epsilon in the denominators
        return (p2 && den == 0) ? 0.0 : num / den;
    };
    // rpf.cpp:461-466
    double W_c_f[MAX_F], W_r_f[MAX_F], W_r_c[NC];
    for (int i = 0; i < NF; ++i) {
        // verbatim reads D_f_ck[i], in bounds only for i<3; beyond that the reference is undefined
        // and this restatement reports NaN in verbatim mode.
        double num = p1 ? D_c_f[i] : (i < NC ? D_f_c[i] : std::nan(""));
        W_c_f[i] = ratio(num, Dfc + Drc + Dpc);
        W_r_f[i] = ratio(D_r_f[i], D_r_f[i] + D_p_f[i]);
    }
    for (int i = 0; i < NC; ++i) W_r_c[i] = ratio(D_r_c[i], D_r_c[i] + D_p_c[i]);  // rpf.cpp:470
    for (int i = 0; i < NC; ++i) alpha[i] = 1 - W_r_c[i];                          // rpf.cpp:475
    for (int i = 0; i < NF; ++i) beta[i] = (1 - W_r_f[i]) * W_c_f[i];              // rpf.cpp:479
    wrc = 0;                                                                       // rpf.cpp:483-487
    for (int i = 0; i < NC; ++i) wrc += W_r_c[i];
    wrc /= NC;
}

}  // namespace

namespace {

struct Frame {
    int W, H, P;
    const float *planes;  // [NCH][S]
    size_t S;
    const double *col_in; // [3][S] current iteration's input colours
    Layout ly;
    double value(int c, size_t idx) const {
        if (c >= CH_COL && c < CH_COL + NC) return col_in[static_cast<size_t>(c - CH_COL) * S + idx];
        return static_cast<double>(planes[static_cast<size_t>(c) * S + idx]);
    }
};

struct PixelScratch {
    std::vector<size_t> nbh;      // accepted sample indices in reference order
    std::vector<double> raw;      // [n][NCH]
    std::vector<double> z;        // [n][NCH]
    std::vector<double> w;        // [P][n]
    ChannelBins cb;
    std::vector<int> joint;
    double *mi_log = nullptr;     // rpf_oracle_pixel_probe: the 96 MI values of this pixel
};

// One pixel of rpf.cpp:556-723. Returns number of NaN colour channels written.
struct Subsample {
    uint32_t keep_q24 = RPF_SUBSAMPLE_ALL;
    uint64_t seed = 0;
    int row0 = 0;
};

// Paper mode (not in the reference): all zero = reference semantics
struct Paper {
    double w_eps = 0, min_sigma = 0, pos_sigmas = 0;
};

int filter_pixel(const Frame &f, int px, int py, int box, double seed, int flags, const Subsample &sub, const Paper &paper, double *col_out,
                 PixelScratch &s, int *n_out, double *alpha_out, double *beta_out, double *wrc_out) {
    const int P = f.P, W = f.W, H = f.H;
    const Layout &ly = f.ly;
    const int NF = ly.NF, NCH = ly.NCH, CH_FEAT = ly.CH_FEAT;
    const size_t base = (static_cast<size_t>(py) * W + px) * P;
    // rpf.cpp:338-347 (FillMeanAndStddev) with ops.h:111-144: sequential sums, population sigma, no clamp
    double mu[MAX_F], sd3[MAX_F];
    for (int k = 0; k < NF; ++k) {
        double sum = 0, sq = 0;
        for (int i = 0; i < P; ++i) {
            const double v = f.value(CH_FEAT + k, base + i);
            sum = sum + v;
            sq = sq + v * v;
        }
        mu[k] = sum / static_cast<double>(P);
        const double sd = std::sqrt(sq / static_cast<double>(P) - mu[k] * mu[k]);
        sd3[k] = sd * 3;  // rpf.cpp:579 multiplyArray(stdDev, 3)
        const bool world_pos = (k >= 3 && k < 6) || (k >= 9 && k < 12);  // p0, p1 of sd.h:62-94
        if (world_pos && paper.pos_sigmas > 0) sd3[k] = sd * paper.pos_sigmas;
        if (paper.min_sigma > 0 && sd3[k] < paper.min_sigma) sd3[k] = paper.min_sigma;  // a NaN sigma stays NaN (accepts)
    }
    // rpf.cpp:558-586: own samples, then xn ascending, yn ascending, sample order
    s.nbh.clear();
    for (int i = 0; i < P; ++i) s.nbh.push_back(base + i);
    const int bd = (box - 1) / 2;  // rpf.cpp:561
    const uint64_t ckey = rpf_subsample_centre_key(sub.seed, box, px, py + sub.row0);
    for (int xn = px - bd; xn <= px + bd; ++xn) {
        for (int yn = py - bd; yn <= py + bd; ++yn) {
            if (xn == px && yn == py) continue;
            if (xn < 0 || xn >= W || yn < 0 || yn >= H) continue;  // rpf.cpp:569
            const size_t nb = (static_cast<size_t>(yn) * W + xn) * P;
            for (int i = 0; i < P; ++i) {
                bool ok = true;  // ops.h:100-107 allLessThan: reject on a >= b (NaN passes)
                for (int k = 0; k < NF; ++k) {
                    const double d = std::fabs(f.value(CH_FEAT + k, nb + i) - mu[k]);
                    if (d >= sd3[k]) { ok = false; break; }
                }
                // optional neighbour subsampling (not in the reference): shared counter-based RNG, global coordinates
                if (ok && sub.keep_q24 < RPF_SUBSAMPLE_ALL) ok = rpf_subsample_keep(ckey, sub.keep_q24, xn, yn + sub.row0, i);
                if (ok) s.nbh.push_back(nb + i);
            }
        }
    }
    const int n = static_cast<int>(s.nbh.size());
    // rpf.cpp:596-601: mean / sigma of all 19 channels over the neighbourhood
    s.raw.resize(static_cast<size_t>(n) * NCH);
    for (int j = 0; j < n; ++j)
        for (int c = 0; c < NCH; ++c) s.raw[static_cast<size_t>(j) * NCH + c] = f.value(c, s.nbh[j]);
    double M[MAX_CH], SD[MAX_CH];
    for (int c = 0; c < NCH; ++c) {
        double sum = 0, sq = 0;
        for (int j = 0; j < n; ++j) {
            const double v = s.raw[static_cast<size_t>(j) * NCH + c];
            sum = sum + v;
            sq = sq + v * v;
        }
        M[c] = sum / static_cast<double>(n);
        double var = sq / static_cast<double>(n) - M[c] * M[c];
        if ((flags & RPF_ORACLE_P3) && !(var > 0)) var = 0.0;  // sqrt(std::max(0.0, var)): a NaN variance gives 0 as well
        SD[c] = std::sqrt(var);
    }
    // rpf.cpp:603-612, sd.h:224-235, ops.h:44-51. The own samples are the first P entries.
    s.z.resize(static_cast<size_t>(n) * NCH);
    for (int j = 0; j < n; ++j)
        for (int c = 0; c < NCH; ++c) {
            const size_t o = static_cast<size_t>(j) * NCH + c;
            s.z[o] = SD[c] == 0 ? 0 : (s.raw[o] - M[c]) / SD[c];
        }
    double alpha[NC], beta[MAX_F], wrc;
    cf_weights(s.z.data(), n, flags, alpha, beta, wrc, s.cb, s.joint, paper.w_eps, s.mi_log, ly);
    if (n_out) *n_out = n;
    if (alpha_out) std::memcpy(alpha_out, alpha, sizeof alpha);
    if (beta_out) std::memcpy(beta_out, beta, NF * sizeof(double));
    if (wrc_out) *wrc_out = wrc;
    // rpf.cpp:531,662-664
    const double sigma_p = static_cast<double>(box / 4);  // integer division
    const double s2 = seed * seed / (1 - wrc) / (1 - wrc);
    const double sp2 = sigma_p * sigma_p;
    s.w.resize(static_cast<size_t>(P) * n);
    for (int i = 0; i < P; ++i) {
        const double *zi = &s.z[static_cast<size_t>(i) * NCH];
        for (int j = 0; j < n; ++j) {
            const double *zj = &s.z[static_cast<size_t>(j) * NCH];
            double sp = 0, sc = 0, sf = 0;  // sumArray: starts at 0, adds in index order
            for (int k = 0; k < NP; ++k) { const double d = zi[CH_POS + k] - zj[CH_POS + k]; sp += d * d; }
            for (int k = 0; k < NC; ++k) { const double d = zi[CH_COL + k] - zj[CH_COL + k]; sc += (d * d) * alpha[k]; }
            for (int k = 0; k < NF; ++k) { const double d = zi[CH_FEAT + k] - zj[CH_FEAT + k]; sf += (d * d) * beta[k]; }
            s.w[static_cast<size_t>(i) * n + j] =
                std::exp(-sp / (2 * sp2)) * std::exp(-sc / (2 * s2)) * std::exp(-sf / (2 * s2));  // rpf.cpp:667-670
        }
    }
    // rpf.cpp:682-717: blend with UN-normalised neighbour colours
    int nans = 0;
    for (int i = 0; i < P; ++i)
        for (int k = 0; k < NC; ++k) {
            double sw = 0, swc = 0;
            for (int j = 0; j < n; ++j) {
                const double wij = s.w[static_cast<size_t>(i) * n + j];
                sw += wij;
                swc += wij * s.raw[static_cast<size_t>(j) * NCH + CH_COL + k];
            }
            const double c = swc / sw;
            if (c != c) ++nans;  // P4: count instead of exit(1) (rpf.cpp:702-705)
            col_out[static_cast<size_t>(k) * f.S + base + i] = c;
        }
    return nans;
}

}  // namespace

extern "C" int rpf_oracle_filter(rpf_oracle_args *a) {
    if (!a || !a->planes || a->width <= 0 || a->height <= 0 || a->spp <= 0 || a->n_iter < 0) return -1;
    const int W = a->width, H = a->height, P = a->spp;
    const size_t S = static_cast<size_t>(W) * H * P;
    std::vector<double> cur(3 * S), nxt(3 * S);
    if (a->colors) std::memcpy(cur.data(), a->colors, 3 * S * sizeof(double));
    else
        for (int k = 0; k < 3; ++k)
            for (size_t i = 0; i < S; ++i) cur[k * S + i] = a->planes[(CH_COL + k) * S + i];
    const int y0 = a->y0 > 0 ? a->y0 : 0, y1 = a->y1 > 0 && a->y1 < H ? a->y1 : H;
    const int x0 = a->x0 > 0 ? a->x0 : 0, x1 = a->x1 > 0 && a->x1 < W ? a->x1 : W;
    int nt = a->n_threads > 0 ? a->n_threads : static_cast<int>(std::thread::hardware_concurrency());
    if (nt < 1) nt = 1;
    std::atomic<long long> nan_total{0};
    Subsample sub;
    if (a->subsample_keep_q24 > 0 && a->subsample_keep_q24 < RPF_SUBSAMPLE_ALL) {
        sub.keep_q24 = a->subsample_keep_q24;
        sub.seed = a->subsample_seed;
        sub.row0 = a->global_row0;
    }
    Paper paper;
    paper.w_eps = a->w_epsilon;
    paper.min_sigma = a->min_sigma;
    paper.pos_sigmas = a->world_pos_sigmas;
    for (int it = 0; it < a->n_iter; ++it) {
        const int box = a->box_sizes[it];
        const double seed = a->sigma_seeds ? a->sigma_seeds[it] : a->sigma_fc_seed;
        const bool last = it == a->n_iter - 1;
        nxt = cur;  // pixels outside the filtered window keep their colours
        Frame f{W, H, P, a->planes, S, cur.data(), Layout(a->n_random, a->n_features)};
        // 16x16 tiles handed out dynamically, like ParallelFor2D at rpf.cpp:542
        const int T = 16, tx = (x1 - x0 + T - 1) / T, ty = (y1 - y0 + T - 1) / T;
        std::atomic<int> next{0};
        auto worker = [&]() {
            PixelScratch s;
            long long nans = 0;
            for (;;) {
                const int t = next.fetch_add(1);
                if (t >= tx * ty) break;
                const int bx = x0 + (t % tx) * T, by = y0 + (t / tx) * T;
                for (int x = bx; x < bx + T && x < x1; ++x)
                    for (int y = by; y < by + T && y < y1; ++y) {
                        const size_t p = static_cast<size_t>(y) * W + x;
                        nans += filter_pixel(f, x, y, box, seed, a->flags, sub, paper, nxt.data(), s,
                                             last && a->nbh_size ? a->nbh_size + p : nullptr,
                                             last && a->alpha ? a->alpha + p * 3 : nullptr,
                                             last && a->beta ? a->beta + p * Layout(a->n_random, a->n_features).NF : nullptr,
                                             last && a->wrc ? a->wrc + p : nullptr);
                    }
            }
            nan_total += nans;
        };
        std::vector<std::thread> th;
        for (int i = 1; i < nt; ++i) th.emplace_back(worker);
        worker();
        for (auto &t : th) t.join();
        cur.swap(nxt);  // rpf.cpp:732
    }
    if (a->hdr_clamp_sigmas > 0 && a->n_iter > 0) rpf_oracle_hdr_clamp(cur.data(), W, H, P, y0, y1, a->hdr_clamp_sigmas);
    if (a->colors) std::memcpy(a->colors, cur.data(), 3 * S * sizeof(double));
    a->nan_count = nan_total.load();
    return 0;
}

extern "C" void rpf_oracle_hdr_clamp(double *colors, int width, int height, int spp, int y0, int y1, double k_sigmas) {
    const size_t S = static_cast<size_t>(width) * height * spp;
    for (int y = y0; y < y1; ++y)
        for (int x = 0; x < width; ++x) {
            const size_t base = (static_cast<size_t>(y) * width + x) * spp;
            double mu[NC], lim[NC];
            for (int k = 0; k < NC; ++k) {
                const double *c = colors + k * S + base;
                double sum = 0, sq = 0;
                for (int i = 0; i < spp; ++i) {
                    sum = sum + c[i];
                    sq = sq + c[i] * c[i];
                }
                mu[k] = sum / static_cast<double>(spp);
                const double var = sq / static_cast<double>(spp) - mu[k] * mu[k];
                lim[k] = k_sigmas * std::sqrt(var > 0 ? var : 0.0);
            }
            for (int i = 0; i < spp; ++i) {
                bool spike = false;
                for (int k = 0; k < NC; ++k)
                    if (std::fabs(colors[k * S + base + i] - mu[k]) > lim[k]) spike = true;
                if (spike)
                    for (int k = 0; k < NC; ++k) colors[k * S + base + i] = mu[k];
            }
        }
}

extern "C" double rpf_oracle_mi(const double *x, const double *y, int n) {
    const int bins = bins_for(n);
    double lx, hx, ly, hy;
    min_max(x, n, lx, hx);
    min_max(y, n, ly, hy);
    std::vector<int> bx(n), by(n), cx(bins, 0), cy(bins, 0), joint;
    for (int i = 0; i < n; ++i) {
        bx[i] = bin_of(x[i], lx, hx, bins);
        by[i] = bin_of(y[i], ly, hy, bins);
        cx[bx[i]]++;
        cy[by[i]]++;
    }
    return mi_from_bins(bx.data(), by.data(), cx.data(), cy.data(), n, bins, joint);
}

extern "C" void rpf_oracle_mean_std(const double *v, int n, double *mean, double *sd) {
    double sum = 0, sq = 0;
    for (int i = 0; i < n; ++i) {
        sum = sum + v[i];
        sq = sq + v[i] * v[i];
    }
    *mean = sum / static_cast<double>(n);
    *sd = std::sqrt(sq / static_cast<double>(n) - *mean * *mean);
}

extern "C" void rpf_oracle_cf_weights(const double *z, int n, int flags, double *alpha, double *beta, double *wrc) {
    ChannelBins cb;
    std::vector<int> joint;
    cf_weights(z, n, flags, alpha, beta, *wrc, cb, joint);
}

// One pixel of one pass (box a->box_sizes[0], input colours a->colors or planes[2..4]): the component-level view the GPU
// exposes through rpf_set_debug_pixel / rpf_download_debug_mi.
extern "C" int rpf_oracle_pixel_probe(const rpf_oracle_args *a, int px, int py, double *mi96, int *n_out, int *bins_out, int *marg,
                                      int bins_cap, double *z_out, int n_cap) {
    if (!a || !a->planes || a->n_iter < 1 || px < 0 || py < 0 || px >= a->width || py >= a->height) return -1;
    const int W = a->width, H = a->height, P = a->spp;
    const size_t S = static_cast<size_t>(W) * H * P;
    std::vector<double> cur(3 * S), out(3 * S);
    if (a->colors) std::memcpy(cur.data(), a->colors, 3 * S * sizeof(double));
    else
        for (int k = 0; k < 3; ++k)
            for (size_t i = 0; i < S; ++i) cur[k * S + i] = a->planes[(CH_COL + k) * S + i];
    Frame f{W, H, P, a->planes, S, cur.data(), Layout(a->n_random, a->n_features)};
    PixelScratch s;
    double mi[1024];
    s.mi_log = mi;
    const Layout ly(a->n_random, a->n_features);
    const int NCH = ly.NCH;
    Subsample sub;
    Paper paper;
    paper.w_eps = a->w_epsilon; paper.min_sigma = a->min_sigma; paper.pos_sigmas = a->world_pos_sigmas;
    int n = 0;
    filter_pixel(f, px, py, a->box_sizes[0], a->sigma_fc_seed, a->flags, sub, paper, out.data(), s, &n, nullptr, nullptr, nullptr);
    const int bins = bins_for(n);
    if (mi96) std::memcpy(mi96, mi, ly.pairs() * sizeof(double));
    if (n_out) *n_out = n;
    if (bins_out) *bins_out = bins;
    if (marg) {
        if (bins_cap < bins) return -2;
        for (int c = 0; c < NCH; ++c)
            for (int i = 0; i < bins_cap; ++i) marg[c * bins_cap + i] = i < bins ? s.cb.hist[static_cast<size_t>(c) * bins + i] : 0;
    }
    if (z_out) {
        if (n_cap < n) return -3;
        std::memcpy(z_out, s.z.data(), static_cast<size_t>(n) * NCH * sizeof(double));
    }
    return 0;
}
