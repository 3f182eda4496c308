/*
 * oracle/oracle.c -- CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 *
 * Plain-C restatement of the ensemble-update hot path of jmatta1/MCMCpp, used
 * ONLY as the checker by tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg.  The product path (mcmcpp_b200/csrc) never links, loads or
 * calls anything in this directory.
 *
 * Parity status: PINNED.  This file is checked (tests/test_oracle_*.py)
 *   - against the pcg-cpp upstream known-answer vectors
 *     (MCMCpp/Utility/pcg-cpp/test-high/expected/check-pcg64.out, check-pcg32.out),
 *   - against golden vectors generated in the build container by running the
 *     UNMODIFIED reference headers (oracle/ref_driver.cpp -> oracle/_ref/,
 *     fixtures + generating script under tests/golden/),
 *   - and, when oracle/_ref/ is present, live against the compiled reference.
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference).  Build: gcc -O2 -ffp-contract=off -fPIC -shared (see Makefile);
 * contraction must stay off, the reference trajectory depends on it.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef unsigned __int128 u128;

/* ------------------------------------------------------------------ pcg64 */
/* MCMCpp/Utility/pcg-cpp/include/pcg_random.hpp:161-162 (128-bit default multiplier) */
#define PCG128_MULT ((((u128)2549297995355413924ULL) << 64) | (u128)4865540595714422341ULL)

typedef struct { u128 state; u128 inc; } pcg64_t;

/* pcg_random.hpp:335-339 (set_stream: inc=(stream<<1)|1), :480-487 (seed ctor:
 * state = bump(seed + increment)), :399-402 (bump = state*mult + inc).
 * MCMCpp passes 64-bit seed/stream zero-extended (Utility/MultiSampler.h:54). */
void oracle_pcg64_seed(pcg64_t* g, uint64_t seed, uint64_t stream)
{
    g->inc = (((u128)stream) << 1) | 1u;
    g->state = ((u128)seed + g->inc) * PCG128_MULT + g->inc;
}

/* pcg_random.hpp:984-1011 (xsl_rr_mixin::output), pcg_extras.hpp:303-312 (rotr) */
static inline uint64_t pcg64_output(u128 s)
{
    unsigned rot = (unsigned)(s >> 122);
    uint64_t x = (uint64_t)(s >> 64) ^ (uint64_t)s;
    return (x >> rot) | (x << ((-rot) & 63u));
}

/* pcg_random.hpp:417-423 with output_previous=false for 128-bit state (:742):
 * advance first, then output the NEW state. */
uint64_t oracle_pcg64_next(pcg64_t* g)
{
    g->state = g->state * PCG128_MULT + g->inc;
    return pcg64_output(g->state);
}

/* pcg_random.hpp:622-651 (engine::advance, square-and-multiply over (mult,plus)) */
void oracle_pcg64_advance(pcg64_t* g, u128 delta)
{
    u128 acc_mult = 1, acc_plus = 0, cur_mult = PCG128_MULT, cur_plus = g->inc;
    while (delta > 0) {
        if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
        cur_plus = (cur_mult + 1) * cur_plus;
        cur_mult *= cur_mult;
        delta >>= 1;
    }
    g->state = acc_mult * g->state + acc_plus;
}
void oracle_pcg64_advance_u64(pcg64_t* g, uint64_t delta) { oracle_pcg64_advance(g, (u128)delta); }
/* backstep(n) == advance(2^128 - n): pcg_random.hpp:445-448 */
void oracle_pcg64_backstep_u64(pcg64_t* g, uint64_t delta) { oracle_pcg64_advance(g, (u128)0 - (u128)delta); }

/* pcg_extras.hpp:514-525 (bounded_rand); *rejections counts redraws */
uint64_t oracle_pcg64_bounded(pcg64_t* g, uint64_t bound, int* rejections)
{
    uint64_t threshold = (0 - bound) % bound;
    for (;;) {
        uint64_t r = oracle_pcg64_next(g);
        if (r >= threshold) return r % bound;
        if (rejections) ++*rejections;
    }
}

/* ------------------------------------------------------------------ pcg32 */
/* pcg_random.hpp:158-159 (64-bit multiplier / default increment), :470-475
 * (one-arg ctor), :808-836 (xsh_rr output on the OLD state). Used by the
 * reference only for test-side initial positions and AutoCorrCalc subsampling. */
typedef struct { uint64_t state; uint64_t inc; } pcg32_t;
#define PCG64_MULT 6364136223846793005ULL
#define PCG64_DEFAULT_INC 1442695040888963407ULL
void oracle_pcg32_seed(pcg32_t* g, uint64_t seed) /* pcg32 eng(seed) */
{
    g->inc = PCG64_DEFAULT_INC;
    g->state = (seed + g->inc) * PCG64_MULT + g->inc;
}
void oracle_pcg32_seed_stream(pcg32_t* g, uint64_t seed, uint64_t stream) /* pcg32 eng(seed, stream) */
{
    g->inc = (stream << 1) | 1u;
    g->state = (seed + g->inc) * PCG64_MULT + g->inc;
}
uint32_t oracle_pcg32_next(pcg32_t* g)
{
    uint64_t old = g->state;
    g->state = old * PCG64_MULT + g->inc;
    unsigned rot = (unsigned)(old >> 59);
    uint32_t x = (uint32_t)(((old >> 18) ^ old) >> 27);
    return (x >> rot) | (x << ((-rot) & 31u));
}

/* ------------------------------------------------- libstdc++ distributions */
/* std::generate_canonical<double,53>(pcg64) (libstdc++ 13 bits/random.tcc:3346-3381):
 * one 64-bit draw, sum = double(r) (RNE), ret = sum / 2^64, clamp to nextafter(1,0).
 * std::uniform_real_distribution<double>(0,1) is the identity on it
 * (Utility/MultiSampler.h:121, bits/random.h:1779+). */
double oracle_u64_to_unit(uint64_t r)
{
    double u = (double)r * 0x1.0p-64;
    if (u >= 1.0) u = 0x1.fffffffffffffp-1;
    return u;
}

/* Utility/GwDistribution.h:45-58: z = (Term1*u + InvSqrtAlpha)^2 with
 * Alpha = num/denom, SqrtAlpha = sqrt(Alpha), InvSqrtAlpha = 1/sqrt(Alpha),
 * Term1 = SqrtAlpha - InvSqrtAlpha. */
void oracle_gw_constants(int alpha_num, int alpha_denom, double* term1, double* inv_sqrt_alpha)
{
    double alpha = (double)alpha_num / (double)alpha_denom;
    double sa = sqrt(alpha);
    double isa = 1.0 / sqrt(alpha);
    *inv_sqrt_alpha = isa;
    *term1 = sa - isa;
}
double oracle_gw_sample(double u, double term1, double inv_sqrt_alpha)
{
    double t = term1 * u + inv_sqrt_alpha;
    return t * t;
}
/* Utility/MultiSampler.h:86 getNegExponentialReal = -expDist(engine);
 * std::exponential_distribution (bits/random.h:4897-4905): -log(1-u)/lambda, lambda=1 */
double oracle_neg_exponential(double u) { return -(-log(1.0 - u) / 1.0); }

/* ---------------------------------------------------------------- targets */
/* Target kinds.  Kind 0 is the reference's only in-repo Calculator; kinds 1-4
 * are the synthetic BASELINE.json configs; their arithmetic
 * order is DEFINED here and in include/mcmcpp_b200/targets.h identically. */
enum { T_SKEWED_GAUSS_2D = 0, T_ROSENBROCK_2D = 1, T_DIAG_GAUSS = 2, T_DENSE_GAUSS = 3, T_LINREG = 4 };

/* params: kind 0: {eps}; 1: {}; 2: iv[D]; 3: P[D*D] row-major; 4: {M, prior_var, X[M*D] row-major, y[M]} */

/* test/sequential/SkewedGaussian/Common/SkewedGaussian.h:52-57 */
static double logp_skewed(const double* x, double eps)
{
    double t1 = x[0] / 2.0 - x[1];
    double t2 = x[0] / 2.0 + x[1];
    return (((t1 * t1) / eps) + (t2 * t2)) / -2.0;
}
/* -(100 (y-x^2)^2 + (1-x)^2)/20 */
static double logp_rosenbrock(const double* x)
{
    double t1 = x[1] - x[0] * x[0];
    double t2 = 1.0 - x[0];
    return ((100.0 * (t1 * t1)) + (t2 * t2)) / -20.0;
}
/* canonical 64-way strided sum + xor-butterfly tree: term d goes to partial
 * d%64 (in increasing d); s[l]=p[2l]+p[2l+1]; then s[l]+=s[l^off] for
 * off=16,8,4,2,1.  This is the order one warp (lane l owns elements 64k+2l,
 * 64k+2l+1) produces, so host and device agree bit for bit. */
static double canonical64_finish(double* p)
{
    double s[32], t[32];
    for (int l = 0; l < 32; ++l) s[l] = p[2 * l] + p[2 * l + 1];
    for (int off = 16; off >= 1; off >>= 1) {
        for (int l = 0; l < 32; ++l) t[l] = s[l] + s[l ^ off];
        memcpy(s, t, sizeof s);
    }
    return s[0];
}
/* -1/2 sum_d x_d^2 iv_d */
static double logp_diag(const double* x, const double* iv, int D)
{
    double p[64];
    for (int l = 0; l < 64; ++l) p[l] = 0.0;
    for (int d = 0; d < D; ++d) p[d & 63] += (x[d] * x[d]) * iv[d];
    return -0.5 * canonical64_finish(p);
}
/* -1/2 x^T P x, y_i = sum_j P_ij x_j then sum_i x_i y_i */
static double logp_dense(const double* x, const double* P, int D)
{
    double q = 0.0;
    for (int i = 0; i < D; ++i) {
        double y = 0.0;
        for (int j = 0; j < D; ++j) y += P[(size_t)i * D + j] * x[j];
        q += x[i] * y;
    }
    return -0.5 * q;
}
/* -1/2 ||y - X th||^2 - 1/2 ||th||^2 / prior_var */
static double logp_linreg(const double* th, const double* p, int D)
{
    long M = (long)p[0];
    double pv = p[1];
    const double* X = p + 2;
    const double* y = X + (size_t)M * D;
    double ss = 0.0, tt = 0.0;
    for (long m = 0; m < M; ++m) {
        double r = 0.0;
        for (int j = 0; j < D; ++j) r += X[(size_t)m * D + j] * th[j];
        r = y[m] - r;
        ss += r * r;
    }
    for (int j = 0; j < D; ++j) tt += th[j] * th[j];
    return -0.5 * ss - 0.5 * (tt / pv);
}

double oracle_logp(int kind, int D, const double* params, const double* x)
{
    switch (kind) {
    case T_SKEWED_GAUSS_2D: return logp_skewed(x, params[0]);
    case T_ROSENBROCK_2D:   return logp_rosenbrock(x);
    case T_DIAG_GAUSS:      return logp_diag(x, params, D);
    case T_DENSE_GAUSS:     return logp_dense(x, params, D);
    case T_LINREG:          return logp_linreg(x, params, D);
    }
    return NAN;
}

/* --------------------------------------------------------- the hot path */
typedef struct {
    int32_t j;          /* complementary index (StretchMove.h:102) */
    int32_t accept;     /* Metropolis decision (StretchMove.h:113) */
    double z;           /* scalingFactor (StretchMove.h:104) */
    double e;           /* log(1-u) = getNegExponentialReal() (StretchMove.h:113) */
    double newp;        /* calcLogPostProb(proposal) (StretchMove.h:111) */
    double diff;        /* logProbDiff (StretchMove.h:112) */
} oracle_trace_t;

/*
 * Sequential-order restatement of EnsembleSampler::runMCMC / performStep
 * (EnsembleSampler.h:284-310, 341-360) with StretchMove::updateWalker
 * (Movers/StretchMove.h:100-123) and Walker bookkeeping (Walker/Walker.h:105,162-179).
 *
 *  pos   [W*D]  AoS [walker][param]; walkers 0..K-1 red, K..W-1 black (EnsembleSampler.h:225-226)
 *  logp  [W]    aux data (current log posterior)
 *  acc/tot [W]  per-walker accepted / total counters.  The caller mirrors
 *               setInitialWalkerPos by starting them at 1 (Walker.h:76,162-170).
 *  nsteps       number of STORED steps; each is `interval` ensemble steps of which
 *               only the last is stored (EnsembleSampler.h:296-308)
 *  chain [nsteps*W*D] AoS [step][walker][param] or NULL (Chain/ChainBlock.h:125-131)
 *  trace [nsteps*interval*W] in update order, or NULL
 *  rng state is carried in *g so that successive calls continue the stream.
 * Returns the number of bounded_rand rejections seen (shifts the 3-draws/update cadence).
 */
int oracle_stretch_run(pcg64_t* g, int kind, const double* tparams, int W, int D,
                       int alpha_num, int alpha_denom,
                       double* pos, double* logp, int64_t* acc, int64_t* tot,
                       int nsteps, int interval, double* chain, oracle_trace_t* trace)
{
    const int K = W / 2;
    double term1, isa;
    oracle_gw_constants(alpha_num, alpha_denom, &term1, &isa);
    double* prop = (double*)malloc(sizeof(double) * (size_t)D);
    int rejections = 0;
    size_t tr = 0;
    for (int s = 0; s < nsteps; ++s) {
        for (int sub = 0; sub < interval; ++sub) {
            int store = (sub == interval - 1);
            for (int h = 0; h < 2; ++h) {
                const int own0 = h * K, oth0 = (1 - h) * K;
                for (int i = 0; i < K; ++i) {
                    const int w = own0 + i;
                    int j = (int)oracle_pcg64_bounded(g, (uint64_t)K, &rejections);
                    const double* sel = pos + (size_t)(oth0 + j) * D;
                    double* cur = pos + (size_t)w * D;
                    double z = oracle_gw_sample(oracle_u64_to_unit(oracle_pcg64_next(g)), term1, isa);
                    for (int d = 0; d < D; ++d) prop[d] = sel[d] + z * (cur[d] - sel[d]);
                    double scaling = log(z) * (double)(D - 1);
                    double newp = oracle_logp(kind, D, tparams, prop);
                    double diff = scaling + newp - logp[w];
                    double e = oracle_neg_exponential(oracle_u64_to_unit(oracle_pcg64_next(g)));
                    int a = (e < diff);
                    if (a) { memcpy(cur, prop, sizeof(double) * (size_t)D); logp[w] = newp; ++acc[w]; }
                    ++tot[w];
                    if (trace) { oracle_trace_t* t = &trace[tr++]; t->j = j; t->accept = a; t->z = z; t->e = e; t->newp = newp; t->diff = diff; }
                    if (store && chain) memcpy(chain + ((size_t)s * W + w) * D, cur, sizeof(double) * (size_t)D);
                }
            }
        }
    }
    free(prop);
    return rejections;
}

/*
 * PER_WALKER rng discipline: walker w owns pcg64(seed, stream = w), i.e. one reference mover per
 * walker re-seeded with setPrng(seed, w) (Movers/StretchMove.h:90, the per-thread streams of
 * Threading/RedBlkUpdater.h:86 taken to one walker per stream).  Same update arithmetic and the
 * same red-then-black order as oracle_stretch_run; rng [W] carries every walker's engine.
 */
void oracle_pcg64_seed_many(pcg64_t* g, uint64_t seed, int W)
{
    for (int w = 0; w < W; ++w) oracle_pcg64_seed(&g[w], seed, (uint64_t)w);
}

int oracle_stretch_run_per_walker(pcg64_t* rng, int kind, const double* tparams, int W, int D,
                                  int alpha_num, int alpha_denom,
                                  double* pos, double* logp, int64_t* acc, int64_t* tot,
                                  int nsteps, int interval, double* chain, oracle_trace_t* trace)
{
    const int K = W / 2;
    double term1, isa;
    oracle_gw_constants(alpha_num, alpha_denom, &term1, &isa);
    double* prop = (double*)malloc(sizeof(double) * (size_t)D);
    int rejections = 0;
    size_t tr = 0;
    for (int s = 0; s < nsteps; ++s) {
        for (int sub = 0; sub < interval; ++sub) {
            int store = (sub == interval - 1);
            for (int h = 0; h < 2; ++h) {
                const int own0 = h * K, oth0 = (1 - h) * K;
                for (int i = 0; i < K; ++i) {
                    const int w = own0 + i;
                    pcg64_t* g = &rng[w];
                    int j = (int)oracle_pcg64_bounded(g, (uint64_t)K, &rejections);
                    const double* sel = pos + (size_t)(oth0 + j) * D;
                    double* cur = pos + (size_t)w * D;
                    double z = oracle_gw_sample(oracle_u64_to_unit(oracle_pcg64_next(g)), term1, isa);
                    for (int d = 0; d < D; ++d) prop[d] = sel[d] + z * (cur[d] - sel[d]);
                    double scaling = log(z) * (double)(D - 1);
                    double newp = oracle_logp(kind, D, tparams, prop);
                    double diff = scaling + newp - logp[w];
                    double e = oracle_neg_exponential(oracle_u64_to_unit(oracle_pcg64_next(g)));
                    int a = (e < diff);
                    if (a) { memcpy(cur, prop, sizeof(double) * (size_t)D); logp[w] = newp; ++acc[w]; }
                    ++tot[w];
                    if (trace) { oracle_trace_t* t = &trace[tr++]; t->j = j; t->accept = a; t->z = z; t->e = e; t->newp = newp; t->diff = diff; }
                    if (store && chain) memcpy(chain + ((size_t)s * W + w) * D, cur, sizeof(double) * (size_t)D);
                }
            }
        }
    }
    free(prop);
    return rejections;
}

/*
 * PER_WALKER32 rng discipline (the fast device engine): walker w owns pcg32(seed, stream = w) -- the reference's
 * facade (Utility/MultiSampler.h:54,60,86,99) with the vendored 32-bit engine (pcg_random.hpp:1568) in place of pcg64:
 *   j    engine(K)                     bounded_rand in 32-bit arithmetic (pcg_extras.hpp:514-525)
 *   u    generate_canonical<double,53> two outputs, low word first: (r0 + r1 * 2^32) / 2^64, below 1
 *                                      (libstdc++ bits/random.tcc:3346-3381 with k = 2)
 *   z    GwDistribution(u), e = log(1 - u) as for pcg64
 * Same update arithmetic and red-then-black order as oracle_stretch_run.  rng [W] carries every walker's engine.
 * Pinned against pcg-cpp's check-pcg32.out and against oracle/ref_driver.cpp:ref_stretch_run_pcg32 (vendored pcg32 +
 * libstdc++ distributions + the reference's GwDistribution), see tests/golden/pcg32_*.npz.
 */
void oracle_pcg32_seed_many(pcg32_t* g, uint64_t seed, int W)
{
    for (int w = 0; w < W; ++w) oracle_pcg32_seed_stream(&g[w], seed, (uint64_t)w);
}

static uint32_t oracle_pcg32_bounded(pcg32_t* g, uint32_t bound)
{
    uint32_t threshold = (0u - bound) % bound;
    for (;;) {
        uint32_t r = oracle_pcg32_next(g);
        if (r >= threshold) return r % bound;
    }
}

static double oracle_pcg32_unit(pcg32_t* g)
{
    uint64_t lo = oracle_pcg32_next(g);
    uint64_t hi = oracle_pcg32_next(g);
    return oracle_u64_to_unit((hi << 32) | lo);
}

int oracle_stretch_run_per_walker32(pcg32_t* rng, int kind, const double* tparams, int W, int D,
                                    int alpha_num, int alpha_denom,
                                    double* pos, double* logp, int64_t* acc, int64_t* tot,
                                    int nsteps, int interval, double* chain, oracle_trace_t* trace)
{
    const int K = W / 2;
    double term1, isa;
    oracle_gw_constants(alpha_num, alpha_denom, &term1, &isa);
    double* prop = (double*)malloc(sizeof(double) * (size_t)D);
    size_t tr = 0;
    for (int s = 0; s < nsteps; ++s) {
        for (int sub = 0; sub < interval; ++sub) {
            int store = (sub == interval - 1);
            for (int h = 0; h < 2; ++h) {
                const int own0 = h * K, oth0 = (1 - h) * K;
                for (int i = 0; i < K; ++i) {
                    const int w = own0 + i;
                    pcg32_t* g = &rng[w];
                    int j = (int)oracle_pcg32_bounded(g, (uint32_t)K);
                    const double* sel = pos + (size_t)(oth0 + j) * D;
                    double* cur = pos + (size_t)w * D;
                    double z = oracle_gw_sample(oracle_pcg32_unit(g), term1, isa);
                    for (int d = 0; d < D; ++d) prop[d] = sel[d] + z * (cur[d] - sel[d]);
                    double scaling = log(z) * (double)(D - 1);
                    double newp = oracle_logp(kind, D, tparams, prop);
                    double diff = scaling + newp - logp[w];
                    double e = oracle_neg_exponential(oracle_pcg32_unit(g));
                    int a = (e < diff);
                    if (a) { memcpy(cur, prop, sizeof(double) * (size_t)D); logp[w] = newp; ++acc[w]; }
                    ++tot[w];
                    if (trace) { oracle_trace_t* t = &trace[tr++]; t->j = j; t->accept = a; t->z = z; t->e = e; t->newp = newp; t->diff = diff; }
                    if (store && chain) memcpy(chain + ((size_t)s * W + w) * D, cur, sizeof(double) * (size_t)D);
                }
            }
        }
    }
    free(prop);
    return 0;
}

/* batched log posterior (tests + timing helpers) */
void oracle_logp_many(int kind, int D, const double* params, const double* x, long n, double* out)
{
    for (long i = 0; i < n; ++i) out[i] = oracle_logp(kind, D, params, x + (size_t)i * D);
}

/*
 * One colour of one ensemble step restricted to slots [slot_begin, slot_begin+slot_count): what ONE RANK of a
 * walker-sharded ensemble does.  Each update positions its own copy of the sampler's stream (seed, stream 0) at
 * draw 3(step*W + half*K + slot) by jump-ahead (pcg_random.hpp:622-651) -- the restatement of the design the
 * survey validated (SURVEY.md s.8(c) "Design validation").  Reads only the other colour; updates pos/logp/acc in place.
 */
void oracle_stretch_half_slots(uint64_t seed, int kind, const double* tparams, int W, int D, int alpha_num, int alpha_denom,
                               double* pos, double* logp, int64_t* acc, long step, int half, int slot_begin, int slot_count)
{
    const int K = W / 2;
    double term1, isa;
    oracle_gw_constants(alpha_num, alpha_denom, &term1, &isa);
    double* prop = (double*)malloc(sizeof(double) * (size_t)D);
    const int own0 = half * K, oth0 = (1 - half) * K;
    for (int i = slot_begin; i < slot_begin + slot_count; ++i) {
        pcg64_t g;
        oracle_pcg64_seed(&g, seed, 0);
        oracle_pcg64_advance(&g, (u128)3 * ((u128)step * (u128)W + (u128)half * (u128)K + (u128)i));
        const int w = own0 + i;
        int j = (int)oracle_pcg64_bounded(&g, (uint64_t)K, NULL);
        const double* sel = pos + (size_t)(oth0 + j) * D;
        double* cur = pos + (size_t)w * D;
        double z = oracle_gw_sample(oracle_u64_to_unit(oracle_pcg64_next(&g)), term1, isa);
        for (int d = 0; d < D; ++d) prop[d] = sel[d] + z * (cur[d] - sel[d]);
        double newp = oracle_logp(kind, D, tparams, prop);
        double diff = log(z) * (double)(D - 1) + newp - logp[w];
        double e = oracle_neg_exponential(oracle_u64_to_unit(oracle_pcg64_next(&g)));
        if (e < diff) { memcpy(cur, prop, sizeof(double) * (size_t)D); logp[w] = newp; if (acc) ++acc[w]; }
    }
    free(prop);
}

/*
 * CovarianceMatrix::calculateCovar (MCMCpp/Analysis/CovarianceMatrix.h:120-141) over a contiguous chain
 * [S][W][D]: rows with (index % sliceInterval) == 0 are used; every point adds x_i to the Kahan-summed averages
 * (addAverage :238-245) and x_i*x_j, j <= i, to the Kahan-summed products, both triangle entries (addProduct :217-235;
 * processPoint :201-214); finalizeMatrix (:143-185): avg /= n; cov_ij = prod_ij / n - avg_i * avg_j;
 * corr = cov / sqrt(var_i) row by row, then / sqrt(var_j) column by column.  mean/cov/corr: D, D*D, D*D doubles.
 * Returns the number of points used (steps used * W).
 */
long oracle_covariance(const double* chain, long S, int W, int D, int sliceInterval, double* mean, double* cov, double* corr)
{
    const size_t DD = (size_t)D * (size_t)D;
    double* prod = (double*)calloc(DD, sizeof(double));
    double* prodC = (double*)calloc(DD, sizeof(double));
    double* avg = (double*)calloc((size_t)D, sizeof(double));
    double* avgC = (double*)calloc((size_t)D, sizeof(double));
    long numPts = 0;
    for (long index = 0; index < S; ++index) {
        if ((index % sliceInterval) != 0) continue;
        for (int w = 0; w < W; ++w) {
            const double* pt = chain + ((size_t)index * W + w) * D;
            for (int i = 0; i < D; ++i) {
                double y = pt[i] - avgC[i];
                double t = avg[i] + y;
                avgC[i] = (t - avg[i]) - y;
                avg[i] = t;
                for (int j = 0; j <= i; ++j) {
                    const double val = pt[i] * pt[j];
                    size_t o = (size_t)i * D + j;
                    y = val - prodC[o]; t = prod[o] + y; prodC[o] = (t - prod[o]) - y; prod[o] = t;
                    if (i != j) {
                        o = (size_t)j * D + i;
                        y = val - prodC[o]; t = prod[o] + y; prodC[o] = (t - prod[o]) - y; prod[o] = t;
                    }
                }
            }
        }
        ++numPts;
    }
    const double ptCount = (double)(numPts * W);
    for (int i = 0; i < D; ++i) { avg[i] /= ptCount; mean[i] = avg[i]; }
    for (int i = 0; i < D; ++i)
        for (int j = 0; j <= i; ++j) {
            const double temp = (prod[(size_t)i * D + j] / ptCount) - (avg[i] * avg[j]);
            cov[(size_t)i * D + j] = temp;
            cov[(size_t)j * D + i] = temp;
        }
    for (int i = 0; i < D; ++i) {
        const double sv = sqrt(cov[(size_t)i * D + i]);
        for (int j = 0; j < D; ++j) corr[(size_t)i * D + j] = cov[(size_t)i * D + j] / sv;
    }
    for (int j = 0; j < D; ++j) {
        const double sv = sqrt(cov[(size_t)j * D + j]);
        for (int i = 0; i < D; ++i) corr[(size_t)i * D + j] /= sv;
    }
    free(prod); free(prodC); free(avg); free(avgC);
    return numPts * W;
}

/* ---- AutoCorrCalc / AutoCov (MCMCpp/Analysis/AutoCorrCalc.h:150-262, Analysis/Detail/AutoCov.h:131-290) -------------
 * Integrated autocorrelation time of every parameter from the walker-averaged, normalised, CIRCULAR autocovariance
 * (FFT length = next power of two >= numSamples, no zero padding: AutoCov.h:285-290).
 *   chain averages: Kahan over the steps, per (walker, param)                          AutoCorrCalc.h:234-262
 *   per walker: centre, bit-reverse, radix-2 FFT, |X|^2 (bit-reversed), radix-2 inverse FFT (no 1/N), divide by
 *   element 0                                                                             AutoCov.h:131-229
 *   average over walkers with the header's compensated sum (value = acov + comp)          AutoCorrCalc.h:206-219
 *   IAT: -rho[0] + Kahan sum of 2 rho[i] until i > windowScaling * sum; negative final sum if it never converges :180-203
 * quirk != 0 reproduces transferWalker's accumulation into the un-cleared buffer (AutoCorrCalc.h:228-232): the series
 * handed to the FFT is chain + (what calcNormAutoCov left there for the previous walker / parameter); the buffer is taken
 * as zero the first time (fresh `new[]` pages).  quirk == 0 is the clean definition (series = chain) the device computes.
 * All W walkers are used (numWalkersToUse == 0).  rho (may be NULL): D * S averaged autocovariance functions.
 */
static unsigned oracle_bitrev(unsigned v, int lg)
{
    v = ((v >> 1) & 0x55555555u) | ((v & 0x55555555u) << 1);
    v = ((v >> 2) & 0x33333333u) | ((v & 0x33333333u) << 2);
    v = ((v >> 4) & 0x0F0F0F0Fu) | ((v & 0x0F0F0F0Fu) << 4);
    v = ((v >> 8) & 0x00FF00FFu) | ((v & 0x00FF00FFu) << 8);
    v = (v >> 16) | (v << 16);
    return v >> (32 - lg);
}

static void oracle_fft_stages(double* re, double* im, const double* wr, const double* wi, int lg, int n)
{
    for (int s = 1; s <= lg; ++s) {
        const int m1 = 1 << s, m2 = m1 >> 1, stride = n >> s;
        for (int k = 0; k < n; k += m1) {
            int tw = 0;
            for (int j = 0; j < m2; ++j) {
                const double xr = re[k + j + m2], xi = im[k + j + m2];
                const double t1r = wr[tw] * xr - wi[tw] * xi;          /* std::complex operator*: (ac - bd, ad + bc) */
                const double t1i = wr[tw] * xi + wi[tw] * xr;
                const double t2r = re[k + j], t2i = im[k + j];
                re[k + j] = t2r + t1r; im[k + j] = t2i + t1i;
                re[k + j + m2] = t2r - t1r; im[k + j + m2] = t2i - t1i;
                tw += stride;
            }
        }
    }
}

int oracle_fft_size_log2(int numSamples) { return (int)ceil(log2((double)numSamples)); }

/* twiddles exactly as AutoCov::calculateRootsOfUnity (AutoCov.h:270-281): cos/sin of TwoPi*i/N, TwoPi from the rational pi */
void oracle_autocov_twiddles(int lg, double* cosPart, double* sinPart)
{
    const int n = 1 << lg;
    const double twoPi = 2.0 * ((double)3141592653589793239ULL / (double)1000000000000000000ULL);
    for (int i = 0; i < n / 2; ++i) {
        const double angle = twoPi * (double)i / (double)n;
        cosPart[i] = cos(angle);
        sinPart[i] = sin(angle);
    }
}

void oracle_autocorr(const double* chain, int S, int W, int D, int windowScaling, int quirk, double* iat, double* rho)
{
    const int lg = oracle_fft_size_log2(S), n = 1 << lg;
    double* cs = (double*)malloc(sizeof(double) * (size_t)(n / 2 + 1));
    double* sn = (double*)malloc(sizeof(double) * (size_t)(n / 2 + 1));
    double* nsn = (double*)malloc(sizeof(double) * (size_t)(n / 2 + 1));
    oracle_autocov_twiddles(lg, cs, sn);
    for (int i = 0; i < n / 2; ++i) nsn[i] = -sn[i];
    double* re1 = (double*)malloc(sizeof(double) * (size_t)n); double* im1 = (double*)malloc(sizeof(double) * (size_t)n);
    double* re2 = (double*)malloc(sizeof(double) * (size_t)n); double* im2 = (double*)malloc(sizeof(double) * (size_t)n);
    double* acov = (double*)calloc((size_t)S, sizeof(double));           /* acovFuncArray */
    double* avgA = (double*)malloc(sizeof(double) * (size_t)S); double* avgC = (double*)malloc(sizeof(double) * (size_t)S);
    double* mean = (double*)calloc((size_t)W * D, sizeof(double)); double* meanC = (double*)calloc((size_t)W * D, sizeof(double));
    for (int t = 0; t < S; ++t)
        for (int j = 0; j < W * D; ++j) {
            const double value = chain[(size_t)t * W * D + j] - meanC[j];
            const double temp = mean[j] + value;
            meanC[j] = (temp - mean[j]) - value;
            mean[j] = temp;
        }
    for (int j = 0; j < W * D; ++j) mean[j] /= (double)S;
    for (int p = 0; p < D; ++p) {
        for (int i = 0; i < S; ++i) { avgA[i] = 0.0; avgC[i] = 0.0; }
        for (int w = 0; w < W; ++w) {
            const int off = w * D + p;
            for (int t = 0; t < S; ++t) acov[t] = chain[(size_t)t * W * D + off] + (quirk ? acov[t] : 0.0);   /* transferWalker */
            for (int i = 0; i < S; ++i) { const unsigned r = oracle_bitrev((unsigned)i, lg); re1[r] = acov[i] - mean[off]; im1[r] = 0.0; }
            for (int i = S; i < n; ++i) { const unsigned r = oracle_bitrev((unsigned)i, lg); re1[r] = 0.0; im1[r] = 0.0; }
            oracle_fft_stages(re1, im1, cs, nsn, lg, n);
            for (int i = 0; i < n; ++i) { const unsigned r = oracle_bitrev((unsigned)i, lg); re2[r] = re1[i] * re1[i] + im1[i] * im1[i]; im2[r] = 0.0; }
            oracle_fft_stages(re2, im2, cs, sn, lg, n);
            const double norm = re2[0];
            for (int i = 0; i < S; ++i) acov[i] = re2[i] / norm;
            for (int i = 0; i < S; ++i) {                                 /* averageAutocovarianceFunctions */
                const double value = acov[i] + avgC[i];
                const double temp = avgA[i] + value;
                avgC[i] = (temp - avgA[i]) - value;
                avgA[i] = temp;
            }
        }
        for (int i = 0; i < S; ++i) avgA[i] /= (double)W;
        if (rho) memcpy(rho + (size_t)p * S, avgA, sizeof(double) * (size_t)S);
        double sum = -avgA[0], comp = 0.0;
        const double factor = (double)windowScaling;
        int converged = 0;
        for (int i = 0; i < S; ++i) {
            const double value = (2.0 * avgA[i]) - comp;
            const double temp = sum + value;
            comp = (temp - sum) - value;
            sum = temp;
            if ((double)i > factor * sum) { converged = 1; break; }
        }
        iat[p] = converged ? sum : -sum;
    }
    free(cs); free(sn); free(nsn); free(re1); free(im1); free(re2); free(im2); free(acov); free(avgA); free(avgC); free(mean); free(meanC);
}

/*
 * DifferentialEvolution::updateWalker (MCMCpp/Movers/DifferentialEvolution.h:113-142) inside the same red/black driver
 * (EnsembleSampler.h:341-360).  Per update: ind1 = bounded(K); ind2 = bounded(K) repeated until != ind1;
 * proposal_i = cur_i + gamma (w1_i - w2_i) + (lowEdge + u_i * width), one uniform per parameter
 * (MultiSampler.h:74); accept iff (newp - logp) > log(1 - u).  gamma = 2.38 / sqrt(2 D) unless overridden (:58,:83);
 * lowEdge = -1e-4, width = 2e-4 (:150-151).  per_walker == 0: ONE stream rng[0] consumed in walker order (the sequential
 * reference sampler); != 0: walker w owns rng[w].  trace.j = ind1, trace.z = (double)ind2, trace.diff = newp - logp.
 */
void oracle_de_run(pcg64_t* rng, int per_walker, int kind, const double* tparams, int W, int D, double gamma, double lowEdge, double width,
                   double* pos, double* logp, int64_t* acc, int64_t* tot, int nsteps, int interval, double* chain, oracle_trace_t* trace)
{
    const int K = W / 2;
    double* prop = (double*)malloc(sizeof(double) * (size_t)D);
    size_t tr = 0;
    for (int s = 0; s < nsteps; ++s)
        for (int sub = 0; sub < interval; ++sub) {
            const int store = (sub == interval - 1);
            for (int h = 0; h < 2; ++h) {
                const int own0 = h * K, oth0 = (1 - h) * K;
                for (int i = 0; i < K; ++i) {
                    const int w = own0 + i;
                    pcg64_t* g = per_walker ? &rng[w] : &rng[0];
                    const int ind1 = (int)oracle_pcg64_bounded(g, (uint64_t)K, NULL);
                    int ind2;
                    do { ind2 = (int)oracle_pcg64_bounded(g, (uint64_t)K, NULL); } while (ind2 == ind1);
                    double* cur = pos + (size_t)w * D;
                    const double* w1 = pos + (size_t)(oth0 + ind1) * D;
                    const double* w2 = pos + (size_t)(oth0 + ind2) * D;
                    for (int d = 0; d < D; ++d)
                        prop[d] = cur[d] + gamma * (w1[d] - w2[d]) + (lowEdge + (oracle_u64_to_unit(oracle_pcg64_next(g)) * width));
                    const double newp = oracle_logp(kind, D, tparams, prop);
                    const double diff = newp - logp[w];
                    const double e = oracle_neg_exponential(oracle_u64_to_unit(oracle_pcg64_next(g)));
                    const int a = diff > e;
                    if (a) { memcpy(cur, prop, sizeof(double) * (size_t)D); logp[w] = newp; ++acc[w]; }
                    ++tot[w];
                    if (trace) { oracle_trace_t* t = &trace[tr++]; t->j = ind1; t->accept = a; t->z = (double)ind2; t->e = e; t->newp = newp; t->diff = diff; }
                    if (store && chain) memcpy(chain + ((size_t)s * W + w) * D, cur, sizeof(double) * (size_t)D);
                }
            }
        }
    free(prop);
}

/*
 * WalkMove::updateWalker (MCMCpp/Movers/WalkMove.h:95-186) inside the same red/black driver.  Per update:
 *   selectWalkers (:117-136): selection sampling of S distinct indices in increasing order, one uniform per examined
 *     walker: skip if (K - examined) * u >= (S - selected), else take it;
 *   calculateProposal (:139-183): S normals r_j (std::normal_distribution, libstdc++ bits/random.tcc: Marsaglia polar,
 *     the second value of a pair is kept for the next call); for every parameter i, a = sum_j r_j v_ji, c = sum_j v_ji
 *     (in order j), b = sum_j r_j (once): proposal_i = cur_i + (a - (b * (c / S)));
 *   accept iff log(1 - u) < newp - logp.
 * One engine = pcg64 + the normal distribution's saved value: nstate[2*e] = saved, nstate[2*e+1] = 1.0 if available.
 * per_walker == 0: engine 0 for everyone in walker order (the sequential reference sampler); else engine w.
 * trace.j = first selected index, trace.z = first normal, trace.diff = newp - logp.
 */
static double oracle_std_normal(pcg64_t* g, double* st)
{
    double ret;
    if (st[1] != 0.0) { st[1] = 0.0; ret = st[0]; }
    else {
        double x, y, r2;
        do {
            x = 2.0 * oracle_u64_to_unit(oracle_pcg64_next(g)) - 1.0;
            y = 2.0 * oracle_u64_to_unit(oracle_pcg64_next(g)) - 1.0;
            r2 = x * x + y * y;
        } while (r2 > 1.0 || r2 == 0.0);
        const double mult = sqrt(-2 * log(r2) / r2);
        st[0] = x * mult; st[1] = 1.0;
        ret = y * mult;
    }
    return ret * 1.0 + 0.0;                               /* stddev 1, mean 0 */
}

void oracle_walk_run(pcg64_t* rng, double* nstate, int per_walker, int kind, const double* tparams, int W, int D, int S,
                     double* pos, double* logp, int64_t* acc, int64_t* tot, int nsteps, int interval, double* chain, oracle_trace_t* trace)
{
    const int K = W / 2;
    const double ptCount = (double)S;
    double* prop = (double*)malloc(sizeof(double) * (size_t)D);
    double* rnd = (double*)malloc(sizeof(double) * (size_t)S);
    int* idx = (int*)malloc(sizeof(int) * (size_t)S);
    size_t tr = 0;
    for (int s = 0; s < nsteps; ++s)
        for (int sub = 0; sub < interval; ++sub) {
            const int store = (sub == interval - 1);
            for (int h = 0; h < 2; ++h) {
                const int own0 = h * K, oth0 = (1 - h) * K;
                for (int i = 0; i < K; ++i) {
                    const int w = own0 + i;
                    pcg64_t* g = per_walker ? &rng[w] : &rng[0];
                    double* st = per_walker ? &nstate[2 * w] : &nstate[0];
                    int numSelected = 0, examined = 0;
                    while (numSelected < S) {
                        const double u = oracle_u64_to_unit(oracle_pcg64_next(g));
                        if (((K - examined) * u) >= (ptCount - numSelected)) ++examined;
                        else { idx[numSelected] = examined; ++examined; ++numSelected; }
                    }
                    double* cur = pos + (size_t)w * D;
                    double i0 = 0.0, i1 = 0.0, i2 = 0.0;
                    for (int j = 0; j < S; ++j) {
                        rnd[j] = oracle_std_normal(g, st);
                        const double value = pos[(size_t)(oth0 + idx[j]) * D];
                        i0 += rnd[j] * value; i1 += rnd[j]; i2 += value;
                    }
                    prop[0] = cur[0] + (i0 - (i1 * (i2 / ptCount)));
                    for (int d = 1; d < D; ++d) {
                        i0 = 0.0; i2 = 0.0;
                        for (int j = 0; j < S; ++j) {
                            const double value = pos[(size_t)(oth0 + idx[j]) * D + d];
                            i0 += rnd[j] * value; i2 += value;
                        }
                        prop[d] = (cur[d] + (i0 - (i1 * (i2 / ptCount))));
                    }
                    const double newp = oracle_logp(kind, D, tparams, prop);
                    const double diff = newp - logp[w];
                    const double e = oracle_neg_exponential(oracle_u64_to_unit(oracle_pcg64_next(g)));
                    const int a = e < diff;
                    if (a) { memcpy(cur, prop, sizeof(double) * (size_t)D); logp[w] = newp; ++acc[w]; }
                    ++tot[w];
                    if (trace) { oracle_trace_t* t = &trace[tr++]; t->j = idx[0]; t->accept = a; t->z = rnd[0]; t->e = e; t->newp = newp; t->diff = diff; }
                    if (store && chain) memcpy(chain + ((size_t)s * W + w) * D, cur, sizeof(double) * (size_t)D);
                }
            }
        }
    free(prop); free(rnd); free(idx);
}
