/* genotype.cu — per-CNV sums and Poisson genotype likelihoods on sm_100a (K4).
 *
 * Replaces find_depths -> calculate_likelihood_CNV (likelihood.c:290-308, :108-169)
 * and lpoisson (likelihood.c:96-105).
 *
 * The reference walks every base of every CNV with three accumulators
 * (likelihood.c:115-124).  Here:
 *   observed_rd  (int)    = D(end) - D(start), D(p) = sum of depth[0, p): the per-chunk /
 *                           per-group / per-super depth sums the record sweep left behind
 *                           (cg_counts) give D at the chunk boundary in three short loads,
 *                           the warp sums the < 2048 int16 depths from there to p.  No
 *                           per-position prefix array exists.
 *   expected_rd  (float)  = the SAME ordered fp32 sum, evaluated window by window:
 *                           inside one 100-position GC window the addend is constant,
 *                           and k repeated fp32 additions of one value are replayed
 *                           exactly in O(1) by add_run_f32 below (no reassociation,
 *                           no tree: the result is bit-identical to the literal loop).
 *   mappability  (double) = ordered fp64 sum of the painted fp32 values, evaluated
 *                           run by run over the flattened track with add_run_f64.
 * Long and short CNVs are mixed, so lanes pull CNVs from a shared counter and refill
 * as soon as their chain ends (no lane idles behind a multi-megabase neighbour).
 *
 * The likelihood kernel gives each CNV four lanes: one per genotype hypothesis
 * (k*log(lambda) - lambda) and one for lgamma(k+1); fp64 throughout, explicit
 * round-to-nearest mul/sub so nothing is contracted into an FMA (the reference is
 * built for baseline x86-64, which has none).
 */
#include "cg_internal.h"
#include <math.h>
#include <stdlib.h>

#define FULL 0xffffffffu

#ifdef CG_DEBUG_COUNTERS                     /* tuning builds only (scripts/build_variant.sh ... -DCG_DEBUG_COUNTERS) */
__device__ unsigned long long g_dbg[16];
#define DBG(i) do { if ((threadIdx.x & 31) == 0) atomicAdd(&g_dbg[i], 1ull); } while (0)
extern "C" void cg_debug_counters(unsigned long long* out, int reset)
{
    cudaMemcpyFromSymbol(out, g_dbg, sizeof(unsigned long long) * 16);
    if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(g_dbg, z, sizeof z); }
}
#else
#define DBG(i) do { } while (0)
#endif

/* ---- k repeated additions of one value, exactly ---------------------------
 * For E > 0 and v > 0 in fp32: while E stays inside one binade [2^e, 2^(e+1)) every
 * rounded addition moves the significand by the same integer d (v = q*ulp + r adds q
 * or q+1; an exact tie settles after one step because the result of a tie step is
 * even).  So: take literal steps until three consecutive sums share an exponent,
 * read d off the last step, jump as many steps as fit below the binade's top, and
 * continue literally across the boundary.  Every path performs the reference's
 * additions or an integer replay proven equal to them; anything unusual (v <= 0,
 * NaN, Inf, subnormal) takes the literal loop. */
__device__ __forceinline__ float add_run_f32(float E, float v, int k)
{
    if (k <= 0) return E;
    if (v == 0.0f) return E;                        /* E is never -0, so E + 0 == E */
    const unsigned vb = __float_as_uint(v);
    if ((vb >> 31) || (vb >> 23) == 0u || (vb >> 23) == 255u || E < 0.0f) {
        for (int i = 0; i < k; i++) E = __fadd_rn(E, v);
        return E;
    }
    while (k > 0) {
        const float E1 = __fadd_rn(E, v);
        if (k < 4) { E = E1; k--; continue; }
        const float E2 = __fadd_rn(E1, v);
        const float E3 = __fadd_rn(E2, v);
        const unsigned b1 = __float_as_uint(E1), b2 = __float_as_uint(E2), b3 = __float_as_uint(E3);
        k -= 3;
        E = E3;
        if ((b1 >> 23) == (b3 >> 23)) {             /* E1, E2, E3 in one binade (monotone) */
            const unsigned d = b3 - b2;
            if (d == 0u) return E;                  /* v is absorbed from here on */
            const unsigned room = 0x7fffffu - (b3 & 0x7fffffu);
            unsigned n;
            if ((unsigned long long)(unsigned)k * d <= (unsigned long long)room) n = (unsigned)k;
            else n = room / d;
            E = __uint_as_float(b3 + n * d);
            k -= (int)n;
        }
    }
    return E;
}

__device__ __forceinline__ double add_run_f64(double S, double v, int k)
{
    if (k <= 0) return S;
    if (v == 0.0) return S;
    const unsigned long long vb = (unsigned long long)__double_as_longlong(v);
    const unsigned ve = (unsigned)(vb >> 52) & 0x7ffu;
    if ((vb >> 63) || ve == 0u || ve == 0x7ffu || S < 0.0) {
        for (int i = 0; i < k; i++) S = __dadd_rn(S, v);
        return S;
    }
    while (k > 0) {
        const double S1 = __dadd_rn(S, v);
        if (k < 4) { S = S1; k--; continue; }
        const double S2 = __dadd_rn(S1, v);
        const double S3 = __dadd_rn(S2, v);
        const unsigned long long b1 = (unsigned long long)__double_as_longlong(S1);
        const unsigned long long b2 = (unsigned long long)__double_as_longlong(S2);
        const unsigned long long b3 = (unsigned long long)__double_as_longlong(S3);
        k -= 3;
        S = S3;
        if ((b1 >> 52) == (b3 >> 52)) {
            const unsigned long long d = b3 - b2;
            if (d == 0ull) return S;
            const unsigned long long room = 0xfffffffffffffull - (b3 & 0xfffffffffffffull);
            unsigned long long n;
            if (__umul64hi((unsigned long long)k, d) == 0ull && (unsigned long long)k * d <= room) n = (unsigned long long)k;
            else n = room / d;
            S = __longlong_as_double((long long)(b3 + n * d));
            k -= (int)n;
        }
    }
    return S;
}

/* ---- the common case without any floating-point instruction -----------------
 * While E sits in the binade with biased exponent e, one rounded addition of v moves
 * E's bit pattern by d(v, e) = q + (rem > half), where v's 24-bit significand mv is
 * split at sh = e - exponent(v):  q = mv >> sh, rem = the sh bits shifted out, half =
 * 2^(sh-1)  (round to nearest; rem == half is the tie, left to add_run_f32).  d does not
 * depend on E's significand, so a run of k additions is `bits += k*d`, and a whole chunk
 * of GC windows is a chain of integer multiply-adds.  All d are >= 0, so if the exponent
 * field is unchanged at the END of the chunk no addition crossed a binade boundary and
 * every step was exactly the reference's fadd; otherwise the chunk is replayed window by
 * window through add_run_f32. */
__device__ __forceinline__ unsigned step_delta_f32(unsigned vb, unsigned e, bool& bad)
{
    if (vb == 0u) return 0u;                         /* + 0.0f changes nothing */
    const unsigned ev = vb >> 23;                    /* sign bit set -> ev >= 256 */
    if (ev == 0u || ev >= 255u) { bad = true; return 0u; }
    const int sh = (int)e - (int)ev;
    if (sh >= 26) return 0u;                         /* v < ulp/2: absorbed */
    if (sh < 1) { bad = true; return 0u; }           /* E < 2v: the sum leaves the binade */
    const unsigned mv = (vb & 0x7fffffu) | 0x800000u;
    const unsigned rem = mv & ((1u << sh) - 1u), half = 1u << (sh - 1);
    if (rem == half) bad = true;                     /* exact tie: parity dependent */
    return (mv >> sh) + (rem > half ? 1u : 0u);
}

__device__ __forceinline__ double add_run_f64_fast(double S, double v, int k)
{
    if (k <= 0 || v == 0.0) return S;
    const unsigned long long sb = (unsigned long long)__double_as_longlong(S);
    const unsigned long long vb = (unsigned long long)__double_as_longlong(v);
    const unsigned e = (unsigned)(sb >> 52), ev = (unsigned)(vb >> 52);   /* sign set -> >= 2048 */
    const int sh = (int)e - (int)ev;
    if (e == 0u || e >= 0x7ffu || ev == 0u || ev >= 0x7ffu || sh < 1) return add_run_f64(S, v, k);
    if (sh >= 55) return S;
    const unsigned long long mv = (vb & 0xfffffffffffffull) | 0x10000000000000ull;
    const unsigned long long rem = mv & ((1ull << sh) - 1ull), half = 1ull << (sh - 1);
    if (rem == half) return add_run_f64(S, v, k);
    const unsigned long long d = (mv >> sh) + (rem > half ? 1ull : 0ull);
    const unsigned long long room = 0xfffffffffffffull - (sb & 0xfffffffffffffull);
    if (__umul64hi((unsigned long long)k, d) != 0ull || (unsigned long long)k * d > room) return add_run_f64(S, v, k);
    return __longlong_as_double((long long)(sb + (unsigned long long)k * d));
}


/* ---- a block of mappability runs without rounding --------------------------------------
 * Every addend of the block is a positive multiple of 2^u, u = the lowest set bit over the
 * accumulator and the block's values.  If accumulator + sum(k_i * v_i) stays below 2^53 * 2^u,
 * every partial sum of the reference's sequential loop is a multiple of 2^u below that bound,
 * hence exactly representable: no addition ever rounds, the order does not matter, and the
 * result is the integer sum scaled by 2^u.  (Mappability values are fp32 in [0,1] and the sums
 * stay far below 2^53 units, so this is the path real tracks take.)  Returns false — nothing
 * changed — when the bound or a precondition fails; the caller then replays the block through
 * the rounding-aware scan. */
__device__ __forceinline__ bool map_block_exact(unsigned long long& MB, int k, double v, int lane)
{
    const unsigned long long vb = (unsigned long long)__double_as_longlong(v);
    const unsigned ev = (unsigned)(vb >> 52);                    /* sign bit set -> >= 2048 */
    const bool active = k > 0 && (vb << 1) != 0ull;
    bool ok = !active || (ev >= 1u && ev <= 2046u);
    int low = 0x7fffffff;                                         /* exponent of the lowest set bit */
    unsigned long long odd = 0ull;
    if (active && ok) {
        const unsigned long long m = (vb & 0xfffffffffffffull) | 0x10000000000000ull;
        const int tz = __ffsll((long long)m) - 1;
        odd = m >> tz;
        low = (int)ev - 1075 + tz;
    }
    const unsigned es = (unsigned)(MB >> 52);
    int low_s = 0x7fffffff;
    unsigned long long odd_s = 0ull;
    if (MB != 0ull) {
        if (es < 1u || es > 2046u) ok = false;                    /* negative, subnormal, inf, nan */
        const unsigned long long m = (MB & 0xfffffffffffffull) | 0x10000000000000ull;
        const int tz = __ffsll((long long)m) - 1;
        odd_s = m >> tz;
        low_s = (int)es - 1075 + tz;
    }
    const int u = min(__reduce_min_sync(FULL, low), low_s);
    unsigned long long units = 0ull;
    if (active && ok) {
        const int sh = low - u;                                   /* >= 0 */
        const unsigned long long prod = (unsigned long long)k * odd;              /* < 2^55 for an fp32 value */
        if ((odd >> 24) != 0ull || sh > 52 || (prod >> (53 - sh)) != 0ull) ok = false;
        else units = prod << sh;                                  /* < 2^53 */
    }
    unsigned long long acc = 0ull;
    if (MB != 0ull && ok) {
        const int sh = low_s - u;
        if (sh > 52 || (odd_s >> (53 - sh)) != 0ull) ok = false;
        else acc = odd_s << sh;                                   /* < 2^53 */
    }
    if (!__all_sync(FULL, ok)) return false;
    if (u == 0x7fffffff) return true;                             /* nothing to add, accumulator zero */
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) units += __shfl_xor_sync(FULL, units, d);     /* 33 terms below 2^53: no overflow */
    const unsigned long long total = acc + units;
    if (total >= (1ull << 53)) return false;
    if (total == 0ull) return true;
    const double t = __ll2double_rn((long long)total);            /* exact: < 2^53 */
    const long long tb = __double_as_longlong(t) + ((long long)u << 52);
    const unsigned et = (unsigned)((unsigned long long)tb >> 52);
    if (et < 1u || et > 2046u) return false;
    MB = (unsigned long long)tb;
    return true;
}

/* ---- Poisson likelihoods, c-score, copy number ------------------------------ */
/* (int)x as x86-64 does it (cvttsd2si): truncate toward zero; NaN or out of range
 * gives INT_MIN.  This is what feeds `int max(int,int)` (common.c:262-268) at
 * likelihood.c:138,160. */
__device__ __forceinline__ int c_double_to_int(double x)
{
    if (!(x > -2147483649.0 && x < 2147483648.0)) return (int)0x80000000;
    return __double2int_rz(x);
}

__device__ __forceinline__ bool near_integer(double x)
{
    const double a = fabs(x), d = fabs(a - rint(a));
    return d <= 1e-9 * fmax(1.0, a);
}


/* ---- double-double evaluation of a Poisson log-likelihood, for the decisions only ----------------
 * likelihood.c takes discontinuous decisions on the three likelihoods: `(int)lhomo`, `(int)lhetero` inside the
 * int-truncating max (:138, :160, common.c:262-268) and `lhomo > lhetero` (:164).  The device's fp64 log / lgamma
 * and glibc's agree to ~1e-13 relative, which is enough for every printed digit unless a likelihood sits on an
 * integer, or the two on each other, closer than that.  A CNV whose likelihoods come within 1e-9 relative of such a
 * boundary (margin_flags) has the three of them evaluated again here in double-double arithmetic (~31 significant
 * digits): k*log(lambda) - lambda - log(k!), log by atanh series after exact range reduction, log(k!) by
 * Stirling's series above 40 with an exact shift below it.  The truncation and the comparison then follow the
 * real-number values; no libm, neither CUDA's nor glibc's, decides.  (If the real value lies within glibc's own
 * rounding error of the boundary, the reference's answer is a property of its libm build, not of its algorithm.) */
struct dd { double hi, lo; };
__device__ __forceinline__ dd dd_quick(double a, double b) { const double s = a + b; dd r; r.hi = s; r.lo = b - (s - a); return r; }
__device__ __forceinline__ dd dd_two_sum(double a, double b) { const double s = a + b, bb = s - a; dd r; r.hi = s; r.lo = (a - (s - bb)) + (b - bb); return r; }
__device__ __forceinline__ dd dd_two_prod(double a, double b) { const double p = a * b; dd r; r.hi = p; r.lo = __fma_rn(a, b, -p); return r; }
__device__ __forceinline__ dd dd_add(dd x, dd y) { dd s = dd_two_sum(x.hi, y.hi); s.lo += x.lo + y.lo; return dd_quick(s.hi, s.lo); }
__device__ __forceinline__ dd dd_neg(dd x) { dd r; r.hi = -x.hi; r.lo = -x.lo; return r; }
__device__ __forceinline__ dd dd_sub(dd x, dd y) { return dd_add(x, dd_neg(y)); }
__device__ __forceinline__ dd dd_mul(dd x, dd y) { dd p = dd_two_prod(x.hi, y.hi); p.lo += x.hi * y.lo + x.lo * y.hi; return dd_quick(p.hi, p.lo); }
__device__ __forceinline__ dd dd_from(double a) { dd r; r.hi = a; r.lo = 0.0; return r; }
__device__ __forceinline__ dd dd_div(dd x, dd y)
{
    const double q1 = x.hi / y.hi;
    dd r = dd_sub(x, dd_mul(y, dd_from(q1)));
    const double q2 = r.hi / y.hi;
    r = dd_sub(r, dd_mul(y, dd_from(q2)));
    const double q3 = r.hi / y.hi;
    dd q = dd_quick(q1, q2);
    return dd_add(q, dd_from(q3));
}

/* log(x) for a finite double x > 0 */
__device__ __noinline__ dd dd_log(double x)
{
    int e;
    double m = frexp(x, &e);                          /* x = m * 2^e, m in [0.5, 1) */
    if (m < 0.70710678118654752) { m *= 2.0; e -= 1; }            /* m in [sqrt(1/2), sqrt(2)): exact */
    const dd num = dd_from(m - 1.0);                  /* exact for m in [0.5, 2] */
    const dd den = dd_two_sum(m, 1.0);
    const dd t = dd_div(num, den);                    /* |t| <= 0.1716 */
    const dd t2 = dd_mul(t, t);
    dd sum = dd_div(dd_from(1.0), dd_from(47.0));     /* sum_k t^(2k) / (2k+1), k = 23 .. 0: (0.0295)^24 < 2e-37 */
    for (int k = 22; k >= 0; k--) sum = dd_add(dd_mul(sum, t2), dd_div(dd_from(1.0), dd_from((double)(2 * k + 1))));
    dd lm = dd_mul(t, sum);
    lm.hi *= 2.0; lm.lo *= 2.0;
    dd ln2; ln2.hi = 0x1.62e42fefa39efp-1; ln2.lo = 0x1.abc9e3b39803fp-56;
    return dd_add(dd_mul(ln2, dd_from((double)e)), lm);
}

/* log(k!) = lgamma(k + 1) for an integer k >= 0 */
__device__ __noinline__ dd dd_log_factorial(int k)
{
    if (k <= 1) return dd_from(0.0);
    double z = (double)k + 1.0;                       /* lgamma's argument */
    dd prod = dd_from(1.0);                           /* lgamma(z) = lgamma(z + m) - log(z (z+1) ... (z+m-1)) */
    while (z < 40.0) { prod = dd_mul(prod, dd_from(z)); z += 1.0; }
    const dd lz = dd_log(z);
    const dd zi = dd_div(dd_from(1.0), dd_from(z)), zi2 = dd_mul(zi, zi);
    /* B_2n / (2n (2n-1)), n = 10 .. 1, as exact integer ratios; the first omitted term is below 4e-33 for z >= 40 */
    const double cn[10] = {1.0, -1.0, 1.0, -1.0, 1.0, -691.0, 1.0, -3617.0, 43867.0, -174611.0};
    const double cd[10] = {12.0, 360.0, 1260.0, 1680.0, 1188.0, 360360.0, 156.0, 122400.0, 244188.0, 125400.0};
    dd ser = dd_from(0.0);
    for (int n = 9; n >= 0; n--) ser = dd_add(dd_mul(ser, zi2), dd_div(dd_from(cn[n]), dd_from(cd[n])));
    ser = dd_mul(ser, zi);
    dd hl2pi; hl2pi.hi = 0x1.d67f1c864beb5p-1; hl2pi.lo = -0x1.65b5a1b7ff5dfp-55;
    dd r = dd_mul(dd_sub(dd_from(z), dd_from(0.5)), lz);
    r = dd_add(dd_sub(r, dd_from(z)), dd_add(hl2pi, ser));
    if (prod.hi != 1.0) {
        dd lp = dd_log(prod.hi);                      /* log(hi + lo) = log(hi) + lo/hi - ...: (lo/hi)^2 < 2e-32 */
        lp = dd_add(lp, dd_from(prod.lo / prod.hi));
        r = dd_sub(r, lp);
    }
    return r;
}

/* lpoisson (likelihood.c:96-105) in double-double; lambda is an exact double (0 already replaced by 0.01) */
__device__ __noinline__ dd dd_lpoisson(int k, double lambda, dd logfact)
{
    return dd_sub(dd_sub(dd_mul(dd_log(lambda), dd_from((double)k)), dd_from(lambda)), logfact);
}

struct dd_decision { int ih, it, homo_gt_het; };
__device__ dd_decision dd_decide(int obs, double lam_homo, double lam_het);

/* (int)x of the real number x = hi + lo as C truncates a double that holds it: toward zero; out of int range as cvttsd2si */
__device__ __forceinline__ int dd_trunc_to_int(dd x)
{
    if (!(x.hi > -2147483649.0 && x.hi < 2147483648.0)) return (int)0x80000000;
    double t = trunc(x.hi);
    if (t == x.hi) {                                  /* hi is an integer: lo decides on which side the value lies */
        if (x.hi > 0.0 && x.lo < 0.0) t -= 1.0;
        else if (x.hi < 0.0 && x.lo > 0.0) t += 1.0;
    }
    return __double2int_rz(t);
}

/* the three decisions of a flagged CNV in one out-of-line call (one call site keeps the spills around it out of the
 * common path of k_genotype) */
__device__ __noinline__ dd_decision dd_decide(int obs, double lam_homo, double lam_het)
{
    const dd lf = dd_log_factorial(obs);
    const dd h = dd_lpoisson(obs, lam_homo, lf), t = dd_lpoisson(obs, lam_het, lf);
    dd_decision r;
    r.ih = dd_trunc_to_int(h); r.it = dd_trunc_to_int(t);
    r.homo_gt_het = lam_homo == lam_het ? 0 : (h.hi > t.hi || (h.hi == t.hi && h.lo > t.lo)) ? 1 : 0;
    return r;
}

/* fp64 twin of step_delta_f32: significand increment of one rounded addition of v while the
 * running sum has biased exponent e. */
__device__ __forceinline__ unsigned long long step_delta_f64(unsigned long long vb, unsigned e, bool& bad)
{
    if ((vb << 1) == 0ull) return 0ull;              /* +-0.0 changes nothing */
    const unsigned ev = (unsigned)(vb >> 52);        /* sign bit set -> ev >= 2048 */
    if (ev == 0u || ev >= 0x7ffu) { bad = true; return 0ull; }
    const int sh = (int)e - (int)ev;
    if (sh >= 55) return 0ull;
    if (sh < 1) { bad = true; return 0ull; }
    const unsigned long long mv = (vb & 0xfffffffffffffull) | 0x10000000000000ull;
    const unsigned long long rem = mv & ((1ull << sh) - 1ull), half = 1ull << (sh - 1);
    if (rem == half) bad = true;
    return (mv >> sh) + (rem > half ? 1ull : 0ull);
}

__device__ __forceinline__ unsigned warp_incl_scan_u32(unsigned v, int lane)
{
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const unsigned o = __shfl_up_sync(FULL, v, d); if (lane >= d) v += o; }
    return v;
}
__device__ __forceinline__ unsigned long long warp_incl_scan_u64(unsigned long long v, int lane)
{
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const unsigned long long o = __shfl_up_sync(FULL, v, d); if (lane >= d) v += o; }
    return v;
}

/* four consecutive GC bytes starting at window `idx` (idx is a multiple of 4) */
__device__ __forceinline__ unsigned load_gc4(const int8_t* __restrict__ gc, int idx, int n_windows)
{
    const int8_t* p = gc + idx;
    if (idx + 3 < n_windows && (reinterpret_cast<uintptr_t>(p) & 3u) == 0u)
        return __ldg(reinterpret_cast<const unsigned*>(p));
    unsigned v = 0u;
#pragma unroll
    for (int j = 0; j < 4; j++) if (idx + j < n_windows) v |= (unsigned)(unsigned char)__ldg(p + j) << (8 * j);
    return v;
}

__device__ __forceinline__ cg_cnv make_cnv(int a, int b) { cg_cnv v; v.start = a; v.end = b; return v; }

/* Sum of depth[s, e) by one warp straight from the depth array: for a CNV of a few thousand positions this is one
 * round of independent 16-byte loads (<= GENO_DIRECT/256 + 1 per lane) instead of two endpoint sums with three
 * count levels each.  Positions outside [s, e) inside the first and last vector are masked out. */
#define GENO_DIRECT 8192
__device__ __forceinline__ int depth_sum_direct(const int16_t* __restrict__ depth, int s, int e, int lane)
{
    const int v0 = s >> 3, v1 = (e + 7) >> 3;                       /* vectors of eight positions */
    const uint4* row = reinterpret_cast<const uint4*>(depth);
    int part = 0;
    CG_CHECK(s >= 0 && e >= s);
    /* four independent loads per lane and round: the sum of a 3 kb CNV is one trip to memory instead of three */
    for (int q0 = v0 + lane; q0 < v1; q0 += 128) {
        uint4 a[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int q = q0 + 32 * u;
            a[u] = q < v1 ? __ldg(row + q) : make_uint4(0u, 0u, 0u, 0u);
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int first = (q0 + 32 * u) << 3;
            uint4 v = a[u];
            if (first < s || first + 8 > e) {                       /* (a vector beyond v1 is all zero already) */
                unsigned w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int p0 = first + 2 * j, p1 = p0 + 1;
                    const unsigned keep = ((p0 >= s && p0 < e) ? 0xffffu : 0u) | ((p1 >= s && p1 < e) ? 0xffff0000u : 0u);
                    w[j] &= keep;
                }
                v = make_uint4(w[0], w[1], w[2], w[3]);
            }
            part = __dp2a_lo((int)v.w, 0x0101, __dp2a_lo((int)v.z, 0x0101, __dp2a_lo((int)v.y, 0x0101, __dp2a_lo((int)v.x, 0x0101, part))));
        }
    }
    return __reduce_add_sync(FULL, part);
}

/* ---- G CNVs per warp ---------------------------------------------------------
 * The work per CNV is irregular (a few windows to several thousand) and the exact fp32 chain
 * has a serial, branchy ramp-up: while E is small every addition changes its exponent.  A warp
 * therefore owns G consecutive CNVs, one per lane, and works in four phases:
 *   0  lane = CNV: record; warp = one CNV at a time: observed_rd from the chunk sums and the depth array
 *   1  lane = CNV: the first RAMP windows of the expected chain and the whole mappability chain,
 *      literally / run by run (add_run_f32, add_run_f64_fast): every lane busy with its own CNV
 *   2  warp = one CNV at a time, for the CNVs longer than RAMP windows: each lane takes 4
 *      consecutive GC windows (one coalesced 32-bit load), turns them into integer increments
 *      under the current exponent, and a warp scan finds how far the chain runs before anything
 *      unusual (binade boundary, tie).  Everything before that point is committed as ONE integer
 *      addition to E's bit pattern — bit-identical to the reference's per-base fadd chain — the
 *      offending lane replays its 4 windows through add_run_f32, and the scan resumes behind it
 *   3  lane = CNV: lgamma, three log-likelihoods, c-score, copy number
 * Phases 0, 1 and 3 run at full SIMT width; only the long, regular part of a long CNV uses the
 * warp cooperatively.  A launch needs n/G warps, which leaves the SMs to the depth pass it runs
 * underneath. */
#define GENO_WARPS 4                        /* warps per CTA */
#define GENO_RAMP 32                        /* windows of the expected chain walked lane-serially */
#define RUN_BUCKET_SHIFT 14                 /* run_idx[b] = first run whose end lies beyond position b << 14 */
__device__ __forceinline__ int cgk_run_index_words_dev(int32_t chrom_len) { return ((chrom_len + (1 << RUN_BUCKET_SHIFT) - 1) >> RUN_BUCKET_SHIFT) + 2; }

/* index of the flattened mappability tracks of a submission: per job one entry per 16384 positions */
__global__ void __launch_bounds__(256)
k_run_index(const __grid_constant__ cg_sub S, const __grid_constant__ cg_dev_job J1)
{
    const cg_dev_job* jobs = S.jobs ? S.jobs : &J1;
    __shared__ int s_blk0[CG_JMAX + 1];
    if (threadIdx.x <= S.n_jobs) s_blk0[threadIdx.x] = threadIdx.x < S.n_jobs ? jobs[threadIdx.x].run_blk0 : S.n_run_blocks;
    __syncthreads();
    int j = 0;
    for (int b = blockIdx.x; b < S.n_run_blocks; b += gridDim.x) {
        while (b >= s_blk0[j + 1]) j++;
        const cg_dev_job& J = jobs[j];
        const cg_maprun* __restrict__ runs = J.runs;
        const int n_runs = J.n_runs, n_buckets = cgk_run_index_words_dev(J.chrom_len) - 1;
        const int r = (b - s_blk0[j]) * 256 + threadIdx.x;
        if (r > n_runs) continue;
        const long long prev_end = r ? runs[r - 1].end : 0;
        const long long end = r < n_runs ? runs[r].end : ((long long)n_buckets << RUN_BUCKET_SHIFT) + 1;
        /* buckets whose first position p satisfies prev_end <= p < end: run r is the first one ending beyond p */
        for (long long q = (prev_end + (1 << RUN_BUCKET_SHIFT) - 1) >> RUN_BUCKET_SHIFT; q <= n_buckets && (q << RUN_BUCKET_SHIFT) < end; q++)
            J.run_idx[q] = r;
    }
}

int cgk_run_index_words(int32_t chrom_len) { return ((chrom_len + (1 << RUN_BUCKET_SHIFT) - 1) >> RUN_BUCKET_SHIFT) + 2; }

cudaError_t cgk_run_index(cudaStream_t s, const cg_sub& S, const cg_dev_job& J1)
{
    if (S.n_run_blocks <= 0) return cudaSuccess;
    int blocks = S.n_run_blocks;
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_run_index<<<blocks, 256, 0, s>>>(S, J1);
    return cudaGetLastError();
}

/* D(p) = sum of depth[0, p) mod 2^32, by one warp (every lane gets the result).  Chunk boundary: the supers
 * in front of p's super, the groups in front of its group, the chunks in front of its chunk (one lane each,
 * the carry of the depth pass's first version).  Inside the chunk: 16-byte loads of eight int16 depths. */
__device__ __forceinline__ int depth_sum_before(const int16_t* __restrict__ depth, const cg_counts& k, int chunk0, int p, int lane)
{
    /* depth points at the job's first position; the count arrays are indexed by chunk of the slot's chunk space */
    const int c = chunk0 + (p >> CG_CHUNK_SHIFT);
    const int g = c >> 5, sp = c >> 10;
    int part = 0;
    {
        const int i1 = (g << 5) + lane, i2 = (sp << 5) + lane;
        if (i1 < c) part += __ldg(&k.cnt1[i1]);
        if (i2 < g) part += __ldg(&k.cnt2[i2]);
        for (int j = (chunk0 >> 10) + lane; j < sp; j += 32) part += __ldg(&k.cnt3[j]);
    }
    const int base = (p >> CG_CHUNK_SHIFT) << CG_CHUNK_SHIFT;
    const uint4* row = reinterpret_cast<const uint4*>(depth + base);
    /* the chunk's vectors in front of p: at most eight per lane, asked for four at a time */
    const int nv = (p - base + 7) >> 3;
    for (int q0 = lane; q0 < nv; q0 += 128) {
        uint4 a[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int q = q0 + 32 * u;
            a[u] = q < nv ? __ldg(row + q) : make_uint4(0u, 0u, 0u, 0u);
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            uint4 v = a[u];
            const int keep = p - (base + (q0 + 32 * u) * 8);           /* positions of this vector in front of p (<= 0: none, zero already) */
            if (keep < 8 && keep > 0) {
                unsigned w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int left = keep - 2 * j;                     /* halves of word j in front of p */
                    w[j] = left >= 2 ? w[j] : left == 1 ? (w[j] & 0xffffu) : 0u;
                }
                v = make_uint4(w[0], w[1], w[2], w[3]);
            }
            part = __dp2a_lo((int)v.w, 0x0101, __dp2a_lo((int)v.z, 0x0101, __dp2a_lo((int)v.y, 0x0101, __dp2a_lo((int)v.x, 0x0101, part))));
        }
    }
    return __reduce_add_sync(FULL, part);
}

/* GENO_G CNVs per warp, one per lane in the lane = CNV phases.  The launch covers every CNV of the submission:
 * the CNVs of a (job, kind) segment are cut into groups of GENO_G, global group index -> segment through the
 * cnv0 prefix (in groups).  With one launch per chromosome (round 1) one CNV per warp was best: the launches were
 * small and latency-bound.  One launch per sample carries 15-40 k CNVs and is bound by issue slots instead (r02b,
 * G = 1: 1.9 IPC at 9.8 active threads per warp, every short CNV a serial walk on a single lane). */
#define GENO_G_MAX 8
/* CNVs per warp for a launch over n CNVs: one per warp while the warps fit the GPU at once (48 per SM) — then the
 * launch lasts as long as its longest CNV — and up to GENO_G_MAX when there are more (then lane efficiency counts) */
int cgk_geno_g(unsigned long long n_cnv, int n_sms)
{
    static int forced = -1;
    if (forced < 0) { const char* e = getenv("CG_GENO_G"); forced = e ? atoi(e) : 0; }     /* tuning only */
    if (forced >= 1 && forced <= 32) return forced;
    const unsigned long long cap = (unsigned long long)n_sms * 48;
    unsigned long long g = (n_cnv + cap - 1) / cap;
    return g < 1 ? 1 : g > GENO_G_MAX ? GENO_G_MAX : (int)g;
}
int cgk_geno_groups(unsigned long long n_cnv, int g) { return (int)((n_cnv + (unsigned)g - 1) / (unsigned)g); }

__global__ void __launch_bounds__(GENO_WARPS * 32, 8)
k_genotype(const __grid_constant__ cg_sub S, const __grid_constant__ cg_dev_job J1)
{
    const cg_dev_job* jobs = S.jobs ? S.jobs : &J1;
    __shared__ unsigned s_expb_all[GENO_WARPS][CG_GC_BINS + 3];   /* bit patterns of the warp's job's expected_read_depth[] */
    __shared__ int s_seg0[2 * CG_JMAX + 1];                       /* first global CNV index of (job, kind) = segment 2*job + kind */
    for (int i = threadIdx.x; i <= 2 * S.n_jobs; i += blockDim.x)
        s_seg0[i] = i < 2 * S.n_jobs ? jobs[i >> 1].cnv0[i & 1] : S.n_cnv;
    __syncthreads();
    const int lane = threadIdx.x & 31, G = S.geno_g;
    unsigned* s_expb = s_expb_all[threadIdx.x >> 5];
    const int n_warps = gridDim.x * GENO_WARPS;
    int seg = 0, cur_job = -1, cur_seg = -1;
    /* what the current segment says, in registers */
    const cg_cnv* __restrict__ cnvs = nullptr; cg_call* out = nullptr;
    const int16_t* __restrict__ depth = nullptr; const int8_t* __restrict__ gc = nullptr;
    const cg_maprun* __restrict__ runs = nullptr; const int32_t* __restrict__ run_idx = nullptr;
    int n_runs = 0, has_map = 0, chrom_len = 0, chunk0 = 0, is_dup = 0, n_windows = 0, n_seg = 0;
    bool use_map = false;
    const cg_counts counts = S.k;

    for (int gi = blockIdx.x * GENO_WARPS + (threadIdx.x >> 5); gi < S.n_cnv; gi += n_warps) {
        while (gi >= s_seg0[seg + 1]) seg++;                     /* a warp's indices ascend */
        if (seg != cur_seg) {
            const cg_dev_job& J = jobs[seg >> 1];
            is_dup = seg & 1;
            cnvs = J.cnv[is_dup]; out = J.calls[is_dup]; n_seg = J.n_cnv[is_dup];
            if ((seg >> 1) != cur_job) {
                depth = S.depth + ((size_t)J.chunk0 << CG_CHUNK_SHIFT); gc = J.gc;
                runs = J.runs; run_idx = J.run_idx; n_runs = J.n_runs; has_map = J.has_map;
                chrom_len = J.chrom_len; chunk0 = J.chunk0;
                n_windows = (chrom_len + 99) / 100;
                use_map = has_map && n_runs > 0;
                __syncwarp();
                for (int i = lane; i < CG_GC_BINS; i += 32) s_expb[i] = __float_as_uint(J.expected[i]);
                __syncwarp();
                cur_job = seg >> 1;
            }
            cur_seg = seg;
        }
        /* ---- phase 0: lane = CNV ---- */
        const int c = (gi - s_seg0[seg]) * G + lane;
        const bool valid = lane < G && c < n_seg;
        cg_cnv v = make_cnv(0, 0);
        CG_CHECK(!valid || (c >= 0 && c < n_seg));
        if (valid) v = cnvs[c];
        int s = v.start, e = v.end;
        if (valid && (s < 0 || e > chrom_len)) { atomicOr(S.sticky, CG_FLAG_RANGE); s = max(s, 0); e = min(e, chrom_len); }
        const bool any = valid && e > s;
        int obs = 0;                                    /* likelihood.c:120, the warp sums for one CNV at a time */
        for (unsigned todo = __ballot_sync(FULL, any); todo; todo &= todo - 1) {
            const int L = __ffs(todo) - 1;
            const int cs = __shfl_sync(FULL, s, L), ce = __shfl_sync(FULL, e, L);
            int d;
            if (ce - cs <= GENO_DIRECT) d = depth_sum_direct(depth, cs, ce, lane);
            else d = depth_sum_before(depth, counts, chunk0, ce, lane) - depth_sum_before(depth, counts, chunk0, cs, lane);
            if (lane == L) obs = d;
        }
        const int w_last = any ? (e - 1) / 100 : -1;
        int w = any ? s / 100 : 0;                      /* next window of the expected chain */
        int r_lo = 0, r_hi = 0;                         /* the first run ending beyond s lies in [r_lo, r_hi] */
        if (use_map && any) {
            CG_CHECK((s >> RUN_BUCKET_SHIFT) + 1 < cgk_run_index_words_dev(chrom_len));
            r_lo = __ldg(&run_idx[s >> RUN_BUCKET_SHIFT]);
            r_hi = __ldg(&run_idx[(s >> RUN_BUCKET_SHIFT) + 1]);
        }

        /* ---- phase 1: expected_rd ramp-up (likelihood.c:115-119), window by window ---- */
        float E = 0.0f;
        for (int i = 0; i < GENO_RAMP && w <= w_last; i++, w++) {
            CG_CHECK(w >= 0 && w < n_windows);
            int bin = (int)__ldg(&gc[w]);                                          /* sonic_interval.c:263-272 */
            if ((unsigned)bin > 100u) bin = 0;
            E = add_run_f32(E, __uint_as_float(s_expb[bin]), min(e, 100 * w + 100) - max(s, 100 * w));
        }

        /* ---- phase 1: first mappability run of the CNV (a crowded index bucket is bisected) ---- */
        if (use_map && any) {
            while (r_hi - r_lo > 8) {
                const int mid = (r_lo + r_hi) >> 1;
                CG_CHECK(mid >= 0 && mid < n_runs);
                if (__ldg(&runs[mid].end) > s) r_hi = mid; else r_lo = mid + 1;
            }
        }

        /* ---- phase 2: mappability_score (likelihood.c:122-123), one CNV at a time: 32 runs of the
         *      flattened track per step, fp64 twin of the expected chain below; the first block of
         *      the next CNV and the next block of this one are asked for ahead of use ---- */
        double S = 0.0;
        {
            unsigned todo = __ballot_sync(FULL, use_map && any);
            const cg_maprun none = {0x7fffffff, 0x7fffffff, 0.0f, 0};
            cg_maprun q_first = none;
            if (todo) {
                const int r0 = __shfl_sync(FULL, r_lo, __ffs(todo) - 1);
                if (r0 + lane < n_runs) q_first = runs[r0 + lane];
            }
            while (todo) {
                const int L = __ffs(todo) - 1;
                todo &= todo - 1;
                const int cs = __shfl_sync(FULL, s, L), ce = __shfl_sync(FULL, e, L), r0 = __shfl_sync(FULL, r_lo, L);
                cg_maprun q = q_first;
                q_first = none;
                if (todo) {
                    const int rn = __shfl_sync(FULL, r_lo, __ffs(todo) - 1);
                    if (rn + lane < n_runs) q_first = runs[rn + lane];
                }
                unsigned long long MB = 0ull;           /* bits of the double accumulator */
                for (int rb = r0; rb < n_runs; rb += 32) {
                    DBG(6);
                    const bool past = q.start >= ce;
                    const bool more = !__any_sync(FULL, past);
                    cg_maprun qn = none;
                    if (more && rb + 32 + lane < n_runs) qn = runs[rb + 32 + lane];
                    const int k = past ? 0 : max(0, min(ce, q.end) - max(cs, q.start));   /* runs ending before s add nothing */
                    const unsigned long long vbits = (unsigned long long)__double_as_longlong((double)q.val);
                    int pos = map_block_exact(MB, k, (double)q.val, lane) ? 32 : 0;
                    while (pos < 32) {
                        const unsigned ee = (unsigned)(MB >> 52);
                        bool bad = false;
                        unsigned long long t = 0ull;
                        if (lane >= pos && k > 0 && (vbits << 1) != 0ull) {
                            if (ee == 0u || ee >= 0x7ffu) {
                                bad = true;             /* zero / negative / non-finite accumulator */
                            } else {
                                const unsigned long long d = step_delta_f64(vbits, ee, bad);
                                t = (__umul64hi((unsigned long long)k, d) != 0ull) ? 0x10000000000000ull
                                                                                  : min((unsigned long long)k * d, 0x10000000000000ull);
                            }
                        }
                        const unsigned long long incl = warp_incl_scan_u64(t, lane);
                        const unsigned long long room = 0xfffffffffffffull - (MB & 0xfffffffffffffull);
                        const unsigned m = __ballot_sync(FULL, lane >= pos && (bad || incl > room));
                        DBG(4); if (__ballot_sync(FULL, lane >= pos && bad)) DBG(5);
                        if (m == 0u) { MB += __shfl_sync(FULL, incl, 31); break; }
                        const int f = __ffs(m) - 1;
                        MB += __shfl_sync(FULL, incl - t, f);
                        unsigned long long Mf = 0ull;
                        if (lane == f) Mf = (unsigned long long)__double_as_longlong(add_run_f64(__longlong_as_double((long long)MB), (double)q.val, k));
                        MB = __shfl_sync(FULL, Mf, f);
                        pos = f + 1;
                        if (pos >= 32) break;
                    }
                    if (!more) break;
                    q = qn;
                }
                if (lane == L) S = __longlong_as_double((long long)MB);
            }
        }

        /* ---- phase 2: the rest of the long expected chains, one CNV at a time across the warp ---- */
        unsigned todo = __ballot_sync(FULL, any && w <= w_last);
        while (todo) {
            const int L = __ffs(todo) - 1;
            todo &= todo - 1;
            const int cs = __shfl_sync(FULL, s, L), ce = __shfl_sync(FULL, e, L);
            const int w_from = __shfl_sync(FULL, w, L), cw_last = (ce - 1) / 100;
            unsigned B = __shfl_sync(FULL, __float_as_uint(E), L);                 /* bits of the accumulator, warp-uniform */
            int wb = w_from & ~3;
            unsigned g4 = 0u;
            if (wb + 4 * lane <= cw_last) g4 = load_gc4(gc, wb + 4 * lane, n_windows);
            for (; wb <= cw_last; wb += 128) {
                const int w4 = wb + 4 * lane;
                unsigned g4n = 0u;                      /* the block after this one */
                if (w4 + 128 <= cw_last) g4n = load_gc4(gc, w4 + 128, n_windows);
                DBG(0);
                unsigned vb[4]; int kk[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int wj = w4 + j;
                    int bin = (int)(signed char)(g4 >> (8 * j));
                    if ((unsigned)bin > 100u) bin = 0;
                    vb[j] = s_expb[bin];
                    kk[j] = (wj < w_from || wj > cw_last) ? 0 : min(ce, 100 * wj + 100) - max(cs, 100 * wj);
                }
                int pos = 0;
                for (;;) {
                    const unsigned ee = B >> 23;
                    /* a zero, negative or non-finite accumulator has no binade to replay in:
                     * any real addend then goes through the literal chain */
                    const bool acc_odd = ee == 0u || ee >= 255u;
                    bool bad = false;
                    unsigned t = 0u;
                    if (lane >= pos) {
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            if (kk[j] > 0 && (vb[j] << 1) != 0u) {
                                if (acc_odd) { bad = true; continue; }
                                const unsigned d = step_delta_f32(vb[j], ee, bad);
                                t += min((unsigned)kk[j] * d, 0x800000u);
                            }
                        }
                    }
                    const unsigned incl = warp_incl_scan_u32(t, lane);
                    const unsigned room = 0x7fffffu - (B & 0x7fffffu);
                    const unsigned m = __ballot_sync(FULL, lane >= pos && (bad || incl > room));
                    DBG(1); if (acc_odd) DBG(2); if (__ballot_sync(FULL, lane >= pos && bad)) DBG(3);
                    if (m == 0u) { B += __shfl_sync(FULL, incl, 31); break; }
                    const int f = __ffs(m) - 1;
                    B += __shfl_sync(FULL, incl - t, f);            /* everything before lane f, exactly */
                    unsigned Bf = 0u;
                    if (lane == f) {
                        float Ef = __uint_as_float(B);
#pragma unroll
                        for (int j = 0; j < 4; j++) Ef = add_run_f32(Ef, __uint_as_float(vb[j]), kk[j]);
                        Bf = __float_as_uint(Ef);
                    }
                    B = __shfl_sync(FULL, Bf, f);
                    pos = f + 1;
                    if (pos >= 32) break;
                }
                g4 = g4n;
            }
            if (lane == L) E = __uint_as_float(B);
        }

        /* ---- phase 3: observed_rd, likelihoods, c-score, copy number; lane = CNV ---- */
        if (valid) {
            const double k = (double)obs;
            const double lg = lgamma((double)(obs + 1));                           /* likelihood.c:104 */
            double lam_homo = is_dup ? (double)__fmul_rn(2.0f, E) : 0.0;           /* :133 / :155 (float product) */
            double lam_het = (is_dup ? 1.5 : 0.5) * (double)E;                     /* :134 / :156 */
            double lam_none = (double)E;                                           /* :135 / :157 */
            if (lam_homo == 0.0) lam_homo = 0.01;                                  /* :100-102 */
            if (lam_het == 0.0) lam_het = 0.01;
            if (lam_none == 0.0) lam_none = 0.01;
            const double lhomo = __dsub_rn(__dsub_rn(__dmul_rn(k, log(lam_homo)), lam_homo), lg);
            const double lhetero = __dsub_rn(__dsub_rn(__dmul_rn(k, log(lam_het)), lam_het), lg);
            const double lnone = __dsub_rn(__dsub_rn(__dmul_rn(k, log(lam_none)), lam_none), lg);
            int ih = c_double_to_int(lhomo), it = c_double_to_int(lhetero);
            bool homo_gt_het = lhomo > lhetero;
            /* a decision within 1e-9 relative of its boundary is taken from the double-double values (see dd_lpoisson);
             * bit-identical likelihoods (E = 0: all three lambdas are 0.01) are equal in any arithmetic */
            unsigned mf = 0;
            if (near_integer(lhomo) || near_integer(lhetero)) mf |= 1u;
            if (lhomo != lhetero && fabs(lhomo - lhetero) <= 1e-9 * fmax(fabs(lhomo), fabs(lhetero))) mf |= 2u;
            if (mf) {
                const dd_decision dcs = dd_decide(obs, lam_homo, lam_het);
                ih = dcs.ih; it = dcs.it; homo_gt_het = dcs.homo_gt_het != 0;
            }
            cg_call o;
            o.observed = obs;
            o.expected = E;
            o.mappability = has_map ? S / (double)(v.end - v.start) : 0.0;         /* :128 */
            o.lhomo = lhomo; o.lhetero = lhetero; o.lnone = lnone;
            o.score = (double)(ih < it ? it : ih) / lnone;                         /* :138 / :160 */
            if (!is_dup) o.copy_number = (__int2float_rn(obs) < __fdiv_rn(E, 4.0f)) ? 2 : 1;   /* :146-149 */
            else         o.copy_number = homo_gt_het ? 2 : 1;                                 /* :164-167 */
            o.margin_flags = mf;
            out[c] = o;
        }
    }
}

cudaError_t cgk_configure_genotype(void)
{
    const int pct = cgk_carveout();                       /* see cgk_configure_depth */
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(k_genotype, cudaFuncAttributePreferredSharedMemoryCarveout, pct)) != cudaSuccess) return e;
    return cudaFuncSetAttribute(k_run_index, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
}

cudaError_t cgk_genotype(cudaStream_t s, const cg_sub& S, const cg_dev_job& J1, int n_sms)
{
    if (S.n_cnv <= 0) return cudaSuccess;
    int blocks = (S.n_cnv + GENO_WARPS - 1) / GENO_WARPS;
    if (blocks > n_sms * 12) blocks = n_sms * 12;
    k_genotype<<<blocks, GENO_WARPS * 32, 0, s>>>(S, J1);
    return cudaGetLastError();
}
