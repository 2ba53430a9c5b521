/* genotype.cu — per-CNV sums and Poisson genotype likelihoods on sm_100a (K4).
 *
 * Replaces find_depths -> calculate_likelihood_CNV (likelihood.c:290-308, :108-169)
 * and lpoisson (likelihood.c:96-105).
 *
 * The reference walks every base of every CNV with three accumulators
 * (likelihood.c:115-124).  Here:
 *   observed_rd  (int)    = P[end-1] - P[start-1] from the int32 prefix the depth
 *                           pass wrote: two loads per CNV.
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

#define FULL 0xffffffffu

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

/* ---- per-CNV sums ----------------------------------------------------------- */
#define WIN_CHUNK 8     /* GC windows a lane advances between refill checks */
#define RUN_CHUNK 4     /* mappability runs a lane advances between refill checks */

__global__ void __launch_bounds__(128)
k_cnv_sums(const cg_cnv* __restrict__ cnvs, int n, const int32_t* __restrict__ prefix,
           const int8_t* __restrict__ gc, const float* __restrict__ expected,
           const cg_maprun* __restrict__ runs, int n_runs, int has_map, int32_t chrom_len,
           unsigned int* flags, cg_call* out, int* work_counter)
{
    __shared__ float s_exp[CG_GC_BINS];
    for (int i = threadIdx.x; i < CG_GC_BINS; i += blockDim.x) s_exp[i] = expected[i];
    __syncthreads();
    const int lane = threadIdx.x & 31;

    int c = -1;                 /* CNV this lane is working on */
    int s = 0, e = 0;           /* clipped interval */
    int w = 0, w_last = -1;     /* next / last GC window of the fp32 chain */
    int r = 0;                  /* next mappability run */
    int phase = 0;              /* 0: expected chain, 1: mappability chain */
    float E = 0.0f;
    double M = 0.0;
    bool more = true;           /* the shared counter may still hold work */

    for (;;) {
        /* refill idle lanes, one counter bump per warp */
        const unsigned need = __ballot_sync(FULL, c < 0 && more);
        if (need) {
            int first = 0;
            if (lane == __ffs(need) - 1) first = atomicAdd(work_counter, __popc(need));
            first = __shfl_sync(FULL, first, __ffs(need) - 1);
            if (c < 0 && more) {
                const int mine = first + __popc(need & ((1u << lane) - 1u));
                if (mine < n) {
                    c = mine;
                    const cg_cnv v = cnvs[c];
                    s = v.start; e = v.end;
                    if (s < 0 || e > chrom_len) { atomicOr(flags, CG_FLAG_RANGE); s = max(s, 0); e = min(e, chrom_len); }
                    E = 0.0f; M = 0.0; phase = 0; r = 0;
                    if (e > s) { w = s / 100; w_last = (e - 1) / 100; } else { w = 0; w_last = -1; }
                } else more = false;
            }
        }
        if (!__any_sync(FULL, c >= 0)) break;
        if (c < 0) continue;

        if (phase == 0) {
            /* likelihood.c:115-119: expected_rd += expected[gc(i)] for i in [s,e), in order */
#pragma unroll 1
            for (int j = 0; j < WIN_CHUNK && w <= w_last; j++, w++) {
                const int k = min(e, 100 * w + 100) - max(s, 100 * w);
                int bin = (int)__ldg(&gc[w]);
                if ((unsigned)bin > 100u) bin = 0;
                E = add_run_f32(E, s_exp[bin], k);
            }
            if (w > w_last) {
                phase = 1;
                if (has_map && e > s && n_runs > 0) {
                    /* first run whose end lies beyond s (runs are disjoint and sorted) */
                    int lo = 0, hi = n_runs;
                    while (lo < hi) { const int mid = (lo + hi) >> 1; if (runs[mid].end > s) hi = mid; else lo = mid + 1; }
                    r = lo;
                } else r = n_runs;
            }
        } else {
            /* likelihood.c:122-123: mappability_score += mappability[i], in order */
#pragma unroll 1
            for (int j = 0; j < RUN_CHUNK && r < n_runs; j++, r++) {
                const cg_maprun q = runs[r];
                if (q.start >= e) { r = n_runs; break; }
                const int k = min(e, q.end) - max(s, q.start);
                M = add_run_f64(M, (double)q.val, k);
            }
            if (r >= n_runs) {
                cg_call* o = &out[c];
                const cg_cnv v = cnvs[c];
                int obs = 0;                                   /* likelihood.c:120 */
                if (e > s) obs = prefix[e - 1] - (s > 0 ? prefix[s - 1] : 0);
                o->observed = obs;
                o->expected = E;
                o->mappability = has_map ? M / (double)(v.end - v.start) : 0.0;   /* likelihood.c:128 */
                c = -1;
            }
        }
    }
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
    return d <= 1e-7 * fmax(1.0, a);
}

__global__ void __launch_bounds__(128)
k_cnv_likelihood(cg_call* out, int n, int is_dup)
{
    const int gid = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31, g = gid & 3, lead = lane & ~3;
    int c = gid >> 2;
    const bool live = c < n;
    if (!live) c = n - 1;
    const int   obs = out[c].observed;
    const float E   = out[c].expected;

    double term;
    if (g == 3) {
        term = lgamma((double)(obs + 1));                          /* likelihood.c:104 */
    } else {
        double lambda;
        if (g == 2)       lambda = (double)E;                      /* lnone:   :135 / :157 */
        else if (g == 1)  lambda = (is_dup ? 1.5 : 0.5) * (double)E;   /* lhetero: :134 / :156 */
        else              lambda = is_dup ? (double)__fmul_rn(2.0f, E) : 0.0;   /* lhomo: :133 / :155 */
        if (lambda == 0.0) lambda = 0.01;                          /* :100-102 */
        term = __dsub_rn(__dmul_rn((double)obs, log(lambda)), lambda);
    }
    const double lg = __shfl_sync(FULL, term, lead + 3);
    const double l  = __dsub_rn(term, lg);
    const double lhomo   = __shfl_sync(FULL, l, lead + 0);
    const double lhetero = __shfl_sync(FULL, l, lead + 1);
    const double lnone   = __shfl_sync(FULL, l, lead + 2);
    if (!live || g != 0) return;

    const int ih = c_double_to_int(lhomo), it = c_double_to_int(lhetero);
    cg_call* o = &out[c];
    o->lhomo = lhomo; o->lhetero = lhetero; o->lnone = lnone;
    o->score = (double)(ih < it ? it : ih) / lnone;                /* :138 / :160 */
    if (!is_dup) o->copy_number = (__int2float_rn(obs) < __fdiv_rn(E, 4.0f)) ? 2 : 1;   /* :146-149 */
    else         o->copy_number = (lhomo > lhetero) ? 2 : 1;                           /* :164-167 */
    unsigned mf = 0;
    if (near_integer(lhomo) || near_integer(lhetero)) mf |= 1u;
    if (fabs(lhomo - lhetero) <= 1e-7 * fmax(fabs(lhomo), fabs(lhetero))) mf |= 2u;
    o->margin_flags = mf;
}

__global__ void k_reset_counter(int* p) { *p = 0; }

cudaError_t cgk_genotype(cudaStream_t s, const cg_cnv* cnvs, int n, int is_dup,
                         const int32_t* prefix, const int8_t* gc, const float* expected,
                         const cg_maprun* runs, int n_runs, int has_map,
                         int32_t chrom_len, unsigned int* flags, cg_call* out, int* work_counter)
{
    if (n <= 0) return cudaSuccess;
    k_reset_counter<<<1, 1, 0, s>>>(work_counter);
    int blocks = (n + 127) / 128;
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_cnv_sums<<<blocks, 128, 0, s>>>(cnvs, n, prefix, gc, expected, runs, n_runs, has_map,
                                      chrom_len, flags, out, work_counter);
    const long long lanes = 4ll * n;
    k_cnv_likelihood<<<(int)((lanes + 127) / 128), 128, 0, s>>>(out, n, is_dup);
    return cudaGetLastError();
}
