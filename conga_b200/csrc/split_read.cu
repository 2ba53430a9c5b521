/* split_read.cu — read-pair support from split reads on sm_100a (K5; `--rp` with `--dups`).
 *
 * Replaces, per chromosome:
 *   build_hash_table + create_hash_table   split_read.c:338-455   10-mer index of the chromosome
 *   almostPerfect_match_seq_ref            split_read.c:76-204    re-mapping of each read half
 *   read_SplitReads' tests + determine_SvType   bam_data.c:29-154 (all but the SONIC satellite test)
 *   count_ReadPairs                        likelihood.c:41-94     rows x CNVs overlap counts
 * The host keeps what needs htslib or SONIC: the gate of bam_data.c:207, the two average-phred tests
 * of find_split_reads (split_read.c:236-246, :301-311), and sonic_is_satellite on the few rows that
 * come back.
 *
 * Index.  The reference chains, for every k-mer value, the positions where it starts, and drops
 * values with >= MAX_SR_HIT (50000) positions.  Here: a counting sort by the 20-bit k-mer code of
 * split_read.c:36-50 (positions whose ten bases are not all ACGT, or run past the sequence, have no
 * k-mer): k_sr_hist counts the positions of every value, k_sr_scan turns the 2^20 counts into each
 * value's span [lo, hi) of one position list, k_sr_scatter writes every position into its value's span
 * (one atomic slot counter per value; the order inside a span is arbitrary, which nothing below minds).
 * Hand-written; round 1 sorted (key, position) pairs with cub::DeviceRadixSort.
 *
 * Search.  One warp per read half.  The lanes walk the seed's span 32 positions at a time and test the
 * window |hit - pos| < 100000 (split_read.c:120,172); for each candidate inside it the lanes compare the bases
 * (hammingDistance, common.c:277-287) and the warp sums mismatches; forward hits are kept (at most 99
 * matter: with >= MAX_MAPPING hits the reference keeps none), reverse-complement hits are only
 * counted — they lower mapq = 60 / n but can never make a row (determine_SvType only pairs
 * FORWARD with FORWARD, bam_data.c:61).  Everything is order-independent, so the unordered row
 * list equals the reference's linked list as a multiset, which is all count_ReadPairs looks at.
 */
#include "cg_internal.h"

#define FULL 0xffffffffu
#define SR_K 10                       /* HASHKMERLEN  split_read.h:18 */
#define SR_BUCKETS (1 << (2 * SR_K))
#define SR_SENTINEL SR_BUCKETS        /* key of a position without a valid k-mer */
#define SR_LOOKAHEAD 100000           /* split_read.c:5 */
#define SR_MAX_MAPPING 100            /* split_read.h:14 */
#define SR_MAX_HIT 50000              /* split_read.h:13 */
#define SR_WARPS 8

__device__ __forceinline__ int base_code(unsigned char c)        /* (c & 0x6) >> 1 for ACGT, -1 otherwise (split_read.c:29-50,64-74) */
{
    return (c == 'A' || c == 'C' || c == 'G' || c == 'T') ? ((c & 0x6) >> 1) : -1;
}

/* the k-mer starting at position i, or -1 if its ten bases are not all ACGT or run past the sequence */
__device__ __forceinline__ int kmer_at(const unsigned char* __restrict__ ref, int i, int n)
{
    if (i + SR_K > n) return -1;
    unsigned v = 0;
    bool ok = true;
#pragma unroll
    for (int j = 0; j < SR_K; j++) {
        const int c = base_code(ref[i + j]);
        ok = ok && c >= 0;
        v = (v << 2) | (unsigned)(c & 3);
    }
    return ok ? (int)v : -1;
}

/* positions per k-mer value */
__global__ void k_sr_hist(const unsigned char* __restrict__ ref, int n, int* cnt)
{
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int k = kmer_at(ref, i, n);
        if (k >= 0) atomicAdd(&cnt[k], 1);
    }
}

/* exclusive scan of the 2^20 counts by ONE CTA of 32 warps: a warp owns 32768 consecutive values and walks them 32 at
 * a time with a shuffle scan and a running carry; the warps' totals are scanned in shared memory.  lo[k] = first slot
 * of value k in the position list, hi[k] = lo[k] + its count; the counts are cleared for k_sr_scatter's slot counters */
__global__ void __launch_bounds__(1024)
k_sr_scan(int* cnt, int* lo, int* hi)
{
    __shared__ int s_tot[32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int per_warp = SR_BUCKETS / 32, base = w * per_warp;
    int carry = 0;
    for (int i = 0; i < per_warp; i += 32) {
        const int v = cnt[base + i + lane];
        int x = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(FULL, x, d); if (lane >= d) x += o; }
        lo[base + i + lane] = carry + x - v;                  /* relative to the warp's first value for now */
        carry += __shfl_sync(FULL, x, 31);
    }
    if (lane == 0) s_tot[w] = carry;
    __syncthreads();
    int off = 0;
    for (int k = 0; k < w; k++) off += s_tot[k];
    for (int i = 0; i < per_warp; i += 32) {
        const int k = base + i + lane;
        const int l = lo[k] + off;
        lo[k] = l;
        hi[k] = l + cnt[k];
        cnt[k] = 0;
    }
}

/* every position into its value's span */
__global__ void k_sr_scatter(const unsigned char* __restrict__ ref, int n, const int* __restrict__ lo, int* fill, int* pos_list)
{
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int k = kmer_at(ref, i, n);
        if (k >= 0) {
            const int slot = lo[k] + atomicAdd(&fill[k], 1);
            CG_CHECK(slot >= 0 && slot < n);
            pos_list[slot] = i;
        }
    }
}

struct sr_search_args {
    const unsigned char* ref; int ref_len;          /* ref_seq and bp_cnt (ref[bp_cnt] is the terminator) */
    const int* pos_list; const int* lo; const int* hi;     /* positions grouped by k-mer value, span of every value */
    const cg_half* halves; long long n_halves;
    const unsigned char* bases;
    int chrom_len, mq_threshold;
    cg_splitrow* rows; unsigned* n_rows; unsigned cap;
    unsigned long long* n_mappings;                 /* diagnostics: (half, mapping) pairs the reference would have listed */
};

/* seed of a string held one base per lane-strided slot: code of its first ten bases or -1 */
__device__ __forceinline__ int seed_of(const unsigned char* s, int len, bool reverse)
{
    if (len < SR_K) return -1;
    unsigned v = 0;
    for (int j = 0; j < SR_K; j++) {
        unsigned char c = reverse ? s[len - 1 - j] : s[j];
        if (reverse) c = c == 'A' ? 'T' : c == 'T' ? 'A' : c == 'G' ? 'C' : c == 'C' ? 'G' : c;   /* split_read.c:138-150 */
        const int b = base_code(c);
        if (b < 0) return -1;
        v = (v << 2) | (unsigned)b;
    }
    return (int)v;
}

__global__ void __launch_bounds__(SR_WARPS * 32)
k_sr_search(const sr_search_args a)
{
    __shared__ int s_hits[SR_WARPS][SR_MAX_MAPPING];
    __shared__ unsigned char s_str[SR_WARPS][1024];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const long long n_warps = (long long)gridDim.x * SR_WARPS;
    for (long long h = (long long)blockIdx.x * SR_WARPS + w; h < a.n_halves; h += n_warps) {
        const cg_half hf = a.halves[h];
        const int len = min((int)hf.len, 1024);
        for (int i = lane; i < len; i += 32) s_str[w][i] = a.bases[(size_t)hf.seq_off + i];
        __syncwarp();
        const unsigned char* str = s_str[w];
        const int dist_max = (int)(0.05 * (double)len);                      /* split_read.c:94 */
        int n_fwd = 0, n_all = 0;
        if (len >= SR_K) {                                                    /* split_read.c:96-97 */
            for (int rev = 0; rev < 2; rev++) {
                if (rev && n_fwd >= SR_MAX_MAPPING) break;                    /* split_read.c:133 */
                const int seed = seed_of(str, len, rev != 0);
                if (seed < 0) continue;
                CG_CHECK(seed >= 0 && seed < SR_BUCKETS);
                const int lo = a.lo[seed], hi = a.hi[seed];
                CG_CHECK(lo >= 0 && hi >= lo && hi <= a.ref_len);
                if (hi - lo >= SR_MAX_HIT) continue;                          /* init_hash_table, split_read.c:440-448 */
                bool done = false;
                for (int c0 = lo; c0 < hi && !done; c0 += 32) {
                    /* 32 positions of the span at a time; the ones inside the window abs(hit - pos) < SR_LOOKAHEAD are candidates */
                    const int mine = c0 + lane < hi ? a.pos_list[c0 + lane] : 0x7fffffff;
                    const long long dlt = (long long)mine - (long long)hf.pos;
                    unsigned cand = __ballot_sync(FULL, c0 + lane < hi && dlt > -(long long)SR_LOOKAHEAD && dlt < (long long)SR_LOOKAHEAD);
                    while (cand) {
                        const int src = __ffs(cand) - 1;
                        cand &= cand - 1;
                        const int hit = __shfl_sync(FULL, mine, src);
                        int d = 0;
                        for (int i = lane; i < len; i += 32) {
                            unsigned char q = rev ? str[len - 1 - i] : str[i];
                            if (rev) q = q == 'A' ? 'T' : q == 'T' ? 'A' : q == 'G' ? 'C' : q == 'C' ? 'G' : q;
                            /* the reference compares against ref_seq[hit + i]; at and beyond the terminator nothing equals a base */
                            const unsigned char t = hit + i < a.ref_len ? a.ref[hit + i] : (unsigned char)0;
                            d += t != q;
                        }
                        d = __reduce_add_sync(FULL, d);
                        if (d <= dist_max) {
                            if (!rev) { if (n_fwd < SR_MAX_MAPPING && lane == 0) s_hits[w][n_fwd] = hit; n_fwd++; }
                            n_all++;
                            if (n_all > SR_MAX_MAPPING + 1) { done = true; break; }   /* already rejected: >= MAX_MAPPING mappings keep none */
                        }
                    }
                }
            }
        }
        __syncwarp();
        if (n_all > 0 && n_all < SR_MAX_MAPPING) {                           /* split_read.c:184 */
            if (lane == 0) atomicAdd(a.n_mappings, (unsigned long long)n_all);
            const int mapq = 60 / n_all;                                      /* split_read.c:193 */
            const int pos = hf.pos;
            const int len_split = hf.read_len / 2, len_read = hf.read_len - len_split;   /* bam_data.c:44-45 */
            for (int k = lane; k < n_fwd; k += 32) {
                const int pm = s_hits[w][k];
                /* read_SplitReads, bam_data.c:104-106 (the satellite term is the host's) */
                if (!((int)hf.mapq > a.mq_threshold && mapq > a.mq_threshold && pos > 0 && pm > 0 && pos < a.chrom_len && pm < a.chrom_len))
                    continue;
                int p11, p12, p21, p22;
                if (pos < pm) { p11 = pos; p12 = pos + len_read; p21 = pm; p22 = pm + len_split; }
                else if (pm < pos) { p11 = pm; p12 = pm + len_split; p21 = pos; p22 = pos + len_read; }
                else continue;
                if (p12 >= p21) continue;                                      /* bam_data.c:62-63 */
                const bool r2 = hf.is_read2 != 0;
                /* both orientations are FORWARD here (bam_data.c:65-78) */
                unsigned char type;
                if ((pos < pm && !r2) || (pos > pm && r2)) type = 'D';
                else type = 'E';
                const unsigned slot = atomicAdd(a.n_rows, 1u);
                if (slot < a.cap) {
                    cg_splitrow row;
                    row.pos = pos; row.pos_map = pm;
                    row.left_start = p11; row.left_end = p12; row.right_start = p21; row.right_end = p22;
                    row.sv_type = type;
                    for (int z = 0; z < 7; z++) row.pad[z] = 0;
                    a.rows[slot] = row;
                }
            }
        }
        __syncwarp();
    }
}

/* count_ReadPairs, likelihood.c:41-94: one thread per row, every CNV of the row's kind */
__global__ void k_sr_count(const cg_splitrow* __restrict__ rows, int n_rows, const cg_cnv* __restrict__ dels, int n_del,
                           const cg_cnv* __restrict__ dups, int n_dup, int* border_rp, int* rp)
{
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_rows; i += stride) {
        const cg_splitrow r = rows[i];
        if (r.sv_type == 'E') {
            for (int k = 0; k < n_dup; k++) {
                const int s = dups[k].start, e = dups[k].end;
                if (r.left_end >= s - 5000 && r.left_end <= e + 5000 && r.right_start <= e + 5000 && r.right_start >= s - 5000)
                    atomicAdd(&rp[k], 1);                                      /* likelihood.c:66-71 */
            }
        } else if (r.sv_type == 'D') {
            for (int k = 0; k < n_del; k++) {
                const int s = dels[k].start, e = dels[k].end;
                if (r.left_end <= s + 100 && r.left_end >= s - 5000 && r.right_start >= e - 100 && r.right_start <= e + 5000)
                    atomicAdd(&border_rp[k], 1);                               /* likelihood.c:81-86 */
            }
        }
    }
}

/* ---- launchers ------------------------------------------------------------------------ */
cudaError_t cgk_sr_index(cudaStream_t s, const unsigned char* ref, int n, int* cnt, int* pos_list, int* lo, int* hi, int n_sms)
{
    cudaError_t e;
    if ((e = cudaMemsetAsync(cnt, 0, sizeof(int) * SR_BUCKETS, s)) != cudaSuccess) return e;
    if (n > 0) k_sr_hist<<<n_sms * 8, 256, 0, s>>>(ref, n, cnt);
    k_sr_scan<<<1, 1024, 0, s>>>(cnt, lo, hi);
    if (n > 0) k_sr_scatter<<<n_sms * 8, 256, 0, s>>>(ref, n, lo, cnt, pos_list);
    return cudaGetLastError();
}

cudaError_t cgk_sr_search(cudaStream_t s, const unsigned char* ref, int ref_len, const int* pos_list, const int* lo, const int* hi,
                          const cg_half* halves, long long n_halves, const unsigned char* bases, int chrom_len, int mq_threshold,
                          cg_splitrow* rows, unsigned* n_rows, unsigned cap, unsigned long long* n_mappings, int n_sms)
{
    if (n_halves <= 0) return cudaSuccess;
    sr_search_args a;
    a.ref = ref; a.ref_len = ref_len; a.pos_list = pos_list; a.lo = lo; a.hi = hi;
    a.halves = halves; a.n_halves = n_halves; a.bases = bases; a.chrom_len = chrom_len; a.mq_threshold = mq_threshold;
    a.rows = rows; a.n_rows = n_rows; a.cap = cap; a.n_mappings = n_mappings;
    long long blocks = (n_halves + SR_WARPS - 1) / SR_WARPS;
    if (blocks > (long long)n_sms * 8) blocks = (long long)n_sms * 8;
    k_sr_search<<<(int)blocks, SR_WARPS * 32, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t cgk_sr_count(cudaStream_t s, const cg_splitrow* rows, int n_rows, const cg_cnv* dels, int n_del, const cg_cnv* dups, int n_dup,
                         int* border_rp, int* rp, int n_sms)
{
    if (n_rows <= 0) return cudaSuccess;
    int blocks = (n_rows + 127) / 128;
    if (blocks > n_sms * 8) blocks = n_sms * 8;
    k_sr_count<<<blocks, 128, 0, s>>>(rows, n_rows, dels, n_del, dups, n_dup, border_rp, rp);
    return cudaGetLastError();
}

cudaError_t cgk_configure_split(void)
{
    const int pct = cgk_carveout();
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(k_sr_hist, cudaFuncAttributePreferredSharedMemoryCarveout, pct)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_sr_scan, cudaFuncAttributePreferredSharedMemoryCarveout, pct)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_sr_scatter, cudaFuncAttributePreferredSharedMemoryCarveout, pct)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_sr_search, cudaFuncAttributePreferredSharedMemoryCarveout, pct)) != cudaSuccess) return e;
    return cudaFuncSetAttribute(k_sr_count, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
}
