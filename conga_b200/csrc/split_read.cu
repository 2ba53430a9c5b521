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
 * values with >= MAX_SR_HIT (50000) positions.  Here: one 21-bit key per position (the 20-bit k-mer
 * code of split_read.c:36-50, or a sentinel where ref[i..i+9] is not all ACGT or runs past the
 * sequence), a stable radix sort of (key, position) — positions stay ascending inside a k-mer — and
 * two int32[2^20] arrays with each k-mer's span in the sorted position list.
 *
 * Search.  One warp per read half.  The window test |hit - pos| < 100000 (split_read.c:120,172) is a
 * binary search into the k-mer's ascending positions; for each candidate the lanes compare the bases
 * (hammingDistance, common.c:277-287) and the warp sums mismatches; forward hits are kept (at most 99
 * matter: with >= MAX_MAPPING hits the reference keeps none), reverse-complement hits are only
 * counted — they lower mapq = 60 / n but can never make a row (determine_SvType only pairs
 * FORWARD with FORWARD, bam_data.c:61).  Everything is order-independent, so the unordered row
 * list equals the reference's linked list as a multiset, which is all count_ReadPairs looks at.
 */
#include "cg_internal.h"
#include <cub/device/device_radix_sort.cuh>

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

/* key of every position: the k-mer starting there, if all ten bases are ACGT and inside the sequence */
__global__ void k_sr_keys(const unsigned char* __restrict__ ref, int n, unsigned* keys, int* vals)
{
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        unsigned key = SR_SENTINEL;
        if (i + SR_K <= n) {
            unsigned v = 0;
            bool ok = true;
#pragma unroll
            for (int j = 0; j < SR_K; j++) {
                const int c = base_code(ref[i + j]);
                ok = ok && c >= 0;
                v = (v << 2) | (unsigned)(c & 3);
            }
            if (ok) key = v;
        }
        keys[i] = key;
        vals[i] = i;
    }
}

/* span of every k-mer value in the sorted list */
__global__ void k_sr_bounds(const unsigned* __restrict__ keys, int n, int* lo, int* hi)
{
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const unsigned k = keys[i];
        if (k >= SR_SENTINEL) continue;
        if (i == 0 || keys[i - 1] != k) lo[k] = i;
        if (i == n - 1 || keys[i + 1] != k) hi[k] = i + 1;
    }
}

struct sr_search_args {
    const unsigned char* ref; int ref_len;          /* ref_seq and bp_cnt (ref[bp_cnt] is the terminator) */
    const int* pos_sorted; const int* lo; const int* hi;
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
                int lo = a.lo[seed], hi = a.hi[seed];
                if (hi - lo >= SR_MAX_HIT) continue;                          /* init_hash_table, split_read.c:440-448 */
                /* first position > pos - SR_LOOKAHEAD */
                const int want = hf.pos - SR_LOOKAHEAD;
                int l = lo, r = hi;
                while (l < r) { const int m = (l + r) >> 1; if (a.pos_sorted[m] > want) r = m; else l = m + 1; }
                for (int c = l; c < hi; c++) {
                    const int hit = a.pos_sorted[c];
                    if (hit >= hf.pos + SR_LOOKAHEAD) break;                  /* abs(hit - pos) < SR_LOOKAHEAD */
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
                        if (n_all > SR_MAX_MAPPING + 1) break;                /* already rejected: >= MAX_MAPPING mappings keep none */
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
size_t cgk_sr_sort_temp_bytes(int n)
{
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const unsigned*)nullptr, (unsigned*)nullptr, (const int*)nullptr, (int*)nullptr, n, 0, 2 * SR_K + 1);
    return bytes;
}

cudaError_t cgk_sr_index(cudaStream_t s, const unsigned char* ref, int n, unsigned* keys_a, unsigned* keys_b, int* vals_a, int* vals_b,
                         void* temp, size_t temp_bytes, int* lo, int* hi, int n_sms)
{
    cudaError_t e;
    if ((e = cudaMemsetAsync(lo, 0, sizeof(int) * SR_BUCKETS, s)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(hi, 0, sizeof(int) * SR_BUCKETS, s)) != cudaSuccess) return e;
    if (n <= 0) return cudaSuccess;
    k_sr_keys<<<n_sms * 8, 256, 0, s>>>(ref, n, keys_a, vals_a);
    if ((e = cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_a, keys_b, vals_a, vals_b, n, 0, 2 * SR_K + 1, s)) != cudaSuccess) return e;
    k_sr_bounds<<<n_sms * 8, 256, 0, s>>>(keys_b, n, lo, hi);
    return cudaGetLastError();
}

cudaError_t cgk_sr_search(cudaStream_t s, const unsigned char* ref, int ref_len, const int* pos_sorted, const int* lo, const int* hi,
                          const cg_half* halves, long long n_halves, const unsigned char* bases, int chrom_len, int mq_threshold,
                          cg_splitrow* rows, unsigned* n_rows, unsigned cap, unsigned long long* n_mappings, int n_sms)
{
    if (n_halves <= 0) return cudaSuccess;
    sr_search_args a;
    a.ref = ref; a.ref_len = ref_len; a.pos_sorted = pos_sorted; a.lo = lo; a.hi = hi;
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
    if ((e = cudaFuncSetAttribute(k_sr_keys, cudaFuncAttributePreferredSharedMemoryCarveout, pct)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_sr_bounds, cudaFuncAttributePreferredSharedMemoryCarveout, pct)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_sr_search, cudaFuncAttributePreferredSharedMemoryCarveout, pct)) != cudaSuccess) return e;
    return cudaFuncSetAttribute(k_sr_count, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
}
