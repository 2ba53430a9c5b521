/* depth_pass.cu — the fused read-depth pass of one chromosome on sm_100a.
 *
 * Replaces, in ONE sweep over the chromosome's positions (no intermediate HBM
 * round trip):
 *   K1  read_depth[pos]++ per passing record        bam_data.c:201-217, read_distribution.c:12-18
 *   K2  chromosome sum + 101 GC-bin sums/counts     read_distribution.c:27-46, :49-73
 *   K3  per-CNV depth sums in O(1): NOT a prefix array — the chunk / group / super depth sums left by the
 *       record sweep plus the depth array itself serve likelihood.c:115-121 (genotype.cu, depth_sum_before)
 *
 * Data layout in HBM (per chromosome slot, reused across chromosomes):
 *   reads      cg_read[N]            8 B records, coordinate sorted
 *   chunk_lo   int32[n_chunks+1]     first record of every 2048-position chunk
 *   cnt1/2/3   int32[...]            sum of the int16 depths per chunk / 32 chunks / 1024 chunks
 *   depth      int16[n_chunks*2048]  the reference's `short read_depth[]` (wraps mod 2^16)
 *   stats      int64[204]            cg_chrom_stats row
 *
 * The scan carry never travels between warps.  A chunk's depth sum is the number of
 * passing records that fall in it (every read adds exactly 1) unless a position
 * collects >= 32768 reads and the int16 wraps; k_mark_batch therefore counts records
 * per chunk / group / super while it checks the pushed batch (the records are read
 * once, 8 B each, against 6 B per POSITION for the pass) and k_rare_paths corrects the
 * counts for wrapped pile-ups.  A chunk's carry is then three short coalesced loads:
 * the supers in front of its super, the groups in front of its group inside the super,
 * the chunks in front of it inside the group.  The pass is embarrassingly parallel
 * over chunks: no look-back, no spinning, no block barrier, no separate scan kernel.
 *
 * Algorithmic bytes per chromosome: 8·N (records) + 2·L (depth, written once) + L/100 (GC bytes); the
 * pass is HBM-write bound.  (SURVEY §8d's plan also wrote a 4·L int32 prefix; rounds r01a-r01g did, at
 * 6 B per position.  It was only ever read at two positions per CNV, see genotype.cu.)
 *
 * Pass kernel shape: persistent warps pulling chunks from a counter; a warp owns 4 KB of
 * packed u16 bins in shared memory.  Per chunk: warp-aggregated shared histogram of the
 * chunk's records -> the eight 256-position rows are summed per lane and their eight
 * warp scans run interleaved (eight independent shuffle chains) -> per row one 16-byte
 * LDS (which is also the 16-byte int16 output vector) and two 16-byte prefix stores,
 * bins zeroed behind -> per-window sums from the prefixes at window ends into 101
 * shared GC accumulators.
 */
#include "cg_internal.h"
#include <cstdlib>

#define FULL 0xffffffffu
#define NT   CG_PASS_THREADS
#define NWARP (NT / 32)
#define ROWS (CG_CHUNK / 256)         /* rows of 256 positions per chunk */
#define MAXWIN 24                     /* GC windows that can overlap one chunk: 2048/100 + 2 */
#define LVL (CG_GROUP_SHIFT - CG_CHUNK_SHIFT)      /* 5: 32 children per level */

static_assert(CG_SUPER_SHIFT - CG_GROUP_SHIFT == LVL && LVL == 5, "32 children per level, one per lane");
static_assert((1 << CG_CHUNK_SHIFT) == CG_CHUNK && ROWS == 8, "chunk geometry");

__device__ __forceinline__ int sext16(unsigned v) { return (int)(short)(v & 0xffffu); }

__device__ __forceinline__ int warp_incl_scan(int v, int lane)
{
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int o = __shfl_up_sync(FULL, v, d);
        if (lane >= d) v += o;
    }
    return v;
}

/* streaming stores: depth/prefix are written once and not re-read by this kernel */
__device__ __forceinline__ void st_cs_v4(void* p, unsigned a, unsigned b, unsigned c, unsigned d)
{
    asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

/* ------------------------------------------------------------------------- */
static inline int n_groups_of(int n_chunks) { return (n_chunks + 31) >> LVL; }
static inline int n_supers_of(int n_chunks) { return (n_chunks + 1023) >> (2 * LVL); }
static inline size_t pad4(size_t n) { return (n + 3) & ~(size_t)3; }

size_t cgk_counts_words(int n_chunks)
{
    return pad4((size_t)n_chunks + 1) + pad4((size_t)n_chunks) + pad4((size_t)n_groups_of(n_chunks)) + pad4((size_t)n_supers_of(n_chunks));
}

cg_counts cgk_counts_layout(int32_t* base, int n_chunks)
{
    cg_counts k;
    k.cnt1 = base;                                            /* the three zeroed arrays first, chunk_lo last */
    k.cnt2 = k.cnt1 + pad4((size_t)n_chunks);
    k.cnt3 = k.cnt2 + pad4((size_t)n_groups_of(n_chunks));
    k.chunk_lo = k.cnt3 + pad4((size_t)n_supers_of(n_chunks));
    return k;
}

/* chromosome start: clear the counts, the stats row and the control words; every chunk
 * bound starts "behind the last record" (the pass clamps it to the record count) */
__global__ void k_chrom_init(int4* base4, int n_zero_vec, int n_vec, long long* stats_row, unsigned int* ctl)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
    for (int j = i; j < n_vec; j += stride)
        base4[j] = j < n_zero_vec ? make_int4(0, 0, 0, 0) : make_int4(0x7fffffff, 0x7fffffff, 0x7fffffff, 0x7fffffff);
    if (i < CG_STATS_WORDS) stats_row[i] = 0;
    if (i < 4) ctl[i] = 0u;
}

cudaError_t cgk_chrom_init(cudaStream_t s, int32_t* counts_base, int n_chunks, long long* stats_row, unsigned int* ctl)
{
    const cg_counts k = cgk_counts_layout(counts_base, n_chunks);
    const int n_zero_vec = (int)((k.chunk_lo - counts_base) / 4);
    const int n_vec = (int)(cgk_counts_words(n_chunks) / 4);
    int blocks = (n_vec + 255) / 256;
    if (blocks < 1) blocks = 1;
    if (blocks > 148 * 2) blocks = 148 * 2;
    k_chrom_init<<<blocks, 256, 0, s>>>(reinterpret_cast<int4*>(counts_base), n_zero_vec, n_vec, stats_row, ctl);
    return cudaGetLastError();
}

/* ------------------------------------------------------------------------- */
/* per pushed batch, one sweep over its records: sortedness and range checks,  */
/* chunk boundaries, passing records per chunk/group/super, pile-up probe      */
__device__ __forceinline__ int chunk_of(int pos, int n_chunks)
{
    int t = pos >> CG_CHUNK_SHIFT;            /* arithmetic shift: negative pos -> negative chunk */
    t = t < 0 ? 0 : t;
    return t >= n_chunks ? n_chunks - 1 : t;
}

/* first lane of a run of equal keys in the warp, and the run's length */
__device__ __forceinline__ bool run_leader(int key, int lane, int& cnt)
{
    const int prev = __shfl_up_sync(FULL, key, 1);
    const bool leader = lane == 0 || key != prev;
    const unsigned lm = __ballot_sync(FULL, leader);
    const unsigned above = lane == 31 ? 0u : (lm & ~((2u << lane) - 1u));
    cnt = (above ? __ffs(above) - 1 : 32) - lane;
    return leader;
}

__global__ void __launch_bounds__(256)
k_mark_batch(const cg_read* __restrict__ reads, long long first, long long n, cg_counts kk,
             int n_chunks, int32_t chrom_len, int mq_threshold, unsigned int* ctl, long long* stats_row)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31;
    unsigned f = 0;
    int n_pass = 0;
    /* whole warps iterate together (k - lane is warp-uniform) so the shuffles below are safe */
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k - lane < n; k += stride) {
        const bool in = k < n;
        const long long i = first + k;
        int p = 0, key = -1 - lane;
        if (in) {
            const uint2 r = __ldg(reinterpret_cast<const uint2*>(reads) + i);
            p = (int)r.x;
            const int mapq = (int)((r.y >> 16) & 0xffu);
            const int cp = chunk_of(p, n_chunks);
            int cq = -1;
            if (i > 0) {
                const int q = reads[i - 1].pos;
                if (p < q) f |= CG_FLAG_UNSORTED;
                cq = chunk_of(q, n_chunks);
            }
            const bool inside = p >= 0 && p < chrom_len;
            if (!inside) f |= CG_FLAG_RANGE;
            for (int t = cq + 1; t <= cp; t++) kk.chunk_lo[t] = (int32_t)i;
            if (mapq > mq_threshold) {                                   /* bam_data.c:205 */
                n_pass++;                                                /* bam_data.c:214 */
                if (inside) key = cp;
            }
            /* a run of >= 32768 equal positions contains two indices that are multiples of the
             * probe distance, so the later one sees the earlier one */
            if ((i & (CG_PILEUP_PROBE - 1)) == 0 && i >= CG_PILEUP_PROBE && reads[i - CG_PILEUP_PROBE].pos == p)
                f |= CG_FLAG_PILEUP;
        }
        /* sorted stream: records of one chunk sit in adjacent lanes -> one atomic per chunk per warp,
         * and the same for the two coarser levels */
        int cnt;
        if (run_leader(key, lane, cnt) && key >= 0) atomicAdd(&kk.cnt1[key], cnt);
        const int gkey = key >= 0 ? key >> LVL : key;
        if (run_leader(gkey, lane, cnt) && gkey >= 0) atomicAdd(&kk.cnt2[gkey], cnt);
        const int skey = key >= 0 ? key >> (2 * LVL) : key;
        if (run_leader(skey, lane, cnt) && skey >= 0) atomicAdd(&kk.cnt3[skey], cnt);
    }
    if (f) atomicOr(&ctl[CG_W_FLAGS], f);
    n_pass = __reduce_add_sync(FULL, n_pass);
    if (lane == 0 && n_pass)
        atomicAdd((unsigned long long*)&stats_row[2 * CG_GC_BINS + 1], (unsigned long long)n_pass);
}

cudaError_t cgk_mark_batch(cudaStream_t s, const cg_read* reads, long long first, long long n,
                           cg_counts k, int n_chunks, int32_t chrom_len, int mq_threshold,
                           unsigned int* ctl, long long* stats_row)
{
    if (n <= 0) return cudaSuccess;
    long long blocks = (n + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_mark_batch<<<(int)blocks, 256, 0, s>>>(reads, first, n, k, n_chunks, chrom_len, mq_threshold, ctl, stats_row);
    return cudaGetLastError();
}

/* ------------------------------------------------------------------------- */
/* rare paths; both kernels return at once for a sorted stream without pile-ups */

/* out-of-order input: global-atomic histogram into u32 words (zero at rest: the
 * pass clears every word it consumes) */
__global__ void k_rare_scatter(const cg_read* __restrict__ reads, long long n_reads, uint32_t* depth32,
                               int32_t chrom_len, int mq_threshold, const unsigned int* ctl)
{
    if (!(ctl[CG_W_FLAGS] & CG_FLAG_UNSORTED)) return;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_reads; i += stride) {
        const uint2 r = *reinterpret_cast<const uint2*>(&reads[i]);
        const int pos = (int)r.x, mapq = (int)((r.y >> 16) & 0xffu);
        if (mapq > mq_threshold && pos >= 0 && pos < chrom_len) atomicAdd(&depth32[pos], 1u);
    }
}

__device__ __forceinline__ void counts_add(const cg_counts& k, int chunk, int delta)
{
    atomicAdd(&k.cnt1[chunk], delta);
    atomicAdd(&k.cnt2[chunk >> LVL], delta);
    atomicAdd(&k.cnt3[chunk >> (2 * LVL)], delta);
}

/* unsorted: chunk sums recomputed from the scattered counts (covers int16 wrap too).
 * sorted with a pile-up flagged: for every run of equal positions longer than the probe
 * distance, count its passing records c and replace the +c by sext16(c). */
__global__ void __launch_bounds__(256)
k_rare_totals(const cg_read* __restrict__ reads, long long n_reads, const uint32_t* __restrict__ depth32,
              cg_counts k, int n_chunks, int32_t chrom_len, int mq_threshold, const unsigned int* ctl)
{
    const unsigned fl = ctl[CG_W_FLAGS];
    if (!(fl & (CG_FLAG_UNSORTED | CG_FLAG_PILEUP))) return;
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long n_warps = ((long long)gridDim.x * blockDim.x) >> 5;
    if (fl & CG_FLAG_UNSORTED) {
        for (long long c = warp; c < n_chunks; c += n_warps) {
            const uint4* src = reinterpret_cast<const uint4*>(depth32 + c * CG_CHUNK);
            int s = 0;
            for (int q = lane; q < CG_CHUNK / 4; q += 32) {
                const uint4 v = src[q];
                s += sext16(v.x) + sext16(v.y) + sext16(v.z) + sext16(v.w);
            }
            s = __reduce_add_sync(FULL, s);
            /* the record count of the chunk is order independent; only a wrap changes it */
            if (lane == 0) {
                const int old = k.cnt1[c];
                if (s != old) counts_add(k, (int)c, s - old);
            }
        }
        return;
    }
    const long long n_groups = (n_reads + 31) >> 5;
    for (long long g = warp; g < n_groups; g += n_warps) {
        const long long i = (g << 5) + lane;
        int p = 0;
        bool start = false;
        if (i < n_reads) {
            p = reads[i].pos;
            start = (i == 0 || reads[i - 1].pos != p) && i + CG_PILEUP_PROBE < n_reads
                    && reads[i + CG_PILEUP_PROBE].pos == p;
        }
        unsigned m = __ballot_sync(FULL, start);
        while (m) {
            const int l = __ffs(m) - 1;
            m &= m - 1;
            const long long i0 = __shfl_sync(FULL, i, l);
            const int p0 = __shfl_sync(FULL, p, l);
            int c = 0;
            for (long long j = i0;; j += 32) {
                const long long jj = j + lane;
                bool inrun = false;
                if (jj < n_reads) {
                    const uint2 r = *reinterpret_cast<const uint2*>(&reads[jj]);
                    inrun = (int)r.x == p0;
                    if (inrun && (int)((r.y >> 16) & 0xffu) > mq_threshold) c++;
                }
                if (!__all_sync(FULL, inrun)) break;
            }
            c = __reduce_add_sync(FULL, c);
            if (lane == 0 && c >= 32768 && p0 >= 0 && p0 < chrom_len)
                counts_add(k, p0 >> CG_CHUNK_SHIFT, sext16((unsigned)c) - c);
        }
    }
}

cudaError_t cgk_rare_paths(cudaStream_t s, const cg_read* reads, long long n_reads, uint32_t* depth32,
                           cg_counts k, int n_chunks, int32_t chrom_len, int mq_threshold,
                           const unsigned int* ctl, int n_sms)
{
    k_rare_scatter<<<n_sms * 4, 256, 0, s>>>(reads, n_reads, depth32, chrom_len, mq_threshold, ctl);
    k_rare_totals<<<n_sms * 4, 256, 0, s>>>(reads, n_reads, depth32, k, n_chunks, chrom_len, mq_threshold, ctl);
    return cudaGetLastError();
}

/* ------------------------------------------------------------------------- */
/* the fused pass                                                            */
struct __align__(16) pass_smem {
    uint4 bins4[NWARP][CG_CHUNK / 8];         /* 8 x 4 KB: eight packed u16 counts per vector, one region per warp */
    int win_end[NWARP][MAXWIN];               /* chunk-local inclusive prefix at each window's last position */
    /* per-CTA rd_per_gc / window_per_gc.  32-bit is exact here: a CTA's share of the depth sum is
     * bounded by the record count (< 2^31) and it sweeps < 2^32 positions; native 32-bit shared
     * atomics instead of the CAS loop a 64-bit shared add compiles to */
    int gc_rd[CG_GC_BINS];
    unsigned gc_cnt[CG_GC_BINS];
    int is_last;
};

#define PASS_CTAS_PER_SM 5            /* register/shared-memory budget the kernel is compiled for */
/* CTAs per SM one launch asks for.  Measured on B200 (C2 chromosomes, pass alone): 1 -> 4.9, 2 -> 6.1,
 * 3 -> 5.6, 4 -> 5.2, 5 -> 5.1 TB/s: more resident warps mean more concurrent write streams and
 * worse HBM efficiency; the rest of the SM stays free for the genotype and preparation kernels. */
#define PASS_GRID_CTAS_PER_SM 2
#ifndef PASS_BATCH
#define PASS_BATCH 4                  /* chunks per visit to the work counter */
#endif

/* expected_read_depth[101], read_distribution.c:75-83:
 *   expected[0] = 0; expected[i] = (float)rd_per_gc[i] / window_per_gc[i]; NaN / Inf -> 0.
 * long->float and int->float round to nearest like x86's cvtsi2ss, and the quotient is one
 * IEEE fp32 division. */
__device__ __forceinline__ float expected_of(long long rd, long long cnt, int i)
{
    float e = 0.0f;
    if (i > 0) {
        e = __fdiv_rn(__ll2float_rn(rd), __int2float_rn((int)cnt));
        if (isnan(e) || isinf(e)) e = 0.0f;
    }
    return e;
}

/* sum of the two sign-extended 16-bit halves of w, added to acc (one IDP.2A) */
__device__ __forceinline__ int add_halves(unsigned w, int acc) { return __dp2a_lo((int)w, 0x0101, acc); }

__global__ void __launch_bounds__(NT, PASS_CTAS_PER_SM)
k_depth_pass(const cg_pass_args a)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    pass_smem& sm = *reinterpret_cast<pass_smem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint4* my4 = sm.bins4[warp];
    unsigned* bins = reinterpret_cast<unsigned*>(my4);                 /* word = positions 2k (low), 2k+1 (high) */
    int* wend = sm.win_end[warp];
    const bool unsorted = (a.ctl[CG_W_FLAGS] & CG_FLAG_UNSORTED) != 0;
    const int n_reads = (int)a.n_reads;

    for (int i = lane; i < CG_CHUNK / 8; i += 32) my4[i] = make_uint4(0, 0, 0, 0);
    if (lane < MAXWIN) wend[lane] = 0;
    for (int i = tid; i < CG_GC_BINS; i += NT) { sm.gc_rd[i] = 0; sm.gc_cnt[i] = 0u; }
    __syncthreads();

    /* Chunks are handed out PASS_BATCH at a time from one counter: at 2 B per position the pass needs
     * 1.5 G chunks/s to saturate HBM, more than one address takes atomics (measured 0.9 G/s, r01g) */
    unsigned cur = 0u;
    if (lane == 0) cur = atomicAdd(&a.ctl[CG_W_TILE], (unsigned)PASS_BATCH);
    int c = (int)__shfl_sync(FULL, cur, 0);
    int left = PASS_BATCH;                            /* chunks of the current batch not yet done, this one included */
    unsigned nxt = 0u;
    int blo = 0;

    while (c < a.n_chunks) {
        /* the batch after this one: asked for at the batch's first chunk, looked at after its last */
        if (left == PASS_BATCH) {
            if (lane == 0) nxt = atomicAdd(&a.ctl[CG_W_TILE], (unsigned)PASS_BATCH);
            /* record bounds of the whole batch in one coalesced load: lane j holds chunk_lo[c + j] */
            blo = lane <= PASS_BATCH ? min(__ldg(&a.k.chunk_lo[min(c + lane, a.n_chunks)]), n_reads) : 0;
        }
        const int base = c << CG_CHUNK_SHIFT;
        const int w0 = base / 100;                    /* first GC window overlapping the chunk */
        /* A window's depth sum is the number of records that fell into it as long as no int16 of the chunk
         * went negative (32768 or more records on one position): counted while the records are binned.
         * `wrapped` chunks (and the out-of-order path) take the sums from a scan of the bins instead. */
        bool big = false;
        /* the GC byte of this lane's window: asked for now, used after the stores */
        const int nwin = (base + CG_CHUNK - 1) / 100 - w0 + 1;
        const int chunk_end = min(base + CG_CHUNK, a.chrom_len);
        const int gw_s = max(base, 100 * (w0 + lane)), gw_e = min(chunk_end, 100 * (w0 + lane) + 100);
        int gbin = 0;
        if (lane < nwin && gw_e > gw_s) gbin = (int)a.gc[w0 + lane];              /* sonic_interval.c:263-272 */

        /* ---- phase A: the chunk's records -> packed shared-memory histogram, records per window ---- */
        if (!unsorted) {
            const int lo = __shfl_sync(FULL, blo, PASS_BATCH - left), hi = __shfl_sync(FULL, blo, PASS_BATCH - left + 1);
            for (int i0 = lo; i0 < hi; i0 += 32) {
                const int i = i0 + lane;
                const bool in = i < hi;
                uint2 r = make_uint2(0u, 0u);
                if (in) r = __ldg(reinterpret_cast<const uint2*>(a.reads) + i);
                const int pos = (int)r.x, mapq = (int)((r.y >> 16) & 0xffu);
                const unsigned off = (unsigned)(pos - base);
                const bool ok = in && mapq > a.mq_threshold                       /* bam_data.c:205 */
                                && off < (unsigned)CG_CHUNK && pos < a.chrom_len;
                /* warp aggregation: the stream is sorted, so equal positions sit in adjacent lanes */
                int cnt;
                const bool leader = run_leader(ok ? (int)off : -1 - lane, lane, cnt);
                if (ok && leader) {
                    unsigned now;                                                 /* the position's count so far, unwrapped */
                    if (off & 1u) {
                        now = (atomicAdd(&bins[off >> 1], (unsigned)cnt << 16) >> 16) + (unsigned)cnt;   /* bam_data.c:213; carry-out = wrap */
                    } else {
                        now = (atomicAdd(&bins[off >> 1], (unsigned)cnt) & 0xffffu) + (unsigned)cnt;
                        /* the low half wrapped into its neighbour's count: take the carry back out */
                        if (now > 0xffffu) atomicSub(&bins[off >> 1], 0x10000u);
                    }
                    big |= now >= 0x8000u;
                    atomicAdd(&wend[pos / 100 - w0], cnt);
                }
            }
        } else {
            uint4* src = reinterpret_cast<uint4*>(a.depth32 + (size_t)base);
            for (int q = lane; q < CG_CHUNK / 8; q += 32) {
                const uint4 x = src[2 * q], y = src[2 * q + 1];
                my4[q] = make_uint4((x.x & 0xffffu) | (x.y << 16), (x.z & 0xffffu) | (x.w << 16),
                                    (y.x & 0xffffu) | (y.y << 16), (y.z & 0xffffu) | (y.w << 16));
                src[2 * q] = make_uint4(0, 0, 0, 0);                              /* zero at rest */
                src[2 * q + 1] = make_uint4(0, 0, 0, 0);
            }
            big = true;
        }
        const bool wrapped = __any_sync(FULL, big);
        __syncwarp();

        if (!wrapped) {
            /* ---- rows -> depth stores: the packed bins ARE the int16 output vectors ---- */
            char* dp = reinterpret_cast<char*>(a.depth) + ((size_t)base + lane * 8) * 2;
#pragma unroll
            for (int r = 0; r < ROWS; r++) {
                const uint4 v = my4[r * 32 + lane];
                my4[r * 32 + lane] = make_uint4(0, 0, 0, 0);          /* clean bins for the next chunk */
                st_cs_v4(dp + r * 512, v.x, v.y, v.z, v.w);
            }
        } else {
            if (lane < MAXWIN) wend[lane] = 0;                        /* the scan below writes window-END sums */
            __syncwarp();
            /* ---- eight row scans at once: lane sums, then eight independent shuffle chains ---- */
            int ex[ROWS];                                 /* -> prefix in front of this lane's 8 positions, per row */
            {
                int tot[ROWS];
    #pragma unroll
                for (int r = 0; r < ROWS; r++) {
                    const uint4 v = my4[r * 32 + lane];
                    tot[r] = add_halves(v.w, add_halves(v.z, add_halves(v.y, add_halves(v.x, 0))));
                    ex[r] = tot[r];
                }
    #pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
    #pragma unroll
                    for (int r = 0; r < ROWS; r++) {
                        const int o = __shfl_up_sync(FULL, ex[r], d);
                        if (lane >= d) ex[r] += o;
                    }
                }
                int run = 0;                                /* sums are relative to the chunk's first position */
    #pragma unroll
                for (int r = 0; r < ROWS; r++) {
                    const int row_total = __shfl_sync(FULL, ex[r], 31);
                    ex[r] = run + ex[r] - tot[r];
                    run += row_total;
                }
            }

            /* ---- rows -> depth stores, window-end sums ---- */
            {
                const int cpos = base + lane * 8;         /* this lane's first position in row 0 */
                int wq = cpos / 100, wm = cpos - wq * 100;                /* its GC window and offset within it */
                char* dp = reinterpret_cast<char*>(a.depth) + (size_t)cpos * 2;
    #pragma unroll
                for (int r = 0; r < ROWS; r++) {
                    const uint4 v = my4[r * 32 + lane];
                    my4[r * 32 + lane] = make_uint4(0, 0, 0, 0);          /* clean bins for the next chunk */
                    const int b = ex[r];
                    const int s0 = sext16(v.x), s1 = add_halves(v.x, 0);
                    const int s2 = s1 + sext16(v.y), s3 = add_halves(v.y, s1);
                    const int s4 = s3 + sext16(v.z), s5 = add_halves(v.z, s3);
                    const int s6 = s5 + sext16(v.w), s7 = add_halves(v.w, s5);
                    st_cs_v4(dp + r * 512, v.x, v.y, v.z, v.w);
                    /* the lane whose 8 positions contain a window's last position (== 99 mod 100) */
                    if (wm >= 92) {
                        const int k = 99 - wm;
                        const int sk = k == 0 ? s0 : k == 1 ? s1 : k == 2 ? s2 : k == 3 ? s3
                                     : k == 4 ? s4 : k == 5 ? s5 : k == 6 ? s6 : s7;
                        wend[wq - w0] = b + sk;
                    }
                    if (r == ROWS - 1 && lane == 31)                      /* the chunk's last position closes a partial window */
                        wend[(base + CG_CHUNK - 1) / 100 - w0] = b + s7;
                    wm += 56; wq += 2;                                    /* 256 = 2*100 + 56 */
                    if (wm >= 100) { wm -= 100; wq++; }
                }
            }
            __syncwarp();

            /* window-end sums -> per-window sums */
            const int nw = (base + CG_CHUNK - 1) / 100 - w0 + 1;
            const int d = lane < nw ? wend[lane] - (lane ? wend[lane - 1] : 0) : 0;
            __syncwarp();
            if (lane < nw) wend[lane] = d;
        }
        __syncwarp();

        /* ---- GC-bin sums from the per-window depth sums (read_distribution.c:63-73) ---- */
        if (lane < nwin && gw_e > gw_s) {
            const int bin = (unsigned)gbin > 100u ? 0 : gbin;
            const int sum = wend[lane];
            if (sum) atomicAdd(&sm.gc_rd[bin], sum);
            atomicAdd(&sm.gc_cnt[bin], (unsigned)(gw_e - gw_s));
        }
        if (lane < MAXWIN) wend[lane] = 0;                            /* clean for the next chunk */
        __syncwarp();
        if (--left) c++;
        else { c = (int)__shfl_sync(FULL, nxt, 0); left = PASS_BATCH; }
    }

    /* ---- CTA epilogue: flush the per-CTA statistics; the last CTA derives the table ---- */
    __syncthreads();
    for (int b = tid; b < CG_GC_BINS; b += NT) {
        if (sm.gc_rd[b])  atomicAdd((unsigned long long*)&a.stats[b], (unsigned long long)(long long)sm.gc_rd[b]);
        if (sm.gc_cnt[b]) atomicAdd((unsigned long long*)&a.stats[CG_GC_BINS + b], (unsigned long long)sm.gc_cnt[b]);
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) sm.is_last = atomicAdd(&a.ctl[CG_W_DONE], 1u) == gridDim.x - 1;
    __syncthreads();
    if (sm.is_last) {
        __threadfence();
        if (tid < CG_GC_BINS) {
            const long long rd = __ldcg(&a.stats[tid]), cnt = __ldcg(&a.stats[CG_GC_BINS + tid]);
            a.expected[tid] = expected_of(rd, cnt, tid);
        }
        if (warp == NWARP - 1) {
            /* calc_mu_per_chr, read_distribution.c:35: |sum| <= number of records < 2^31 */
            int tot = 0;
            const int n_supers = (a.n_chunks + 1023) >> (2 * LVL);
            for (int j = lane; j < n_supers; j += 32) tot += a.k.cnt3[j];
            tot = __reduce_add_sync(FULL, tot);
            if (lane == 0) {
                a.stats[2 * CG_GC_BINS] = (long long)tot;
                if (a.ctl[CG_W_FLAGS] & CG_FLAG_RANGE) atomicOr(a.sticky, CG_FLAG_RANGE);
            }
        }
        if (a.host_row) {                     /* asynchronous HOST mode: the row goes straight to page-locked host memory */
            __syncthreads();
            if (tid < CG_STATS_WORDS) a.host_row[tid] = __ldcg(&a.stats[tid]);
        }
    }
}

int cgk_depth_pass_ctas_per_sm(void) { return PASS_CTAS_PER_SM; }

/* Two kernels share an SM only if they ask for the same L1/shared-memory split; otherwise the
 * CTAs of the second wait until the SM has drained.  Every kernel of the library therefore
 * states the same carve-out, so that the preparation and genotype kernels of one chromosome
 * really run on the SMs a neighbour's depth pass occupies. */
int cgk_carveout(void)
{
    static int pct = -1;
    if (pct < 0) {
        const char* e = getenv("CG_CARVEOUT");            /* tuning only */
        pct = e ? atoi(e) : 100;
        if (pct < 0 || pct > 100) pct = 100;
    }
    return pct;
}

cudaError_t cgk_configure_depth(void)
{
    const int pct = cgk_carveout();
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(k_depth_pass, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(pass_smem))) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_depth_pass, cudaFuncAttributePreferredSharedMemoryCarveout, pct)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_chrom_init, cudaFuncAttributePreferredSharedMemoryCarveout, pct)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_mark_batch, cudaFuncAttributePreferredSharedMemoryCarveout, pct)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_rare_scatter, cudaFuncAttributePreferredSharedMemoryCarveout, pct)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_rare_totals, cudaFuncAttributePreferredSharedMemoryCarveout, pct)) != cudaSuccess) return e;
    return cudaSuccess;
}

/* Diagnostic (cg_download_prefix): the int32 inclusive prefix of the depth array, P[i] = sum of depth[0..i]
 * mod 2^32, rebuilt from the depth array and the chunk / group / super sums.  One warp per chunk. */
__global__ void __launch_bounds__(256)
k_prefix_from_depth(const int16_t* __restrict__ depth, cg_counts k, int n_chunks, int32_t* __restrict__ prefix)
{
    const int lane = threadIdx.x & 31;
    const int n_warps = gridDim.x * (blockDim.x >> 5);
    for (int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); c < n_chunks; c += n_warps) {
        int part = 0;
        const int g = c >> LVL, sp = c >> (2 * LVL);
        const int i1 = (g << LVL) + lane, i2 = (sp << LVL) + lane;
        if (i1 < c) part += k.cnt1[i1];
        if (i2 < g) part += k.cnt2[i2];
        for (int j = lane; j < sp; j += 32) part += k.cnt3[j];
        int run = __reduce_add_sync(FULL, part);
        const size_t base = (size_t)c << CG_CHUNK_SHIFT;
        const uint4* src = reinterpret_cast<const uint4*>(depth + base);
        for (int r = 0; r < ROWS; r++) {
            const uint4 v = src[r * 32 + lane];
            const int s0 = sext16(v.x), s1 = add_halves(v.x, 0);
            const int s2 = s1 + sext16(v.y), s3 = add_halves(v.y, s1);
            const int s4 = s3 + sext16(v.z), s5 = add_halves(v.z, s3);
            const int s6 = s5 + sext16(v.w), s7 = add_halves(v.w, s5);
            const int incl = warp_incl_scan(s7, lane);
            const int b = run + incl - s7;
            int4* dst = reinterpret_cast<int4*>(prefix + base + (size_t)(r * 32 + lane) * 8);
            dst[0] = make_int4(b + s0, b + s1, b + s2, b + s3);
            dst[1] = make_int4(b + s4, b + s5, b + s6, b + s7);
            run += __shfl_sync(FULL, incl, 31);
        }
    }
}

cudaError_t cgk_prefix_from_depth(cudaStream_t s, const int16_t* depth, cg_counts k, int n_chunks, int32_t* prefix)
{
    if (n_chunks <= 0) return cudaSuccess;
    int blocks = (n_chunks + 7) / 8;
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_prefix_from_depth<<<blocks, 256, 0, s>>>(depth, k, n_chunks, prefix);
    return cudaGetLastError();
}

cudaError_t cgk_depth_pass(cudaStream_t s, const cg_pass_args& a, int n_sms)
{
    const int smem = (int)sizeof(pass_smem);           /* opted in by cgk_configure_depth */
    static int ctas_per_sm = 0;               /* CG_PASS_CTAS: CTAs per SM one launch asks for (tuning) */
    if (!ctas_per_sm) {
        const char* e = getenv("CG_PASS_CTAS");
        ctas_per_sm = e ? atoi(e) : PASS_GRID_CTAS_PER_SM;
        if (ctas_per_sm < 1 || ctas_per_sm > PASS_CTAS_PER_SM) ctas_per_sm = PASS_CTAS_PER_SM;
    }
    int grid = n_sms * ctas_per_sm;
    const int want = (a.n_chunks + NWARP - 1) / NWARP;
    if (grid > want) grid = want;
    if (grid < 1) grid = 1;
    k_depth_pass<<<grid, NT, smem, s>>>(a);
    return cudaGetLastError();
}

/* ------------------------------------------------------------------------- */
/* the table alone (cg_set_expected replays, tests) */
__global__ void k_expected_table(const long long* __restrict__ stats_row, float* expected)
{
    const int i = threadIdx.x;
    if (i < CG_GC_BINS) expected[i] = expected_of(stats_row[i], stats_row[CG_GC_BINS + i], i);
}

cudaError_t cgk_expected_table(cudaStream_t s, const long long* stats_row, float* expected)
{
    k_expected_table<<<1, 128, 0, s>>>(stats_row, expected);
    return cudaGetLastError();
}
