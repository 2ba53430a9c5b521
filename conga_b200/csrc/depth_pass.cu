/* depth_pass.cu — the fused read-depth pass of one chromosome on sm_100a.
 *
 * Replaces, in ONE sweep over the chromosome's positions (no intermediate HBM
 * round trip):
 *   K1  read_depth[pos]++ per passing record        bam_data.c:201-217, read_distribution.c:12-18
 *   K2  chromosome sum + 101 GC-bin sums/counts     read_distribution.c:27-46, :49-73
 *   K3  int32 inclusive prefix of the depth array   (serves likelihood.c:115-121 as O(1) queries)
 *
 * Data layout in HBM (per context, reused across chromosomes):
 *   reads      cg_read[N]            8 B records, coordinate sorted
 *   tile_lo    int32[n_tiles+1]      first record of every 16384-position tile
 *   chunk_cnt  int32[n_tiles*8]      sum of the int16 depths of every 2048-position chunk
 *   tile_carry u32[n_tiles]          prefix in front of every tile (mod 2^32)
 *   depth      int16[n_tiles*16384]  the reference's `short read_depth[]` (wraps mod 2^16)
 *   prefix     int32[n_tiles*16384]  P[i] = sum_{j<=i} depth[j]   (mod 2^32)
 *   stats      int64[204]            cg_chrom_stats row
 *
 * The scan carry never travels between CTAs.  A chunk's depth sum is the number of
 * passing records that fall in it (every read adds exactly 1) unless a position
 * collects >= 32768 reads and the int16 wraps; k_mark_batch therefore counts records
 * per chunk while it checks the pushed batch (the records are read once, 8 B each,
 * against 6 B per POSITION for the pass), k_rare_paths corrects the count for wrapped
 * pile-ups, and k_tile_scan turns the counts into per-tile carries.  The pass itself
 * is then embarrassingly parallel over tiles: no look-back, no spinning.
 *
 * Algorithmic bytes per chromosome: 8·N (records) + 2·L (depth) +
 * 4·L (prefix) + L/100 (GC bytes); the pass is HBM-write bound.
 *
 * Pass kernel shape: persistent CTAs pulling tiles from a counter, 256 threads,
 * 32 KB of packed u16 bins in shared memory.  Per tile: warp-aggregated shared
 * histogram of the tile's records -> barrier -> every warp scans its own 2048
 * positions in 8 rows (lane = 8 consecutive positions = one 16-byte LDS, which is
 * also the 16-byte int16 output vector; one warp scan per row; two 16-byte prefix
 * stores) and zeroes the bins behind it -> barrier -> per-window sums from the
 * prefixes at window ends into 101 shared int64 GC accumulators.
 */
#include "cg_internal.h"

#define FULL 0xffffffffu
#define TILE CG_TILE
#define NT   CG_PASS_THREADS
#define NWARP (NT / 32)
#define ROWS (CG_CHUNK / 256)         /* rows of 256 positions per warp chunk */
#define MAXWIN (TILE / 100 + 2)       /* GC windows that can overlap one tile */

static_assert(NWARP == CG_CHUNKS_PER_TILE, "one warp per chunk");
static_assert((1 << CG_TILE_SHIFT) == TILE && (1 << CG_CHUNK_SHIFT) == CG_CHUNK, "shifts");

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
/* chromosome start: clear the chunk counts, the stats row and the control words */
__global__ void k_chrom_init(int4* chunk_cnt4, int n_vec, long long* stats_row, unsigned int* ctl)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    for (int j = i; j < n_vec; j += gridDim.x * blockDim.x) chunk_cnt4[j] = make_int4(0, 0, 0, 0);
    if (i < CG_STATS_WORDS) stats_row[i] = 0;
    if (i < 4) ctl[i] = 0u;
}

cudaError_t cgk_chrom_init(cudaStream_t s, int32_t* chunk_cnt, int n_chunks, long long* stats_row, unsigned int* ctl)
{
    const int n_vec = n_chunks / 4;                      /* n_chunks is a multiple of 8 */
    int blocks = (n_vec + 255) / 256;
    if (blocks < 1) blocks = 1;
    if (blocks > 148) blocks = 148;
    k_chrom_init<<<blocks, 256, 0, s>>>(reinterpret_cast<int4*>(chunk_cnt), n_vec, stats_row, ctl);
    return cudaGetLastError();
}

/* ------------------------------------------------------------------------- */
/* per pushed batch, one sweep over its records: sortedness and range checks,  */
/* tile boundaries, passing records per chunk, pile-up probe                   */
__device__ __forceinline__ int tile_of(int pos, int n_tiles)
{
    int t = pos >> CG_TILE_SHIFT;             /* arithmetic shift: negative pos -> negative tile */
    t = t < 0 ? 0 : t;
    return t >= n_tiles ? n_tiles - 1 : t;
}

__global__ void __launch_bounds__(256)
k_mark_batch(const cg_read* __restrict__ reads, long long first, long long n, int32_t* tile_lo,
             int32_t* chunk_cnt, int n_tiles, int32_t chrom_len, int mq_threshold,
             unsigned int* ctl, long long* stats_row)
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
            const int tp = tile_of(p, n_tiles);
            int tq = -1;
            if (i > 0) {
                const int q = reads[i - 1].pos;
                if (p < q) f |= CG_FLAG_UNSORTED;
                tq = tile_of(q, n_tiles);
            }
            const bool inside = p >= 0 && p < chrom_len;
            if (!inside) f |= CG_FLAG_RANGE;
            for (int t = tq + 1; t <= tp; t++) tile_lo[t] = (int32_t)i;
            if (mapq > mq_threshold) {                                   /* bam_data.c:205 */
                n_pass++;                                                /* bam_data.c:214 */
                if (inside) key = p >> CG_CHUNK_SHIFT;
            }
            /* a run of >= 32768 equal positions contains two indices that are multiples of the
             * probe distance, so the later one sees the earlier one */
            if ((i & (CG_PILEUP_PROBE - 1)) == 0 && i >= CG_PILEUP_PROBE && reads[i - CG_PILEUP_PROBE].pos == p)
                f |= CG_FLAG_PILEUP;
        }
        /* sorted stream: records of one chunk sit in adjacent lanes -> one atomic per chunk per warp */
        const int prev = __shfl_up_sync(FULL, key, 1);
        const bool leader = lane == 0 || key != prev;
        const unsigned lm = __ballot_sync(FULL, leader);
        if (leader && key >= 0) {
            const unsigned above = lane == 31 ? 0u : (lm & ~((2u << lane) - 1u));
            const int cnt = (above ? __ffs(above) - 1 : 32) - lane;
            atomicAdd(&chunk_cnt[key], cnt);
        }
    }
    if (f) atomicOr(&ctl[CG_W_FLAGS], f);
    n_pass = __reduce_add_sync(FULL, n_pass);
    if (lane == 0 && n_pass)
        atomicAdd((unsigned long long*)&stats_row[2 * CG_GC_BINS + 1], (unsigned long long)n_pass);
}

cudaError_t cgk_mark_batch(cudaStream_t s, const cg_read* reads, long long first, long long n,
                           int32_t* tile_lo, int32_t* chunk_cnt, int n_tiles, int32_t chrom_len,
                           int mq_threshold, unsigned int* ctl, long long* stats_row)
{
    if (n <= 0) return cudaSuccess;
    long long blocks = (n + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_mark_batch<<<(int)blocks, 256, 0, s>>>(reads, first, n, tile_lo, chunk_cnt, n_tiles, chrom_len,
                                             mq_threshold, ctl, stats_row);
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

/* unsorted: chunk sums recomputed from the scattered counts (covers int16 wrap too).
 * sorted with a pile-up flagged: for every run of equal positions longer than the probe
 * distance, count its passing records c and replace the chunk's +c by sext16(c). */
__global__ void __launch_bounds__(256)
k_rare_totals(const cg_read* __restrict__ reads, long long n_reads, const uint32_t* __restrict__ depth32,
              int32_t* chunk_cnt, int n_chunks, int32_t chrom_len, int mq_threshold, const unsigned int* ctl)
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
            if (lane == 0) chunk_cnt[c] = s;
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
                atomicAdd(&chunk_cnt[p0 >> CG_CHUNK_SHIFT], sext16((unsigned)c) - c);
        }
    }
}

cudaError_t cgk_rare_paths(cudaStream_t s, const cg_read* reads, long long n_reads, uint32_t* depth32,
                           int32_t* chunk_cnt, int n_tiles, int32_t chrom_len, int mq_threshold,
                           const unsigned int* ctl, int n_sms)
{
    k_rare_scatter<<<n_sms * 4, 256, 0, s>>>(reads, n_reads, depth32, chrom_len, mq_threshold, ctl);
    k_rare_totals<<<n_sms * 4, 256, 0, s>>>(reads, n_reads, depth32, chunk_cnt, n_tiles * CG_CHUNKS_PER_TILE,
                                           chrom_len, mq_threshold, ctl);
    return cudaGetLastError();
}

/* ------------------------------------------------------------------------- */
/* one CTA: tile bounds behind the last record, per-tile carries, chromosome sum.
 * |sum of any set of depths| <= number of records < 2^31, so int32 arithmetic is exact. */
#define SCAN_THREADS 1024
__global__ void __launch_bounds__(SCAN_THREADS)
k_tile_scan(const cg_read* __restrict__ reads, long long n_reads, int32_t* tile_lo,
            const int32_t* __restrict__ chunk_cnt, uint32_t* tile_carry, int n_tiles, long long* stats_row)
{
    __shared__ int wsum[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int last = n_reads > 0 ? tile_of(reads[n_reads - 1].pos, n_tiles) : -1;
    for (int t = last + 1 + tid; t <= n_tiles; t += SCAN_THREADS) tile_lo[t] = (int32_t)n_reads;

    int run = 0;
    for (int b = 0; b < n_tiles; b += SCAN_THREADS) {
        const int t = b + tid;
        int tot = 0;
        if (t < n_tiles) {
            const int4 x = __ldg(reinterpret_cast<const int4*>(chunk_cnt) + 2 * t);
            const int4 y = __ldg(reinterpret_cast<const int4*>(chunk_cnt) + 2 * t + 1);
            tot = x.x + x.y + x.z + x.w + y.x + y.y + y.z + y.w;
        }
        const int incl = warp_incl_scan(tot, lane);
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        if (warp == 0) wsum[lane] = warp_incl_scan(wsum[lane], lane);
        __syncthreads();
        const int woff = warp ? wsum[warp - 1] : 0;
        if (t < n_tiles) tile_carry[t] = (uint32_t)(run + woff + incl - tot);
        run += wsum[31];
        __syncthreads();
    }
    if (tid == 0) stats_row[2 * CG_GC_BINS] = (long long)run;          /* calc_mu_per_chr, read_distribution.c:35 */
}

cudaError_t cgk_tile_scan(cudaStream_t s, const cg_read* reads, long long n_reads, int32_t* tile_lo,
                          const int32_t* chunk_cnt, uint32_t* tile_carry, int n_tiles, long long* stats_row)
{
    k_tile_scan<<<1, SCAN_THREADS, 0, s>>>(reads, n_reads, tile_lo, chunk_cnt, tile_carry, n_tiles, stats_row);
    return cudaGetLastError();
}

/* ------------------------------------------------------------------------- */
/* the fused pass                                                            */
struct __align__(16) pass_smem {
    uint4 bins4[TILE / 8];                    /* 32 KB: eight packed u16 counts per vector */
    /* per-CTA rd_per_gc / window_per_gc.  32-bit is exact here: a CTA's share of the depth sum is
     * bounded by the record count (< 2^31) and it sweeps < 2^32 positions; native 32-bit shared
     * atomics instead of the CAS loop a 64-bit shared add compiles to */
    int gc_rd[CG_GC_BINS];
    unsigned gc_cnt[CG_GC_BINS];
    int win_end[MAXWIN];                      /* tile-local inclusive prefix at each window's last position */
    int next_tile;
    int is_last;
};

#define PASS_CTAS_PER_SM 5

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

__global__ void __launch_bounds__(NT, PASS_CTAS_PER_SM)
k_depth_pass(const cg_pass_args a)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    pass_smem& sm = *reinterpret_cast<pass_smem*>(smem_raw);
    unsigned* bins = reinterpret_cast<unsigned*>(sm.bins4);            /* word = positions 2k (low), 2k+1 (high) */
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool unsorted = (a.ctl[CG_W_FLAGS] & CG_FLAG_UNSORTED) != 0;

    for (int i = tid; i < TILE / 8; i += NT) sm.bins4[i] = make_uint4(0, 0, 0, 0);
    for (int i = tid; i < CG_GC_BINS; i += NT) { sm.gc_rd[i] = 0; sm.gc_cnt[i] = 0u; }
    if (tid == 0) sm.next_tile = (int)atomicAdd(&a.ctl[CG_W_TILE], 1u);
    __syncthreads();

    for (;;) {
        const int t = sm.next_tile;
        if (t >= a.n_tiles) break;
        const int base = t << CG_TILE_SHIFT;
        /* carry in front of this warp's chunk: tile carry + the chunks before it */
        const int4 c0 = __ldg(reinterpret_cast<const int4*>(a.chunk_cnt) + 2 * t);
        const int4 c1 = __ldg(reinterpret_cast<const int4*>(a.chunk_cnt) + 2 * t + 1);
        const unsigned tile_carry = __ldg(&a.tile_carry[t]);

        /* ---- phase A: the tile's records -> packed shared-memory histogram ---- */
        if (!unsorted) {
            const int lo = __ldg(&a.tile_lo[t]), hi = __ldg(&a.tile_lo[t + 1]);
            for (int i0 = lo + warp * 32; i0 < hi; i0 += NT) {
                const int i = i0 + lane;
                const bool in = i < hi;
                uint2 r = make_uint2(0u, 0u);
                if (in) r = __ldg(reinterpret_cast<const uint2*>(a.reads) + i);
                const int pos = (int)r.x, mapq = (int)((r.y >> 16) & 0xffu);
                const unsigned off = (unsigned)(pos - base);
                const bool ok = in && mapq > a.mq_threshold                       /* bam_data.c:205 */
                                && off < (unsigned)TILE && pos < a.chrom_len;
                /* warp aggregation: the stream is sorted, so equal positions sit in adjacent lanes */
                const int key = ok ? (int)off : -1 - lane;
                const int prev = __shfl_up_sync(FULL, key, 1);
                const bool leader = (lane == 0) || key != prev;
                const unsigned lm = __ballot_sync(FULL, leader);
                if (ok && leader) {
                    const unsigned above = (lane == 31) ? 0u : (lm & ~((2u << lane) - 1u));
                    const unsigned cnt = (unsigned)((above ? __ffs(above) - 1 : 32) - lane);
                    if (off & 1u) {
                        atomicAdd(&bins[off >> 1], cnt << 16);                    /* bam_data.c:213; carry-out = wrap */
                    } else {
                        const unsigned old = atomicAdd(&bins[off >> 1], cnt);
                        /* the low half wrapped into its neighbour's count: take the carry back out */
                        if ((old & 0xffffu) + cnt > 0xffffu) atomicSub(&bins[off >> 1], 0x10000u);
                    }
                }
            }
        } else {
            uint4* src = reinterpret_cast<uint4*>(a.depth32 + (size_t)base);
            for (int q = tid; q < TILE / 8; q += NT) {
                const uint4 x = src[2 * q], y = src[2 * q + 1];
                sm.bins4[q] = make_uint4((x.x & 0xffffu) | (x.y << 16), (x.z & 0xffffu) | (x.w << 16),
                                         (y.x & 0xffffu) | (y.y << 16), (y.z & 0xffffu) | (y.w << 16));
                src[2 * q] = make_uint4(0, 0, 0, 0);                              /* zero at rest */
                src[2 * q + 1] = make_uint4(0, 0, 0, 0);
            }
        }
        __syncthreads();

        /* everyone has read sm.next_tile; fetch the tile after this one while the scan runs */
        unsigned nxt = 0u;
        if (tid == 0) nxt = atomicAdd(&a.ctl[CG_W_TILE], 1u);

        /* ---- scan: bins -> depth + prefix, one warp per chunk, 8 rows of 256 positions ---- */
        const int w0 = base / 100;                     /* first GC window overlapping the tile */
        {
            const int cc[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
            int woff = 0;
#pragma unroll
            for (int j = 0; j < 7; j++) woff += (j < warp) ? cc[j] : 0;
            unsigned run = tile_carry + (unsigned)woff;
            const int cpos = base + warp * CG_CHUNK + lane * 8;       /* this lane's first position in row 0 */
            int wq = cpos / 100, wm = cpos - wq * 100;                /* its GC window and offset within it */
            int vi = warp * (CG_CHUNK / 8) + lane;                    /* uint4 index into the bins */
            char* dp = reinterpret_cast<char*>(a.depth) + (size_t)cpos * 2;
            char* pp = reinterpret_cast<char*>(a.prefix) + (size_t)cpos * 4;
#pragma unroll 2
            for (int r = 0; r < ROWS; r++) {
                const uint4 v = sm.bins4[vi];
                sm.bins4[vi] = make_uint4(0, 0, 0, 0);                /* clean bins for the next tile */
                const int s0 = sext16(v.x), s1 = s0 + ((int)v.x >> 16);
                const int s2 = s1 + sext16(v.y), s3 = s2 + ((int)v.y >> 16);
                const int s4 = s3 + sext16(v.z), s5 = s4 + ((int)v.z >> 16);
                const int s6 = s5 + sext16(v.w), s7 = s6 + ((int)v.w >> 16);
                const int incl = warp_incl_scan(s7, lane);
                const unsigned b = run + (unsigned)(incl - s7);
                st_cs_v4(dp, v.x, v.y, v.z, v.w);
                st_cs_v4(pp, b + s0, b + s1, b + s2, b + s3);
                st_cs_v4(pp + 16, b + s4, b + s5, b + s6, b + s7);
                /* the lane whose 8 positions contain a window's last position (== 99 mod 100) */
                if (wm >= 92) {
                    const int k = 99 - wm;
                    const int sk = k == 0 ? s0 : k == 1 ? s1 : k == 2 ? s2 : k == 3 ? s3
                                 : k == 4 ? s4 : k == 5 ? s5 : k == 6 ? s6 : s7;
                    sm.win_end[wq - w0] = (int)(b - tile_carry) + sk;
                }
                if (r == ROWS - 1 && tid == NT - 1)                   /* the tile's last position closes a partial window */
                    sm.win_end[(base + TILE - 1) / 100 - w0] = (int)(b - tile_carry) + s7;
                run += (unsigned)__shfl_sync(FULL, incl, 31);
                vi += 32; dp += 512; pp += 1024;
                wm += 56; wq += 2;                                    /* 256 = 2*100 + 56 */
                if (wm >= 100) { wm -= 100; wq++; }
            }
        }
        if (tid == 0) sm.next_tile = (int)nxt;
        __syncthreads();

        /* ---- GC-bin sums from window-end prefixes (read_distribution.c:63-73) ---- */
        {
            const int nwin = (base + TILE - 1) / 100 - w0 + 1;
            const int tile_end = min(base + TILE, a.chrom_len);
            if (tid < nwin) {
                const int w = w0 + tid;
                const int ws = max(base, 100 * w), we = min(tile_end, 100 * w + 100);
                if (we > ws) {
                    int bin = (int)a.gc[w];                           /* sonic_interval.c:263-272 */
                    if ((unsigned)bin > 100u) bin = 0;
                    const int sum = sm.win_end[tid] - (tid ? sm.win_end[tid - 1] : 0);
                    if (sum) atomicAdd(&sm.gc_rd[bin], sum);
                    atomicAdd(&sm.gc_cnt[bin], (unsigned)(we - ws));
                }
            }
        }
        /* win_end is next written after the next tile's first barrier, bins after this loop-top read */
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
    if (sm.is_last && tid < CG_GC_BINS) {
        __threadfence();
        const long long rd = __ldcg(&a.stats[tid]), cnt = __ldcg(&a.stats[CG_GC_BINS + tid]);
        a.expected[tid] = expected_of(rd, cnt, tid);
        if (tid == 0 && (a.ctl[CG_W_FLAGS] & CG_FLAG_RANGE)) atomicOr(a.sticky, CG_FLAG_RANGE);
    }
}

int cgk_depth_pass_ctas_per_sm(void) { return PASS_CTAS_PER_SM; }

cudaError_t cgk_depth_pass(cudaStream_t s, const cg_pass_args& a, int n_sms)
{
    static bool configured = false;
    const int smem = (int)sizeof(pass_smem);
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(k_depth_pass, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    int grid = n_sms * PASS_CTAS_PER_SM;
    if (grid > a.n_tiles) grid = a.n_tiles;
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
