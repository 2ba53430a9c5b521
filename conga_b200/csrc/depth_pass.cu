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
 *   depth      int16[n_tiles*16384]  the reference's `short read_depth[]` (wraps mod 2^16)
 *   prefix     int32[n_tiles*16384]  P[i] = sum_{j<=i} depth[j]   (mod 2^32)
 *   tile_state u64[n_tiles]          decoupled look-back: flag<<32 | value
 *   stats      int64[204]            cg_chrom_stats row
 *
 * Algorithmic bytes per chromosome: 8·N (records) + 2·L (depth) +
 * 4·L (prefix) + L/100 (GC bytes); the pass is HBM-write bound.
 *
 * Kernel shape: persistent CTAs (3 per SM × 148 SMs), 256 threads, 64 KB of u32
 * bins in shared memory per CTA.  Per tile: warp-aggregated shared-memory
 * histogram of the tile's records -> tile total published for the single-pass
 * look-back -> 8 block-scan iterations that read the bins with conflict-free
 * 16-byte LDS, zero them for the next tile, and write depth (8-byte) and prefix
 * (16-byte) vectors fully coalesced -> per-window sums from prefix differences
 * into 101 shared int64 GC accumulators.
 */
#include "cg_internal.h"

#define FULL 0xffffffffu
#define TILE CG_TILE
#define NT   CG_PASS_THREADS
#define NWARP (NT / 32)
#define SPAN (NT * 8)                 /* positions per block-scan iteration */
#define ITERS (TILE / SPAN)
#define MAXWIN (TILE / 100 + 2)       /* GC windows that can overlap one tile */

static_assert(TILE % SPAN == 0, "tile must be a whole number of scan iterations");
static_assert((1 << CG_TILE_SHIFT) == TILE, "tile shift");

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

__device__ __forceinline__ unsigned long long ld_state(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_state(unsigned long long* p, unsigned long long v)
{
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}
/* streaming stores: depth/prefix are written once and not re-read by this kernel */
__device__ __forceinline__ void st_cs_v4(int32_t* p, int a, int b, int c, int d)
{
    asm volatile("st.global.cs.v4.s32 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void st_cs_v2(int16_t* p, unsigned a, unsigned b)
{
    asm volatile("st.global.cs.v2.u32 [%0], {%1,%2};" :: "l"(p), "r"(a), "r"(b) : "memory");
}

#define ST_AGG  1ull
#define ST_INCL 2ull

/* ------------------------------------------------------------------------- */
/* chromosome start: clear look-back words, tile counter and the stats row   */
__global__ void k_chrom_init(unsigned long long* tile_state, int n_tiles, unsigned int* tile_counter,
                             long long* stats_row, unsigned int* flags)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    for (int j = i; j < n_tiles; j += gridDim.x * blockDim.x) tile_state[j] = 0ull;
    if (i < CG_STATS_WORDS) stats_row[i] = 0;
    if (i == 0) { *tile_counter = 0u; *flags = 0u; }
}

cudaError_t cgk_chrom_init(cudaStream_t s, unsigned long long* tile_state, int n_tiles,
                           unsigned int* tile_counter, long long* stats_row, unsigned int* flags)
{
    int blocks = (n_tiles + 255) / 256;
    if (blocks < 1) blocks = 1;
    if (blocks > 148) blocks = 148;
    k_chrom_init<<<blocks, 256, 0, s>>>(tile_state, n_tiles, tile_counter, stats_row, flags);
    return cudaGetLastError();
}

/* ------------------------------------------------------------------------- */
/* per pushed batch: sortedness check + tile boundaries in one sweep          */
__device__ __forceinline__ int tile_of(int pos, int n_tiles)
{
    int t = pos >> CG_TILE_SHIFT;             /* arithmetic shift: negative pos -> negative tile */
    t = t < 0 ? 0 : t;
    return t >= n_tiles ? n_tiles - 1 : t;
}

__global__ void k_mark_batch(const cg_read* __restrict__ reads, long long first, long long n,
                             int32_t* tile_lo, int n_tiles, int32_t chrom_len, unsigned int* flags)
{
    long long stride = (long long)gridDim.x * blockDim.x;
    unsigned f = 0;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
        long long i = first + k;
        int p = reads[i].pos;
        int tp = tile_of(p, n_tiles), tq = -1;
        if (i > 0) {
            int q = reads[i - 1].pos;
            if (p < q) f |= CG_FLAG_UNSORTED;
            tq = tile_of(q, n_tiles);
        }
        if (p < 0 || p >= chrom_len) f |= CG_FLAG_RANGE;
        for (int t = tq + 1; t <= tp; t++) tile_lo[t] = (int32_t)i;
    }
    if (f) atomicOr(flags, f);
}

cudaError_t cgk_mark_batch(cudaStream_t s, const cg_read* reads, long long first, long long n,
                           int32_t* tile_lo, int n_tiles, int32_t chrom_len, unsigned int* flags)
{
    if (n <= 0) return cudaSuccess;
    long long blocks = (n + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_mark_batch<<<(int)blocks, 256, 0, s>>>(reads, first, n, tile_lo, n_tiles, chrom_len, flags);
    return cudaGetLastError();
}

/* tiles after the last record's tile start at N */
__global__ void k_seal_bounds(const cg_read* __restrict__ reads, long long n_reads, int32_t* tile_lo, int n_tiles)
{
    int last = n_reads > 0 ? tile_of(reads[n_reads - 1].pos, n_tiles) : -1;
    for (int t = last + 1 + blockIdx.x * blockDim.x + threadIdx.x; t <= n_tiles; t += gridDim.x * blockDim.x)
        tile_lo[t] = (int32_t)n_reads;
}

cudaError_t cgk_seal_bounds(cudaStream_t s, const cg_read* reads, long long n_reads, int32_t* tile_lo, int n_tiles)
{
    k_seal_bounds<<<8, 256, 0, s>>>(reads, n_reads, tile_lo, n_tiles);
    return cudaGetLastError();
}

/* ------------------------------------------------------------------------- */
/* out-of-order input: global-atomic histogram into u32 words; both kernels   */
/* return at once when the batch checks found the stream sorted               */
__global__ void k_unsorted_zero(uint4* depth32, long long n_vec, const unsigned int* flags)
{
    if (!(*flags & CG_FLAG_UNSORTED)) return;
    long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += stride)
        depth32[i] = make_uint4(0, 0, 0, 0);
}

__global__ void k_unsorted_scatter(const cg_read* __restrict__ reads, long long n_reads, uint32_t* depth32,
                                   int32_t chrom_len, int mq_threshold, const unsigned int* flags,
                                   long long* stats_row)
{
    if (!(*flags & CG_FLAG_UNSORTED)) return;
    long long stride = (long long)gridDim.x * blockDim.x;
    int pass = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_reads; i += stride) {
        uint2 r = *reinterpret_cast<const uint2*>(&reads[i]);
        int pos = (int)r.x, mapq = (int)((r.y >> 16) & 0xffu);
        if (mapq > mq_threshold) {
            pass++;
            if (pos >= 0 && pos < chrom_len) atomicAdd(&depth32[pos], 1u);
        }
    }
    pass = __reduce_add_sync(FULL, pass);
    if ((threadIdx.x & 31) == 0 && pass)
        atomicAdd((unsigned long long*)&stats_row[2 * CG_GC_BINS + 1], (unsigned long long)pass);
}

cudaError_t cgk_unsorted_scatter(cudaStream_t s, const cg_read* reads, long long n_reads,
                                 uint32_t* depth32, long long n_words, int32_t chrom_len,
                                 int mq_threshold, const unsigned int* flags, long long* stats_row, int n_sms)
{
    k_unsorted_zero<<<n_sms * 4, 256, 0, s>>>(reinterpret_cast<uint4*>(depth32), n_words / 4, flags);
    k_unsorted_scatter<<<n_sms * 4, 256, 0, s>>>(reads, n_reads, depth32, chrom_len, mq_threshold, flags, stats_row);
    return cudaGetLastError();
}

/* ------------------------------------------------------------------------- */
/* the fused pass                                                            */
struct __align__(16) pass_smem {
    uint4 bins4[TILE / 4];                    /* 64 KB: one u32 count per position of the tile */
    unsigned long long gc_rd[CG_GC_BINS];     /* per-CTA rd_per_gc (two's complement int64) */
    unsigned long long gc_cnt[CG_GC_BINS];    /* per-CTA window_per_gc */
    int win_end[MAXWIN];                      /* tile-local inclusive prefix at each window's last position */
    int wtot[2][NWARP];
    int tile;
    int tile_total;
    unsigned carry;
};

__global__ void __launch_bounds__(NT, 3)
k_depth_pass(const cg_pass_args a)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    pass_smem& sm = *reinterpret_cast<pass_smem*>(smem_raw);
    unsigned* bins = reinterpret_cast<unsigned*>(sm.bins4);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool unsorted = (*a.flags & CG_FLAG_UNSORTED) != 0;

    for (int i = tid; i < TILE / 4; i += NT) sm.bins4[i] = make_uint4(0, 0, 0, 0);
    for (int i = tid; i < CG_GC_BINS; i += NT) { sm.gc_rd[i] = 0ull; sm.gc_cnt[i] = 0ull; }
    int my_pass = 0;            /* records of this thread that passed the mapq test */
    long long cta_rd = 0;       /* thread 0: sum of this CTA's tile totals (rd_cnt share) */

    for (;;) {
        if (tid == 0) { sm.tile = (int)atomicAdd(a.tile_counter, 1u); sm.tile_total = 0; }
        __syncthreads();
        const int t = sm.tile;
        if (t >= a.n_tiles) break;
        const int base = t << CG_TILE_SHIFT;

        /* ---- phase A: the tile's records -> shared-memory histogram ---------- */
        int contrib = 0;        /* this thread's share of sum(sign-extended int16 depth) */
        if (!unsorted) {
            const int lo = a.tile_lo[t], hi = a.tile_lo[t + 1];
            for (int i0 = lo + warp * 32; i0 < hi; i0 += NT) {
                const int i = i0 + lane;
                const bool in = i < hi;
                uint2 r = make_uint2(0u, 0u);
                if (in) r = __ldg(reinterpret_cast<const uint2*>(a.reads) + i);
                const int pos = (int)r.x, mapq = (int)((r.y >> 16) & 0xffu);
                const bool pass = in && mapq > a.mq_threshold;                 /* bam_data.c:205 */
                my_pass += pass;
                const unsigned off = (unsigned)(pos - base);
                const bool ok = pass && off < (unsigned)TILE && pos < a.chrom_len;
                /* warp aggregation: the stream is sorted, so equal positions sit in adjacent lanes */
                const int key = ok ? (int)off : -1 - lane;
                const int prev = __shfl_up_sync(FULL, key, 1);
                const bool leader = (lane == 0) || key != prev;
                const unsigned lm = __ballot_sync(FULL, leader);
                if (ok && leader) {
                    const unsigned above = (lane == 31) ? 0u : (lm & ~((2u << lane) - 1u));
                    const int cnt = (above ? __ffs(above) - 1 : 32) - lane;
                    const unsigned old = atomicAdd(&bins[off], (unsigned)cnt);     /* bam_data.c:213 */
                    contrib += sext16(old + cnt) - sext16(old);
                }
            }
        } else {
            const uint4* src = reinterpret_cast<const uint4*>(a.depth32 + (size_t)base);
            for (int q = tid; q < TILE / 4; q += NT) {
                uint4 v = src[q];
                sm.bins4[q] = v;
                contrib += sext16(v.x) + sext16(v.y) + sext16(v.z) + sext16(v.w);
            }
        }
        contrib = __reduce_add_sync(FULL, contrib);
        if (lane == 0 && contrib) atomicAdd(&sm.tile_total, contrib);
        __syncthreads();

        /* ---- single-pass look-back for the tile's starting prefix ------------ */
        if (warp == 0) {
            const unsigned agg = (unsigned)sm.tile_total;
            unsigned excl = 0;
            if (t == 0) {
                if (lane == 0) st_state(&a.tile_state[0], (ST_INCL << 32) | agg);
            } else {
                if (lane == 0) st_state(&a.tile_state[t], (ST_AGG << 32) | agg);
                int idx = t - 1 - lane;
                for (;;) {
                    unsigned long long w = (ST_INCL << 32);
                    if (idx >= 0) {
                        do { w = ld_state(&a.tile_state[idx]); } while ((w >> 32) == 0ull);
                    }
                    const unsigned incl = __ballot_sync(FULL, (w >> 32) == ST_INCL);
                    unsigned v = (unsigned)w;
                    if (incl) {
                        const int first = __ffs(incl) - 1;       /* nearest tile with a full prefix */
                        if (lane > first) v = 0;
                        excl += __reduce_add_sync(FULL, v);
                        break;
                    }
                    excl += __reduce_add_sync(FULL, v);
                    idx -= 32;
                }
                if (lane == 0) st_state(&a.tile_state[t], (ST_INCL << 32) | (unsigned)(excl + agg));
            }
            if (lane == 0) { sm.carry = excl; cta_rd += (long long)(int)agg; }
        }

        /* ---- block-scan iterations: bins -> depth + prefix -------------------- */
        const int w0 = base / 100;                 /* first GC window overlapping the tile */
        unsigned run = 0;                          /* prefix at the start of this iteration's span */
        unsigned carry0 = 0;
#pragma unroll 1
        for (int it = 0; it < ITERS; it++) {
            /* warp `warp` owns 256 positions; lane owns quad A (4 positions) and quad B 128 further on */
            const int sp = it * SPAN + warp * 256;
            const int qa = (sp >> 2) + lane, qb = qa + 32;
            const uint4 va = sm.bins4[qa], vb = sm.bins4[qb];
            sm.bins4[qa] = make_uint4(0, 0, 0, 0);             /* clean bins for the next tile */
            sm.bins4[qb] = make_uint4(0, 0, 0, 0);
            const int a0 = sext16(va.x), a1 = sext16(va.y), a2 = sext16(va.z), a3 = sext16(va.w);
            const int b0 = sext16(vb.x), b1 = sext16(vb.y), b2 = sext16(vb.z), b3 = sext16(vb.w);
            const int sa = a0 + a1 + a2 + a3, sb = b0 + b1 + b2 + b3;
            const int ia = warp_incl_scan(sa, lane), ib = warp_incl_scan(sb, lane);
            const int ta = __shfl_sync(FULL, ia, 31), tb = __shfl_sync(FULL, ib, 31);
            if (lane == 31) sm.wtot[it & 1][warp] = ta + tb;
            __syncthreads();
            if (it == 0) { carry0 = sm.carry; run = carry0; }
            int woff = 0, btot = 0;
#pragma unroll
            for (int j = 0; j < NWARP; j++) {
                const int v = sm.wtot[it & 1][j];
                btot += v;
                if (j < warp) woff += v;
            }
            const int pa0 = (int)run + woff + (ia - sa) + a0, pa1 = pa0 + a1, pa2 = pa1 + a2, pa3 = pa2 + a3;
            const int pb0 = (int)run + woff + ta + (ib - sb) + b0, pb1 = pb0 + b1, pb2 = pb1 + b2, pb3 = pb2 + b3;
            const size_t ga = (size_t)base + (size_t)(qa << 2), gb = ga + 128;
            st_cs_v2(a.depth + ga, (va.x & 0xffffu) | (va.y << 16), (va.z & 0xffffu) | (va.w << 16));
            st_cs_v2(a.depth + gb, (vb.x & 0xffffu) | (vb.y << 16), (vb.z & 0xffffu) | (vb.w << 16));
            st_cs_v4(a.prefix + ga, pa0, pa1, pa2, pa3);
            st_cs_v4(a.prefix + gb, pb0, pb1, pb2, pb3);
            /* a quad starting at global position == 96 (mod 100) ends a GC window */
            const unsigned pga = (unsigned)ga, pgb = (unsigned)gb;
            if (pga % 100u == 96u) sm.win_end[pga / 100u - w0] = pa3 - (int)carry0;
            if (pgb % 100u == 96u) sm.win_end[pgb / 100u - w0] = pb3 - (int)carry0;
            if (it == ITERS - 1 && tid == NT - 1)
                sm.win_end[(unsigned)(base + TILE - 1) / 100u - w0] = pb3 - (int)carry0;
            run += (unsigned)btot;
        }
        __syncthreads();

        /* ---- GC-bin sums from window-end prefixes (read_distribution.c:63-73) -- */
        {
            const int nwin = (base + TILE - 1) / 100 - w0 + 1;
            const int tile_end = min(base + TILE, a.chrom_len);
            if (tid < nwin) {
                const int w = w0 + tid;
                const int ws = max(base, 100 * w), we = min(tile_end, 100 * w + 100);
                if (we > ws) {
                    int bin = (int)a.gc[w];                       /* sonic_interval.c:263-272 */
                    if ((unsigned)bin > 100u) bin = 0;
                    const int sum = sm.win_end[tid] - (tid ? sm.win_end[tid - 1] : 0);
                    if (sum) atomicAdd(&sm.gc_rd[bin], (unsigned long long)(long long)sum);
                    atomicAdd(&sm.gc_cnt[bin], (unsigned long long)(we - ws));
                }
            }
        }
        /* the loop-top barrier orders these reads before the next tile's writes */
    }

    /* ---- CTA epilogue: flush the per-CTA statistics ---------------------------- */
    for (int b = tid; b < CG_GC_BINS; b += NT) {
        if (sm.gc_rd[b])  atomicAdd((unsigned long long*)&a.stats[b], sm.gc_rd[b]);
        if (sm.gc_cnt[b]) atomicAdd((unsigned long long*)&a.stats[CG_GC_BINS + b], sm.gc_cnt[b]);
    }
    if (tid == 0 && cta_rd) atomicAdd((unsigned long long*)&a.stats[2 * CG_GC_BINS], (unsigned long long)cta_rd);
    my_pass = __reduce_add_sync(FULL, my_pass);
    if (lane == 0 && my_pass)
        atomicAdd((unsigned long long*)&a.stats[2 * CG_GC_BINS + 1], (unsigned long long)my_pass);
}

int cgk_depth_pass_ctas_per_sm(void) { return 3; }

cudaError_t cgk_depth_pass(cudaStream_t s, const cg_pass_args& a, int n_sms)
{
    static bool configured = false;
    const int smem = (int)sizeof(pass_smem);
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(k_depth_pass, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    int grid = n_sms * cgk_depth_pass_ctas_per_sm();
    if (grid > a.n_tiles) grid = a.n_tiles;
    if (grid < 1) grid = 1;
    k_depth_pass<<<grid, NT, smem, s>>>(a);
    return cudaGetLastError();
}

/* ------------------------------------------------------------------------- */
/* expected_read_depth[101], read_distribution.c:75-83:                       */
/*   expected[0] = 0; expected[i] = (float)rd_per_gc[i] / window_per_gc[i];    */
/*   NaN / Inf -> 0.  long->float and int->float round to nearest like x86's   */
/*   cvtsi2ss, and the quotient is one IEEE fp32 division.                     */
__global__ void k_expected_table(const long long* __restrict__ stats_row, float* expected)
{
    int i = threadIdx.x;
    if (i >= CG_GC_BINS) return;
    float e = 0.0f;
    if (i > 0) {
        const float num = __ll2float_rn(stats_row[i]);
        const float den = __int2float_rn((int)stats_row[CG_GC_BINS + i]);
        e = __fdiv_rn(num, den);
        if (isnan(e) || isinf(e)) e = 0.0f;
    }
    expected[i] = e;
}

cudaError_t cgk_expected_table(cudaStream_t s, const long long* stats_row, float* expected)
{
    k_expected_table<<<1, 128, 0, s>>>(stats_row, expected);
    return cudaGetLastError();
}
