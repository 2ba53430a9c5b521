/* cg_internal.h — shared between the translation units of libconga_gpu.so.
 * Not part of the ABI (include/conga_gpu.h is). */
#ifndef CG_INTERNAL_H
#define CG_INTERNAL_H

#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/conga_gpu.h"

/* ---- geometry of the fused depth pass (K1+K2+K3) ------------------------- */
#define CG_TILE        16384          /* positions per tile: 32 KB of packed u16 bins in smem */
#define CG_PASS_THREADS  256
#define CG_TILE_SHIFT  14
#define CG_CHUNK       2048           /* positions one warp scans; unit of the record pre-count */
#define CG_CHUNK_SHIFT 11
#define CG_CHUNKS_PER_TILE (CG_TILE / CG_CHUNK)
#define CG_PILEUP_PROBE 16384         /* k_mark_batch looks this many records back for an equal pos */
#define CG_STATS_WORDS 204            /* int64 words per chromosome row = sizeof(cg_chrom_stats)/8 */

/* Device-side error/condition bits, one u32 word per context. */
#define CG_FLAG_UNSORTED   1u         /* a pushed record had pos < its predecessor */
#define CG_FLAG_RANGE      2u         /* a record or CNV outside [0, chrom_len) */
#define CG_FLAG_PILEUP     4u         /* > CG_PILEUP_PROBE records share one pos: int16 wrap possible */
/* words of the per-context u32 control block (d_flags) */
#define CG_W_FLAGS   0
#define CG_W_TILE    1                /* dynamic tile counter of the depth pass */
#define CG_W_DONE    2                /* CTAs of the depth pass that have flushed their statistics */

struct cg_pass_args {
    const cg_read* reads;             /* all records of the chromosome, device */
    long long      n_reads;
    const int32_t* tile_lo;           /* [n_tiles+1] first record index of each tile (sorted input) */
    const int32_t* chunk_cnt;         /* [n_tiles*8] sum of the (int16) depths of each 2048-position chunk */
    const uint32_t* tile_carry;       /* [n_tiles] prefix (mod 2^32) in front of each tile */
    const int8_t*  gc;                /* SONIC GC byte per 100-position window */
    int32_t        chrom_len;
    int32_t        n_tiles;
    int            mq_threshold;
    int16_t*       depth;             /* [n_tiles*CG_TILE] */
    int32_t*       prefix;            /* [n_tiles*CG_TILE] inclusive int32 prefix of depth */
    uint32_t*      depth32;           /* unsorted fallback: per-position u32 counts (zero at rest) */
    unsigned int*  ctl;               /* CG_W_* control words */
    long long*     stats;             /* this chromosome's int64[204] row */
    float*         expected;          /* [101] written by the last CTA to finish */
    unsigned int*  sticky;            /* context-wide word: CG_FLAG_RANGE is copied here */
};

#ifdef __cplusplus
#include <vector>
extern "C++" {
/* conga_gpu.cu / flatten.cu */
int  cg_fail(cg_ctx* c, int code, const char* fmt, ...);
void cg_flatten_rows(const cg_mapiv* rows, size_t n, int32_t chrom_len, std::vector<cg_maprun>& out);
/* depth_pass.cu */
cudaError_t cgk_chrom_init(cudaStream_t s, int32_t* chunk_cnt, int n_chunks, long long* stats_row, unsigned int* ctl);
cudaError_t cgk_mark_batch(cudaStream_t s, const cg_read* reads, long long first, long long n,
                           int32_t* tile_lo, int32_t* chunk_cnt, int n_tiles, int32_t chrom_len,
                           int mq_threshold, unsigned int* ctl, long long* stats_row);
cudaError_t cgk_rare_paths(cudaStream_t s, const cg_read* reads, long long n_reads, uint32_t* depth32,
                           int32_t* chunk_cnt, int n_tiles, int32_t chrom_len, int mq_threshold,
                           const unsigned int* ctl, int n_sms);
cudaError_t cgk_tile_scan(cudaStream_t s, const cg_read* reads, long long n_reads, int32_t* tile_lo,
                          const int32_t* chunk_cnt, uint32_t* tile_carry, int n_tiles, long long* stats_row);
cudaError_t cgk_depth_pass(cudaStream_t s, const cg_pass_args& a, int n_sms);
cudaError_t cgk_expected_table(cudaStream_t s, const long long* stats_row, float* expected);
int         cgk_depth_pass_ctas_per_sm(void);

/* genotype.cu */
cudaError_t cgk_genotype(cudaStream_t s, const cg_cnv* cnvs, int n, int is_dup,
                         const int32_t* prefix, const int8_t* gc, const float* expected,
                         const cg_maprun* runs, int n_runs, int has_map,
                         int32_t chrom_len, unsigned int* flags, cg_call* out,
                         unsigned int* work_counter, unsigned int* work_base, int n_sms);
}
#endif

#endif /* CG_INTERNAL_H */
