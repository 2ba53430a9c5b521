/* conga_oracle.c — CPU restatement of CONGA's read-depth genotyping hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the checker the CUDA path is compared
 * against; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may build, load or call it.  Nothing under conga_b200/
 * links or imports it, and the product has no CPU fallback.
 *
 * Parity status: CONGA ships no tests, golden vectors or example data
 * (SURVEY.md §4), so upstream pins nothing.  This restatement is instead pinned
 * against the UNMODIFIED reference objects compiled from /root/reference
 * (oracle/Makefile target `ref` -> oracle/_ref/conga_dump, a --wrap shim that
 * dumps the reference's own intermediates); the resulting vectors are committed
 * under tests/golden/ and tests/test_oracle_golden.py replays them here.
 * libm (log, lgamma) comes from the host glibc, exactly as in the reference.
 *
 * Written from the behaviour of the reference, one loop per reference loop, with
 * the same C types so the same integer wrap and fp32/fp64 rounding happen:
 *   orc_count_reads          bam_data.c:201-217        (count_reads_bam body)
 *   orc_calc_mu              read_distribution.c:27-46 (calc_mu_per_chr)
 *   orc_calc_mean_per_chr    read_distribution.c:49-84 (calc_mean_per_chr)
 *   orc_gc_bin               sonic/sonic_interval.c:245-273 as called with a
 *                            <=100-wide interval, + (int)round() at the call sites
 *   orc_paint_mappability    svs.c:317-377 + read_distribution.c:20-24
 *   orc_lpoisson             likelihood.c:96-105
 *   orc_imax                 common.c:262-268 (int max(int,int) fed doubles)
 *   orc_calculate_likelihood likelihood.c:108-169
 *   orc_compare_start_pos    common.c:199-215
 *   orc_gc_profile           sonic/sonic_reference.c:57-104 (the SONIC builder's GC bytes of one record)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/conga_gpu.h"

/* bam_data.c:201-217.  `depth` plays in_bam->read_depth (short, zeroed by
 * init_rd_per_chr, read_distribution.c:12-18).  The increment is done in
 * unsigned 16-bit arithmetic and stored back, which is what gcc emits for
 * `short++` (wraps 32767 -> -32768). */
void orc_count_reads(const cg_read* recs, size_t n, int mq_threshold,
                     int16_t* depth, int32_t chrom_len, int32_t* total_read_count)
{
    int32_t cnt = *total_read_count;
    for (size_t i = 0; i < n; i++) {
        if ((int)recs[i].mapq > mq_threshold) {            /* bam_data.c:205 */
            int32_t p = recs[i].pos;
            if (p >= 0 && p < chrom_len)                   /* iterator guarantees it, A3 */
                depth[p] = (int16_t)(uint16_t)((uint16_t)depth[p] + 1u);   /* :213 */
            cnt++;                                          /* :214 */
        }
    }
    *total_read_count = cnt;
}

/* read_distribution.c:27-46 */
void orc_calc_mu(const int16_t* depth, int32_t chrom_len, int64_t* rd_cnt_out, float* mean_out)
{
    long rd_cnt = 0, window_total = 0;
    for (int i = 0; i < chrom_len; i++) {
        rd_cnt += (long)depth[i];
        window_total++;
    }
    float mean = (double)rd_cnt / (double)window_total;    /* :39, stored to a float field */
    *rd_cnt_out = rd_cnt;
    *mean_out = mean;
}

/* sonic_get_gc_content(sonic, chr, i, end) with end - i <= 100 runs its loop
 * exactly once (sonic/sonic_interval.c:263-267) and returns
 * (float)gc_profile[i/100] / 1, or 0.0 when that is 0; the callers apply
 * (int)round() (read_distribution.c:70, likelihood.c:117). */
static inline int orc_gc_bin_inl(const int8_t* gc, int pos)
{
    float gc_content = 0.0f;
    gc_content += (float)gc[pos / 100];
    float r = (gc_content != 0.0f) ? (gc_content / 1) : 0.0f;
    return (int)round(r);
}
int orc_gc_bin(const int8_t* gc, int pos) { return orc_gc_bin_inl(gc, pos); }

/* read_distribution.c:49-84 */
void orc_calc_mean_per_chr(const int16_t* depth, int32_t chrom_len, const int8_t* gc,
                           int64_t rd_per_gc[CG_GC_BINS], int64_t window_per_gc_out[CG_GC_BINS],
                           float expected[CG_GC_BINS])
{
    int window_per_gc[CG_GC_BINS];
    long rd_per_gc_unfiltered[CG_GC_BINS];
    for (int i = 0; i < CG_GC_BINS; i++) { rd_per_gc_unfiltered[i] = 0; window_per_gc[i] = 0; }

    for (int i = 0; i < chrom_len; i++) {
        int gc_val = orc_gc_bin_inl(gc, i);                /* :70 */
        rd_per_gc_unfiltered[gc_val] += (long)depth[i];    /* :71 */
        window_per_gc[gc_val]++;                           /* :72 */
    }
    expected[0] = 0.0;                                     /* :75 */
    for (int i = 1; i < CG_GC_BINS; i++) {
        expected[i] = (float)rd_per_gc_unfiltered[i] / (window_per_gc[i]);   /* :78 */
        if (isnan(expected[i]) || isinf(expected[i]))      /* :80-82 */
            expected[i] = 0;
    }
    for (int i = 0; i < CG_GC_BINS; i++) {
        rd_per_gc[i] = rd_per_gc_unfiltered[i];
        window_per_gc_out[i] = window_per_gc[i];
    }
}

/* read_distribution.c:20-24 + svs.c:350-371.  Rows are the current chromosome's
 * rows in file order; inclusive end; later rows overwrite.  The reference paints
 * unchecked (A10); rows are clipped to the array here so the oracle cannot
 * corrupt memory — generators never emit out-of-range rows. */
void orc_paint_mappability(float* mapp, int32_t chrom_len, const cg_mapiv* rows, size_t n)
{
    memset(mapp, 0, (size_t)chrom_len * sizeof(float));
    for (size_t r = 0; r < n; r++) {
        int s = rows[r].start, e = rows[r].end;
        if (s < 0) s = 0;
        if (e > chrom_len - 1) e = chrom_len - 1;
        for (int i = s; i <= e; i++) mapp[i] = rows[r].val;
    }
}

/* likelihood.c:96-105 */
double orc_lpoisson(int observed, double lambda)
{
    double penalty = 0.01;
    if (lambda == 0.0) lambda = penalty;
    return observed * log(lambda) - lambda - lgamma(observed + 1);
}

/* common.c:262-268 — `int max(int,int)`; the call sites pass doubles, which C
 * converts by truncation toward zero. */
static int orc_imax_i(int x, int y) { return (x < y) ? y : x; }
int orc_imax(double x, double y) { return orc_imax_i((int)x, (int)y); }

/* likelihood.c:108-169.  mapp == NULL <=> params->mappability_file == NULL. */
void orc_calculate_likelihood(const int16_t* depth, const int8_t* gc,
                              const float expected[CG_GC_BINS], const float* mapp,
                              int32_t start, int32_t end, int is_dup, cg_call* out)
{
    float expected_rd = 0;
    int observed_rd = 0;
    double mappability_score = 0;

    for (int i = start; i < end; i++) {                    /* :115 */
        int gc_val = orc_gc_bin_inl(gc, i);                /* :117 */
        expected_rd += expected[gc_val];                   /* :119  fp32, in order */
        observed_rd += depth[i];                           /* :120 */
        if (mapp != NULL) mappability_score += mapp[i];    /* :122-123 fp64 += fp32 */
    }
    memset(out, 0, sizeof(*out));
    if (mapp != NULL) out->mappability = mappability_score / (double)(end - start);   /* :128 */

    if (!is_dup) {
        out->lhomo   = orc_lpoisson(observed_rd, 0.0);                 /* :133 */
        out->lhetero = orc_lpoisson(observed_rd, 0.5 * expected_rd);   /* :134 */
        out->lnone   = orc_lpoisson(observed_rd, expected_rd);         /* :135 */
        out->score = orc_imax_i(out->lhomo, out->lhetero) / out->lnone;   /* :138 */
        out->observed = observed_rd;
        out->expected = expected_rd;
        out->copy_number = (observed_rd < (expected_rd / 4)) ? 2 : 1;  /* :146-149 */
    } else {
        out->lhomo   = orc_lpoisson(observed_rd, 2 * expected_rd);     /* :155 (float product) */
        out->lhetero = orc_lpoisson(observed_rd, 1.5 * expected_rd);   /* :156 */
        out->lnone   = orc_lpoisson(observed_rd, expected_rd);         /* :157 */
        out->score = orc_imax_i(out->lhomo, out->lhetero) / out->lnone;   /* :160 */
        out->observed = observed_rd;
        out->expected = expected_rd;
        out->copy_number = (out->lhomo > out->lhetero) ? 2 : 1;        /* :164-167 */
    }
}

/* find_depths, likelihood.c:290-308: every CNV of one type in array order. */
void orc_find_depths(const int16_t* depth, const int8_t* gc, const float expected[CG_GC_BINS],
                     const float* mapp, const cg_cnv* cnvs, size_t n, int is_dup, cg_call* out)
{
    for (size_t c = 0; c < n; c++)
        orc_calculate_likelihood(depth, gc, expected, mapp, cnvs[c].start, cnvs[c].end, is_dup, &out[c]);
}

/* common.c:199-215 on (start,end) pairs; used with qsort like likelihood.c:324-328. */
int orc_compare_start_pos(const void* a, const void* b)
{
    const cg_cnv* x = (const cg_cnv*)a; const cg_cnv* y = (const cg_cnv*)b;
    if (x->start > y->start) return 1;
    else if (x->start == y->start) return x->end - y->end;
    else return -1;
}
void orc_sort_cnvs(cg_cnv* cnvs, size_t n) { qsort(cnvs, n, sizeof(cg_cnv), orc_compare_start_pos); }

/* The whole per-chromosome path in one call: the CPU baseline bench.py times
 * (kind "port").  depth and mapp are caller-provided scratch of chrom_len. */
void orc_chromosome(const cg_read* recs, size_t n_reads, int mq_threshold,
                    int32_t chrom_len, const int8_t* gc,
                    const cg_mapiv* rows, size_t n_rows,
                    const cg_cnv* dels, size_t n_dels, cg_call* del_out,
                    const cg_cnv* dups, size_t n_dups, cg_call* dup_out,
                    int16_t* depth, float* mapp, cg_chrom_stats* stats, float expected[CG_GC_BINS])
{
    int32_t cnt = 0; float mean;
    memset(depth, 0, (size_t)chrom_len * sizeof(int16_t));        /* read_distribution.c:17 */
    orc_count_reads(recs, n_reads, mq_threshold, depth, chrom_len, &cnt);
    orc_calc_mu(depth, chrom_len, &stats->rd_cnt, &mean);
    orc_calc_mean_per_chr(depth, chrom_len, gc, stats->rd_per_gc, stats->window_per_gc, expected);
    stats->reads_passed = cnt;
    const float* m = NULL;
    if (n_rows > 0 && mapp != NULL) { orc_paint_mappability(mapp, chrom_len, rows, n_rows); m = mapp; }
    orc_find_depths(depth, gc, expected, m, dels, n_dels, 0, del_out);
    orc_find_depths(depth, gc, expected, m, dups, n_dups, 1, dup_out);
}

/* ======================================================================================
 * Split-read support (--rp with --dups): restatement of the part of split_read.c /
 * bam_data.c / likelihood.c that the CUDA path replaces.  Same caveat as above: TEST
 * INFRASTRUCTURE ONLY.  Pinned against the reference binary through the per-CNV rp /
 * border_rp values its --wrap dump holds (tests/golden/rp/ref_rp.npz) and, on the GPU box,
 * through the command-line parity tests.
 *   orc_sr_rows    build_hash_table x2 + init_hash_table     split_read.c:338-455
 *                  almostPerfect_match_seq_ref               split_read.c:76-204
 *                  read_SplitReads' tests (satellite term excluded) + determine_SvType
 *                                                            bam_data.c:29-154
 *   orc_sr_count   count_ReadPairs                           likelihood.c:41-94
 * ====================================================================================== */
#include <stdio.h>

#define ORC_K 10                 /* HASHKMERLEN   split_read.h:18 */
#define ORC_MAX_SR_HIT 50000     /* split_read.h:13 */
#define ORC_MAX_MAPPING 100      /* split_read.h:14 */
#define ORC_LOOKAHEAD 100000     /* split_read.c:5 */

static int orc_is_dna(char c) { return c == 'A' || c == 'C' || c == 'G' || c == 'T'; }       /* common.c:303-314 */
static int orc_kmer_valid(const char* s)                                                       /* split_read.c:60-74 */
{
    for (int i = 0; i < ORC_K; i++) { if (s[i] == 0) return 0; if (!orc_is_dna(s[i])) return 0; }
    return 1;
}
static unsigned orc_hash_ref(const char* s)                                                    /* split_read.c:36-50 */
{
    unsigned val = 0;
    for (int i = 0; i < ORC_K; i++) val = (val << 2) | (unsigned)((s[i] & 0x6) >> 1);
    return val;
}

/* one scan of build_hash_table: mode 0 counts, mode 1 fills (split_read.c:338-421); the rolling
 * update is hash_function_next (split_read.c:29-34) */
static void orc_build(const char* ref, int len, int hash_size, int mode, int* count, int* iter, int** array)
{
    int i = 0, j;
    const unsigned mask = (unsigned)(hash_size - 1) >> 2;
    const int len_limit = len - ORC_K + 1;
    unsigned h;
    if (len < ORC_K) return;
    while (i < len && !orc_is_dna(ref[i])) i++;
    while (i < len_limit && !orc_kmer_valid(ref + i)) i++;
    if (i >= len_limit) return;
    h = orc_hash_ref(ref + i);
#define ORC_ADD(pos) do { if (mode == 0) count[h]++; else if (count[h] != 0) { array[h][iter[h]] = (pos); iter[h]++; } } while (0)
    ORC_ADD(i);
    j = i + ORC_K;
    while (i < len_limit) {
        i++;
        if (orc_is_dna(ref[j])) {
            h = ((h & mask) << 2) | (unsigned)((ref[j++] & 0x6) >> 1);
            ORC_ADD(i);
        } else {
            while (!orc_is_dna(ref[j]) && i < len_limit) { i++; j++; }
            if (i >= len_limit) break;
            while (i < len_limit && !orc_kmer_valid(ref + i)) i++;
            if (i >= len_limit) break;             /* the reference would run off the sequence here */
            j = i + ORC_K;
            h = orc_hash_ref(ref + i);
            ORC_ADD(i);
        }
    }
#undef ORC_ADD
}

/* rows in the order the reference's lists would yield them is not reproduced: callers compare multisets */
long orc_sr_rows(const char* ref, int ref_len, const cg_half* halves, size_t n_halves, const char* bases,
                 int chrom_len, int mq_threshold, cg_splitrow* out, size_t cap, uint64_t* n_mappings)
{
    const int hash_size = 1 << (2 * ORC_K);
    int* count = (int*)calloc((size_t)hash_size, sizeof(int));
    int* iter = (int*)calloc((size_t)hash_size, sizeof(int));
    int** array = (int**)calloc((size_t)hash_size, sizeof(int*));
    char* refz = (char*)malloc((size_t)ref_len + 2048);
    long n_rows = 0;
    uint64_t n_map = 0;
    memcpy(refz, ref, (size_t)ref_len);
    memset(refz + ref_len, 0, 2048);                /* the terminator; what lies behind it is not a base */
    orc_build(refz, ref_len, hash_size, 0, count, iter, array);
    for (int i = 0; i < hash_size; i++) {           /* init_hash_table, split_read.c:431-449 */
        if (count[i] != 0 && count[i] < ORC_MAX_SR_HIT) array[i] = (int*)malloc(sizeof(int) * (size_t)count[i]);
        else count[i] = 0;
    }
    orc_build(refz, ref_len, hash_size, 1, count, iter, array);

    int* posMap = (int*)malloc(sizeof(int) * 400000);
    char* orient = (char*)malloc(400000);
    for (size_t k = 0; k < n_halves; k++) {
        const cg_half* hf = &halves[k];
        char str[2048], rev[2048];
        const int len = hf->len < 2047 ? hf->len : 2047;
        memcpy(str, bases + hf->seq_off, (size_t)len);
        str[len] = 0;
        const int dist_max = (int)(0.05 * len);
        int size = 0;
        if (len < ORC_K) continue;
        if (orc_kmer_valid(str)) {
            const unsigned idx = orc_hash_ref(str);
            for (int c = 0; c < count[idx]; c++) {
                const int hit = array[idx][c];
                if (abs(hit - hf->pos) < ORC_LOOKAHEAD) {
                    int d = 0;
                    for (int i = 0; i < len; i++) if (refz[hit + i] != str[i]) d++;          /* hammingDistance, common.c:277-287 */
                    if (d <= dist_max && size < 400000) { posMap[size] = hit; orient[size] = 'F'; size++; }
                }
            }
        }
        if (size < ORC_MAX_MAPPING) {
            for (int i = 0; i < len; i++) {
                const char c = str[i];
                rev[len - i - 1] = c == 'A' ? 'T' : c == 'T' ? 'A' : c == 'G' ? 'C' : c == 'C' ? 'G' : c;
            }
            rev[len] = 0;
            if (orc_kmer_valid(rev)) {
                const unsigned idx = orc_hash_ref(rev);
                for (int c = 0; c < count[idx]; c++) {
                    const int hit = array[idx][c];
                    if (abs(hit - hf->pos) < ORC_LOOKAHEAD) {
                        int d = 0;
                        for (int i = 0; i < len; i++) if (refz[hit + i] != rev[i]) d++;
                        if (d <= dist_max) { posMap[size] = hit; orient[size] = 'R'; size++; }
                    }
                    if (size > ORC_MAX_MAPPING) break;
                }
            }
        }
        if (!(size > 0 && size < ORC_MAX_MAPPING)) continue;
        n_map += (uint64_t)size;
        const int mapq = 60 / size;
        for (int m = 0; m < size; m++) {
            const int pos = hf->pos, pm = posMap[m];
            /* read_SplitReads, bam_data.c:104-106, without the satellite term */
            if (!((int)hf->mapq > mq_threshold && mapq > mq_threshold && pos > 0 && pm > 0 && pos < chrom_len && pm < chrom_len)) continue;
            /* determine_SvType, bam_data.c:29-78 */
            const int lengthSplit = hf->read_len / 2, lengthRead = hf->read_len - lengthSplit;
            int p11, p12, p21, p22;
            if (pos < pm) { p11 = pos; p12 = pos + lengthRead; p21 = pm; p22 = pm + lengthSplit; }
            else if (pm < pos) { p11 = pm; p12 = pm + lengthSplit; p21 = pos; p22 = pos + lengthRead; }
            else continue;
            if (p12 >= p21) continue;
            if (orient[m] != 'F') continue;                      /* splitRead.orient is always FORWARD */
            char type;
            if ((pos < pm && !hf->is_read2) || (pos > pm && hf->is_read2)) type = 'D';
            else if ((pm < pos && !hf->is_read2) || (pm > pos && hf->is_read2)) type = 'E';
            else continue;
            if ((size_t)n_rows < cap) {
                cg_splitrow r;
                memset(&r, 0, sizeof r);
                r.pos = pos; r.pos_map = pm; r.left_start = p11; r.left_end = p12; r.right_start = p21; r.right_end = p22;
                r.sv_type = (uint8_t)type;
                out[n_rows] = r;
            }
            n_rows++;
        }
    }
    for (int i = 0; i < hash_size; i++) free(array[i]);
    free(count); free(iter); free(array); free(refz); free(posMap); free(orient);
    if (n_mappings) *n_mappings = n_map;
    return n_rows;
}

/* count_ReadPairs, likelihood.c:41-94; rows carry the widened coordinates */
void orc_sr_count(const cg_splitrow* rows, size_t n, const cg_cnv* dels, size_t n_del, const cg_cnv* dups, size_t n_dup,
                  int32_t* border_rp, int32_t* rp)
{
    for (size_t i = 0; i < n_del; i++) border_rp[i] = 0;
    for (size_t i = 0; i < n_dup; i++) rp[i] = 0;
    for (size_t k = 0; k < n; k++) {
        const cg_splitrow* r = &rows[k];
        if (r->sv_type == 'E') {
            for (size_t i = 0; i < n_dup; i++)
                if (r->left_end >= dups[i].start - 5000 && r->left_end <= dups[i].end + 5000
                    && r->right_start <= dups[i].end + 5000 && r->right_start >= dups[i].start - 5000) rp[i]++;
        } else if (r->sv_type == 'D') {
            for (size_t i = 0; i < n_del; i++)
                if (r->left_end <= dels[i].start + 100 && r->left_end >= dels[i].start - 5000
                    && r->right_start >= dels[i].end - 100 && r->right_start <= dels[i].end + 5000) border_rp[i]++;
        }
    }
}

/* ---- SONIC's GC profile of one FASTA record ----
 * sonic/sonic_reference.c:57-104, the body of sonic_write_gc_profile's character loop for the text between one
 * '>' line and the next: white space skipped (:86), everything else is a base (:88), only G/C after
 * toupper count (:87,:90), char_count starts at 1 (:82) so the first byte closes after 99 bases, a byte is
 * written per complete window as (char)(100.0 * gc / SONIC_GC_WINDOW) (:93-99; SONIC_GC_WINDOW = 100,
 * sonic/sonic.h).  text: n characters (line feeds allowed); out: capacity `cap`.  Returns the number of
 * bytes the builder writes for the record (all of them stored when cap allows). */
#include <ctype.h>
long orc_gc_profile(const char* text, long n, signed char* out, long cap)
{
    int char_count = 1, gc = 0;
    long written = 0;
    for (long i = 0; i < n; i++) {
        char ch = text[i];
        if (isspace((unsigned char)ch)) continue;
        ch = (char)toupper((unsigned char)ch);
        char_count++;
        if (ch == 'G' || ch == 'C') gc++;
        if (char_count % 100 == 0) {
            const char gc_content = (char)(100.0 * gc / 100);
            if (written < cap) out[written] = gc_content;
            written++;
            gc = 0;
        }
    }
    return written;
}
