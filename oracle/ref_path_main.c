/* ref_path_main.c — drives the UNMODIFIED reference hot-path functions on
 * in-memory read records, so the reference's own CPU implementation of the path
 * can be timed and diffed WITHOUT htslib BAM decoding in the way.
 *
 * TEST / BASELINE INFRASTRUCTURE ONLY (see oracle/conga_oracle.c header).  Built
 * only where /root/reference exists (oracle/Makefile target `ref`) into
 * oracle/_ref/ref_path; bench.py runs that prebuilt binary for `--impl reference`
 * and for the cpu_baseline of kind "reference".
 *
 * What is reference code here: init_rd_per_chr, count_reads_bam, calc_mean_per_chr,
 * find_SVs (load_known_SVs, qsort, load_mappability_regions, find_depths,
 * calculate_likelihood_CNV, lpoisson, output_SVs) and sonic_load — the objects are
 * compiled from /root/reference untouched.  What is ours: this main(), which
 * replays the chromosome loop of read_bam (bam_data.c:269-339) and, through
 * -Wl,--wrap=hts_itr_next, feeds count_reads_bam's `sam_itr_next` loop
 * (bam_data.c:201; sam_itr_next is a static inline over hts_itr_next,
 * htslib/htslib/sam.h:1186-1200) from an array of cg_read records instead of a
 * BGZF stream.  The split-read (--rp) side path is left off (no_sr = 1).
 *
 * usage: ref_path --sonic F --reads-dir D --out PREFIX [--dels F] [--dups F]
 *                 [--mappability F] [--min-mapq N] [--c-score X] [--min-sv-size N]
 *                 [--first-chr A] [--last-chr B]
 *   D/reads.<sonic chromosome index>.bin holds that chromosome's cg_read records.
 * Prints one JSON line with per-phase seconds (CLOCK_MONOTONIC).
 */
#define _GNU_SOURCE
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#include <time.h>

#include "common.h"
#include "bam_data.h"
#include "read_distribution.h"
#include "likelihood.h"
#include "free.h"
#include "../include/conga_gpu.h"

/* defined in svdepth.c in the reference (svdepth.c:11-13); that TU holds main() and is not linked */
FILE* logFile = NULL;
int total_dels = 0;
int total_dups = 0;

void count_reads_bam(bam_info* in_bam, parameters* params, int chr_index, int* base_count_bam);  /* bam_data.c:192 */

static const cg_read* g_recs = NULL;
static size_t g_n = 0, g_i = 0;

int __wrap_hts_itr_next(BGZF* fp, hts_itr_t* iter, void* r, void* data)
{
    (void)fp; (void)iter; (void)data;
    if (g_i >= g_n) return -1;
    bam1_t* b = (bam1_t*)r;
    const cg_read* q = &g_recs[g_i++];
    b->core.pos = q->pos;
    b->core.qual = q->mapq;
    b->core.l_qseq = q->len;
    b->core.flag = 0;
    return 1;
}

static double now_s(void)
{
    struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

static cg_read* load_records(const char* dir, int idx, size_t* n)
{
    char path[2048];
    snprintf(path, sizeof path, "%s/reads.%d.bin", dir, idx);
    FILE* f = fopen(path, "rb");
    *n = 0;
    if (!f) return NULL;
    fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
    cg_read* r = (cg_read*)malloc(sz > 0 ? sz : 1);
    if (sz > 0 && fread(r, 1, sz, f) != (size_t)sz) { perror(path); exit(2); }
    fclose(f);
    *n = sz / sizeof(cg_read);
    return r;
}

int main(int argc, char** argv)
{
    parameters* params; bam_info* in_bam;
    const char* reads_dir = NULL; char* out = NULL;
    init_params(&params);                                   /* common.c:16 */
    params->no_sr = 1; params->rp_support = 10; params->mq_threshold = -1;
    params->c_score = 0.5; params->min_sv_size = 1000; params->min_read_length = 60;
    for (int i = 1; i + 1 < argc; i += 2) {
        const char* k = argv[i]; char* v = argv[i + 1];
        if      (!strcmp(k, "--sonic"))       params->sonic_file = strdup(v);
        else if (!strcmp(k, "--reads-dir"))   reads_dir = v;
        else if (!strcmp(k, "--out"))         out = v;
        else if (!strcmp(k, "--dels"))        params->del_file = strdup(v);
        else if (!strcmp(k, "--dups"))        params->dup_file = strdup(v);
        else if (!strcmp(k, "--mappability")) params->mappability_file = strdup(v);
        else if (!strcmp(k, "--min-mapq"))    params->mq_threshold = atoi(v);
        else if (!strcmp(k, "--c-score"))     params->c_score = atof(v);
        else if (!strcmp(k, "--min-sv-size")) params->min_sv_size = atoi(v);
        else if (!strcmp(k, "--first-chr"))   params->first_chrom = atoi(v);
        else if (!strcmp(k, "--last-chr"))    params->last_chrom = atoi(v);
        else { fprintf(stderr, "ref_path: unknown option %s\n", k); return 2; }
    }
    if (!params->sonic_file || !reads_dir || !out) { fprintf(stderr, "ref_path: --sonic, --reads-dir, --out required\n"); return 2; }

    char path[2048];
    snprintf(path, sizeof path, "%s.log", out);
    logFile = fopen(path, "w");
    params->outdir = strdup(""); params->outprefix = strdup(out);
    params->this_sonic = sonic_load(params->sonic_file);    /* svdepth.c:47 */
    if (!params->this_sonic) { fprintf(stderr, "ref_path: cannot load SONIC\n"); return 2; }
    if (params->last_chrom < params->first_chrom)           /* svdepth.c:49-50 */
        params->last_chrom = params->this_sonic->number_of_chromosomes - 1;

    in_bam = (bam_info*)calloc(1, sizeof(bam_info));
    in_bam->sample_name = strdup("synth");
    htsFile fake_file; memset(&fake_file, 0, sizeof fake_file); fake_file.is_bgzf = 1;
    in_bam->bam_file = &fake_file;

    /* output files + headers exactly as read_bam opens them (bam_data.c:232-250) */
    FILE *fpSVs, *fpDel = NULL, *fpDup = NULL;
    snprintf(path, sizeof path, "%s_svs.bed", out); fpSVs = fopen(path, "w");
    fprintf(fpSVs, "#CHR\tSTART_SV\tEND_SV\tSV_TYPE\tCOPY_NUMBER\tLIKELIHOOD\tREAD_PAIR\tMAPPABILITY\n");
    if (params->del_file) {
        snprintf(path, sizeof path, "%s_dels.bed", out); fpDel = fopen(path, "w");
        fprintf(fpDel, "#CHR\tSTART_SV\tEND_SV\tCOPY_NUMBER\tLIKELIHOOD\tREAD_PAIR\tMAPPABILITY\tOBSERVED_READS\tEXPECTED_READS\n");
    }
    if (params->dup_file) {
        snprintf(path, sizeof path, "%s_dups.bed", out); fpDup = fopen(path, "w");
        fprintf(fpDup, "#CHR\tSTART_SV\tEND_SV\tCOPY_NUMBER\tLIKELIHOOD\tREAD_PAIR\tMAPPABILITY\tOBSERVED_READS\tEXPECTED_READS\n");
    }

    double t_count = 0, t_mean = 0, t_svs = 0, t_init = 0; long n_reads = 0; int n_chr = 0;
    double t_all0 = now_s();
    for (int c = params->first_chrom; c <= params->last_chrom && c < params->this_sonic->number_of_chromosomes; c++) {
        char* name = params->this_sonic->chromosome_names[c];
        if (strstr(name, "X") != NULL || strstr(name, "Y") != NULL) continue;   /* bam_data.c:280 */
        size_t n; cg_read* recs = load_records(reads_dir, c, &n);
        if (!recs) continue;                                /* "not in BAM" (bam_data.c:286-291) */
        g_recs = recs; g_n = n; g_i = 0;
        hts_itr_t* itr = (hts_itr_t*)calloc(1, sizeof(hts_itr_t));
        in_bam->iter = itr;
        int dummy = 0;
        double t0 = now_s();
        init_rd_per_chr(in_bam, params, c);                 /* bam_data.c:304 */
        double t1 = now_s();
        count_reads_bam(in_bam, params, c, &dummy);         /* bam_data.c:316 */
        double t2 = now_s();
        calc_mean_per_chr(params, in_bam, c);               /* bam_data.c:322 */
        double t3 = now_s();
        find_SVs(in_bam, params, fpDel, fpDup, fpSVs, name, c);   /* bam_data.c:338 */
        double t4 = now_s();
        t_init += t1 - t0; t_count += t2 - t1; t_mean += t3 - t2; t_svs += t4 - t3;
        n_reads += (long)n; n_chr++;
        free(itr); free(recs);
    }
    double t_all = now_s() - t_all0;
    if (fpDel) fclose(fpDel);
    if (fpDup) fclose(fpDup);
    fclose(fpSVs); fclose(logFile);
    printf("{\"chromosomes\": %d, \"reads\": %ld, \"total_s\": %.6f, \"init_s\": %.6f, \"count_reads_s\": %.6f, "
           "\"calc_mean_s\": %.6f, \"find_svs_s\": %.6f}\n", n_chr, n_reads, t_all, t_init, t_count, t_mean, t_svs);
    return 0;
}
