/* conga_host.cpp — `conga-b200`: CONGA's command line over libconga_gpu.so.
 *
 * Same flags, same input files, byte-identical <out>_svs.bed / _dels.bed / _dups.bed and
 * conga.log as the reference program (asylvz/CONGA @58d4fe4), with the four numeric stages of
 * read_bam() routed through the C-ABI (include/conga_gpu.h) to the sm_100a kernels:
 *
 *   reference                                   here
 *   main                  svdepth.c:16-74       main()
 *   parse_cmd_line        cmdline.c:12-207      parse_cmd_line()        same option table and defaults
 *   read_bam              bam_data.c:224-359    run()                   chromosome loop in SONIC order
 *   init_rd_per_chr       bam_data.c:304        cg_chrom_begin
 *   count_reads_bam       bam_data.c:316        decode threads -> cg_push_reads
 *   calc_mean_per_chr     bam_data.c:322        cg_chrom_finish (+ the conga.log line)
 *   find_SVs              bam_data.c:338        CNV selection/sort, cg_set_mappability, cg_genotype x2
 *   output_SVs            likelihood.c:172-288  write_rows()            same printf formats
 *
 * What the host does differently — none of it visible in the outputs:
 *   * every BED is parsed once (the reference re-reads the del/dup BEDs twice and the mappability
 *     BED once PER CHROMOSOME, svs.c:7-240, :317-377);
 *   * htslib decoding runs in worker threads, one chromosome per task with its own htsFile and
 *     iterator (the index makes chromosomes independently seekable), ahead of the GPU;
 *   * there is no short read_depth[L] and no float mappability[L] on the host at all;
 *   * CONGA_GPUS=N spreads the chromosomes over N GPUs of the box (longest-processing-time by
 *     length, one host thread per GPU) and ends with the path's one collective, an NCCL all-reduce
 *     of the per-chromosome statistics table over NVLink; the outputs are written in SONIC order
 *     and do not depend on N.
 *
 * There is no CPU implementation of the numeric stages in this program: without the library and
 * a B200 it stops with an error.  With --rp and --dups the split-read side path (split_read.c) runs too:
 * the gate and the average-phred tests here (they need the BAM record and SONIC), the k-mer index, the
 * re-mapping of the read halves and the overlap counts on the GPU.
 *
 * htslib and libsonic are third-party inputs (the same two libraries the reference links).
 */
#include <getopt.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cctype>
#include <cmath>
#include <chrono>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <htslib/faidx.h>
#include <htslib/hts.h>
#include <htslib/sam.h>
#include "sonic.h"

#include <cuda_runtime.h>

#include "conga_gpu.h"

#ifndef CONGA_VERSION
#define CONGA_VERSION "1.0-b200"
#endif
#ifndef CONGA_UPDATE
#define CONGA_UPDATE "November 24, 2022"
#endif
#ifndef BUILD_DATE
#define BUILD_DATE __DATE__
#endif

namespace {

/* parameters, common.h:62-86 (only what the genotyping path reads) */
struct Params {
    std::string ref_genome, outdir, outprefix, low_map_regions, dup_file, del_file, bam_file, sonic_file,
                mappability_file, sonic_info;
    bool has_ref = false, has_out = false, has_dup = false, has_del = false, has_map = false, has_sonic = false,
         has_lowmap = false;
    int min_sv_size = 0, min_read_length = 0, first_chrom = 0, last_chrom = -1;
    float c_score = 0.5f;
    int mq_threshold = -1, rp_support = 10, no_sr = 1;
    sonic* this_sonic = nullptr;
};

FILE* g_log = nullptr;

/* CONGA_TIMING=1: phase times on stderr */
const auto g_t0 = std::chrono::steady_clock::now();
void tick(const char* what)
{
    static const bool on = getenv("CONGA_TIMING") != nullptr;
    if (on) fprintf(stderr, "[timing] %8.3f s  %s\n", std::chrono::duration<double>(std::chrono::steady_clock::now() - g_t0).count(), what);
}

[[noreturn]] void die(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vfprintf(stderr, fmt, ap);
    va_end(ap);
    fputc('\n', stderr);
    exit(1);
}

FILE* open_or_die(const std::string& path, const char* mode)
{
    FILE* f = fopen(path.c_str(), mode);
    if (!f) {   /* common.c:102-119: message + exit(1) */
        fprintf(stderr, "\n[CONGA INPUT ERROR] Unable to open file %s in %s mode.\n", path.c_str(), mode[0] == 'w' ? "write" : "read");
        fprintf(stderr, "Invoke parameter -h for help.\n");
        exit(1);
    }
    return f;
}

void print_help()
{
    fprintf(stdout, "\n\t... CONGA (COpy Number variation Genotyping in Ancient genomes) — B200 host ...\n\n");
    fprintf(stdout, "\tVersion %s\n\tLast update: %s, build date: %s\n\n", CONGA_VERSION, CONGA_UPDATE, BUILD_DATE);
    fprintf(stdout, "\tParameters (identical to conga):\n");
    fprintf(stdout, "\t--input [BAM file]        : Input file in sorted and indexed BAM format (required).\n");
    fprintf(stdout, "\t--out [output prefix]     : Prefix for the output file names (required).\n");
    fprintf(stdout, "\t--ref [reference genome]  : Reference genome in FASTA format (required).\n");
    fprintf(stdout, "\t--sonic [sonic file]      : SONIC file that contains assembly annotations (required).\n");
    fprintf(stdout, "\t--dels [BED file]         : Known deletion SVs in BED format\n");
    fprintf(stdout, "\t--dups [BED file]         : Known duplication SVs in BED format\n");
    fprintf(stdout, "\t--first-chr [chr index]   : The index of the first chromosome for genotyping in your BAM\n");
    fprintf(stdout, "\t--last-chr [chr index]    : The index of the last chromosome for genotyping in your BAM\n");
    fprintf(stdout, "\t--mappability [BED file]  : Mappability file in BED format\n");
    fprintf(stdout, "\t--min-read-length [INT]   : Minimum length of a read to be processed for RP (default: 60 bps)\n");
    fprintf(stdout, "\t--min-sv-size [INT]       : Minimum length of a CNV (default: 1000 bps)\n");
    fprintf(stdout, "\t--min-mapq [INT]          : Minimum mapping quality filter for reads (default: no-filter)\n");
    fprintf(stdout, "\t--c-score [FLOAT]         : Minimum c-score to filter variants (default: 0.5).\n");
    fprintf(stdout, "\t--rp [INT]                : Enable split-read and set minimum read-pair support for a duplication (Suggested for >5x only).\n");
    fprintf(stdout, "\n\tEnvironment: CONGA_GPUS (GPUs to spread the chromosomes over, default 1), CONGA_GPU_DEVICE (first CUDA device,\n"
                    "\tdefault 0), CONGA_DECODE_THREADS (default: cores, at most 16)\n");
    fprintf(stdout, "\n\tInformation:\n\t--version                 : Print version and exit.\n\t--help                    : Print this help screen and exit.\n\n");
}

/* cmdline.c:12-207.  Returns 3 = parameter error (exit 1), 0 = help/version shown (exit 0), 1 = go on. */
int parse_cmd_line(int argc, char** argv, Params& p)
{
    enum { FIRST_CHROM = 10001, LAST_CHROM = 10002 };
    static int no_sr_flag = 1;
    static struct option long_options[] = {
        {"c-score", required_argument, 0, 'a'},      {"min-read-length", required_argument, 0, 'b'},
        {"dels", required_argument, 0, 'd'},         {"min-mapq", required_argument, 0, 'e'},
        {"ref", required_argument, 0, 'f'},          {"help", no_argument, 0, 'h'},
        {"input", required_argument, 0, 'i'},        {"rp", required_argument, 0, 'j'},
        {"min-sv-size", required_argument, 0, 'l'},  {"mappability", required_argument, 0, 'm'},
        {"sonic-info", required_argument, 0, 'n'},   {"out", required_argument, 0, 'o'},
        {"sonic", required_argument, 0, 's'},        {"dups", required_argument, 0, 'u'},
        {"version", no_argument, 0, 'v'},            {"exclude", required_argument, 0, 'x'},
        {"no-sr", no_argument, &no_sr_flag, 1},      {"first-chr", required_argument, 0, FIRST_CHROM},
        {"last-chr", required_argument, 0, LAST_CHROM}, {0, 0, 0, 0}};
    if (argc == 1) { print_help(); return 0; }
    bool load_sonic = false, has_c = false, has_mq = false, has_rp = false;
    std::string c_score, min_mapq, min_rp;
    int o, index;
    /* the reference's short-option string, cmdline.c:54 (note: "x" takes no argument there) */
    while ((o = getopt_long(argc, argv, "hvb:i:f:g:d:r:o:m:c:s:a:e:n:j:k:u:x", long_options, &index)) != -1) {
        switch (o) {
        case 'a': c_score = optarg; has_c = true; break;
        case 'b': p.min_read_length = atoi(optarg); break;
        case 'd': p.del_file = optarg; p.has_del = true; break;
        case 'e': min_mapq = optarg; has_mq = true; break;
        case 'f': p.ref_genome = optarg; p.has_ref = true; break;
        case 'h': print_help(); return 0;
        case 'i': p.bam_file = optarg; break;
        case 'j': min_rp = optarg; has_rp = true; break;
        case 'l': p.min_sv_size = atoi(optarg); break;
        case 'm': p.mappability_file = optarg; p.has_map = true; break;
        case 'n': p.sonic_info = optarg; break;
        case 'o': p.outprefix = optarg; p.has_out = true; break;
        case 's': p.sonic_file = optarg; p.has_sonic = true; load_sonic = true; break;
        case 'u': p.dup_file = optarg; p.has_dup = true; break;
        case 'v':
            fprintf(stderr, "\n\tCONGA Version %s\n\tLast update: %s, build date: %s\n", CONGA_VERSION, CONGA_UPDATE, BUILD_DATE);
            fprintf(stderr, "\tFor more information, check https://github.com/asylvz/CONGA\n\n");
            return 0;
        case 'x': if (optarg) { p.low_map_regions = optarg; p.has_lowmap = true; } break;
        case FIRST_CHROM: p.first_chrom = atoi(optarg); break;
        case LAST_CHROM: p.last_chrom = atoi(optarg); break;
        default: break;
        }
    }
    if (!p.has_out) { fprintf(stderr, "[CONGA CMDLINE ERROR] Please enter the output file name prefix using the --out option.\n"); return 3; }
    if (!p.has_ref) { fprintf(stderr, "[CONGA CMDLINE ERROR] Please enter reference genome file (FASTA) using the --ref option.\n"); return 3; }
    if (!p.has_sonic && load_sonic) { fprintf(stderr, "[CONGA CMDLINE ERROR] Please enter the SONIC file (BED) using the --sonic option.\n"); return 3; }
    if (p.min_sv_size <= 0) { p.min_sv_size = 1000; fprintf(stderr, "Minimum size of an SV is set to %d\n", p.min_sv_size); }
    if (p.min_read_length <= 0) { p.min_read_length = 60; fprintf(stderr, "Minimum size of a read is set to %d\n", p.min_read_length); }
    p.c_score = has_c ? (float)atof(c_score.c_str()) : 0.5f;                 /* float c_score, common.h:74 */
    if (has_rp) { p.rp_support = atoi(min_rp.c_str()); p.no_sr = 0; } else { p.rp_support = 10; p.no_sr = 1; }
    p.mq_threshold = has_mq ? atoi(min_mapq.c_str()) : -1;
    if (p.sonic_info.empty()) p.sonic_info = p.ref_genome;
    /* get_working_directory, common.c:45-83: --out dir/pre splits at the last '/' */
    const size_t slash = p.outprefix.rfind('/');
    if (slash == std::string::npos) {
        p.outdir = "";
    } else {
        const std::string prefix = p.outprefix.substr(slash + 1);
        fprintf(stderr, "prefix: %s\n", prefix.c_str());
        p.outdir = p.outprefix.substr(0, slash) + "/";
        fprintf(stderr, "prefix2: %s\n", prefix.c_str());
        p.outprefix = prefix;
    }
    fprintf(stderr, "[CONGA INFO] Working directory: %s\n", p.outdir.c_str());
    return 1;
}

/* print_params, common.c:86-99 */
void print_params(const Params& p)
{
    printf("\n");
    printf("%-30s%s\n", "BAM input:", p.bam_file.c_str());
    fprintf(g_log, "%-30s%s\n", "BAM input:", p.bam_file.c_str());
    printf("%-30s%s\n", "Reference genome:", p.ref_genome.c_str());
    printf("%-30s%s\n", "SONIC file:", p.has_sonic ? p.sonic_file.c_str() : "(null)");
    fprintf(g_log, "%-30s%s\n", "Reference genome:", p.ref_genome.c_str());
    fprintf(g_log, "%-30s%s\n", "SONIC file:", p.has_sonic ? p.sonic_file.c_str() : "(null)");
    fprintf(g_log, "%-30s%d\n", "First chrom:", p.first_chrom);
    fprintf(g_log, "%-30s%d\n", "Last chrom:", p.last_chrom);
}

/* ---- BED files, each read once --------------------------------------------------------- */
struct BedRow { int start, end; };

/* One pass with the reference's line discipline (svs.c:26-65): fgets into 512 bytes, lines of
 * white space skipped, tokens split on " \t\r\n", atoi.  Rows are grouped by column 1; the
 * per-chromosome filter (name == BAM target name, end-start >= min_sv_size; svs.c:55) is applied
 * when the chromosome comes up. */
std::map<std::string, std::vector<BedRow>> read_cnv_bed(const std::string& path)
{
    std::map<std::string, std::vector<BedRow>> out;
    FILE* f = open_or_die(path, "r");
    char line[512];
    while (fgets(line, sizeof line, f)) {
        bool blank = true;
        for (const char* c = line; *c; c++) if (!isspace((unsigned char)*c)) { blank = false; break; }
        if (blank) continue;
        char* save = nullptr;
        const char* chr = strtok_r(line, " \t\r\n", &save);
        const char* a = strtok_r(nullptr, " \t\r\n", &save);
        const char* b = strtok_r(nullptr, " \t\r\n", &save);
        if (!chr || !a || !b) die("[CONGA INPUT ERROR] %s: a row needs at least three columns", path.c_str());
        out[chr].push_back(BedRow{atoi(a), atoi(b)});
    }
    fclose(f);
    return out;
}

/* Mappability BED (svs.c:329-374): chr start end value, rows of one chromosome kept in FILE ORDER
 * because the reference paints them in that order and later rows win. */
std::map<std::string, std::vector<cg_mapiv>> read_mappability_bed(const std::string& path)
{
    std::map<std::string, std::vector<cg_mapiv>> out;
    FILE* f = open_or_die(path, "r");
    char line[512];
    uint32_t line_no = 0;
    while (fgets(line, sizeof line, f)) {
        bool blank = true;
        for (const char* c = line; *c; c++) if (!isspace((unsigned char)*c)) { blank = false; break; }
        if (blank) continue;
        char* save = nullptr;
        const char* chr = strtok_r(line, " \t\r\n", &save);
        const char* a = strtok_r(nullptr, " \t\r\n", &save);
        const char* b = strtok_r(nullptr, " \t\r\n", &save);
        const char* v = strtok_r(nullptr, " \t\r\n", &save);
        if (!chr || !a || !b || !v) die("[CONGA INPUT ERROR] %s: a mappability row needs four columns", path.c_str());
        cg_mapiv r;
        r.start = atoi(a); r.end = atoi(b);
        r.val = (float)atof(v);                                    /* float mappability = atof(...), svs.c:365 */
        r.line = line_no++;
        out[chr].push_back(r);
    }
    fclose(f);
    return out;
}

/* ---- decoding ahead of the GPU ---------------------------------------------------------- */
/* readReferenceSeq, common.c:421-458, fetches the chromosome in 1000-base steps while `max < length` — so a tail
 * shorter than the last step is never loaded and the sequence the k-mer index sees can be shorter than the
 * SONIC length.  Returns bp_cnt, the number of bases it ends up with. */
long reference_loaded_length(int L, int fa_len)
{
    long bp = 0;
    for (long mn = 0, mx = 999; mx < L;) {
        const long e = std::min<long>(mx, (long)fa_len - 1);
        const long loc = e >= mn ? e - mn + 1 : 0;
        if (loc <= 0) break;
        bp += loc;
        if (bp >= L) break;
        mn += loc; mx += loc;
    }
    return std::min<long>(bp, L);
}

/* A chromosome is decoded in PARTS of CHROM_PART positions, each by whichever thread is free (the index
 * makes any region seekable): part [beg, end) keeps the records whose pos lies inside it, so the parts in
 * order are exactly the records the reference's iterator over [0, length) returns (bam_data.c:293). */
#define CHROM_PART (4 << 20)
struct Part {
    int32_t beg = 0, end = 0;
    std::vector<cg_read> reads;
    std::vector<cg_half> halves;          /* --rp: the read halves find_split_reads would list, and their bases */
    std::vector<char> bases;
    std::vector<size_t> half_off;         /* offset of each half's bases */
};
struct ChromTask {
    int chr_index = -1, tid = -1;
    int32_t length = 0;
    std::vector<Part> parts;
    std::vector<char> ref;                /* --rp: the chromosome as readReferenceSeq leaves it, filled part by part */
    int parts_left = 0;
    bool done = false, failed = false;
    size_t n_halves() const { size_t n = 0; for (const Part& q : parts) n += q.halves.size(); return n; }
};

struct Decoder {
    const Params& p;
    std::vector<ChromTask>& tasks;
    hts_idx_t* idx;                       /* shared, read-only */
    std::atomic<size_t> next{0};
    std::mutex mu;
    std::condition_variable cv;
    size_t consumed = 0;                  /* tasks the consumers are done with (any order) */
    size_t window;                        /* decode at most this many tasks ahead of the consumer */
    std::vector<std::thread> threads;
    std::vector<std::pair<size_t, size_t>> items;   /* (task, part) in decoding order */
    const bool sr_on;                     /* split-read machinery: --rp together with --dups (bam_data.c:207) */

    /* The gate of bam_data.c:207 and the part of find_split_reads (split_read.c:205-336) that needs the
     * BAM record: which halves exist (average-phred tests) and their bases.  The re-mapping itself
     * (almostPerfect_match_seq_ref) is the GPU's. */
    void split_halves(const ChromTask& ct, Part& t, const bam1_t* b) const
    {
        const bam1_core_t& c = b->core;
        const int l = c.l_qseq;
        if (!(l > p.min_read_length)) return;
        if ((c.flag & (BAM_FSECONDARY | BAM_FSUPPLEMENTARY | BAM_FDUP | BAM_FQCFAIL)) != 0) return;         /* is_proper, common.c:317-323 */
        /* libsonic caches the chromosome it looked at last inside the sonic struct (sonic_refind_chromosome_index writes
         * this_sonic->last_chromosome_index, sonic_reference.c:124-131) and this runs on up to 16 decode threads over parts
         * of different chromosomes: every thread asks through its own shallow copy of the struct (the tables behind the
         * pointers are read-only), so no thread can return the index another one has just stored */
        static thread_local sonic local_sonic;
        static thread_local const sonic* local_of = nullptr;
        if (local_of != p.this_sonic) { local_sonic = *p.this_sonic; local_of = p.this_sonic; }
        if (sonic_is_satellite(&local_sonic, local_sonic.chromosome_names[ct.chr_index], (int)c.pos, (int)c.pos + 20) != 0) return;
        if (c.pos == 0) return;                                                                               /* split_read.c:215-216 */
        const uint8_t* q = bam_get_qual(b);
        const uint8_t* sq = bam_get_seq(b);
        const int h = l / 2;
        auto emit = [&](int from, int to, int32_t pos, uint8_t read2) {
            cg_half hf;
            hf.pos = pos; hf.seq_off = (uint32_t)0; hf.len = (uint16_t)(to - from); hf.read_len = (uint16_t)l;
            hf.mapq = c.qual; hf.is_read2 = read2; hf.pad = 0;
            t.half_off.push_back(t.bases.size());
            for (int i = from; i < to; i++) {
                const int code = bam_seqi(sq, i);
                t.bases.push_back(code == 1 ? 'A' : code == 2 ? 'C' : code == 4 ? 'G' : code == 8 ? 'T' : 'N');
            }
            t.halves.push_back(hf);
        };
        /* the running average is NOT reset between the two halves (split_read.c:212, :236-239, :301-304) */
        float avg = 0;
        for (int i = h; i < l; i++) avg = avg + q[i];
        avg = (float)avg / (float)(l - h);
        if ((int)floorf(avg) < p.mq_threshold) return;
        emit(h, l, (int32_t)c.pos, 0);
        for (int i = 0; i < h; i++) avg = avg + q[i];
        avg = (float)avg / (float)h;
        if ((int)floorf(avg) < p.mq_threshold) return;
        emit(0, h, (int32_t)(c.pos + h), 1);
    }

    Decoder(const Params& p_, std::vector<ChromTask>& t, hts_idx_t* i, int n_threads, size_t win)
        : p(p_), tasks(t), idx(i), window(win), sr_on(!p_.no_sr && p_.has_dup)
    {
        faidx_t* fai = nullptr;
        if (sr_on && !(fai = fai_load(p.ref_genome.c_str()))) die("Error: cannot load the FASTA index of %s", p.ref_genome.c_str());
        for (size_t k = 0; k < tasks.size(); k++) {
            ChromTask& ct = tasks[k];
            if (sr_on) {
                const char* name = p.this_sonic->chromosome_names[ct.chr_index];
                const int fa_len = faidx_seq_len(fai, name);
                if (fa_len < 0) die("Error: chromosome %s is not in %s", name, p.ref_genome.c_str());
                ct.ref.resize((size_t)reference_loaded_length(ct.length, fa_len));
            }
            for (int64_t beg = 0; beg < std::max<int64_t>(1, ct.length); beg += CHROM_PART) {
                Part q;
                q.beg = (int32_t)beg; q.end = (int32_t)std::min<int64_t>(beg + CHROM_PART, ct.length);
                ct.parts.push_back(std::move(q));
            }
            ct.parts_left = (int)ct.parts.size();
            for (size_t j = 0; j < ct.parts.size(); j++) items.emplace_back(k, j);
        }
        if (fai) fai_destroy(fai);
        for (int k = 0; k < n_threads; k++) threads.emplace_back([this] { work(); });
    }
    ~Decoder() { for (auto& t : threads) t.join(); }

    void work()
    {
        htsFile* fp = hts_open(p.bam_file.c_str(), "r");
        bam_hdr_t* hdr = fp ? sam_hdr_read(fp) : nullptr;
        bam1_t* b = bam_init1();
        faidx_t* fai = sr_on ? fai_load(p.ref_genome.c_str()) : nullptr;     /* one handle per thread: fetching seeks */
        for (;;) {
            const size_t w = next.fetch_add(1);
            if (w >= items.size()) break;
            const size_t k = items[w].first;
            {   /* bounded look-ahead: the reads of whole chromosomes are held in memory.  Every consumer
                 * takes its tasks in increasing order, so the smallest unreleased task is always
                 * inside the window and progress is guaranteed. */
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return k < consumed + window; });
            }
            ChromTask& t = tasks[k];
            Part& q = t.parts[items[w].second];
            bool ok = fp && hdr;
            if (ok) {
                hts_itr_t* it = sam_itr_queryi(idx, t.tid, q.beg, q.end);
                if (!it) ok = false;
                else {
                    while (sam_itr_next(fp, it, b) > 0) {
                        const bam1_core_t& c = b->core;
                        if (c.pos < q.beg || c.pos >= q.end) continue;     /* starts in a neighbouring part */
                        cg_read r;
                        r.pos = (int32_t)c.pos;
                        r.len = c.l_qseq > 65535 ? 65535 : (uint16_t)c.l_qseq;
                        r.mapq = c.qual;
                        r.fl = (uint8_t)(((c.flag & BAM_FSECONDARY) ? 1 : 0) | ((c.flag & BAM_FQCFAIL) ? 2 : 0) | ((c.flag & BAM_FDUP) ? 4 : 0)
                             | ((c.flag & BAM_FSUPPLEMENTARY) ? 8 : 0) | ((c.flag & BAM_FUNMAP) ? 16 : 0) | ((c.flag & BAM_FREVERSE) ? 32 : 0));
                        q.reads.push_back(r);
                        if (sr_on && (int)c.qual > p.mq_threshold) split_halves(t, q, b);
                    }
                    sam_itr_destroy(it);
                }
            }
            if (ok && sr_on) {   /* this part's slice of the reference sequence, upper case (common.c:441-446) */
                const int64_t e = std::min<int64_t>(q.end, (int64_t)t.ref.size());
                if (e > q.beg) {
                    int got = 0;
                    char* raw = fai ? faidx_fetch_seq(fai, p.this_sonic->chromosome_names[t.chr_index], q.beg, (int)e - 1, &got) : nullptr;
                    if (!raw || got < e - q.beg) ok = false;
                    else for (int64_t i = 0; i < e - q.beg; i++) t.ref[(size_t)(q.beg + i)] = (char)toupper((unsigned char)raw[i]);
                    free(raw);
                }
            }
            {
                std::lock_guard<std::mutex> lk(mu);
                if (!ok) t.failed = true;
                if (--t.parts_left == 0) t.done = true;
            }
            cv.notify_all();
        }
        bam_destroy1(b);
        if (fai) fai_destroy(fai);
        if (hdr) bam_hdr_destroy(hdr);
        if (fp) hts_close(fp);
    }

    ChromTask& wait_for(size_t k)
    {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return tasks[k].done; });
        return tasks[k];
    }
    void release(size_t k)
    {
        {
            std::lock_guard<std::mutex> lk(mu);
            std::vector<Part>().swap(tasks[k].parts);
            std::vector<char>().swap(tasks[k].ref);
            consumed++;
        }
        cv.notify_all();
    }
};

/* ---- output_SVs, likelihood.c:172-288 ---------------------------------------------------- */
struct Totals {
    int dels = 0, dups = 0;
    /* CONGA_STATS side file */
    long long reads = 0, positions = 0, cnvs = 0, split_rows = 0;
    int chromosomes = 0, samples = 0;
};

void write_rows(const Params& p, const char* chr, const std::vector<cg_cnv>& dels, const std::vector<cg_call>& dcall,
                const std::vector<cg_cnv>& dups, const std::vector<cg_call>& ucall, const std::vector<int32_t>& del_border_rp,
                const std::vector<int32_t>& dup_rp, FILE* fp_svs, FILE* fp_del, FILE* fp_dup, Totals& tot)
{
    int n_del = 0, n_dup = 0;
    const float c_score = p.c_score;
    const float inv_c = 1 / c_score;                               /* `1 / params->c_score`: a float quotient */
    if (p.has_del) {
        for (size_t i = 0; i < dels.size(); i++) {
            const cg_call& r = dcall[i];
            const char* gt = r.score < c_score ? (r.copy_number == 2 ? "1/1" : "0/1") : (r.score <= inv_c ? "N/A" : "0/0");
            const int border_rp = del_border_rp[i];                /* 0 unless --rp and --dups are both given (A19) */
            if (p.has_map) {
                fprintf(fp_del, "%s\t%d\t%d\t%s\t%.2f\t%d\t%.2lf\t%d\t%.1f\n", chr, dels[i].start, dels[i].end, gt, r.score, border_rp,
                        r.mappability, r.observed, r.expected);
                if (r.score < c_score && r.mappability > 0.5) {
                    fprintf(fp_svs, "%s\t%d\t%d\tDEL\t%s\t%.2f\t%d\t%.2lf\n", chr, dels[i].start, dels[i].end,
                            r.copy_number == 2 ? "1/1" : "0/1", r.score, border_rp, r.mappability);
                    n_del++;
                }
            } else {
                fprintf(fp_del, "%s\t%d\t%d\t%s\t%.2f\t%d\tN/A\t%d\t%.1f\n", chr, dels[i].start, dels[i].end, gt, r.score, border_rp,
                        r.observed, r.expected);
                if (r.score < c_score) {
                    fprintf(fp_svs, "%s\t%d\t%d\tDEL\t%s\t%.2f\t%d\n", chr, dels[i].start, dels[i].end,
                            r.copy_number == 2 ? "1/1" : "0/1", r.score, border_rp);
                    n_del++;
                }
            }
        }
    }
    if (p.has_dup) {
        for (size_t i = 0; i < dups.size(); i++) {
            const cg_call& r = ucall[i];
            const char* gt = r.copy_number == 2 ? "1/1" : "0/1";
            const int rp = dup_rp[i];
            /* with --rp a duplication also qualifies through its read pairs (likelihood.c:245,265) */
            const bool take = !p.no_sr ? (rp > p.rp_support || r.score < c_score) : (r.score < c_score);
            if (p.has_map) {
                fprintf(fp_dup, "%s\t%d\t%d\t%s\t%.2lf\t%d\t%.2lf\t%d\t%.1f\n", chr, dups[i].start, dups[i].end, gt, r.score, rp,
                        r.mappability, r.observed, r.expected);
                if (take && r.mappability > 0.5) {
                    fprintf(fp_svs, "%s\t%d\t%d\tDUP\t%s\t%.2lf\t%d\t%.2lf\n", chr, dups[i].start, dups[i].end, gt, r.score, rp, r.mappability);
                    n_dup++;
                }
            } else {
                fprintf(fp_dup, "%s\t%d\t%d\t%s\t%.2lf\t%d\tN/A\t%d\t%.1f\n", chr, dups[i].start, dups[i].end, gt, r.score, rp,
                        r.observed, r.expected);
                if (take) {
                    fprintf(fp_svs, "%s\t%d\t%d\tDUP\t%s\t%.2lf\t%d\tN/A\n", chr, dups[i].start, dups[i].end, gt, r.score, rp);
                    n_dup++;
                }
            }
        }
    }
    fprintf(stderr, "\nFound %d DELs - %d DUPs\n\n", n_del, n_dup);
    tot.dels += n_del;
    tot.dups += n_dup;
}

/* CNVs of one chromosome as find_SVs sees them: rows whose column 1 equals the BAM target name and whose
 * length reaches min_sv_size (svs.c:55), sorted by start then end (common.c:199-215; stable, like glibc's
 * merge-sorting qsort, for rows that compare equal). */
std::vector<cg_cnv> select_cnvs(const std::map<std::string, std::vector<BedRow>>& bed, const char* chr, int min_sv_size)
{
    std::vector<cg_cnv> v;
    auto it = bed.find(chr);
    if (it == bed.end()) return v;
    for (const BedRow& r : it->second)
        if (r.end - r.start >= min_sv_size) { cg_cnv c; c.start = r.start; c.end = r.end; v.push_back(c); }
    std::stable_sort(v.begin(), v.end(), [](const cg_cnv& a, const cg_cnv& b) { return a.start != b.start ? a.start < b.start : a.end < b.end; });
    return v;
}

#define CG(ctx, call) do { if ((call) != CG_OK) die("Error: %s: %s", #call, cg_last_error(ctx)); } while (0)

/* everything one chromosome contributes to the outputs */
struct ChromResult {
    bool in_bam = false, has_cnvs = false;
    std::string chr;                      /* BAM target name: what column 1 of the BEDs is compared with */
    int32_t length = 0;
    cg_chrom_stats st;
    std::vector<cg_cnv> dels, dups;
    std::vector<cg_call> dcall, ucall;
    std::vector<int32_t> border_rp, rp;   /* svs.border_rp of the deletions, svs.rp of the duplications */
    long n_split = 0, n_paired = 0;
    bool done = false;
};

struct Inputs {
    std::map<std::string, std::vector<BedRow>> del_bed, dup_bed;
    std::map<std::string, std::vector<cg_mapiv>> map_bed;
    /* the painted mappability track of a chromosome as disjoint runs (cg_flatten_mappability), made on first
     * use and kept: in batch mode every sample reuses it */
    std::mutex mu;
    std::map<std::string, std::pair<cg_maprun*, size_t>> runs;

    const std::pair<cg_maprun*, size_t>& runs_of(const std::string& chr, int32_t chrom_len)
    {
        std::lock_guard<std::mutex> lk(mu);
        auto it = runs.find(chr);
        if (it != runs.end()) return it->second;
        std::pair<cg_maprun*, size_t> r(nullptr, 0);
        auto rows = map_bed.find(chr);
        if (rows != map_bed.end() && !rows->second.empty()
            && cg_flatten_mappability(rows->second.data(), rows->second.size(), chrom_len, &r.first, &r.second) != CG_OK)
            die("Error: cg_flatten_mappability: %s", cg_last_error(nullptr));
        return runs.emplace(chr, r).first->second;
    }
};

/* The split-read side path of one chromosome (bam_data.c:306-335): reference k-mer index and re-mapping
 * of every half on the GPU, then read_SplitReads' satellite test and +-50 widening on the rows. */
std::vector<cg_splitrow> split_read_rows(const Params& p, cg_ctx* ctx, const ChromTask& t, ChromResult& r)
{
    sonic* sn = p.this_sonic;
    const char* name = sn->chromosome_names[t.chr_index];
    const std::vector<char>& ref = t.ref;                       /* fetched by the decode threads, part by part */
    CG(ctx, cg_sr_reference(ctx, ref.empty() ? "" : ref.data(), (int32_t)ref.size()));
    tick("split reads: k-mer index built");
    const size_t MAX_BASES = (size_t)1 << 28;
    std::vector<cg_half> batch;
    for (const Part& q : t.parts) {
        for (size_t i = 0; i < q.halves.size();) {
            const size_t base0 = q.half_off[i];
            size_t j = i;
            batch.clear();
            while (j < q.halves.size() && q.half_off[j] + q.halves[j].len - base0 <= MAX_BASES) {
                cg_half hf = q.halves[j];
                hf.seq_off = (uint32_t)(q.half_off[j] - base0);
                batch.push_back(hf);
                j++;
            }
            const size_t nb = (j < q.halves.size() ? q.half_off[j] : q.bases.size()) - base0;
            CG(ctx, cg_sr_push(ctx, batch.data(), batch.size(), q.bases.data() + base0, nb));
            i = j;
        }
    }
    r.n_split = (long)t.n_halves();
    tick("split reads: halves re-mapped");
    size_t n = 0;
    std::vector<cg_splitrow> rows((size_t)1 << 16);
    CG(ctx, cg_sr_rows(ctx, rows.data(), rows.size(), &n, nullptr));
    if (n > rows.size()) { rows.resize(n); CG(ctx, cg_sr_rows(ctx, rows.data(), rows.size(), &n, nullptr)); }
    rows.resize(n);
    std::vector<cg_splitrow> kept;
    for (cg_splitrow row : rows) {
        /* bam_data.c:101-104: both ends outside satellites */
        const float sat = sonic_is_satellite(sn, name, row.pos, row.pos + 1) + sonic_is_satellite(sn, name, row.pos_map, row.pos_map + 1);
        if (sat != 0) continue;
        row.left_start -= 50; row.left_end -= 50; row.right_start += 50; row.right_end += 50;   /* SOFTCLIP_WRONGMAP_WINDOW, :112-125 */
        kept.push_back(row);
    }
    r.n_paired = (long)kept.size();
    return kept;
}

/* The body of the reference's chromosome loop (bam_data.c:304-338) for one chromosome on one context. */
void process_chromosome(const Params& p, Inputs& in, cg_ctx* ctx, const ChromTask& t, ChromResult& r)
{
    sonic* sn = p.this_sonic;
    /* seams 1-3 */
    CG(ctx, cg_chrom_begin(ctx, t.chr_index, t.length, (const int8_t*)sn->chromosome_gc_profile[t.chr_index], t.length / 100 + 1, p.mq_threshold));
    const size_t BATCH = (size_t)1 << 22;
    for (const Part& q : t.parts)
        for (size_t i = 0; i < q.reads.size(); i += BATCH)
            CG(ctx, cg_push_reads(ctx, q.reads.data() + i, std::min(BATCH, q.reads.size() - i)));
    CG(ctx, cg_chrom_finish(ctx, &r.st));
    std::vector<cg_splitrow> rows;
    if (!p.no_sr && p.has_dup) rows = split_read_rows(p, ctx, t, r);
    /* find_SVs, likelihood.c:311-371 */
    if (p.has_del) r.dels = select_cnvs(in.del_bed, r.chr.c_str(), p.min_sv_size);
    if (p.has_dup) r.dups = select_cnvs(in.dup_bed, r.chr.c_str(), p.min_sv_size);
    r.has_cnvs = !(r.dels.empty() && r.dups.empty());
    if (!r.has_cnvs) return;                                        /* likelihood.c:332-336 */
    if (p.has_map) {
        /* no run at all (no row on this chromosome, or only zeros): mappability 0.0 for every CNV */
        const auto& runs = in.runs_of(r.chr, t.length);
        CG(ctx, cg_set_mappability_runs(ctx, runs.first, runs.second, 1));
    }
    r.dcall.resize(r.dels.size());
    r.ucall.resize(r.dups.size());
    r.border_rp.assign(r.dels.size(), 0);
    r.rp.assign(r.dups.size(), 0);
    if (!p.no_sr && p.has_dup)                                       /* count_ReadPairs, likelihood.c:344-348 */
        CG(ctx, cg_sr_count(ctx, rows.data(), rows.size(), r.dels.data(), r.dels.size(), r.dups.data(), r.dups.size(),
                            r.border_rp.data(), r.rp.data()));
    if (!r.dels.empty()) CG(ctx, cg_genotype(ctx, r.dels.data(), r.dels.size(), 0, r.dcall.data()));
    if (!r.dups.empty()) CG(ctx, cg_genotype(ctx, r.dups.data(), r.dups.size(), 1, r.ucall.data()));
}

/* What the reference prints and writes for one chromosome, in its order. */
void emit_chromosome(const Params& p, const std::string& sample, const char* sonic_name, const ChromResult& r,
                     FILE* fp_svs, FILE* fp_del, FILE* fp_dup, Totals& tot)
{
    if (!r.in_bam) {
        fprintf(stderr, "\nCannot find chromosome name %s in BAM/CRAM %s", sonic_name, sample.c_str());
        return;
    }
    const char* chr = r.chr.c_str();
    tot.reads += (long long)r.st.reads_passed; tot.positions += (long long)r.length; tot.chromosomes++;
    tot.cnvs += (long long)(r.dels.size() + r.dups.size()); tot.split_rows += r.n_paired;
    fprintf(stderr, "\nReading BAM [%s] - Chromosome: %s", sample.c_str(), chr);
    fprintf(stderr, "\n-->counting reads");
    fprintf(stderr, " (%d reads, %ld split-reads)\n", (int)r.st.reads_passed, r.n_split);
    if (!p.no_sr && p.has_dup) fprintf(stderr, "Loading Split-Reads\n\nCONGA paired %ld single-end reads\n", r.n_paired);
    {   /* calc_mu_per_chr, read_distribution.c:39-41 */
        const float mean = (float)((double)r.st.rd_cnt / (double)(long)r.length);
        fprintf(g_log, "Read Count:%li  Window count:%li mean=%f\n", (long)r.st.rd_cnt, (long)r.length, mean);
    }
    fprintf(stderr, "\nLoading known SVs");
    fprintf(stderr, "(%d DELS, %d DUPS in chromosome %s - larger than the threshold %d)\n", (int)r.dels.size(), (int)r.dups.size(), chr, p.min_sv_size);
    if (!r.has_cnvs) return;
    fprintf(stderr, "Finding mappability for each region\n");
    fprintf(stderr, "\nCalculating Likelihoods\n");
    write_rows(p, chr, r.dels, r.dcall, r.dups, r.ucall, r.border_rp, r.rp, fp_svs, fp_del, fp_dup, tot);
}

/* Longest-processing-time assignment by chromosome length: the depth pass streams
 * 6 bytes per position, so length is the cost.  Same rule as conga_b200/multi.py. */
std::vector<int> lpt_assign(const std::vector<ChromTask>& todo, int n_gpus)
{
    std::vector<size_t> order(todo.size());
    for (size_t i = 0; i < order.size(); i++) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return todo[a].length > todo[b].length; });
    std::vector<long long> load(n_gpus, 0);
    std::vector<int> gpu_of(todo.size(), 0);
    for (size_t i : order) {
        int g = 0;
        for (int k = 1; k < n_gpus; k++) if (load[k] < load[g]) g = k;
        gpu_of[i] = g;
        load[g] += todo[i].length;
    }
    return gpu_of;
}

/* CUDA contexts brought up on side threads from the moment the command line is known, and the NCCL
 * communicators over them (created on first use, kept for the following samples of a batch) */
struct Gpus {
    int first_dev = 0;
    std::vector<cg_ctx*> ctx;
    std::vector<std::thread> threads;
    void start()
    {
        first_dev = getenv("CONGA_GPU_DEVICE") ? atoi(getenv("CONGA_GPU_DEVICE")) : 0;
        const int n = getenv("CONGA_GPUS") ? std::max(1, atoi(getenv("CONGA_GPUS"))) : 1;
        if (!getenv("CUDA_VISIBLE_DEVICES")) {
            /* the driver initialises every visible GPU (about 6 s on an 8-GPU box against 1 s for one): show it
             * only the ones this run uses; they become devices 0..n-1 */
            std::string list;
            for (int g = 0; g < n; g++) list += (g ? "," : "") + std::to_string(first_dev + g);
            setenv("CUDA_VISIBLE_DEVICES", list.c_str(), 1);
            first_dev = 0;
        }
        ctx.assign(n, nullptr);
        for (int g = 0; g < n; g++) threads.emplace_back([this, g] { ctx[g] = cg_create(first_dev + g); });
    }
    void wait()
    {
        for (auto& t : threads) t.join();
        threads.clear();
        for (cg_ctx* c : ctx)
            if (!c) die("Error: %s (conga-b200 has no CPU implementation of the genotyping path)", cg_last_error(nullptr));
    }
    /* the path's one collective, issued by the library: ncclAllReduce(int64, sum) of the statistics tables, one
     * communicator per GPU of this process (cg_comm_init_all on first use, kept for the following samples of a batch) */
    bool have_comm = false;
    void allreduce(int n_rows)
    {
        const int n = (int)ctx.size();
        if (!have_comm) {
            if (cg_comm_init_all(ctx.data(), n) != CG_OK) die("Error: %s", cg_last_error(ctx[0]));
            have_comm = true;
        }
        if (cg_stats_allreduce_all(ctx.data(), n, n_rows) != CG_OK) die("Error: %s", cg_last_error(ctx[0]));
        for (int g = 0; g < n; g++) if (cg_sync(ctx[g]) != CG_OK) die("Error: cg_sync: %s", cg_last_error(ctx[g]));
    }
    void shutdown()
    {
        for (cg_ctx* c : ctx) cg_destroy(c);                 /* destroys the context's communicator too */
        ctx.clear();
    }
};

/* One BAM being worked on: its header and index, the chromosomes to visit, and the threads decoding them. */
struct Sample {
    Params p;                             /* the run's parameters with this sample's BAM and output prefix */
    htsFile* fp = nullptr;
    bam_hdr_t* hdr = nullptr;
    hts_idx_t* idx = nullptr;
    std::string name;
    std::vector<ChromTask> tasks, todo;
    std::vector<size_t> todo_of, task_of_todo;
    std::vector<ChromResult> results;
    Decoder* dec = nullptr;

    Sample(const Params& p_, int n_gpus) : p(p_)
    {
        fp = hts_open(p.bam_file.c_str(), "r");
        if (!fp) die("Error opening BAM file %s", p.bam_file.c_str());
        hdr = sam_hdr_read(fp);
        if (!hdr) die("Error: cannot read the BAM header of %s", p.bam_file.c_str());
        idx = sam_index_load(fp, p.bam_file.c_str());
        if (!idx) die("Error: Sam Index cannot be loaded (sam_index_load)");
        /* get_sample_name, common.c:325-352: first header token that starts with "SM" */
        std::string text(hdr->text ? hdr->text : "");
        for (size_t i = 0; i < text.size();) {
            size_t j = text.find_first_of("\t\n", i);
            if (j == std::string::npos) j = text.size();
            if (j - i >= 2 && text[i] == 'S' && text[i + 1] == 'M') { name = j - i > 3 ? text.substr(i + 3, j - i - 3) : ""; break; }
            i = j + 1;
        }
        /* the chromosome loop's selection logic, bam_data.c:269-291 */
        sonic* sn = p.this_sonic;
        for (int ci = 0; ci < sn->number_of_chromosomes; ci++) {
            if (ci < p.first_chrom) ci = p.first_chrom;
            if (ci > p.last_chrom || ci >= sn->number_of_chromosomes) break;
            const char* cname = sn->chromosome_names[ci];
            if (strstr(cname, "X") || strstr(cname, "Y")) continue;
            int tid = -1;
            for (int t = 0; t < hdr->n_targets; t++) if (strcmp(cname, hdr->target_name[t]) == 0) { tid = t; break; }
            ChromTask t;
            t.chr_index = ci; t.tid = tid; t.length = sn->chromosome_lengths[ci];
            tasks.push_back(std::move(t));
        }
        /* tasks for targets missing from the BAM are never decoded */
        todo_of.assign(tasks.size(), (size_t)-1);
        results.resize(tasks.size());
        for (size_t k = 0; k < tasks.size(); k++) {
            if (tasks[k].tid < 0) continue;
            todo_of[k] = todo.size();
            task_of_todo.push_back(k);
            todo.push_back(tasks[k]);
            results[k].in_bam = true;
            results[k].chr = hdr->target_name[tasks[k].tid];
            results[k].length = tasks[k].length;
        }
        int n_threads = (int)std::min<size_t>(std::max<unsigned>(1u, std::thread::hardware_concurrency()), 16);
        if (const char* e = getenv("CONGA_DECODE_THREADS")) n_threads = std::max(1, atoi(e));
        size_t n_parts = 0;
        for (const ChromTask& t : todo) n_parts += ((size_t)std::max<int32_t>(1, t.length) + CHROM_PART - 1) / CHROM_PART;
        n_threads = (int)std::min<size_t>((size_t)n_threads, std::max<size_t>(1, n_parts));
        dec = new Decoder(p, todo, idx, n_threads, (size_t)n_threads + 2 * (size_t)n_gpus);
    }
    ~Sample()
    {
        delete dec;
        if (fp && hts_close(fp) != 0) die("Error closing BAM file");
        if (hdr) bam_hdr_destroy(hdr);
        if (idx) hts_idx_destroy(idx);
    }
};

/* read_bam, bam_data.c:224-359, for one sample whose decoding is already under way */
void run_sample(Sample& sm, Inputs& in, Gpus& gpus, Totals& tot)
{
    const Params& p = sm.p;
    const std::string f_svs = p.outdir + p.outprefix + "_svs.bed";
    fprintf(stderr, "\nOutput SV file: %s\n", f_svs.c_str());
    FILE* fp_svs = open_or_die(f_svs, "w");
    fprintf(fp_svs, "#CHR\tSTART_SV\tEND_SV\tSV_TYPE\tCOPY_NUMBER\tLIKELIHOOD\tREAD_PAIR\tMAPPABILITY\n");
    FILE *fp_del = nullptr, *fp_dup = nullptr;
    if (p.has_del) {
        const std::string f = p.outdir + p.outprefix + "_dels.bed";
        fprintf(stderr, "Output Del file: %s\n", f.c_str());
        fp_del = open_or_die(f, "w");
        fprintf(fp_del, "#CHR\tSTART_SV\tEND_SV\tCOPY_NUMBER\tLIKELIHOOD\tREAD_PAIR\tMAPPABILITY\tOBSERVED_READS\tEXPECTED_READS\n");
    }
    if (p.has_dup) {
        const std::string f = p.outdir + p.outprefix + "_dups.bed";
        fprintf(stderr, "Output DUP file: %s\n", f.c_str());
        fp_dup = open_or_die(f, "w");
        fprintf(fp_dup, "#CHR\tSTART_SV\tEND_SV\tCOPY_NUMBER\tLIKELIHOOD\tREAD_PAIR\tMAPPABILITY\tOBSERVED_READS\tEXPECTED_READS\n");
    }

    sonic* sn = p.this_sonic;
    std::vector<cg_ctx*>& ctx = gpus.ctx;
    const int n_gpus = (int)std::min<size_t>(ctx.size(), std::max<size_t>(1, sm.todo.size()));
    Decoder& dec = *sm.dec;
    if (n_gpus == 1) {
        /* streaming, like the reference: each chromosome is written as soon as it is genotyped */
        for (size_t k = 0; k < sm.tasks.size(); k++) {
            if (sm.results[k].in_bam) {
                ChromTask& t = dec.wait_for(sm.todo_of[k]);
                if (t.failed) die("Error: Iterator cannot be loaded (bam_itr_queryi)");
                process_chromosome(p, in, ctx[0], t, sm.results[k]);
                dec.release(sm.todo_of[k]);
            }
            emit_chromosome(p, sm.name, sn->chromosome_names[sm.tasks[k].chr_index], sm.results[k], fp_svs, fp_del, fp_dup, tot);
        }
    } else {
        /* one host thread per GPU, each walking its own chromosomes in SONIC order */
        const std::vector<int> gpu_of = lpt_assign(sm.todo, n_gpus);
        for (int g = 0; g < (int)ctx.size(); g++) CG(ctx[g], cg_stats_reset(ctx[g]));      /* batch mode: the previous sample's rows */
        std::vector<std::thread> workers;
        for (int g = 0; g < n_gpus; g++)
            workers.emplace_back([&, g] {
                for (size_t j = 0; j < sm.todo.size(); j++) {
                    if (gpu_of[j] != g) continue;
                    ChromTask& t = dec.wait_for(j);
                    if (t.failed) die("Error: Iterator cannot be loaded (bam_itr_queryi)");
                    process_chromosome(p, in, ctx[g], t, sm.results[sm.task_of_todo[j]]);
                    dec.release(j);
                }
            });
        for (auto& w : workers) w.join();
        /* the path's one collective: every GPU's int64[n_chrom][204] statistics table summed over NVLink, so
         * that the writer of conga.log (GPU 0's host side) holds every chromosome's row; a chromosome's row is
         * zero everywhere but on the GPU that processed it */
        gpus.allreduce(sn->number_of_chromosomes);
        for (size_t k = 0; k < sm.tasks.size(); k++) {
            if (sm.results[k].in_bam) {
                cg_chrom_stats st;
                CG(ctx[0], cg_stats_row(ctx[0], sm.tasks[k].chr_index, &st));
                if (memcmp(&st, &sm.results[k].st, sizeof st) != 0) die("Error: statistics row %d differs after the all-reduce", sm.tasks[k].chr_index);
                sm.results[k].st = st;
            }
            emit_chromosome(p, sm.name, sn->chromosome_names[sm.tasks[k].chr_index], sm.results[k], fp_svs, fp_del, fp_dup, tot);
        }
    }
    fprintf(stderr, "\n");
    if (fp_del) fclose(fp_del);
    if (fp_dup) fclose(fp_dup);
    fclose(fp_svs);
}

/* CONGA_BATCH=<file>: one "BAM <white space> output prefix" per line.  Every sample is genotyped against the
 * same --dels/--dups/--mappability/--sonic with the same flags and gets its own three BED files plus
 * <prefix>_conga.log — the files separate runs would write — while the CUDA contexts, the SONIC tables, the
 * parsed BEDs and the flattened mappability tracks are set up once, and sample k+1 is decoded while k is on
 * the GPUs (BASELINE.json configs[4]). */
std::vector<std::pair<std::string, std::string>> read_batch_list(const char* path)
{
    std::vector<std::pair<std::string, std::string>> v;
    FILE* f = open_or_die(path, "r");
    char line[8192];
    while (fgets(line, sizeof line, f)) {
        char* save = nullptr;
        const char* bam = strtok_r(line, " \t\r\n", &save);
        const char* out = bam ? strtok_r(nullptr, " \t\r\n", &save) : nullptr;
        if (!bam || bam[0] == '#') continue;
        if (!out) die("[CONGA INPUT ERROR] %s: a batch row needs a BAM and an output prefix", path);
        v.emplace_back(bam, out);
    }
    fclose(f);
    return v;
}

}  // namespace

/* --out dir/pre -> outdir "dir/", outprefix "pre" (get_working_directory, common.c:45-83) */
void split_outprefix(Params& p, bool chatty)
{
    const size_t slash = p.outprefix.rfind('/');
    if (slash == std::string::npos) { p.outdir = ""; return; }
    const std::string prefix = p.outprefix.substr(slash + 1);
    if (chatty) fprintf(stderr, "prefix: %s\n", prefix.c_str());
    p.outdir = p.outprefix.substr(0, slash) + "/";
    if (chatty) fprintf(stderr, "prefix2: %s\n", prefix.c_str());
    p.outprefix = prefix;
}

void log_header(FILE* f)
{
    time_t raw;
    time(&raw);
    const struct tm* ti = localtime(&raw);
    fprintf(f, "#CreationDate=%d.%d.%d\n\n", ti->tm_year + 1900, ti->tm_mon + 1, ti->tm_mday);   /* svdepth.c:31 */
}

int main(int argc, char** argv)
{
    g_log = open_or_die("conga.log", "w");                          /* svdepth.c:30: always in the CWD */
    log_header(g_log);

    Params p;
    const int rc = parse_cmd_line(argc, argv, p);
    if (rc == 3) return EXIT_FAILURE;
    if (rc == 0) return EXIT_SUCCESS;
    print_params(p);
    if (!p.has_sonic) die("[CONGA CMDLINE ERROR] Please enter the SONIC file using the --sonic option.");

    tick("start");
    Gpus gpus;
    gpus.start();
    p.this_sonic = sonic_load(const_cast<char*>(p.sonic_file.c_str()));
    tick("SONIC loaded");
    if (!p.this_sonic) die("Error: cannot load SONIC file %s", p.sonic_file.c_str());
    const int first_as_given = p.first_chrom, last_as_given = p.last_chrom;
    if (p.last_chrom < p.first_chrom) p.last_chrom = p.this_sonic->number_of_chromosomes - 1;   /* svdepth.c:49-50 */

    /* the samples: the command line's one BAM, or the rows of CONGA_BATCH */
    std::vector<Params> samples;
    const char* batch = getenv("CONGA_BATCH");
    if (batch) {
        for (const auto& row : read_batch_list(batch)) {
            Params q = p;
            q.bam_file = row.first;
            q.outprefix = row.second;
            split_outprefix(q, false);
            samples.push_back(q);
        }
        if (samples.empty()) die("[CONGA INPUT ERROR] %s lists no sample", batch);
    } else {
        samples.push_back(p);
    }

    Totals tot;
    Inputs in;
    Sample* cur = new Sample(samples[0], (int)gpus.ctx.size());       /* decoding starts here */
    tick("decoders started");
    if (p.has_del) in.del_bed = read_cnv_bed(p.del_file);
    if (p.has_dup) in.dup_bed = read_cnv_bed(p.dup_file);
    if (p.has_map) in.map_bed = read_mappability_bed(p.mappability_file);
    tick("BEDs parsed");
    gpus.wait();
    tick("CUDA contexts up");
    /* CONGA_STATS=<file>: an opt-in machine-readable summary of the run (metrics); nothing else changes */
    const char* stats_path = getenv("CONGA_STATS");
    const auto t_work0 = std::chrono::steady_clock::now();
    if (getenv("CONGA_TIMING") || stats_path) for (cg_ctx* c : gpus.ctx) cg_set_profiling(c, 1);
    FILE* main_log = g_log;
    for (size_t i = 0; i < samples.size(); i++) {
        Sample* next = i + 1 < samples.size() ? new Sample(samples[i + 1], (int)gpus.ctx.size()) : nullptr;   /* decoded under sample i */
        if (batch) {   /* what a separate run would have left in ./conga.log */
            g_log = open_or_die(cur->p.outdir + cur->p.outprefix + "_conga.log", "w");
            log_header(g_log);
            Params shown = cur->p;
            shown.first_chrom = first_as_given; shown.last_chrom = last_as_given;
            fprintf(g_log, "%-30s%s\n", "BAM input:", shown.bam_file.c_str());
            fprintf(g_log, "%-30s%s\n", "Reference genome:", shown.ref_genome.c_str());
            fprintf(g_log, "%-30s%s\n", "SONIC file:", shown.sonic_file.c_str());
            fprintf(g_log, "%-30s%d\n", "First chrom:", shown.first_chrom);
            fprintf(g_log, "%-30s%d\n", "Last chrom:", shown.last_chrom);
        }
        run_sample(*cur, in, gpus, tot);
        if (batch) { fclose(g_log); g_log = main_log; fprintf(g_log, "sample %zu: %s -> %s%s\n", i, cur->p.bam_file.c_str(), cur->p.outdir.c_str(), cur->p.outprefix.c_str()); }
        delete cur;
        cur = next;
        tot.samples++;
        tick("sample done");
    }
    if (getenv("CONGA_TIMING") && !stats_path)             /* with CONGA_STATS the kernel times go into the side file */
        for (cg_ctx* c : gpus.ctx) {
            double ms[CG_NSLOTS]; int64_t n[CG_NSLOTS];
            if (cg_kernel_times(c, ms, n) == CG_OK)
                fprintf(stderr, "[timing] kernels: prep %.3f ms (%ld), depth pass %.3f ms (%ld), genotype %.3f ms (%ld), split reads %.3f ms (%ld)\n",
                        ms[0], (long)n[0], ms[1], (long)n[1], ms[2], (long)n[2], ms[3], (long)n[3]);
        }
    if (stats_path) {
        double ms[CG_NSLOTS] = {0, 0, 0, 0}; int64_t n[CG_NSLOTS] = {0, 0, 0, 0};
        for (cg_ctx* c : gpus.ctx) {
            double m[CG_NSLOTS]; int64_t k[CG_NSLOTS];
            if (cg_kernel_times(c, m, k) == CG_OK) for (int i = 0; i < CG_NSLOTS; i++) { ms[i] += m[i]; n[i] += k[i]; }
        }
        const double work_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_work0).count();
        const double total_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - g_t0).count();
        /* bytes the depth pass moves by construction: 8 per record, 2 per position, 1 per 100 positions */
        const double pass_bytes = 8.0 * (double)tot.reads + 2.01 * (double)tot.positions;
        FILE* f = fopen(stats_path, "w");
        if (f) {
            fprintf(f, "{\"samples\": %d, \"gpus\": %zu, \"chromosomes\": %d, \"positions\": %lld, \"reads\": %lld, \"cnvs\": %lld, \"split_read_rows\": %lld,\n"
                       " \"seconds_total\": %.4f, \"seconds_after_contexts_up\": %.4f, \"reads_per_s\": %.1f, \"cnvs_per_s\": %.1f,\n"
                       " \"kernel_ms\": {\"prep\": %.4f, \"depth_pass\": %.4f, \"genotype\": %.4f, \"split_reads\": %.4f},\n"
                       " \"kernel_launches\": {\"prep\": %ld, \"depth_pass\": %ld, \"genotype\": %ld, \"split_reads\": %ld},\n"
                       " \"depth_pass_gb_per_s\": %.1f, \"genotype_cnvs_per_s\": %.1f}\n",
                    tot.samples, gpus.ctx.size(), tot.chromosomes, tot.positions, tot.reads, tot.cnvs, tot.split_rows,
                    total_s, work_s, work_s > 0 ? (double)tot.reads / work_s : 0.0, work_s > 0 ? (double)tot.cnvs / work_s : 0.0,
                    ms[0], ms[1], ms[2], ms[3], (long)n[0], (long)n[1], (long)n[2], (long)n[3],
                    ms[1] > 0 ? pass_bytes / (ms[1] * 1e-3) / 1e9 : 0.0, ms[2] > 0 ? (double)tot.cnvs / (ms[2] * 1e-3) : 0.0);
            fclose(f);
        }
    }
    gpus.shutdown();
    tick("contexts destroyed");
    for (auto& kv : in.runs) cg_free_runs(kv.second.first);

    char user[256] = "";
    if (getlogin_r(user, sizeof user - 1) != 0) snprintf(user, sizeof user, "you");
    fprintf(stderr, "\nThank you %s. I found %d DELs and %d DUPs. Hope to see you again...\n", user, tot.dels, tot.dups);
    fclose(g_log);
    return EXIT_SUCCESS;
}
