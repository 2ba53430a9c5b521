// sonic-b200 — the SONIC builder's command line (sonic/sonic_exe.c) with the GC profile computed on the GPU.
//
// Everything except the GC profile is the builder's own library code: this program calls libsonic's
// sonic_build (sonic/sonic.c:11-150) and is linked with -Wl,--wrap=sonic_write_gc_profile, so the one call
// sonic_build makes to sonic_write_gc_profile (sonic/sonic.c:130) lands in __wrap_sonic_write_gc_profile below.
// That function keeps the framing of the original (sonic/sonic_reference.c:30-106: chromosome index, GC
// bytes, SONIC_END_OF_CHR per record, SONIC_END_OF_GC at the end — including that a record without a
// complete window gets no terminator, :61-64, and that a record whose name is not in the .fai gets no index,
// :76-78) and replaces the per-character loop by cg_gc_profile (include/conga_gpu.h) on the record's bases.
//
// Not reproduced: text before the first '>' (the original counts it with char_count starting at 0 and
// writes its bytes without an index, which its own reader cannot parse) — reported as an error instead.
#include <fcntl.h>
#include <getopt.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <future>
#include <string>
#include <thread>
#include <vector>

#include "conga_gpu.h"
extern "C" {
#include "sonic.h"
}

static double now_s()
{
    using namespace std::chrono;
    return duration<double>(steady_clock::now().time_since_epoch()).count();
}
static const bool g_timing = getenv("CONGA_TIMING") != nullptr;
static double g_t0 = now_s();
static void mark(const char* what)
{
    if (g_timing) fprintf(stderr, "[timing] %8.3f s  %s\n", now_s() - g_t0, what);
}

static bool g_space[256];

// bases of text[a, b) with white space removed -> out (capacity >= b - a); `threads` workers, two passes
static size_t strip_space(const unsigned char* text, size_t a, size_t b, char* out, int threads)
{
    const size_t n = b - a;
    if (threads < 1) threads = 1;
    if (n < (1u << 20)) threads = 1;
    std::vector<size_t> cnt(threads + 1, 0);
    auto seg = [&](int t) { return a + n * (size_t)t / (size_t)threads; };
    auto count = [&](int t) {
        size_t k = 0;
        for (size_t i = seg(t), e = seg(t + 1); i < e; i++) k += !g_space[text[i]];
        cnt[t + 1] = k;
    };
    auto copy = [&](int t) {
        // a white-space byte is stored into the slot of the next base and overwritten by it; stopping after
        // the segment's last base keeps every store inside the worker's own output range
        size_t e = seg(t + 1);
        const size_t i0 = seg(t);
        while (e > i0 && g_space[text[e - 1]]) e--;
        char* o = out + cnt[t];
        for (size_t i = i0; i < e; i++) { *o = (char)text[i]; o += !g_space[text[i]]; }
    };
    if (threads == 1) {
        count(0);
        copy(0);
        return cnt[1];
    }
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; t++) pool.emplace_back(count, t);
    for (auto& th : pool) th.join();
    pool.clear();
    for (int t = 0; t < threads; t++) cnt[t + 1] += cnt[t];
    for (int t = 0; t < threads; t++) pool.emplace_back(copy, t);
    for (auto& th : pool) th.join();
    return cnt[threads];
}

// The CUDA context (about a second to bring up) is created on a side thread from main(), behind
// sonic_build's parsing of the .fai, the BED files and the RepeatMasker file.
// The thread is detached and hands the context over through a future, so an exit() inside libsonic
// (missing file, malformed RepeatMasker line) ends the program the way it ends the original.
static std::shared_future<cg_ctx*> g_ctx_ready;
static cg_ctx* create_context()
{
    const char* dev_env = getenv("CONGA_GPU_DEVICE");
    return cg_create(dev_env ? atoi(dev_env) : 0);
}

static void die(const char* msg)
{
    fprintf(stderr, "[sonic-b200] %s\n", msg);
    exit(1);
}

extern "C" void __wrap_sonic_write_gc_profile(gzFile sonic_file, FILE* ref_file, int number_of_chromosomes, char** chromosome_names)
{
    mark("gc profile: start");
    for (int c = 0; c < 256; c++) g_space[c] = c == ' ' || (c >= '\t' && c <= '\r');      // isspace, C locale
    struct stat sb;
    const int fd = fileno(ref_file);
    if (fstat(fd, &sb) != 0) die("cannot stat the reference genome");
    size_t size = (size_t)sb.st_size;
    const unsigned char* text = nullptr;
    if (size) {
        void* m = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
        if (m == MAP_FAILED) die("cannot map the reference genome");
        madvise(m, size, MADV_SEQUENTIAL);
        text = static_cast<const unsigned char*>(m);
        // `char ch = fgetc(); if (ch == EOF) break;` (sonic_reference.c:57-59): a 0xFF byte ends the scan
        if (const void* ff = memchr(text, 0xFF, size)) size = (size_t)(static_cast<const unsigned char*>(ff) - text);
    }

    cg_ctx* ctx = g_ctx_ready.valid() ? g_ctx_ready.get() : create_context();
    if (!ctx) { fprintf(stderr, "[sonic-b200] %s\n", cg_last_error(nullptr)); exit(1); }
    mark("gc profile: GPU context ready");
    int threads = (int)std::thread::hardware_concurrency();
    if (const char* e = getenv("CONGA_DECODE_THREADS")) threads = atoi(e);
    if (threads < 1) threads = 1;
    if (threads > 64) threads = 64;

    char* bases = nullptr; size_t bases_cap = 0;
    int8_t* gc = nullptr; size_t gc_cap = 0;
    const char end_of_chr = (char)SONIC_END_OF_CHR;
    const int end_of_gc = SONIC_END_OF_GC;
    int written_gc = 0;
    size_t p = 0;

    // text before the first '>'
    {
        const void* gt = size ? memchr(text, '>', size) : nullptr;
        const size_t first = gt ? (size_t)(static_cast<const unsigned char*>(gt) - text) : size;
        for (size_t i = 0; i < first; i++)
            if (!g_space[text[i]]) die("the reference genome does not start with a '>' line");
        p = first;
    }
    while (p < size) {                                     /* text[p] == '>' */
        p++;
        if (written_gc != 0) gzwrite(sonic_file, &end_of_chr, 1);                        /* :61-64 */
        while (p < size && g_space[text[p]]) p++;                                       /* fscanf("%s") :66 */
        if (p >= size) break;
        const size_t name0 = p;
        while (p < size && !g_space[text[p]]) p++;
        const std::string name(reinterpret_cast<const char*>(text) + name0, p - name0);
        if (p >= size) break;                                                           /* fgets == NULL :70-72 */
        for (int k = 0; k < MAX_LENGTH - 1 && p < size; k++) if (text[p++] == '\n') break;   /* fgets :70 */
        int chromosome_index = -1;
        for (int i = 0; i < number_of_chromosomes; i++)
            if (name == chromosome_names[i]) { chromosome_index = i; break; }
        if (chromosome_index != -1) gzwrite(sonic_file, &chromosome_index, sizeof(chromosome_index));   /* :76-78 */
        written_gc = 0;
        const void* gt = p < size ? memchr(text + p, '>', size - p) : nullptr;
        const size_t end = gt ? (size_t)(static_cast<const unsigned char*>(gt) - text) : size;
        if (end - p + 64 > bases_cap) {
            if (bases) cg_free_pinned(ctx, bases);
            bases_cap = (end - p) + (end - p) / 8 + 4096;
            bases = static_cast<char*>(cg_alloc_pinned(ctx, bases_cap));
            if (!bases) die(cg_last_error(ctx));
        }
        const size_t n = strip_space(text, p, end, bases, threads);
        if (n > 0x7fffffffull) die("a chromosome of more than 2^31 - 1 bases");
        const size_t n_w = (n + 1) / 100;
        if (n_w + 64 > gc_cap) {
            if (gc) cg_free_pinned(ctx, gc);
            gc_cap = n_w + n_w / 8 + 4096;
            gc = static_cast<int8_t*>(cg_alloc_pinned(ctx, gc_cap));
            if (!gc) die(cg_last_error(ctx));
        }
        if (cg_gc_profile(ctx, bases, (int64_t)n, gc, (int32_t)n_w) != CG_OK) die(cg_last_error(ctx));
        if (n_w) gzwrite(sonic_file, gc, (unsigned)n_w);
        written_gc = (int)n_w;
        p = end;
    }
    gzwrite(sonic_file, &end_of_chr, 1);                                                /* :103 */
    gzwrite(sonic_file, &end_of_gc, sizeof(end_of_gc));                                 /* :104 */
    if (bases) cg_free_pinned(ctx, bases);
    if (gc) cg_free_pinned(ctx, gc);
    cg_destroy(ctx);
    if (text) munmap(const_cast<unsigned char*>(text), (size_t)sb.st_size);
    mark("gc profile: done");
}

static int print_help(const char* prog)
{
    fprintf(stderr, "\nSONIC: Some Organism's Nucleotide Information Container.\nVersion: %s (GC profile on the GPU)\n\n", SONIC_VERSION);
    fprintf(stderr, "Usage: %s [arguments]\n\n", prog);
    fprintf(stderr, "COMPULSORY PARAMETERS FOR CREATING SONIC:\n");
    fprintf(stderr, "\t--ref        [in.fasta]        : Reference genome FASTA file.\n");
    fprintf(stderr, "\t--gaps       [gap.bed]         : Reference genome assembly gap coodinates (BED).\n");
    fprintf(stderr, "\t--reps       [repeatmasker.out]: RepeatMasker output files (.out).\n");
    fprintf(stderr, "\t--dups       [dups.bed]        : Segmental duplication coordinates in the reference (BED).\n");
    fprintf(stderr, "\t--make-sonic [ref.sonic]       : SONIC file to be created.\n");
    fprintf(stderr, "VALIDATING EXISTING SONIC FILE:\n");
    fprintf(stderr, "\t--test-sonic [ref.sonic]       : SONIC file to be checked for validity.\n");
    fprintf(stderr, "OPTIONAL PARAMETERS:\n");
    fprintf(stderr, "\t--info       [info_string]     : Information string to annotate the SONIC file with, such as version.\n");
    fprintf(stderr, "\t--help                         : Print this help information and exit.\n\n");
    fprintf(stderr, "\t--version                      : Print version information and exit.\n\n");
    return 0;
}

int main(int argc, char** argv)
{
    // option table of sonic_exe.c:56-69 (the two "testing purposes only" entries left out)
    static struct option long_options[] = {
        {"ref", required_argument, 0, 'f'},  {"gaps", required_argument, 0, 'g'}, {"dups", required_argument, 0, 'd'},
        {"reps", required_argument, 0, 'r'}, {"mei", required_argument, 0, 'm'},  {"info", required_argument, 0, 'i'},
        {"make-sonic", required_argument, 0, 'c'}, {"test-sonic", required_argument, 0, 't'},
        {"help", no_argument, 0, 'h'}, {"version", no_argument, 0, 'v'}, {0, 0, 0, 0}};
    std::string ref, gaps, dups, reps, info, sonic_path;
    bool have_info = false, do_make = true, do_test = false;
    if (argc == 1) return print_help(argv[0]);
    int o, index;
    while ((o = getopt_long(argc, argv, "f:i:g:d:r:m:c:ht:v", long_options, &index)) != -1) {
        switch (o) {
        case 'f': ref = optarg; break;
        case 'g': gaps = optarg; break;
        case 'd': dups = optarg; break;
        case 'r': reps = optarg; break;
        case 'm': break;                                       /* accepted and unused, as in sonic_build */
        case 'i': info = optarg; have_info = true; break;
        case 'c': sonic_path = optarg; do_make = true; break;
        case 't': sonic_path = optarg; do_test = true; do_make = false; break;
        case 'h': return print_help(argv[0]);
        case 'v': fprintf(stderr, "\nSONIC: Some Organism's Nucleotide Information Container.\nVersion: %s\n\n", SONIC_VERSION); return 0;
        }
    }
    if (do_test) {
        sonic* s = sonic_load(const_cast<char*>(sonic_path.c_str()));
        if (!s) { fprintf(stderr, "[SONIC] The SONIC file %s is invalid.\n", sonic_path.c_str()); exit(-1); }
        fprintf(stderr, "[SONIC] Number of chromosomes: %d\n", s->number_of_chromosomes);
        fprintf(stderr, "[SONIC] Genome length: %ld\n", s->genome_length);
        fprintf(stderr, "[SONIC] The SONIC file %s seems to be valid.\n", sonic_path.c_str());
        return 0;
    }
    if (!do_make) return 0;
    if (ref.empty()) { fprintf(stderr, "[SONIC CMDLINE ERROR] Please enter reference genome file (FASTA) using the --ref option.\n"); return 0; }
    if (gaps.empty()) { fprintf(stderr, "[SONIC CMDLINE ERROR] Please enter the assembly gaps file (BED) using the --gaps option.\n"); return 0; }
    if (reps.empty()) { fprintf(stderr, "[SONIC CMDLINE ERROR] Please enter the repeats file (RepeaMasker) using the --reps option.\n"); return 0; }
    if (dups.empty()) { fprintf(stderr, "[SONIC CMDLINE ERROR] Please enter the segmental duplications file (BED) using the --dups option.\n"); return 0; }
    if (sonic_path.empty()) { fprintf(stderr, "[SONIC CMDLINE ERROR] Please enter the SONIC file to create using the --make-sonic option.\n"); return 0; }
    // sonic_exe.c hands sonic_build an uninitialised buffer when --info is absent; sonic_build's documented
    // default for "no info" is the reference genome's path (sonic.c:58-59): pass NULL to get it
    std::vector<char> a_ref(ref.begin(), ref.end()), a_gaps(gaps.begin(), gaps.end()), a_reps(reps.begin(), reps.end()),
        a_dups(dups.begin(), dups.end()), a_info(info.begin(), info.end()), a_sonic(sonic_path.begin(), sonic_path.end());
    for (auto* v : {&a_ref, &a_gaps, &a_reps, &a_dups, &a_info, &a_sonic}) v->push_back(0);
    mark("start");
    {
        std::packaged_task<cg_ctx*()> task(create_context);
        g_ctx_ready = task.get_future().share();
        std::thread(std::move(task)).detach();
    }
    sonic_build(a_ref.data(), a_gaps.data(), a_reps.data(), a_dups.data(), have_info ? a_info.data() : nullptr, a_sonic.data());
    mark("done");
    return 0;
}
