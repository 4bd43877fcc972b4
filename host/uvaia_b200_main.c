/* uvaia_b200 -- C host program of the B200-native n-best search: the `uvaia` command line (src/nearest.c:57-84,
 * 159-353 of the reference) on top of the C ABI of libuvaia_b200.so (include/uvaia_b200.h).
 *
 * The host does what the reference's host code does around the scoring seam and nothing else:
 *   - FASTA reader with the reference's hygiene rules (fastaseq.c:410-486, nexus_common.c:133-145), codec by content
 *     (file_compression.c:21-47): raw and gzip through zlib, xz / bzip2 through the command line tools;
 *   - query quality filter (utils.c:10-48), ordering (fastaseq.c:779-795), trim clamp (fastaseq.c:708-712);
 *   - the reference stream in batches: -x name lookup (nearest.c:255-261), length check (nearest.c:263-278);
 *   - <prefix>.csv.xz (nearest.c:513-547) and <prefix>.aln.xz (nearest.c:308-314, align.c:346-355).
 * Everything that scores a pair or keeps a neighbour runs on the GPU: without the library and a B200 the program
 * fails, there is no CPU path.  -k/--keep_resolved prunes redundant queries through uvg_prune_queries (fastaseq.c:797-841).
 */
#define _GNU_SOURCE
#include <ctype.h>
#include <fcntl.h>
#include <getopt.h>
#include <pthread.h>
#include <sched.h>
#include <semaphore.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>
#include <zlib.h>

#include "uvaia_b200.h"

#define PROGRAM_VERSION "uvaia_b200 0.1 (drop-in for uvaia 2.0.2)"

static void die(const char* fmt, ...)
{   /* biomcmc_error (lowlevel.c:39-52): message to stderr, exit(EXIT_FAILURE) */
    va_list ap;
    va_start(ap, fmt);
    fprintf(stderr, "uvaia_b200 error: ");
    vfprintf(stderr, fmt, ap);
    fprintf(stderr, "\n");
    va_end(ap);
    exit(EXIT_FAILURE);
}
static void* xmalloc(size_t n) { void* p = malloc(n ? n : 1); if (!p) die("out of memory (%zu bytes)", n); return p; }
static void* xrealloc(void* q, size_t n) { void* p = realloc(q, n ? n : 1); if (!p) die("out of memory (%zu bytes)", n); return p; }
static double now_s(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }

/* ---------------------------------------------------------------------------------------------------------------
 * input: codec by content, lines ending at \n or \r
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct {
    gzFile gz; FILE* pipe;
    unsigned char* buf; size_t cap, beg, end; int eof;
    unsigned char* line; size_t line_cap;
    int mem;                    /* buf is a borrowed memory range (a slice of a mapped raw file): nothing to read, nothing to free */
} reader_t;

static char* shell_quote(const char* s)
{
    size_t n = strlen(s), k = 0;
    char* q = xmalloc(4 * n + 3);
    q[k++] = '\'';
    for (size_t i = 0; i < n; ++i) { if (s[i] == '\'') { memcpy(q + k, "'\\''", 4); k += 4; } else q[k++] = s[i]; }
    q[k++] = '\''; q[k] = 0;
    return q;
}

static void reader_open(reader_t* r, const char* path)
{
    memset(r, 0, sizeof *r);
    unsigned char magic[6] = {0};
    FILE* f = fopen(path, "rb");
    if (!f) die("cannot open file %s", path);
    size_t got = fread(magic, 1, 6, f);
    fclose(f);
    const char* tool = NULL;
    if (got >= 6 && !memcmp(magic, "\xfd" "7zXZ\0", 6)) tool = "xz -dc -T0 -- ";          /* xz >= 5.4 decodes the blocks of a multi-block file (what `xz -T0` writes) on all cores */
    else if (got >= 3 && !memcmp(magic, "BZh", 3)) tool = "bzip2 -dc -- ";
    if (tool) {
        char* q = shell_quote(path);
        char* cmd = xmalloc(strlen(tool) + strlen(q) + 1);
        sprintf(cmd, "%s%s", tool, q);
        r->pipe = popen(cmd, "r");
        if (!r->pipe) die("cannot run '%s'", cmd);
        free(cmd); free(q);
    } else {
        r->gz = gzopen(path, "rb");                         /* transparent for uncompressed files */
        if (!r->gz) die("cannot open file %s", path);
        gzbuffer(r->gz, 1 << 20);
    }
    r->cap = 4 << 20;
    r->buf = xmalloc(r->cap);
}

static void reader_open_mem(reader_t* r, const unsigned char* base, size_t len)
{
    memset(r, 0, sizeof *r);
    r->buf = (unsigned char*)base; r->end = r->cap = len; r->eof = 1; r->mem = 1;
}

static void reader_close(reader_t* r)
{
    if (r->gz) gzclose(r->gz);
    if (r->pipe) pclose(r->pipe);
    if (!r->mem) free(r->buf);
    free(r->line);
    memset(r, 0, sizeof *r);
}

static int reader_fill(reader_t* r)
{
    if (r->eof) return 0;
    if (r->beg > 0) { memmove(r->buf, r->buf + r->beg, r->end - r->beg); r->end -= r->beg; r->beg = 0; }
    if (r->end == r->cap) { r->cap *= 2; r->buf = xrealloc(r->buf, r->cap); }
    long n = r->gz ? gzread(r->gz, r->buf + r->end, (unsigned)(r->cap - r->end)) : (long)fread(r->buf + r->end, 1, r->cap - r->end, r->pipe);
    if (n < 0) die("read error");
    if (n == 0) { r->eof = 1; return 0; }
    r->end += (size_t)n;
    return 1;
}

/* next line without its terminator (\n or \r, file_compression.c:216); returns 0 at end of input */
static int reader_line(reader_t* r, const unsigned char** line, size_t* len)
{
    size_t scan = r->beg;
    for (;;) {
        for (; scan < r->end; ++scan) {
            const unsigned char* p = memchr(r->buf + scan, '\n', r->end - scan);
            size_t stop = p ? (size_t)(p - r->buf) : r->end;
            const unsigned char* c = memchr(r->buf + scan, '\r', stop - scan);
            if (c) stop = (size_t)(c - r->buf);
            else if (!p) { scan = r->end; break; }
            *line = r->buf + r->beg; *len = stop - r->beg;
            r->beg = stop + 1;
            return 1;
        }
        const size_t done = scan - r->beg;
        if (!reader_fill(r)) {
            if (r->end > r->beg) { *line = r->buf + r->beg; *len = r->end - r->beg; r->beg = r->end; return 1; }
            return 0;
        }
        scan = r->beg + done;
    }
}

/* ---------------------------------------------------------------------------------------------------------------
 * FASTA records with the reference's hygiene (hazard H5): lines shorter than 2 bytes and lines whose first non-blank
 * byte is # or ; are dropped (nexus_common.c:133-145); a line containing '>' starts a record named by everything after
 * the first '>' (fastaseq.c:432-442); sequence lines lose white space and are upper-cased (fastaseq.c:451-452).
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct {
    reader_t rd;
    char* next_name;            /* name of the record whose header line has been read already */
    char* name; unsigned char* seq; size_t len, cap;
    int accumulate; size_t start;   /* accumulate: records are laid one behind the other in seq; the current one is seq[start, start+len) */
} fasta_t;

static int is_blank(unsigned char c) { return c == ' ' || (c >= '\t' && c <= '\r'); }

static void fasta_open(fasta_t* f, const char* path) { memset(f, 0, sizeof *f); reader_open(&f->rd, path); }
static void fasta_open_mem(fasta_t* f, const unsigned char* base, size_t len) { memset(f, 0, sizeof *f); reader_open_mem(&f->rd, base, len); f->accumulate = 1; }
static void fasta_close(fasta_t* f) { reader_close(&f->rd); free(f->next_name); free(f->name); free(f->seq); memset(f, 0, sizeof *f); }

/* 1 = a record is in f->name / f->seq / f->len (owned by f until the next call) */
static int fasta_next(fasta_t* f)
{
    const unsigned char* ln; size_t n;
    free(f->name); f->name = NULL;
    if (f->accumulate) f->start += f->len;
    f->len = 0;
    for (;;) {
        int have = reader_line(&f->rd, &ln, &n);
        if (!have) {
            if (f->next_name) { f->name = f->next_name; f->next_name = NULL; return 1; }
            return 0;
        }
        if (n < 2) continue;
        size_t i = 0;
        while (i < n && is_blank(ln[i])) ++i;
        if (i == n || ln[i] == '#' || ln[i] == ';') continue;
        const unsigned char* gt = memchr(ln, '>', n);
        if (gt) {
            const size_t k = (size_t)(gt - ln) + 1;
            char* nm = xmalloc(n - k + 1);
            memcpy(nm, ln + k, n - k); nm[n - k] = 0;
            if (f->next_name) { f->name = f->next_name; f->next_name = nm; return 1; }
            f->next_name = nm;
            continue;
        }
        if (!f->next_name) continue;                        /* sequence text before the first header */
        if (f->start + f->len + n > f->cap) { f->cap = 2 * (f->start + f->len + n) + 1024; f->seq = xrealloc(f->seq, f->cap); }
        {   /* upper-case copy (branch-free, vectorised by the compiler); a line with blanks inside is compacted afterwards */
            unsigned char* dst = f->seq + f->start + f->len;
            unsigned any_blank = 0;
            for (i = 0; i < n; ++i) {
                const unsigned char ch = ln[i];
                any_blank |= (unsigned)(ch == ' ') | (unsigned)((unsigned char)(ch - 9) < 5);
                dst[i] = (unsigned char)(ch - (((unsigned char)(ch - 'a') < 26) << 5));
            }
            if (!any_blank) f->len += n;
            else { size_t k = 0; for (i = 0; i < n; ++i) if (!is_blank(ln[i])) dst[k++] = dst[i]; f->len += k; }
        }
    }
}

/* ---------------------------------------------------------------------------------------------------------------
 * byte classes
 * ------------------------------------------------------------------------------------------------------------- */
static unsigned char g_degen[256], g_validtxt[256], g_isacgt[256];
static void init_tables(void)
{   /* degeneracy of the IUPAC codes (biomcmc alignment.c:906-931); indel bytes of quick_count_sequence_non_N (fastaseq.c:642-648) */
    const char* d1 = "ACGTU"; const char* d2 = "MRWSYK"; const char* d3 = "VHDB";
    for (const char* p = d1; *p; ++p) g_degen[(unsigned char)*p] = 1;
    for (const char* p = d2; *p; ++p) g_degen[(unsigned char)*p] = 2;
    for (const char* p = d3; *p; ++p) g_degen[(unsigned char)*p] = 3;
    memset(g_validtxt, 1, 256);
    for (const char* p = "NnXx-?Oo."; *p; ++p) g_validtxt[(unsigned char)*p] = 0;
    for (const char* p = "ACGTacgt"; *p; ++p) g_isacgt[(unsigned char)*p] = 1;
}

/* ---------------------------------------------------------------------------------------------------------------
 * queries: read, filter (utils.c:10-48), order (fastaseq.c:779-795)
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct { char** name; unsigned char* text; int n, nchar; } queries_t;

typedef struct { long cnt; int idx; } order_key;
static int cmp_order(const void* a, const void* b)
{
    const order_key* x = a; const order_key* y = b;
    if (x->cnt != y->cnt) return x->cnt < y->cnt ? -1 : 1;
    return x->idx - y->idx;
}

static queries_t read_queries(const char* path, double ambiguity, int trim_arg, int acgt, int* n_read)
{
    fasta_t f;
    fasta_open(&f, path);
    char** name = NULL; unsigned char** seq = NULL; size_t* len = NULL; int n = 0, cap = 0;
    *n_read = 0;
    while (fasta_next(&f)) {
        ++*n_read;
        const size_t L = f.len;
        if (L < 5) continue;
        size_t d0 = 0, d1 = 0;
        for (size_t i = 0; i < L; ++i) { const unsigned char d = g_degen[f.seq[i]]; d0 += d == 0; d1 += d == 1; }
        if ((double)d0 / (double)L > ambiguity) continue;
        if ((double)d1 / (double)L < 1.0 - 1.1 * ambiguity) continue;
        if (n == cap) { cap = cap ? 2 * cap : 256; name = xrealloc(name, cap * sizeof *name); seq = xrealloc(seq, cap * sizeof *seq); len = xrealloc(len, cap * sizeof *len); }
        name[n] = f.name; f.name = NULL;
        seq[n] = xmalloc(L); memcpy(seq[n], f.seq, L); len[n] = L;
        ++n;
    }
    fasta_close(&f);
    queries_t q = {0};
    if (n < 1) return q;
    for (int i = 1; i < n; ++i) if (len[i] != len[0]) die("query sequences are not aligned");                  /* utils.c:38-42 */
    const int L = (int)len[0];
    int trim = trim_arg < 0 ? 0 : trim_arg;                                                                   /* fastaseq.c:708-712 */
    if (trim > L / 2.1) trim = (int)(L / 2.1);
    order_key* key = xmalloc(n * sizeof *key);
    for (int i = 0; i < n; ++i) {
        long c = 0;
        for (int s = trim; s < L - trim; ++s) c += acgt ? g_isacgt[seq[i][s]] : g_validtxt[seq[i][s]];
        key[i].cnt = c; key[i].idx = i;
    }
    qsort(key, n, sizeof *key, cmp_order);
    q.n = n; q.nchar = L;
    q.name = xmalloc(n * sizeof *q.name);
    q.text = xmalloc((size_t)n * L);
    for (int i = 0; i < n; ++i) {
        q.name[i] = name[key[i].idx];
        memcpy(q.text + (size_t)i * L, seq[key[i].idx], L);
    }
    for (int i = 0; i < n; ++i) free(seq[i]);
    free(seq); free(len); free(name); free(key);
    return q;
}

/* -x: names of the kept queries (open addressing over FNV-1a) */
typedef struct { char** slot; size_t mask; } nameset_t;
static uint64_t fnv(const char* s) { uint64_t h = 1469598103934665603ull; for (; *s; ++s) { h ^= (unsigned char)*s; h *= 1099511628211ull; } return h; }
static nameset_t nameset_build(char** name, int n)
{
    nameset_t s; size_t cap = 16;
    while (cap < (size_t)n * 2 + 2) cap *= 2;
    s.mask = cap - 1; s.slot = calloc(cap, sizeof *s.slot);
    if (!s.slot) die("out of memory");
    for (int i = 0; i < n; ++i) { size_t h = fnv(name[i]) & s.mask; while (s.slot[h]) h = (h + 1) & s.mask; s.slot[h] = name[i]; }
    return s;
}
static int nameset_has(const nameset_t* s, const char* nm)
{
    for (size_t h = fnv(nm) & s->mask; s->slot[h]; h = (h + 1) & s->mask) if (!strcmp(s->slot[h], nm)) return 1;
    return 0;
}

/* ---------------------------------------------------------------------------------------------------------------
 * output: xz through the command line tool when there is one, else gzip under the same name (what the reference's
 * zlib-only build does, file_compression.c:288-317)
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct { FILE* pipe; gzFile gz; } writer_t;
static writer_t writer_open(const char* path)
{
    writer_t w = {0};
    if (system("command -v xz > /dev/null 2>&1") == 0) {
        char* q = shell_quote(path);
        char* cmd = xmalloc(strlen(q) + 32);
        sprintf(cmd, "xz -c -T0 -2 > %s", q);
        w.pipe = popen(cmd, "w");
        free(cmd); free(q);
    }
    if (!w.pipe) { w.gz = gzopen(path, "wb"); if (!w.gz) die("cannot write %s", path); }
    return w;
}
static void writer_put(writer_t* w, const void* p, size_t n)
{
    if (!n) return;
    if (w->pipe) { if (fwrite(p, 1, n, w->pipe) != n) die("write error"); }
    else if (gzwrite(w->gz, p, (unsigned)n) != (int)n) die("write error");
}
static void writer_close(writer_t* w) { if (w->pipe && pclose(w->pipe) != 0) die("xz failed"); if (w->gz) gzclose(w->gz); }

/* ---------------------------------------------------------------------------------------------------------------
 * reference stream: the reader thread (nearest.c:245-286 without the scoring)
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct { unsigned char* text; char** name; long n; int last; long fend[1025]; int n_fend; } batch_t;   /* fend: rows of the batch before each file end */
typedef struct {
    const char** file; int n_files, L, min_non_n; long batch_rows;
    const nameset_t* qset; double t0;
    batch_t batch[2];
    sem_t free_slots, full_slots;
    uint8_t* rec_state; int64_t n_rec, rec_cap;          /* per record: 0 = skipped, 2 = pushed (in stream order) */
    long count, n_invalid, same_name;
} stream_t;

/* one parsed record enters the stream: counters, -x, the length check, the copy into the pinned batch.  `name` is owned
 * by the callee (kept for accepted rows, freed otherwise). */
typedef struct {
    stream_t* sp; batch_t* bt; int k;                        /* first pass: rows into the pinned batches */
    uint64_t digest; long n_digest;                          /* --parse_only */
    writer_t* aln; const uint8_t *rec_ok, *best; int64_t n_rec, rec, ord; long n_output;   /* second pass: the .aln.xz dump */
} stream_pos_t;
static void stream_take(stream_pos_t* ps, char* name, const unsigned char* seq, size_t len)
{
    stream_t* sp = ps->sp;
    if (ps->aln) {                                           /* second pass: write the records that ever entered a heap */
        if (ps->rec >= ps->n_rec) die("a reference file changed while running");
        if (ps->rec_ok[ps->rec++] == 1 && ps->best[ps->ord++]) {
            writer_put(ps->aln, ">", 1); writer_put(ps->aln, name, strlen(name)); writer_put(ps->aln, "\n", 1);
            writer_put(ps->aln, seq, len); writer_put(ps->aln, "\n", 1);
            ++ps->n_output;
        }
        free(name);
        return;
    }
    if (!sp) {                                               /* --parse_only: a digest of what the parser delivers, in order */
        uint64_t h = ps->digest ? ps->digest : 1469598103934665603ull;
        for (const char* c = name; *c; ++c) { h ^= (unsigned char)*c; h *= 1099511628211ull; }
        h ^= 0xff; h *= 1099511628211ull;
        size_t i = 0;
        for (; i + 8 <= len; i += 8) { uint64_t w; memcpy(&w, seq + i, 8); h ^= w; h *= 1099511628211ull; h ^= h >> 29; }
        for (; i < len; ++i) { h ^= seq[i]; h *= 1099511628211ull; }
        ps->digest = h; ++ps->n_digest;
        free(name);
        return;
    }
    ++sp->count;
    if (sp->n_rec == sp->rec_cap) { sp->rec_cap = sp->rec_cap ? 2 * sp->rec_cap : 1 << 16; sp->rec_state = xrealloc(sp->rec_state, sp->rec_cap); }
    sp->rec_state[sp->n_rec++] = 0;
    if (sp->qset && nameset_has(sp->qset, name)) { ++sp->same_name; free(name); return; }        /* nearest.c:255-261 */
    if ((int64_t)len != sp->L) {
        /* nearest.c:263-278: a poor-quality reference of the wrong length is skipped, a good one is fatal */
        long nn = 0;
        for (size_t i = 0; i < len; ++i) nn += g_validtxt[seq[i]];
        if (nn < sp->min_non_n) { ++sp->n_invalid; free(name); return; }
        die("all sequences must be aligned: reference %s has %zu sites, queries have %d", name, len, sp->L);
    }
    batch_t* bt = ps->bt;
    memcpy(bt->text + (size_t)bt->n * sp->L, seq, sp->L);
    bt->name[bt->n++] = name;
    sp->rec_state[sp->n_rec - 1] = 2;
    if (bt->n == sp->batch_rows) {
        sem_post(&sp->full_slots);
        ps->k ^= 1;
        sem_wait(&sp->free_slots);
        ps->bt = &sp->batch[ps->k];
        ps->bt->n = 0; ps->bt->last = 0; ps->bt->n_fend = 0;
    }
}

/* Raw (uncompressed) FASTA files are mapped and parsed by several threads: a line that holds '>' starts a record whatever
 * came before it, so a slice that begins at such a line parses on its own.  The file goes through in windows; every
 * worker parses one slice of the window into its own buffer, then this thread takes the records in file order. */
static size_t env_size(const char* name, size_t dflt) { const char* e = getenv(name); return e && atoll(e) > 0 ? (size_t)atoll(e) : dflt; }
typedef struct { char* name; size_t off, len; } prec_t;
typedef struct { const unsigned char* base; size_t len; prec_t* rec; size_t n, cap; unsigned char* seq; } pslice_t;

static void* parse_slice(void* arg)
{
    pslice_t* ps = arg;
    fasta_t f;
    fasta_open_mem(&f, ps->base, ps->len);
    while (fasta_next(&f)) {
        if (ps->n == ps->cap) { ps->cap = ps->cap ? 2 * ps->cap : 1024; ps->rec = xrealloc(ps->rec, ps->cap * sizeof *ps->rec); }
        ps->rec[ps->n].name = f.name; ps->rec[ps->n].off = f.start; ps->rec[ps->n].len = f.len; ++ps->n;
        f.name = NULL;
    }
    ps->seq = f.seq; f.seq = NULL;
    fasta_close(&f);
    return NULL;
}
/* offset of the first line at or after `from` that starts a record (same hygiene as fasta_next), or `size` */
static size_t next_header(const unsigned char* m, size_t size, size_t from)
{
    size_t p = from;
    if (p > 0) { while (p < size && m[p - 1] != '\n' && m[p - 1] != '\r') ++p; }       /* to the start of a line */
    while (p < size) {
        size_t e = p;
        while (e < size && m[e] != '\n' && m[e] != '\r') ++e;
        const size_t n = e - p;
        if (n >= 2) {
            size_t i = p;
            while (i < e && is_blank(m[i])) ++i;
            if (i < e && m[i] != '#' && m[i] != ';' && memchr(m + p, '>', n)) return p;
        }
        p = e + 1;
    }
    return size;
}

static int parse_threads(void)
{
    const char* e = getenv("UVG_PARSE_THREADS");
    long t = e ? atol(e) : sysconf(_SC_NPROCESSORS_ONLN) - 2;
    return t < 1 ? 1 : t > 32 ? 32 : (int)t;
}
/* 1 = the file was raw text and has been streamed; 0 = compressed (or not mappable): the caller reads it serially */
static int stream_file_parallel(stream_pos_t* pos, const char* path)
{
    const int T = parse_threads();
    if (T < 2) return 0;
    int fd = open(path, O_RDONLY);
    if (fd < 0) die("cannot open file %s", path);
    struct stat sb;
    if (fstat(fd, &sb) || !S_ISREG(sb.st_mode) || sb.st_size < (off_t)env_size("UVG_PARSE_MIN_BYTES", 1 << 20)) { close(fd); return 0; }
    const size_t size = (size_t)sb.st_size, window = env_size("UVG_PARSE_WINDOW", (size_t)256 << 20);
    unsigned char* m = mmap(NULL, size, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (m == MAP_FAILED) return 0;
    if ((m[0] == 0x1f && m[1] == 0x8b) || !memcmp(m, "\xfd" "7zXZ", 5) || !memcmp(m, "BZh", 3)) { munmap(m, size); return 0; }
    madvise(m, size, MADV_SEQUENTIAL);
    /* two sets of slices: the parsers of the next window run while this thread takes the records of the current one */
    pslice_t* sl[2]; pthread_t* th[2];
    for (int k = 0; k < 2; ++k) { sl[k] = xmalloc(T * sizeof **sl); th[k] = xmalloc(T * sizeof **th); }
    double t_join = 0, t_take = 0;
    size_t w0 = 0;                                           /* start of the window: offset 0 or a header line */
    int cur = 0, have = 0;
    for (;;) {
        /* launch the window that starts at w0 into the free set */
        const int nxt = have ? cur ^ 1 : cur;
        int launched = 0;
        if (w0 < size) {
            const size_t w1 = w0 + window >= size ? size : next_header(m, size, w0 + window);
            size_t cut[33];
            cut[0] = w0; cut[T] = w1;
            for (int t = 1; t < T; ++t) {
                cut[t] = next_header(m, w1, w0 + (w1 - w0) / T * t);
                if (cut[t] < cut[t - 1]) cut[t] = cut[t - 1];
            }
            for (int t = 0; t < T; ++t) {
                memset(&sl[nxt][t], 0, sizeof sl[nxt][t]);
                sl[nxt][t].base = m + cut[t]; sl[nxt][t].len = cut[t + 1] - cut[t];
                if (pthread_create(&th[nxt][t], NULL, parse_slice, &sl[nxt][t])) die("cannot start a parser thread");
            }
            w0 = w1; launched = 1;
        }
        if (!have) { if (!launched) break; have = 1; continue; }
        for (int t = 0; t < T; ++t) {
            const double ta = now_s();
            pthread_join(th[cur][t], NULL);
            const double tb = now_s();
            pslice_t* ps = &sl[cur][t];
            for (size_t i = 0; i < ps->n; ++i) stream_take(pos, ps->rec[i].name, ps->seq + ps->rec[i].off, ps->rec[i].len);
            free(ps->rec); free(ps->seq);
            t_join += tb - ta; t_take += now_s() - tb;
        }
        if (!launched) break;
        cur ^= 1;
    }
    if (getenv("UVG_PARSE_TIMING")) fprintf(stderr, "parallel parse of %s: %d threads, waited %.3f s for the parsers, %.3f s taking the records\n", path, T, t_join, t_take);
    for (int k = 0; k < 2; ++k) { free(sl[k]); free(th[k]); }
    munmap(m, size);
    return 1;
}

static void* stream_reader(void* arg)
{
    stream_t* sp = arg;
    stream_pos_t pos = {sp, NULL, 0};
    sem_wait(&sp->free_slots);
    pos.bt = &sp->batch[0];
    pos.bt->n = 0; pos.bt->last = 0; pos.bt->n_fend = 0;
    for (int j = 0; j < sp->n_files; ++j) {
        const double t1 = now_s();
        if (!stream_file_parallel(&pos, sp->file[j])) {
            fasta_t f;
            fasta_open(&f, sp->file[j]);
            while (fasta_next(&f)) { char* nm = f.name; f.name = NULL; stream_take(&pos, nm, f.seq, f.len); }
            fasta_close(&f);
        }
        pos.bt->fend[pos.bt->n_fend++] = pos.bt->n;          /* the reference's batch ends with the file (nearest.c:245-286) */
        fprintf(stderr, "Finished reading file %s in %.3lf secs;\n", sp->file[j], now_s() - t1);
    }
    pos.bt->last = 1;
    sem_post(&sp->full_slots);
    return NULL;
}

/* ---------------------------------------------------------------------------------------------------------------
 * --cache FILE: the packed references of a run (uvg_export_packed blobs + names + which FASTA records were accepted),
 * so that the next run over the same reference files skips reading, parsing and encoding them.  Layout (host endian):
 *   "UVGCACHE" u32 version, i32 nchar, i32 acgt, i32 min_non_n, i32 n_files, i32 pad; per file: u64 size, i64 mtime
 *   records: u64 tag (1 blob, 2 record map, 3 end, 4 file end: u64 accepted references up to here) ...
 *     blob:       u64 blob_bytes, u64 n_refs, u64 names_bytes, blob, names (NUL terminated each), padding to 8
 *     record map: u64 n_rec, i64 count, i64 n_invalid, rec_ok[n_rec], padding to 8
 * A cache is used only if nchar, mode, the -A threshold and the size and mtime of every -r file match; not with -x (the
 * set of pushed references then depends on the query names).
 * ------------------------------------------------------------------------------------------------------------- */
#define CACHE_VERSION 2u
typedef struct { FILE* fp; char* path; char* tmp; void* buf; size_t buf_cap; int64_t exported, base; } cache_writer_t;

static void cache_put(FILE* fp, const void* p, size_t n) { if (n && fwrite(p, 1, n, fp) != n) die("cannot write the cache file"); }
static void cache_pad8(FILE* fp, size_t n) { static const char z[8] = {0}; if (n % 8) cache_put(fp, z, 8 - n % 8); }
static int cache_get(FILE* fp, void* p, size_t n) { return !n || fread(p, 1, n, fp) == n; }

static void cache_header(FILE* fp, int nchar, int acgt, int min_non_n, const char** file, int n_files)
{
    const uint32_t ver = CACHE_VERSION; const int32_t v[5] = {nchar, acgt, min_non_n, n_files, 0};
    cache_put(fp, "UVGCACHE", 8); cache_put(fp, &ver, 4); cache_put(fp, v, sizeof v);
    for (int j = 0; j < n_files; ++j) {
        struct stat sb;
        if (stat(file[j], &sb)) die("cannot stat %s", file[j]);
        const uint64_t size = (uint64_t)sb.st_size; const int64_t mtime = (int64_t)sb.st_mtime;
        cache_put(fp, &size, 8); cache_put(fp, &mtime, 8);
    }
}
/* -> open file positioned at the first record, or NULL when there is no usable cache */
static FILE* cache_open_valid(const char* path, int nchar, int acgt, int min_non_n, const char** file, int n_files)
{
    FILE* fp = fopen(path, "rb");
    if (!fp) return NULL;
    char magic[8]; uint32_t ver = 0; int32_t v[5];
    int ok = cache_get(fp, magic, 8) && !memcmp(magic, "UVGCACHE", 8) && cache_get(fp, &ver, 4) && ver == CACHE_VERSION && cache_get(fp, v, sizeof v)
             && v[0] == nchar && v[1] == acgt && v[2] == min_non_n && v[3] == n_files;
    for (int j = 0; ok && j < n_files; ++j) {
        struct stat sb; uint64_t size; int64_t mtime;
        ok = cache_get(fp, &size, 8) && cache_get(fp, &mtime, 8) && !stat(file[j], &sb) && (uint64_t)sb.st_size == size && (int64_t)sb.st_mtime == mtime;
    }
    if (!ok) { fclose(fp); return NULL; }
    return fp;
}
/* export the resident references that are not in the file yet: whole groups of 32, everything when `final` (before a
 * drop and at the end of the stream; the reader then drops at the same place) */
static void cache_flush(cache_writer_t* cw, uvg_ctx* gpu, char** rname, int final)
{
    if (!cw->fp) return;
    const int64_t db_n = uvg_db_count(gpu), hi = final ? db_n : db_n / 32 * 32;
    if (hi > cw->exported) {
        const int64_t first = cw->exported, n = hi - first;
        size_t bytes = 0;
        if (uvg_packed_bytes(gpu, first, n, &bytes)) die("%s", uvg_last_error(gpu));
        if (bytes > cw->buf_cap) {
            uvg_host_free_pinned(cw->buf);
            cw->buf_cap = bytes + bytes / 8; cw->buf = uvg_host_alloc_pinned(cw->buf_cap);
            if (!cw->buf) die("cannot allocate %zu bytes of pinned memory", cw->buf_cap);
        }
        if (uvg_export_packed(gpu, first, n, cw->buf, cw->buf_cap)) die("%s", uvg_last_error(gpu));
        uint64_t names_bytes = 0;
        for (int64_t i = 0; i < n; ++i) names_bytes += strlen(rname[cw->base + first + i]) + 1;
        const uint64_t rec[4] = {1, bytes, (uint64_t)n, names_bytes};
        cache_put(cw->fp, rec, sizeof rec); cache_put(cw->fp, cw->buf, bytes);
        for (int64_t i = 0; i < n; ++i) { const char* nm = rname[cw->base + first + i]; cache_put(cw->fp, nm, strlen(nm) + 1); }
        cache_pad8(cw->fp, names_bytes);
        cw->exported = hi;
    }
    if (final) { cw->base += db_n; cw->exported = 0; }
}

/* ---------------------------------------------------------------------------------------------------------------
 * main: nearest.c:159-353
 * ------------------------------------------------------------------------------------------------------------- */
static void usage(const char* prog, int full)
{
    printf("%s\nFor every query sequence, finds the closest neighbours in the reference alignment (n-best search on a B200 GPU)\n", PROGRAM_VERSION);
    printf("Usage: %s [-hvxk] [--acgt] [-n <int>] [--trim=<int>] [-A <double>] [-a <double>] [-p <int>] -r <ref.fa(.gz,.xz)> [-r ...] "
           "[-t <int>] [-o <without suffix>] <seqs.fa(.gz,.xz)>\n", prog);
    if (!full) return;
    printf("  -n, --nbest            number of best reference sequences per query to store (default=100)\n"
           "      --trim             number of sites to trim from both ends (default=0, suggested for sarscov2=230)\n"
           "  -a, --query_ambiguity  maximum allowed ambiguity for QUERY sequence to be excluded (default=0.5)\n"
           "  -A, --ref_ambiguity    maximum allowed ambiguity for REFERENCE sequence to be excluded (default=0.5)\n"
           "      --acgt             considers only ACGT sites in query sequences (mismatch-based)\n"
           "  -k, --keep_resolved    keep more resolved and exclude redundant query seqs (default is to keep all)\n"
           "  -x, --exclude_self     exclude reference sequences with same name as a query sequence\n"
           "  -p, --pool             batch size of the reference program (default 64 per thread); its table can depend on it and\n"
           "                         the search reproduces the table of this batch size\n"
           "  -r, --reference        aligned reference sequences (can be several files)\n"
           "  -t, --nthreads         accepted for compatibility (scoring runs on the GPU)\n"
           "  -o, --output           prefix of xzipped output alignment and table with nearest neighbour sequences\n"
           "      --device           CUDA device ordinal (default 0);  --batch: reference rows per push (default 65536)\n"
           "      --devices          several GPUs of the box, e.g. 0-7 or 0,2,5: the references are row-sharded over them (the\n"
           "                         tables do not depend on it; --cache and --max_resident do not apply)\n"
           "      --cache            file that keeps the packed references of this run; a later run over the same reference\n"
           "                         files (same size and mtime, same -A, same mode) reads it instead of parsing them\n");
}

/* ---------------------------------------------------------------------------------------------------------------
 * `uvaiaball` (src/ball.c:138-291) over uvg_ball_*: keeps the reference sequences within a radius of some query.
 * Invoked as `uvaia_b200 ball ...` or through a program name that starts with "uvaiaball".
 * ------------------------------------------------------------------------------------------------------------- */
static void ball_usage(const char* prog, int full)
{
    printf("%s ball  [-h|--help] [-v|--version] [-x|--acgt] [-k|--keep_resolved] [-d|--distance=<int>] [--trim=<int>] [-A|--ref_ambiguity=<double>] "
           "[-a|--query_ambiguity=<double>] [-p|--pool=<int>] -r|--reference=<ref.fa(.gz,.xz)> [-r ...]... <seqs.fa(.gz,.xz)> [-o|--output=<without suffix>]\n", prog);
    if (!full) return;
    printf("Finds all reference neighbours to query sequences within a ball radius (i.e. close to any query sequence), on a B200 GPU\n"
           "  -x, --acgt             considers only ACGT sites (i.e. unambiguous SNP differences)\n"
           "  -k, --keep_resolved    when excluding redundant query seqs, keep the more resolved (instead of default, less resolved)\n"
           "  -d, --distance         ball radius, i.e. refs within this distance to any query seq are kept (default=5)\n"
           "      --trim             number of sites to trim from both ends (default=0, suggested for sarscov2=230)\n"
           "  -a, --query_ambiguity  maximum allowed ambiguity for QUERY sequence to be excluded (default=0.5)\n"
           "  -A, --ref_ambiguity    maximum allowed ambiguity for REFERENCE sequence to be excluded (default=0.5)\n"
           "  -p, --pool             accepted for compatibility\n"
           "  -r, --reference        aligned reference sequences (can be several files)\n"
           "  -o, --output           prefix of xzipped output alignment with subset of ref sequences (default ball_uvaia)\n"
           "      --device           CUDA device ordinal (default 0);  --batch: reference rows per push (default 65536)\n");
}

static int ball_main(int argc, char** argv)
{
    int dist = 5, trim = 0, acgt = 0, hires = 0, device = 0;                    /* ball.c:63-66 */
    long batch_rows = 65536;
    double ambig_q = 0.5, ambig_r = 0.5;
    const char* out_prefix = "ball_uvaia";
    const char* ref_file[1024]; int n_ref_files = 0;
    static struct option lo[] = {
        {"help", 0, 0, 'h'}, {"version", 0, 0, 'v'}, {"acgt", 0, 0, 'x'}, {"keep_resolved", 0, 0, 'k'}, {"distance", 1, 0, 'd'}, {"trim", 1, 0, 1001},
        {"query_ambiguity", 1, 0, 'a'}, {"ref_ambiguity", 1, 0, 'A'}, {"pool", 1, 0, 'p'}, {"reference", 1, 0, 'r'}, {"output", 1, 0, 'o'},
        {"device", 1, 0, 1002}, {"batch", 1, 0, 1003}, {0, 0, 0, 0}};
    int ch;
    while ((ch = getopt_long(argc, argv, "hvxkd:a:A:p:r:o:", lo, NULL)) != -1) {
        switch (ch) {
        case 'h': ball_usage(argv[0], 1); return EXIT_SUCCESS;
        case 'v': printf("%s\n", PROGRAM_VERSION); return EXIT_SUCCESS;
        case 'x': acgt = 1; break;
        case 'k': hires = 1; break;
        case 'd': dist = atoi(optarg); break;
        case 1001: trim = atoi(optarg); break;
        case 'a': ambig_q = atof(optarg); break;
        case 'A': ambig_r = atof(optarg); break;
        case 'p': break;
        case 'r': if (n_ref_files == 1024) die("at most 1024 reference files"); ref_file[n_ref_files++] = optarg; break;
        case 'o': out_prefix = optarg; break;
        case 1002: device = atoi(optarg); break;
        case 1003: batch_rows = atol(optarg); break;
        default: ball_usage(argv[0], 0); return EXIT_FAILURE;
        }
    }
    if (n_ref_files < 1 || argc - optind != 1) { ball_usage(argv[0], 0); return EXIT_FAILURE; }
    const char* query_file = argv[optind];
    if (ambig_q < 0.001) ambig_q = 0.001;                                       /* ball.c:154-157 */
    if (ambig_q > 1.) ambig_q = 1.;
    if (ambig_r < 0.001) ambig_r = 0.001;
    if (ambig_r > 1.) ambig_r = 1.;
    if (batch_rows < 1) batch_rows = 1;
    init_tables();
    const double t0 = now_s();
    fprintf(stderr, "Experimental program: uvaia_b200 ball package: %s\n", PROGRAM_VERSION);
    int n_read = 0;
    queries_t q = read_queries(query_file, ambig_q, trim, acgt, &n_read);
    fprintf(stderr, "Finished reading %d query references in %lf secs;\n", n_read, now_s() - t0);
    fprintf(stderr, "Query database composed of %d valid references, after excluding low quality.\n", q.n);
    if (q.n < 1) die("No valid reference sequences found. Please check file %s.", query_file);
    const int L = q.nchar;
    fprintf(stderr, acgt ? "Only ACGT differences are taken into account. \n" : "Partial matches are taken into account (excluding only gaps and Ns).\n");
    {   /* ball.c:190: redundant queries always go; -k keeps the more resolved one */
        uint8_t* keep = xmalloc(q.n);
        int n_kept = 0;
        if (uvg_prune_queries(q.text, q.n, L, trim, acgt, hires, keep, &n_kept)) die("uvg_prune_queries failed");
        int w = 0;
        for (int i = 0; i < q.n; ++i) {
            if (!keep[i]) { free(q.name[i]); continue; }
            if (w != i) { memmove(q.text + (size_t)w * L, q.text + (size_t)i * L, L); q.name[w] = q.name[i]; }
            ++w;
        }
        q.n = n_kept;
        free(keep);
        fprintf(stderr, "Query database now composed of %d valid references, after removing redundant (%s resolved) sequences.\n", q.n, hires ? "less" : "more");
    }
    uvg_ball_params prm;
    memset(&prm, 0, sizeof prm);
    prm.abi_version = UVG_ABI_VERSION; prm.nchar = L; prm.trim = trim; prm.dist = dist; prm.acgt_mode = acgt; prm.device = device;
    uvg_ball* gpu = NULL;
    uvg_query_info qi;
    if (uvg_ball_create(&prm, &gpu)) die("%s", uvg_ball_last_error(NULL));
    if (uvg_ball_set_queries(gpu, q.text, q.n, &qi)) die("%s", uvg_ball_last_error(gpu));
    fprintf(stderr, "Query sequence alignment: %d segregating, %d non-segregating sites with indels, and %d constant sites (all are used in comparisons)\n",
            qi.n_idx, qi.n_idx_m, qi.n_idx_c);
    fprintf(stderr, "radius distance is %d (refs more distant than this are excluded)\n", (int)uvg_ball_radius(gpu));

    const int min_non_n = (int)(L * ambig_r);                                   /* ball.c:201: A, not 1 - A */
    if (batch_rows > ((long)1 << 30) / L) batch_rows = (((long)1 << 30) / L) > 0 ? ((long)1 << 30) / L : 1;
    size_t plen = strlen(out_prefix);
    char* outname = xmalloc(plen + 16);
    sprintf(outname, "%s.aln.xz", out_prefix);
    writer_t w = writer_open(outname);
    stream_t sp;
    memset(&sp, 0, sizeof sp);
    sp.file = ref_file; sp.n_files = n_ref_files; sp.L = L; sp.min_non_n = min_non_n; sp.batch_rows = batch_rows; sp.qset = NULL; sp.t0 = t0;
    for (int k = 0; k < 2; ++k) {
        sp.batch[k].text = uvg_host_alloc_pinned((size_t)batch_rows * L);
        if (!sp.batch[k].text) die("cannot allocate %zu bytes of pinned memory", (size_t)batch_rows * L);
        sp.batch[k].name = xmalloc(batch_rows * sizeof(char*));
    }
    sem_init(&sp.free_slots, 0, 2);
    sem_init(&sp.full_slots, 0, 0);
    pthread_t reader;
    if (pthread_create(&reader, NULL, stream_reader, &sp)) die("cannot start the reader thread");
    uint8_t* keep = xmalloc(batch_rows);
    long n_output = 0, n_invalid = 0;
    for (int k = 0;; k ^= 1) {
        sem_wait(&sp.full_slots);
        batch_t* bt = &sp.batch[k];
        if (bt->n > 0) {
            if (uvg_ball_push_refs(gpu, bt->text, bt->n, min_non_n, keep, NULL)) die("%s", uvg_ball_last_error(gpu));
            for (long i = 0; i < bt->n; ++i) {
                if (keep[i] == 2) {                                             /* ball.c:246-252: within the radius, in file order */
                    writer_put(&w, ">", 1); writer_put(&w, bt->name[i], strlen(bt->name[i])); writer_put(&w, "\n", 1);
                    writer_put(&w, bt->text + (size_t)i * L, L); writer_put(&w, "\n", 1);
                    ++n_output;
                } else if (keep[i] == 0) ++n_invalid;
                free(bt->name[i]);
            }
        }
        const int last = bt->last;
        sem_post(&sp.free_slots);
        if (last) break;
    }
    pthread_join(reader, NULL);
    n_invalid += sp.n_invalid;
    writer_close(&w);
    fprintf(stderr, "Total of %ld sequences read, %ld sequences within radius (kept), %ld too ambiguous (excluded)\n", sp.count, n_output, n_invalid);
    fprintf(stderr, "Saved %ld sequences to file %s, %.3lf secs elapsed.\n", n_output, outname, now_s() - t0);
    uvg_ball_destroy(gpu);
    return EXIT_SUCCESS;
}

/* ---------------------------------------------------------------------------------------------------------------
 * --devices: the database row-sharded over several GPUs of the box inside ONE process -- one host thread and one
 * uvg_ctx per GPU.  The main thread cuts every batch of the reader into runs of rows that belong to one
 * rank (uniform rounds of MG_ROUND accepted references per rank, dealt in file order: uvg_shard_of with n_total = -1)
 * and hands them to the rank threads, which copy, filter and encode them concurrently; the search runs when the
 * last file has been read (the contexts run the pool-independent closed form and count hazard-H1 flags).
 * ------------------------------------------------------------------------------------------------------------- */
static long MG_ROUND = 16384;             /* accepted references per GPU and round; UVG_MG_ROUND overrides (multiple of 64: the tests use small rounds) */
enum { MG_QUIT = 0, MG_PUSH, MG_MARK, MG_SEARCH, MG_EVER };
typedef struct {
    int rank, world;
    uvg_params prm;
    const unsigned char* qtext; int nq;
    unsigned char nccl_id[128];
    uvg_ctx* ctx; uvg_query_info qi;
    pthread_t th; sem_t go, done;
    int job, busy;
    const unsigned char* text; long n; int min_non_n; int64_t count, n_acc; uint8_t* flags;
    char err[600]; int failed;
} mg_rank_t;

static void mg_fail(mg_rank_t* r, const char* what) { snprintf(r->err, sizeof r->err, "GPU %d (rank %d): %s: %s", r->prm.device, r->rank, what, uvg_last_error(r->ctx)); r->failed = 1; }
static void* mg_thread(void* arg)
{
    mg_rank_t* r = arg;
    if (uvg_create(&r->prm, &r->ctx)) { snprintf(r->err, sizeof r->err, "GPU %d: %s", r->prm.device, uvg_last_error(NULL)); r->failed = 1; }
    else if (uvg_nccl_init(r->ctx, r->nccl_id)) mg_fail(r, "uvg_nccl_init");
    else if (uvg_set_queries(r->ctx, r->qtext, r->nq, &r->qi)) mg_fail(r, "uvg_set_queries");      /* rank 0 prepares, the others receive by broadcast */
    sem_post(&r->done);
    for (;;) {
        sem_wait(&r->go);
        if (r->job == MG_QUIT) break;
        if (!r->failed) {
            if (r->job == MG_PUSH) { r->n_acc = 0; if (uvg_push_refs(r->ctx, r->text, r->n, r->min_non_n, NULL, NULL, &r->n_acc)) mg_fail(r, "uvg_push_refs"); }
            else if (r->job == MG_MARK) { if (uvg_mark_file_end(r->ctx, r->count)) mg_fail(r, "uvg_mark_file_end"); }
            else if (r->job == MG_SEARCH) { if (uvg_set_global_count(r->ctx, r->count) || uvg_search(r->ctx)) mg_fail(r, "uvg_search"); }
            else if (r->job == MG_EVER) { if (r->count && uvg_get_ever_best(r->ctx, 0, r->count, r->flags)) mg_fail(r, "uvg_get_ever_best"); }
        }
        sem_post(&r->done);
    }
    return NULL;
}
static void mg_wait(mg_rank_t* r) { if (r->busy) { sem_wait(&r->done); r->busy = 0; if (r->failed) die("%s", r->err); } }
static void mg_submit(mg_rank_t* r, int job) { mg_wait(r); r->job = job; r->busy = 1; sem_post(&r->go); }
static void mg_all(mg_rank_t* rk, int G, int job, int64_t count)
{
    for (int g = 0; g < G; ++g) { mg_wait(&rk[g]); rk[g].count = count; mg_submit(&rk[g], job); }
    for (int g = 0; g < G; ++g) mg_wait(&rk[g]);
}
/* valid_ref_sites of the rows of a batch on host threads (quick_count_sequence_non_N, fastaseq.c:642-648): the host has to
 * know which rows the filter accepts before it can deal them out by accepted ordinal */
typedef struct { const unsigned char* text; long a, b; int L; int* out; } nn_job_t;
static void* nn_worker(void* arg)
{
    nn_job_t* j = arg;
    for (long i = j->a; i < j->b; ++i) {
        const unsigned char* row = j->text + (size_t)i * j->L;
        int c = 0;
        for (int k = 0; k < j->L; ++k) c += g_validtxt[row[k]];
        j->out[i] = c;
    }
    return NULL;
}
static void batch_non_n(const unsigned char* text, long n, int L, int* out)
{
    long T = sysconf(_SC_NPROCESSORS_ONLN) - 1;
    if (T > 32) T = 32;
    if (T > n / 64) T = n / 64;
    if (T < 2) { nn_job_t j = {text, 0, n, L, out}; nn_worker(&j); return; }
    pthread_t th[32]; nn_job_t job[32];
    for (long t = 0; t < T; ++t) {
        job[t] = (nn_job_t){text, n * t / T, n * (t + 1) / T, L, out};
        if (pthread_create(&th[t], NULL, nn_worker, &job[t])) die("cannot start a worker thread");
    }
    for (long t = 0; t < T; ++t) pthread_join(th[t], NULL);
}

/* what omp_get_max_threads() answers in the reference program before -t is applied (nearest.c:81): OMP_NUM_THREADS when it
 * is set, else the processors this process may run on */
static int omp_max_threads_like(void)
{
    const char* e = getenv("OMP_NUM_THREADS");
    if (e && atoi(e) > 0) return atoi(e);
    cpu_set_t set;
    if (!sched_getaffinity(0, sizeof set, &set) && CPU_COUNT(&set) > 0) return CPU_COUNT(&set);
    const long n = sysconf(_SC_NPROCESSORS_ONLN);
    return n > 0 ? (int)n : 1;
}

int main(int argc, char** argv)
{
    {   /* `uvaia_b200 ball ...` or a program name that starts with "uvaiaball": the radius search */
        const char* base = strrchr(argv[0], '/');
        base = base ? base + 1 : argv[0];
        if (!strncmp(base, "uvaiaball", 9)) return ball_main(argc, argv);
        if (argc > 1 && !strcmp(argv[1], "ball")) { argv[1] = argv[0]; return ball_main(argc - 1, argv + 1); }
    }
    int nbest = 100, trim = 0, acgt = 0, xclude = 0, pool = 0, device = 0, hires = 0;
    int devs[8], n_dev = 0;                                                     /* --devices: row sharding over several GPUs */
    long batch_rows = 65536, max_resident = 1 << 20;
    const char* cache_path = NULL;
    int parse_only = 0;
    double ambig_q = 0.5, ambig_r = 0.5;
    const char* out_prefix = NULL;
    const char* ref_file[1024]; int n_ref_files = 0;
    static struct option lo[] = {
        {"help", 0, 0, 'h'}, {"version", 0, 0, 'v'}, {"acgt", 0, 0, 1000}, {"keep_resolved", 0, 0, 'k'}, {"exclude_self", 0, 0, 'x'},
        {"nbest", 1, 0, 'n'}, {"trim", 1, 0, 1001}, {"query_ambiguity", 1, 0, 'a'}, {"ref_ambiguity", 1, 0, 'A'}, {"pool", 1, 0, 'p'},
        {"reference", 1, 0, 'r'}, {"nthreads", 1, 0, 't'}, {"output", 1, 0, 'o'}, {"device", 1, 0, 1002}, {"batch", 1, 0, 1003},
        {"max_resident", 1, 0, 1004}, {"cache", 1, 0, 1005}, {"parse_only", 0, 0, 1006}, {"devices", 1, 0, 1007}, {0, 0, 0, 0}};
    int ch;
    while ((ch = getopt_long(argc, argv, "hvkxn:a:A:p:r:t:o:", lo, NULL)) != -1) {
        switch (ch) {
        case 'h': usage(argv[0], 1); return EXIT_SUCCESS;                       /* nearest.c:113-157: help and version exit 0 */
        case 'v': printf("%s\n", PROGRAM_VERSION); return EXIT_SUCCESS;
        case 1000: acgt = 1; break;
        case 'k': hires = 1; break;
        case 'x': xclude = 1; break;
        case 'n': nbest = atoi(optarg); break;
        case 1001: trim = atoi(optarg); break;
        case 'a': ambig_q = atof(optarg); break;
        case 'A': ambig_r = atof(optarg); break;
        case 'p': pool = atoi(optarg); break;
        case 'r': if (n_ref_files == 1024) die("at most 1024 reference files"); ref_file[n_ref_files++] = optarg; break;
        case 't': break;
        case 'o': out_prefix = optarg; break;
        case 1002: device = atoi(optarg); break;
        case 1003: batch_rows = atol(optarg); break;
        case 1004: max_resident = atol(optarg); break;
        case 1005: cache_path = optarg; break;
        case 1006: parse_only = 1; break;
        case 1007: {                                                            /* "0,1,2" or "0-7" */
            n_dev = 0;
            for (const char* c = optarg; *c && n_dev < 8;) {
                const int a = atoi(c);
                while (*c >= '0' && *c <= '9') ++c;
                int b = a;
                if (*c == '-') { b = atoi(++c); while (*c >= '0' && *c <= '9') ++c; }
                for (int d = a; d <= b && n_dev < 8; ++d) devs[n_dev++] = d;
                if (*c == ',') ++c; else if (*c) { usage(argv[0], 0); return EXIT_FAILURE; }
            }
            break;
        }
        default: usage(argv[0], 0); return EXIT_FAILURE;
        }
    }
    if (parse_only) {
        /* diagnostic: run the reference files through the FASTA reader (parallel for raw files, UVG_PARSE_THREADS) and
         * print how many records came out and a digest of their names and sequences; needs no GPU */
        init_tables();
        for (int j = 0; j < n_ref_files; ++j) {
            stream_pos_t pos;
            memset(&pos, 0, sizeof pos);
            const double t1 = now_s();
            const int par = stream_file_parallel(&pos, ref_file[j]);
            if (!par) {
                fasta_t f;
                fasta_open(&f, ref_file[j]);
                while (fasta_next(&f)) { char* nm = f.name; f.name = NULL; stream_take(&pos, nm, f.seq, f.len); }
                fasta_close(&f);
            }
            printf("%s records=%ld digest=%016llx parser=%s secs=%.3f\n", ref_file[j], pos.n_digest, (unsigned long long)pos.digest,
                   par ? "parallel" : "serial", now_s() - t1);
        }
        return EXIT_SUCCESS;
    }
    if (n_ref_files < 1 || argc - optind != 1) { usage(argv[0], 0); return EXIT_FAILURE; }
    const char* query_file = argv[optind];
    if (ambig_q < 0.001) ambig_q = 0.001;                                       /* nearest.c:175-179 */
    if (ambig_q > 1.) ambig_q = 1.;
    if (ambig_r < 0.001) ambig_r = 0.001;
    if (ambig_r > 1.) ambig_r = 1.;
    if (nbest < 1) nbest = 1;
    if (pool < 1) pool = 64 * omp_max_threads_like();                            /* nearest.c:81: 64 x omp_get_max_threads(), set before -t is read */
    if (batch_rows < 1) batch_rows = 1;
    if (!out_prefix) out_prefix = acgt ? "nn_uvaia_acgt" : "nn_uvaia";          /* nearest.c:198-200 */
    init_tables();
    const double t0 = now_s();
    fprintf(stderr, "program: uvaia_b200 package: %s\n", PROGRAM_VERSION);
    fprintf(stderr, "Creating a queue of %d sequences; for each query, the %d closest sequences will be stored\n", pool, nbest);   /* nearest.c:192 */

    /* 1. queries */
    int n_read = 0;
    queries_t q = read_queries(query_file, ambig_q, trim, acgt, &n_read);
    fprintf(stderr, "Finished reading %d query references in %lf secs;\n", n_read, now_s() - t0);
    fprintf(stderr, "Query database composed of %d valid references, after excluding low quality.\n", q.n);
    if (q.n < 1) die("No valid reference sequences found. Please check file %s.", query_file);
    const int L = q.nchar, K = nbest < 2 ? 2 : nbest;
    uvg_params prm;
    memset(&prm, 0, sizeof prm);
    prm.abi_version = UVG_ABI_VERSION; prm.nchar = L; prm.trim = trim; prm.nbest = nbest; prm.acgt_mode = acgt; prm.pool = pool; prm.device = device;
    uvg_ctx* gpu = NULL;
    uvg_query_info qi;
    const int multi = n_dev > 1;
    if (multi && getenv("UVG_MG_ROUND") && atol(getenv("UVG_MG_ROUND")) >= 64) MG_ROUND = atol(getenv("UVG_MG_ROUND")) / 64 * 64;
    if (n_dev == 1) prm.device = device = devs[0];
    if (multi && cache_path) { fprintf(stderr, "Note: --cache is ignored with --devices.\n"); cache_path = NULL; }
    if (!multi && uvg_create(&prm, &gpu)) die("%s", uvg_last_error(NULL));
    if (hires) {                                                                /* nearest.c:219-224 */
        uint8_t* keep = xmalloc(q.n);
        int n_kept = 0;
        if (uvg_prune_queries(q.text, q.n, L, trim, acgt, 1, keep, &n_kept)) die("uvg_prune_queries failed");
        int w = 0;
        for (int i = 0; i < q.n; ++i) {
            if (!keep[i]) { free(q.name[i]); continue; }
            if (w != i) { memmove(q.text + (size_t)w * L, q.text + (size_t)i * L, L); q.name[w] = q.name[i]; }
            ++w;
        }
        q.n = n_kept;
        free(keep);
        fprintf(stderr, "Updated query database composed of %d valid references, after removing redundant (less resolved) sequences.\n", q.n);
    }
    mg_rank_t rk[8];
    memset(rk, 0, sizeof rk);
    if (multi) {
        /* one thread + one context per GPU; the query planes are prepared by rank 0 and broadcast over NVLink */
        unsigned char id[128];
        if (uvg_nccl_unique_id(id)) die("%s", uvg_last_error(NULL));
        for (int g = 0; g < n_dev; ++g) {
            rk[g].rank = g; rk[g].world = n_dev; rk[g].prm = prm;
            rk[g].prm.device = devs[g]; rk[g].prm.rank = g; rk[g].prm.world = n_dev;
            rk[g].prm.first_round = (int32_t)MG_ROUND; rk[g].prm.max_round = (int32_t)MG_ROUND; rk[g].prm.nominal_tail = 1;
            rk[g].qtext = q.text; rk[g].nq = q.n; memcpy(rk[g].nccl_id, id, 128);
            sem_init(&rk[g].go, 0, 0); sem_init(&rk[g].done, 0, 0);
            rk[g].busy = 1;
            if (pthread_create(&rk[g].th, NULL, mg_thread, &rk[g])) die("cannot start the thread of GPU %d", devs[g]);
        }
        for (int g = 0; g < n_dev; ++g) mg_wait(&rk[g]);
        gpu = rk[0].ctx; qi = rk[0].qi;
        fprintf(stderr, "The reference sequences are row-sharded over %d GPUs (rounds of %ld accepted references per GPU).\n", n_dev, MG_ROUND);
    } else if (uvg_set_queries(gpu, q.text, q.n, &qi)) die("%s", uvg_last_error(gpu));
    fprintf(stderr, "Query sequence alignment: %d segregating, %d non-segregating sites with indels, and %d constant sites (all are used in comparisons)\n",
            qi.n_idx, qi.n_idx_m, qi.n_idx_c);                                  /* fastaseq.c:774 */
    fprintf(stderr, acgt ? "Considering ACGT differences only (excluding all other characters). \n"
                         : "Considering text match and partially ambiguous (excluding only gaps and Ns).\n");
    nameset_t qset = {0};
    if (xclude) {
        fprintf(stderr, "Reference sequences with same name as query sequences will be excluded from the comparison.\n");
        qset = nameset_build(q.name, q.n);
    }

    /* 2. reference stream.  A reader thread parses the files into two batches of rows in pinned memory; this thread hands
     * every full batch to uvg_push_refs (copy + encoder) and uvg_search_async: reading and parsing batch i+1, the copy of
     * batch i and the scoring rounds of batch i-1 overlap. */
    const int min_non_n = (int)(L * (1.0 - ambig_r));                           /* nearest.c:237 */
    if (batch_rows > ((long)1 << 30) / L) batch_rows = (((long)1 << 30) / L) > 0 ? ((long)1 << 30) / L : 1;
    uint8_t* accepted = xmalloc(batch_rows);
    char** rname = NULL; int64_t n_acc_total = 0, rname_cap = 0;                /* names of the accepted references, by ordinal */
    uint8_t* push_ok = NULL; int64_t n_pushed = 0, push_cap = 0;                /* per pushed row, in stream order: accepted by the filter */
    long n_invalid = 0, count = 0, same_name = 0;
    uint8_t* rec_ok = NULL; int64_t n_rec = 0;
    int* mg_nonn = NULL; int64_t mg_mark[1100]; int mg_n_mark = 0;              /* --devices: valid_ref_sites of a batch; accepted counts at the file ends */
    void* pinned_text[2] = {NULL, NULL}; char** batch_names[2] = {NULL, NULL};
    FILE* cache_in = NULL;
    cache_writer_t cw;
    memset(&cw, 0, sizeof cw);
    if (cache_path && xclude) { fprintf(stderr, "Note: --cache is ignored with -x (the references that are pushed depend on the query names).\n"); cache_path = NULL; }
    if (cache_path) {
        cache_in = cache_open_valid(cache_path, L, acgt, min_non_n, ref_file, n_ref_files);
        if (cache_in) fprintf(stderr, "Packed references are read from the cache %s; the reference files are not parsed.\n", cache_path);
        else {
            cw.path = xmalloc(strlen(cache_path) + 1); strcpy(cw.path, cache_path);
            cw.tmp = xmalloc(strlen(cache_path) + 8); sprintf(cw.tmp, "%s.tmp", cache_path);
            cw.fp = fopen(cw.tmp, "wb");
            if (!cw.fp) die("cannot write the cache file %s", cw.tmp);
            setvbuf(cw.fp, NULL, _IOFBF, 8 << 20);
            cache_header(cw.fp, L, acgt, min_non_n, ref_file, n_ref_files);
            fprintf(stderr, "Packed references will be saved to the cache %s.\n", cache_path);
        }
    }
    if (cache_in) {
        /* cached run: blob -> pinned buffer -> uvg_import_packed -> rounds; the next blob is read under them */
        void* buf = NULL; size_t buf_cap = 0;
        for (;;) {
            uint64_t tag = 0;
            if (!cache_get(cache_in, &tag, 8)) die("the cache file %s is truncated (delete it)", cache_path);
            if (tag == 3) break;
            if (tag == 1) {
                uint64_t h[3];
                if (!cache_get(cache_in, h, sizeof h)) die("the cache file %s is truncated (delete it)", cache_path);
                const size_t bytes = h[0], names_bytes = h[2]; const int64_t n = (int64_t)h[1];
                if (bytes > buf_cap) {
                    uvg_host_free_pinned(buf);
                    buf_cap = bytes + bytes / 8; buf = uvg_host_alloc_pinned(buf_cap);
                    if (!buf) die("cannot allocate %zu bytes of pinned memory", buf_cap);
                }
                char* names = xmalloc(names_bytes + 8);
                if (!cache_get(cache_in, buf, bytes) || !cache_get(cache_in, names, (names_bytes + 7) / 8 * 8)) die("the cache file %s is truncated (delete it)", cache_path);
                if (n_acc_total + n > rname_cap) { rname_cap = 2 * (n_acc_total + n) + 1024; rname = xrealloc(rname, rname_cap * sizeof *rname); }
                const char* nm = names;
                for (int64_t i = 0; i < n; ++i) {
                    const size_t len = strlen(nm);
                    rname[n_acc_total] = xmalloc(len + 1); memcpy(rname[n_acc_total++], nm, len + 1);
                    nm += len + 1;
                }
                free(names);
                int64_t got = 0;
                if (uvg_import_packed(gpu, buf, bytes, &got) || got != n) die("%s", uvg_last_error(gpu));
                if (uvg_search_async(gpu)) die("%s", uvg_last_error(gpu));
                if (n % 32 || uvg_db_count(gpu) >= max_resident) { if (uvg_wait(gpu) || uvg_drop_refs(gpu)) die("%s", uvg_last_error(gpu)); }
            } else if (tag == 4) {
                uint64_t e = 0;
                if (!cache_get(cache_in, &e, 8)) die("the cache file %s is truncated (delete it)", cache_path);
                if (uvg_mark_file_end(gpu, (int64_t)e)) die("%s", uvg_last_error(gpu));
            } else if (tag == 2) {
                uint64_t nr = 0; int64_t v[2];
                if (!cache_get(cache_in, &nr, 8) || !cache_get(cache_in, v, sizeof v)) die("the cache file %s is truncated (delete it)", cache_path);
                n_rec = (int64_t)nr; count = (long)v[0]; n_invalid = (long)v[1];
                rec_ok = xmalloc((nr + 7) / 8 * 8 + 8);
                if (!cache_get(cache_in, rec_ok, (nr + 7) / 8 * 8)) die("the cache file %s is truncated (delete it)", cache_path);
            } else die("the cache file %s is corrupt (delete it)", cache_path);
        }
        fclose(cache_in);
        uvg_host_free_pinned(buf);
        fprintf(stderr, "Finished reading the cache %s in %.3lf secs;\n", cache_path, now_s() - t0);
    } else {
        stream_t sp;
        memset(&sp, 0, sizeof sp);
        sp.file = ref_file; sp.n_files = n_ref_files; sp.L = L; sp.min_non_n = min_non_n; sp.batch_rows = batch_rows;
        sp.qset = xclude ? &qset : NULL; sp.t0 = t0;
        for (int k = 0; k < 2; ++k) {
            sp.batch[k].text = uvg_host_alloc_pinned((size_t)batch_rows * L);
            if (!sp.batch[k].text) die("cannot allocate %zu bytes of pinned memory", (size_t)batch_rows * L);
            sp.batch[k].name = xmalloc(batch_rows * sizeof(char*));
        }
        sem_init(&sp.free_slots, 0, 2);
        sem_init(&sp.full_slots, 0, 0);
        pthread_t reader;
        if (pthread_create(&reader, NULL, stream_reader, &sp)) die("cannot start the reader thread");
        for (int k = 0;; k ^= 1) {
            sem_wait(&sp.full_slots);
            batch_t* bt = &sp.batch[k];
            if (multi && bt->n > 0) {
                /* which rows the filter accepts (host threads), then runs of rows per rank: accepted ordinal o belongs to rank
                 * (o / MG_ROUND) % G (uniform rounds in file order); rows the filter drops travel with the run they fall into
                 * and are dropped again by the rank's encoder (same rule) */
                if (!mg_nonn) mg_nonn = xmalloc(batch_rows * sizeof *mg_nonn);
                batch_non_n(bt->text, bt->n, L, mg_nonn);
                if (n_acc_total + bt->n > rname_cap) { rname_cap = 2 * (n_acc_total + bt->n) + 1024; rname = xrealloc(rname, rname_cap * sizeof *rname); }
                if (n_pushed + bt->n > push_cap) { push_cap = 2 * (n_pushed + bt->n) + 1024; push_ok = xrealloc(push_ok, push_cap); }
                long run0 = 0; int run_rank = -1; int64_t run_acc = 0;
                int fe = 0;
                for (long i = 0; i <= bt->n; ++i) {
                    while (fe < bt->n_fend && bt->fend[fe] == i) { mg_mark[mg_n_mark++] = n_acc_total; ++fe; }      /* a file ends before row i */
                    const int ok = i < bt->n && mg_nonn[i] >= min_non_n;
                    const int r = ok ? (int)((n_acc_total / MG_ROUND) % n_dev) : run_rank;
                    if (i == bt->n || (ok && run_rank >= 0 && r != run_rank)) {
                        if (run_rank >= 0 && i > run0) {                       /* close the run [run0, i) */
                            mg_rank_t* t = &rk[run_rank];
                            mg_wait(t);
                            if (t->n_acc != t->count) die("GPU %d accepted %lld of a run of references, the host counted %lld", t->prm.device, (long long)t->n_acc, (long long)t->count);
                            t->text = bt->text + (size_t)run0 * L; t->n = i - run0; t->min_non_n = min_non_n; t->count = run_acc;
                            mg_submit(t, MG_PUSH);
                        }
                        run0 = i; run_acc = 0; run_rank = -1;
                        if (i == bt->n) break;
                    }
                    if (i < bt->n) {
                        push_ok[n_pushed++] = (uint8_t)ok;
                        if (ok) { if (run_rank < 0) run_rank = r; rname[n_acc_total++] = bt->name[i]; ++run_acc; } else { free(bt->name[i]); ++n_invalid; }
                    }
                }
                for (int g = 0; g < n_dev; ++g) {                              /* the batch buffer goes back to the reader: everything pushed */
                    mg_wait(&rk[g]);
                    if (rk[g].n_acc != rk[g].count) die("GPU %d accepted %lld of a run of references, the host counted %lld", rk[g].prm.device, (long long)rk[g].n_acc, (long long)rk[g].count);
                    rk[g].n_acc = rk[g].count = 0;
                }
                bt->n = 0;
            }
            /* the rows of a batch go to the device in pieces that end where a file ends: uvg_mark_file_end tells the library
             * where the reference's --pool batches restart (nearest.c:245-286) */
            long row0 = 0;
            for (int fe = 0; !multi && fe <= bt->n_fend; ++fe) {
                const long row1 = fe < bt->n_fend ? bt->fend[fe] : bt->n, m = row1 - row0;
                if (m > 0) {
                    int64_t na = 0;
                    if (uvg_push_refs(gpu, bt->text + (size_t)row0 * L, m, min_non_n, NULL, accepted, &na)) die("%s", uvg_last_error(gpu));
                    if (n_acc_total + na > rname_cap) { rname_cap = 2 * (n_acc_total + na) + 1024; rname = xrealloc(rname, rname_cap * sizeof *rname); }
                    if (n_pushed + m > push_cap) { push_cap = 2 * (n_pushed + m) + 1024; push_ok = xrealloc(push_ok, push_cap); }
                    for (long i = 0; i < m; ++i) {
                        push_ok[n_pushed++] = accepted[i];
                        if (accepted[i]) rname[n_acc_total++] = bt->name[row0 + i]; else { free(bt->name[row0 + i]); ++n_invalid; }
                    }
                    if (uvg_search_async(gpu)) die("%s", uvg_last_error(gpu));
                    const int drop = uvg_db_count(gpu) >= max_resident;
                    cache_flush(&cw, gpu, rname, drop);
                    if (drop) { if (uvg_wait(gpu) || uvg_drop_refs(gpu)) die("%s", uvg_last_error(gpu)); }
                }
                if (fe < bt->n_fend) {
                    if (uvg_mark_file_end(gpu, n_acc_total)) die("%s", uvg_last_error(gpu));
                    if (cw.fp) { const uint64_t rec[2] = {4, (uint64_t)n_acc_total}; cache_put(cw.fp, rec, sizeof rec); }
                }
                row0 = row1;
            }
            const int last = bt->last;
            sem_post(&sp.free_slots);
            if (last) break;
        }
        pthread_join(reader, NULL);
        n_invalid += sp.n_invalid;
        count = sp.count; same_name = sp.same_name;
        /* per streamed record: 1 = accepted reference (the second pass for the .aln dump needs it) */
        rec_ok = sp.rec_state; n_rec = sp.n_rec;
        { int64_t k = 0; for (int64_t i = 0; i < n_rec; ++i) rec_ok[i] = rec_ok[i] == 2 ? push_ok[k++] : 0; }
        pinned_text[0] = sp.batch[0].text; pinned_text[1] = sp.batch[1].text; batch_names[0] = sp.batch[0].name; batch_names[1] = sp.batch[1].name;
        if (cw.fp) {
            cache_flush(&cw, gpu, rname, 1);
            const uint64_t rec[2] = {2, (uint64_t)n_rec}; const int64_t v[2] = {count, n_invalid};
            cache_put(cw.fp, rec, sizeof rec); cache_put(cw.fp, v, sizeof v); cache_put(cw.fp, rec_ok, (size_t)n_rec); cache_pad8(cw.fp, (size_t)n_rec);
            const uint64_t end = 3;
            cache_put(cw.fp, &end, 8);
            if (fclose(cw.fp) || rename(cw.tmp, cw.path)) die("cannot finish the cache file %s", cw.path);
            uvg_host_free_pinned(cw.buf); free(cw.tmp); free(cw.path);
            fprintf(stderr, "Saved the packed references to the cache %s, %.3lf secs elapsed.\n", cache_path, now_s() - t0);
        }
    }
    fprintf(stderr, "Total of %ld sequences read; %ld too ambiguous (excluded). %.3lf secs elapsed. \n", count, n_invalid, now_s() - t0);
    if (xclude) fprintf(stderr, " %ld reference sequences already present in query alignment (based on name only).\n", same_name);
    uint8_t* best = xmalloc((size_t)n_acc_total + 1);
    long long mg_flags = 0;
    if (multi) {
        uint8_t* scratch = xmalloc((size_t)n_acc_total + 1);
        for (int k = 0; k < mg_n_mark; ++k) mg_all(rk, n_dev, MG_MARK, mg_mark[k]);
        mg_all(rk, n_dev, MG_SEARCH, n_acc_total);                              /* collective: rounds over all shards, tables to every rank */
        for (int g = 0; g < n_dev; ++g) rk[g].flags = g ? scratch : best;       /* collective: the ranks' ever-best flags are combined */
        mg_all(rk, n_dev, MG_EVER, n_acc_total);
        for (int g = 0; g < n_dev; ++g) { uvg_stats sg; if (!uvg_get_stats(rk[g].ctx, &sg)) mg_flags += sg.n_h1_flags; }
        fprintf(stderr, "Search over %d GPUs finished, %.3lf secs elapsed.\n", n_dev, now_s() - t0);
        free(scratch);
    } else if (uvg_wait(gpu)) die("%s", uvg_last_error(gpu));

    /* 3. <prefix>.aln.xz: every reference that ever entered a heap, in file order (nearest.c:303-314).  The text is not
     * kept on the host: the reference files are streamed a second time and the flagged records are written. */
    uvg_stats st;
    if (uvg_get_stats(gpu, &st)) die("%s", uvg_last_error(gpu));
    if (!multi && n_acc_total && uvg_get_ever_best(gpu, 0, n_acc_total, best)) die("%s", uvg_last_error(gpu));
    size_t plen = strlen(out_prefix);
    char* outname = xmalloc(plen + 16);
    sprintf(outname, "%s.aln.xz", out_prefix);
    long n_output = 0;
    {
        writer_t w = writer_open(outname);
        stream_pos_t pos;
        memset(&pos, 0, sizeof pos);
        pos.aln = &w; pos.rec_ok = rec_ok; pos.best = best; pos.n_rec = n_rec;
        for (int j = 0; j < n_ref_files; ++j) {
            if (stream_file_parallel(&pos, ref_file[j])) continue;
            fasta_t f;
            fasta_open(&f, ref_file[j]);
            while (fasta_next(&f)) { char* nm = f.name; f.name = NULL; stream_take(&pos, nm, f.seq, f.len); }
            fasta_close(&f);
        }
        n_output = pos.n_output;
        writer_close(&w);
    }
    fprintf(stderr, "Total of %ld sequences read; %ld saved sequences include closest neighbours and intermediate, %ld too ambiguous (excluded). "
                    "Highest number of ACGT mismatches = %lld in current neighbourhood.\n", count, n_output, n_invalid, (long long)st.max_mismatch);
    fprintf(stderr, "Saved %ld sequences to file %s , %.3lf secs elapsed.\n", n_output, outname, now_s() - t0);
    if (multi && mg_flags)
        fprintf(stderr, "Note: %lld admissions may depend on --pool in the reference program (its consensus stage truncates per batch); sharded searches give "
                        "the pool-independent table (= --pool 1).  Run on one GPU for the table of --pool %d.\n", mg_flags, pool);
    if (!multi && st.n_h1_flags)
        fprintf(stderr, "Note: %lld pairs were scored with consensus counts truncated at the batch-wide limit, as the reference program does "
                        "with --pool %d (its table depends on the pool size for this input).\n", (long long)st.n_h1_flags, pool);

    /* 4. <prefix>.csv.xz (save_distance_table, nearest.c:513-547) */
    sprintf(outname, "%s.csv.xz", out_prefix);
    {
        writer_t w = writer_open(outname);
        const char* hdr = acgt ? "query,reference,rank,ACGT_matches,valid_ACGT_comparisons,ACGT_matches_unique,valid_ref_sites,dist_consensus,dist_unique\n"
                               : "query,reference,rank,ACGT_matches,text_matches,partial_matches,valid_pair_comparisons,ACGT_matches_unique,valid_ref_sites\n";
        writer_put(&w, hdr, strlen(hdr));
        uvg_item* rows = xmalloc((size_t)K * sizeof *rows);
        size_t lcap = 1 << 16; char* line = xmalloc(lcap);
        for (int iq = 0; iq < q.n; ++iq) {
            int32_t n = 0;
            if (uvg_get_heap(gpu, iq, rows, &n)) die("%s", uvg_last_error(gpu));
            for (int i = 0; i < n; ++i) {
                const char* rn = rname[rows[i].ref];
                const size_t need = strlen(q.name[iq]) + strlen(rn) + 128;
                if (need > lcap) { lcap = 2 * need; line = xrealloc(line, lcap); }
                const int k = snprintf(line, lcap, "%s,%s,%d,%d,%d,%d,%d,%d,%d\n", q.name[iq], rn, i + 1, rows[i].score[0], rows[i].score[1],
                                       rows[i].score[2], rows[i].score[3], rows[i].score[4], rows[i].score[5]);
                writer_put(&w, line, (size_t)k);
            }
        }
        writer_close(&w);
        free(rows); free(line);
    }
    fprintf(stderr, "Saved distance table to file %s , %.3lf secs elapsed.\n", outname, now_s() - t0);

    if (multi) {
        for (int g = 0; g < n_dev; ++g) { mg_submit(&rk[g], MG_QUIT); pthread_join(rk[g].th, NULL); }
        for (int g = 0; g < n_dev; ++g) uvg_destroy(rk[g].ctx);
    } else uvg_destroy(gpu);
    free(mg_nonn);
    for (int k = 0; k < 2; ++k) { uvg_host_free_pinned(pinned_text[k]); free(batch_names[k]); }
    for (int64_t i = 0; i < n_acc_total; ++i) free(rname[i]);
    for (int i = 0; i < q.n; ++i) free(q.name[i]);
    free(rname); free(rec_ok); free(push_ok); free(accepted); free(best); free(outname); free(q.name); free(q.text); free(qset.slot);
    return EXIT_SUCCESS;
}
