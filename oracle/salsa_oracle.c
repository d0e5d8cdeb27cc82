/*
 * salsa_oracle.c -- CPU restatement of SALSA's Hi-C link-scoring path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in salsa_b200/ (the product) may import,
 * link or execute this file.  Allowed users: tests/, __graft_entry__.smoke()
 * and the cpu_baseline / --impl reference legs of bench.py.
 *
 * PARITY STATUS
 *   - make_links / RE_sites / fast_scaled_scores: the reference is Python-2
 *     only (no python2 and no lib2to3 in this image) and its tree holds no
 *     tests or golden vectors.  PINNED against vectors produced by the
 *     reference's OWN scripts, executed from /root/reference by
 *     tests/golden/ref_py2emu.py, which restores the few Python-2 semantics
 *     they rely on (print statements, long, iteritems, int/int floor
 *     division, 12-digit str(float)) mechanically and leaves the program
 *     logic to the reference's source text: tests/golden/links_ref.json
 *     (+ make_golden.py, the generating script).  Also: the two sample lines
 *     in fast_scaled_scores.py:39-40, hand-derived known-answer tests
 *     (tests/test_oracle.py) and the independent line-by-line py3 twin
 *     (oracle/py_twin.py).  What the emulation cannot restore is Python-2
 *     dict order, so link files compare as sorted lines.
 *   - coverage (break_contigs_start / break_contigs): pinned against the
 *     UNMODIFIED reference sources compiled into oracle/_ref/ (oracle/Makefile)
 *     and against fixtures generated with those binaries (tests/golden/).
 *
 * Every function cites the reference file:line it restates
 * (paths relative to the marbl/SALSA tree).
 */
#define _GNU_SOURCE
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <errno.h>
#include <pthread.h>

/* ------------------------------------------------------------------ */
/* error reporting                                                      */
/* ------------------------------------------------------------------ */
static __thread char g_err[256];
const char *oracle_last_error(void) { return g_err; }
#define FAIL(code, ...) do { snprintf(g_err, sizeof g_err, __VA_ARGS__); return (code); } while (0)

/* error codes mirror the Python exceptions the reference would raise */
enum {
    ORC_OK = 0,
    ORC_INDEX_ERROR = -1,   /* IndexError: too few fields (make_links.py:71,76-77) */
    ORC_VALUE_ERROR = -2,   /* ValueError: long()/float()/int() of bad text      */
    ORC_KEY_ERROR = -3,     /* KeyError: unknown contig (make_links.py:101-113)  */
    ORC_ZERO_DIV = -4,      /* ZeroDivisionError (make_links.py:106-107)         */
    ORC_NOMEM = -5,
    ORC_BAD_INPUT = -6,
    ORC_EXIT1 = -7          /* sys.exit(1) paths of RE_sites.py                  */
};

/* ------------------------------------------------------------------ */
/* Python-2 str(float): '%.12g', plus '.0' when the text looks integral */
/* (float_str -> PyOS_double_to_string(x,'g',12,Py_DTSF_ADD_DOT_0)).    */
/* Evidence of 12 digits: fast_scaled_scores.py:39-40.                  */
/* ------------------------------------------------------------------ */
int oracle_py2_str_float(double x, char *buf, size_t cap)
{
    if (isnan(x)) return snprintf(buf, cap, "nan");
    if (isinf(x)) return snprintf(buf, cap, x < 0 ? "-inf" : "inf");
    int n = snprintf(buf, cap, "%.12g", x);
    int plain = 1;
    for (int i = 0; i < n; i++)
        if (buf[i] == '.' || buf[i] == 'e' || buf[i] == 'E') { plain = 0; break; }
    if (plain && (size_t)n + 2 < cap) { buf[n++] = '.'; buf[n++] = '0'; buf[n] = 0; }
    return n;
}

/* float(str(x)): the double the next program sees after the 12-digit text
 * round trip (make_links.py:169 -> fast_scaled_scores.py:48,71). */
double oracle_round12(double x)
{
    char b[64];
    oracle_py2_str_float(x, b, sizeof b);
    return strtod(b, NULL);
}

/* ------------------------------------------------------------------ */
/* str.split() with no argument: runs of ' ' \t \n \r \v \f separate    */
/* ------------------------------------------------------------------ */
typedef struct { const char *p; size_t n; } tok_t;

static inline int is_pyspace(unsigned char c)
{
    return c == ' ' || (c >= 9 && c <= 13);
}

static int split_ws(const char *s, size_t n, tok_t *t, int maxtok)
{
    int k = 0;
    size_t i = 0;
    while (i < n) {
        while (i < n && is_pyspace((unsigned char)s[i])) i++;
        if (i >= n) break;
        size_t j = i;
        while (j < n && !is_pyspace((unsigned char)s[j])) j++;
        if (k < maxtok) { t[k].p = s + i; t[k].n = j - i; }
        k++;
        i = j;
    }
    return k;
}

static inline int tok_eq(tok_t a, tok_t b)
{
    return a.n == b.n && memcmp(a.p, b.p, a.n) == 0;
}

/* Python-2 byte-string '<' (make_links.py:90): memcmp, shorter prefix first */
static inline int tok_lt(tok_t a, tok_t b)
{
    size_t m = a.n < b.n ? a.n : b.n;
    int c = memcmp(a.p, b.p, m);
    if (c) return c < 0;
    return a.n < b.n;
}

/* long(text): strict [+-]?[0-9]+ (the only forms a BED carries); anything
 * else is reported as ValueError.  Values beyond int64 are rejected too. */
static int parse_long(tok_t t, int64_t *out)
{
    size_t i = 0;
    int neg = 0;
    if (t.n == 0) return -1;
    if (t.p[0] == '+' || t.p[0] == '-') { neg = t.p[0] == '-'; i = 1; }
    if (i >= t.n || t.n - i > 18) return -1;
    int64_t v = 0;
    for (; i < t.n; i++) {
        unsigned d = (unsigned char)t.p[i] - '0';
        if (d > 9) return -1;
        v = v * 10 + d;
    }
    *out = neg ? -v : v;
    return 0;
}

/* ------------------------------------------------------------------ */
/* tiny string -> int map (FNV-1a, open addressing)                     */
/* ------------------------------------------------------------------ */
typedef struct { const char *p; uint32_t n; int32_t v; } sm_slot;
typedef struct { sm_slot *s; size_t cap; } strmap;

static uint64_t fnv1a(const char *p, size_t n)
{
    uint64_t h = 1469598103934665603ull;
    for (size_t i = 0; i < n; i++) { h ^= (unsigned char)p[i]; h *= 1099511628211ull; }
    return h;
}
static int sm_init(strmap *m, size_t n)
{
    size_t cap = 16;
    while (cap < 2 * n + 2) cap <<= 1;
    m->s = calloc(cap, sizeof(sm_slot));
    m->cap = cap;
    return m->s ? 0 : -1;
}
static void sm_free(strmap *m) { free(m->s); m->s = NULL; }
static int32_t sm_get(const strmap *m, const char *p, size_t n)
{
    size_t i = fnv1a(p, n) & (m->cap - 1);
    while (m->s[i].p) {
        if (m->s[i].n == n && memcmp(m->s[i].p, p, n) == 0) return m->s[i].v;
        i = (i + 1) & (m->cap - 1);
    }
    return -1;
}
/* insert or overwrite (dict assignment semantics: later wins) */
static void sm_put(strmap *m, const char *p, size_t n, int32_t v)
{
    size_t i = fnv1a(p, n) & (m->cap - 1);
    while (m->s[i].p) {
        if (m->s[i].n == n && memcmp(m->s[i].p, p, n) == 0) { m->s[i].v = v; return; }
        i = (i + 1) & (m->cap - 1);
    }
    m->s[i].p = p; m->s[i].n = (uint32_t)n; m->s[i].v = v;
}

/* u64 -> i64 counter map */
typedef struct { uint64_t k; int64_t c; } cm_slot;
typedef struct { cm_slot *s; size_t cap, used; } cntmap;
#define CM_EMPTY 0xffffffffffffffffull
static int cm_init(cntmap *m, size_t cap0)
{
    size_t cap = 1024;
    while (cap < cap0) cap <<= 1;
    m->s = malloc(cap * sizeof(cm_slot));
    if (!m->s) return -1;
    for (size_t i = 0; i < cap; i++) { m->s[i].k = CM_EMPTY; m->s[i].c = 0; }
    m->cap = cap; m->used = 0;
    return 0;
}
static inline uint64_t mix64(uint64_t x)
{
    x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
    return x;
}
static int cm_add(cntmap *m, uint64_t k, int64_t d);
static int cm_grow(cntmap *m)
{
    cntmap n;
    if (cm_init(&n, m->cap * 2)) return -1;
    for (size_t i = 0; i < m->cap; i++)
        if (m->s[i].k != CM_EMPTY) cm_add(&n, m->s[i].k, m->s[i].c);
    free(m->s);
    *m = n;
    return 0;
}
static int cm_add(cntmap *m, uint64_t k, int64_t d)
{
    if ((m->used + 1) * 2 > m->cap && cm_grow(m)) return -1;
    size_t i = mix64(k) & (m->cap - 1);
    while (m->s[i].k != CM_EMPTY && m->s[i].k != k) i = (i + 1) & (m->cap - 1);
    if (m->s[i].k == CM_EMPTY) { m->s[i].k = k; m->used++; }
    m->s[i].c += d;
    return 0;
}

/* ================================================================== */
/* make_links.py                                                        */
/* ================================================================== */
/*
 * Contig table handed in by the caller, in ANY order:
 *   names_blob/name_off[ncontig+1] : names from scaffold_length_iteration_N
 *   length[i]   : float(attrs[1])               make_links.py:37-41
 *   has_re[i], re_left[i], re_right[i] : ints   make_links.py:43-48
 * avoid pairs are indices into that table (both orders are implied,
 * make_links.py:29-35).  A contig that appears in links_avoid but not in the
 * length table cannot be expressed by index; such pairs can never matter for
 * counting (a counted pair needs both lengths) but they DO matter for the
 * stale-prev rule, so avoid pairs are given by NAME as well: avoid_blob/off.
 *
 * Link key = ((a*ncontig + b) << 2) | ends, a/b table indices of first/second
 * (first < second as byte strings, make_links.py:90-96), ends bit1 = first is
 * 'E', bit0 = second is 'E'.
 *
 * The scan [lo,hi) variant evaluates exactly the adjacent-line pairs whose
 * SECOND line starts in [lo,hi) -- used for threaded CPU baselines; with an
 * avoid list only the whole-file call is valid (stale prev_line,
 * make_links.py:72-74).
 */
typedef struct {
    const char *bed; size_t nbytes, lo, hi;
    int32_t ncontig; const strmap *names;
    const double *length; const uint8_t *has_re; const int64_t *re_left, *re_right;
    const strmap *avoid; /* key "a$b" strings, or NULL */
    cntmap links;
    int64_t nlines, npairs_counted;
    int rc; int64_t err_line; char err[200];
} ml_job;

static size_t line_end(const char *b, size_t n, size_t p)
{
    const char *q = memchr(b + p, '\n', n - p);
    return q ? (size_t)(q - b) + 1 : n;   /* one past the '\n' (Python keeps it) */
}

static void ml_scan(ml_job *J)
{
    const char *b = J->bed;
    size_t n = J->nbytes;
    J->rc = ORC_OK; J->nlines = 0; J->npairs_counted = 0; J->err_line = -1;
    if (cm_init(&J->links, 1 << 16)) { J->rc = ORC_NOMEM; return; }
    if (n == 0) return;

    /* find the first line that starts at or after lo, and the line before it */
    size_t cur = J->lo;
    size_t prev = (size_t)-1;
    if (cur > 0) {
        /* snap cur forward to a line start */
        if (b[cur - 1] != '\n') cur = line_end(b, n, cur);
        /* previous line start */
        if (cur > 0) {
            size_t q = cur - 1;           /* the '\n' ending the previous line */
            while (q > 0 && b[q - 1] != '\n') q--;
            prev = q;
        }
    }
    /* make_links.py:60-68: the very first line only ever plays "previous" */
    if (prev == (size_t)-1) {
        if (cur >= n) return;
        prev = cur;
        cur = line_end(b, n, cur);
        J->nlines++;
    }
    size_t prev_end = line_end(b, n, prev);
    char keybuf[1024];

    while (cur < n && cur < J->hi) {
        size_t e = line_end(b, n, cur);
        J->nlines++;
        tok_t a[4], pa[4];
        int na = split_ws(b + cur, e - cur, a, 4);         /* make_links.py:62 */
        int np = split_ws(b + prev, prev_end - prev, pa, 4); /* make_links.py:70 */
        if (np < 1 || na < 1) {                              /* make_links.py:71 */
            J->rc = ORC_INDEX_ERROR; J->err_line = J->nlines;
            snprintf(J->err, sizeof J->err, "IndexError: blank BED line near byte %zu", cur);
            return;
        }
        if (!tok_eq(pa[0], a[0])) {                          /* make_links.py:71 */
            if (J->avoid) {                                  /* make_links.py:72-74 */
                size_t kl = pa[0].n + 1 + a[0].n;
                if (kl < sizeof keybuf) {
                    memcpy(keybuf, pa[0].p, pa[0].n);
                    keybuf[pa[0].n] = '$';
                    memcpy(keybuf + pa[0].n + 1, a[0].p, a[0].n);
                    if (sm_get(J->avoid, keybuf, kl) >= 0) {
                        /* `continue` WITHOUT prev_line = line */
                        cur = e;
                        continue;
                    }
                }
            }
            if (np < 4 || na < 4) {                          /* make_links.py:76-77 */
                J->rc = ORC_INDEX_ERROR; J->err_line = J->nlines;
                snprintf(J->err, sizeof J->err, "IndexError: BED line with < 4 fields near byte %zu", cur);
                return;
            }
            /* make_links.py:78  read.split('/')[0] */
            tok_t r1 = pa[3], r2 = a[3];
            const char *s1 = memchr(r1.p, '/', r1.n), *s2 = memchr(r2.p, '/', r2.n);
            if (s1) r1.n = (size_t)(s1 - r1.p);
            if (s2) r2.n = (size_t)(s2 - r2.p);
            if (tok_eq(r1, r2)) {
                int64_t ps, pe, cs, ce;
                if (parse_long(pa[1], &ps) || parse_long(pa[2], &pe) ||
                    parse_long(a[1], &cs) || parse_long(a[2], &ce)) { /* :81-82 */
                    J->rc = ORC_VALUE_ERROR; J->err_line = J->nlines;
                    snprintf(J->err, sizeof J->err, "ValueError: bad start/end near byte %zu", cur);
                    return;
                }
                double pos1 = (double)(ps + pe) / 2.0, pos2 = (double)(cs + ce) / 2.0;
                tok_t first = pa[0], second = a[0];
                if (!tok_lt(pa[0], a[0])) {                  /* make_links.py:90-96 */
                    first = a[0]; second = pa[0];
                    double t = pos1; pos1 = pos2; pos2 = t;
                }
                int32_t i1 = sm_get(J->names, first.p, first.n);
                int32_t i2 = sm_get(J->names, second.p, second.n);
                if (i1 < 0 || i2 < 0) {                      /* make_links.py:101-102 */
                    J->rc = ORC_KEY_ERROR; J->err_line = J->nlines;
                    snprintf(J->err, sizeof J->err, "KeyError: contig not in length table near byte %zu", cur);
                    return;
                }
                double len1 = J->length[i1], len2 = J->length[i2];
                double l1 = len1 / 2, l2 = len2 / 2;         /* make_links.py:104-105 */
                if (len1 == 0 || len2 == 0) {                /* make_links.py:106-107 */
                    J->rc = ORC_ZERO_DIV; J->err_line = J->nlines;
                    snprintf(J->err, sizeof J->err, "ZeroDivisionError: zero-length contig near byte %zu", cur);
                    return;
                }
                /* make_links.py:108-139: the four mutually exclusive tests */
                int e1 = -1, e2 = -1;
                if (pos1 <= l1) e1 = 0;
                if (pos1 > len1 - l1) e1 = (e1 == 0) ? 2 : 1;
                if (pos2 <= l2) e2 = 0;
                if (pos2 > len2 - l2) e2 = (e2 == 0) ? 2 : 1;
                if (e1 == 2 || e2 == 2 || e1 < 0 || e2 < 0) {
                    /* cannot happen for finite lengths: len - len/2 == len/2 */
                    J->rc = ORC_BAD_INPUT; J->err_line = J->nlines;
                    snprintf(J->err, sizeof J->err, "non-exclusive end classification near byte %zu", cur);
                    return;
                }
                if (!J->has_re[i1] || !J->has_re[i2]) {      /* make_links.py:113.. KeyError */
                    J->rc = ORC_KEY_ERROR; J->err_line = J->nlines;
                    snprintf(J->err, sizeof J->err, "KeyError: contig not in re_counts near byte %zu", cur);
                    return;
                }
                uint64_t key = (((uint64_t)i1 * (uint64_t)J->ncontig + (uint64_t)i2) << 2) | (uint64_t)(e1 * 2 + e2);
                if (cm_add(&J->links, key, 1)) { J->rc = ORC_NOMEM; return; }   /* :140-144 */
                J->npairs_counted++;
            }
        }
        prev = cur; prev_end = e;                            /* make_links.py:146 */
        cur = e;
    }
}

static void *ml_thread(void *p) { ml_scan((ml_job *)p); return NULL; }

typedef struct {
    int64_t n;            /* number of distinct link keys                      */
    int32_t *first, *second;   /* table indices                                 */
    uint8_t *ends;        /* bit1: first is E, bit0: second is E                */
    int64_t *count;
    double *norm;         /* make_links.py:113-137 */
    double *score;        /* make_links.py:165-168 (0.0 stands for the int 0)  */
    uint8_t *score_is_int0;
    int64_t nlines, npairs_counted;
} oracle_links;

void oracle_links_free(oracle_links *L)
{
    if (!L) return;
    free(L->first); free(L->second); free(L->ends); free(L->count);
    free(L->norm); free(L->score); free(L->score_is_int0);
    memset(L, 0, sizeof *L);
}

typedef struct { uint64_t k; int64_t c; } kc_t;
static const char *g_sort_blob; static const int64_t *g_sort_off; static int32_t g_sort_nc;
static int name_cmp(int32_t a, int32_t b)
{
    size_t na = (size_t)(g_sort_off[a + 1] - g_sort_off[a]), nb = (size_t)(g_sort_off[b + 1] - g_sort_off[b]);
    size_t m = na < nb ? na : nb;
    int c = memcmp(g_sort_blob + g_sort_off[a], g_sort_blob + g_sort_off[b], m);
    if (c) return c;
    return (na > nb) - (na < nb);
}
static int kc_cmp(const void *x, const void *y)
{
    const kc_t *a = x, *b = y;
    int32_t a1 = (int32_t)((a->k >> 2) / (uint64_t)g_sort_nc), a2 = (int32_t)((a->k >> 2) % (uint64_t)g_sort_nc);
    int32_t b1 = (int32_t)((b->k >> 2) / (uint64_t)g_sort_nc), b2 = (int32_t)((b->k >> 2) % (uint64_t)g_sort_nc);
    int c = name_cmp(a1, b1);
    if (c) return c;
    c = name_cmp(a2, b2);
    if (c) return c;
    return (int)(a->k & 3) - (int)(b->k & 3);
}

/*
 * oracle_make_links: restates make_links.py:11-172 on an in-memory BED.
 * Output rows are sorted by (first name, second name, ends) because the
 * reference's own order is Python-2 dict order (make_links.py:161) which no
 * consumer depends on (fast_scaled_scores re-loads into a dict; `sort` makes
 * the final order total).
 * nthreads > 1 splits the BED into byte ranges (only valid with navoid == 0).
 */
int oracle_make_links(const char *bed, size_t nbytes,
                      int32_t ncontig, const char *names_blob, const int64_t *name_off,
                      const double *length, const uint8_t *has_re,
                      const int64_t *re_left, const int64_t *re_right,
                      int32_t navoid, const char *avoid_blob, const int64_t *avoid_off,
                      int nthreads, oracle_links *out)
{
    memset(out, 0, sizeof *out);
    strmap names, avoid;
    if (sm_init(&names, (size_t)ncontig)) FAIL(ORC_NOMEM, "out of memory");
    for (int32_t i = 0; i < ncontig; i++)   /* later duplicate wins, like dict assignment */
        sm_put(&names, names_blob + name_off[i], (size_t)(name_off[i + 1] - name_off[i]), i);
    int have_avoid = navoid > 0;
    char *akeys = NULL;
    if (have_avoid) {
        /* make_links.py:29-35: both "a$b" and "b$a" */
        size_t tot = 0;
        for (int32_t i = 0; i < navoid; i++)
            tot += (size_t)(avoid_off[2 * i + 2] - avoid_off[2 * i]) + 1;
        akeys = malloc(2 * tot + 16);
        if (!akeys || sm_init(&avoid, (size_t)navoid * 2)) FAIL(ORC_NOMEM, "out of memory");
        char *w = akeys;
        for (int32_t i = 0; i < navoid; i++) {
            const char *a = avoid_blob + avoid_off[2 * i]; size_t an = (size_t)(avoid_off[2 * i + 1] - avoid_off[2 * i]);
            const char *b2 = avoid_blob + avoid_off[2 * i + 1]; size_t bn = (size_t)(avoid_off[2 * i + 2] - avoid_off[2 * i + 1]);
            char *k1 = w; memcpy(w, a, an); w[an] = '$'; memcpy(w + an + 1, b2, bn); w += an + 1 + bn;
            sm_put(&avoid, k1, an + 1 + bn, 1);
            char *k2 = w; memcpy(w, b2, bn); w[bn] = '$'; memcpy(w + bn + 1, a, an); w += an + 1 + bn;
            sm_put(&avoid, k2, an + 1 + bn, 1);
        }
        nthreads = 1;
    }
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    ml_job *jobs = calloc((size_t)nthreads, sizeof(ml_job));
    pthread_t *th = calloc((size_t)nthreads, sizeof(pthread_t));
    for (int t = 0; t < nthreads; t++) {
        ml_job *J = &jobs[t];
        J->bed = bed; J->nbytes = nbytes;
        J->lo = nbytes / (size_t)nthreads * (size_t)t;
        J->hi = (t == nthreads - 1) ? nbytes : nbytes / (size_t)nthreads * (size_t)(t + 1);
        /* a range owns the lines that START in it */
        J->ncontig = ncontig; J->names = &names; J->length = length; J->has_re = has_re;
        J->re_left = re_left; J->re_right = re_right; J->avoid = have_avoid ? &avoid : NULL;
    }
    /* range t owns pairs whose second line starts in [lo,hi): ml_scan snaps lo
     * forward to a line start and stops at the first line starting >= hi. */
    if (nthreads == 1) ml_scan(&jobs[0]);
    else {
        for (int t = 0; t < nthreads; t++) pthread_create(&th[t], NULL, ml_thread, &jobs[t]);
        for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
    }
    int rc = ORC_OK;
    cntmap all;
    cm_init(&all, 1 << 16);
    for (int t = 0; t < nthreads; t++) {
        ml_job *J = &jobs[t];
        if (J->rc != ORC_OK && rc == ORC_OK) { rc = J->rc; snprintf(g_err, sizeof g_err, "%s", J->err); }
        out->nlines += J->nlines; out->npairs_counted += J->npairs_counted;
        if (J->links.s) {
            for (size_t i = 0; i < J->links.cap; i++)
                if (J->links.s[i].k != CM_EMPTY) cm_add(&all, J->links.s[i].k, J->links.s[i].c);
            free(J->links.s);
        }
    }
    /* threads other than 0 counted their "previous" line too */
    free(jobs); free(th);
    if (rc == ORC_OK) {
        size_t n = all.used;
        kc_t *v = malloc((n + 1) * sizeof(kc_t));
        size_t j = 0;
        for (size_t i = 0; i < all.cap; i++)
            if (all.s[i].k != CM_EMPTY) { v[j].k = all.s[i].k; v[j].c = all.s[i].c; j++; }
        g_sort_blob = names_blob; g_sort_off = name_off; g_sort_nc = ncontig;
        qsort(v, n, sizeof(kc_t), kc_cmp);
        out->n = (int64_t)n;
        out->first = malloc((n + 1) * sizeof(int32_t)); out->second = malloc((n + 1) * sizeof(int32_t));
        out->ends = malloc(n + 1); out->count = malloc((n + 1) * sizeof(int64_t));
        out->norm = malloc((n + 1) * sizeof(double)); out->score = malloc((n + 1) * sizeof(double));
        out->score_is_int0 = malloc(n + 1);
        for (size_t i = 0; i < n; i++) {
            int32_t a = (int32_t)((v[i].k >> 2) / (uint64_t)ncontig), b2 = (int32_t)((v[i].k >> 2) % (uint64_t)ncontig);
            int e = (int)(v[i].k & 3);
            out->first[i] = a; out->second[i] = b2; out->ends[i] = (uint8_t)e; out->count[i] = v[i].c;
            /* make_links.py:104-107,113-137: r1 = (len/2)/len == 0.5 exactly */
            double len1 = length[a], len2 = length[b2];
            double r1 = (len1 / 2) / len1, r2 = (len2 / 2) / len2;
            double x = (double)((e & 2) ? re_right[a] : re_left[a]);
            double y = (double)((e & 1) ? re_right[b2] : re_left[b2]);
            double norm = x * r1 + y * r2;
            out->norm[i] = norm;
            if (norm == 0) { out->score[i] = 0.0; out->score_is_int0[i] = 1; }       /* :165-166 */
            else { out->score[i] = (double)v[i].c * 1.0 / norm; out->score_is_int0[i] = 0; } /* :168 */
        }
        free(v);
    }
    free(all.s);
    sm_free(&names);
    if (have_avoid) { sm_free(&avoid); free(akeys); }
    return rc;
}

/* Text of one contig_links_iteration_N line (make_links.py:169). */
int oracle_format_link_line(const char *first, int first_n, const char *second, int second_n,
                            int ends, double score, int score_is_int0, int64_t count,
                            char *buf, size_t cap)
{
    char s[64];
    if (score_is_int0) snprintf(s, sizeof s, "0");
    else oracle_py2_str_float(score, s, sizeof s);
    return snprintf(buf, cap, "%.*s:%c\t%.*s:%c\t%s\t%lld\n", first_n, first, (ends & 2) ? 'E' : 'B',
                    second_n, second, (ends & 1) ? 'E' : 'B', s, (long long)count);
}

/* ================================================================== */
/* fast_scaled_scores.py                                                */
/* ================================================================== */
/*
 * Edges are given as node indices (a node = one contig end, e.g. "ctg:B"),
 * weights are float(text) of column 3 and links int(column 4)
 * (fast_scaled_scores.py:35-60).  The graph must be simple (make_links never
 * emits a key twice); duplicate edges are rejected rather than emulating
 * networkx's attribute overwrite.
 *
 * Outputs per edge, in input order (fast_scaled_scores.py:64-86):
 *   best_alt[i]     : the float printed in column 4, or 1.0 with
 *   best_alt_is_int1[i] = 1 when the reference prints the int literal `1`
 *   ratio[i]        : w / bestAlt
 */
int oracle_scaled_scores(int64_t nedges, int32_t nnodes, const int32_t *u, const int32_t *v,
                         const double *w, double *best_alt, uint8_t *best_alt_is_int1, double *ratio)
{
    /* node2weight: fast_scaled_scores.py:47-57 (max via >=) */
    double *nw = malloc(((size_t)nnodes + 1) * sizeof(double));
    uint8_t *seen = calloc((size_t)nnodes + 1, 1);
    int64_t *deg = calloc((size_t)nnodes + 2, sizeof(int64_t));
    if (!nw || !seen || !deg) FAIL(ORC_NOMEM, "out of memory");
    for (int64_t i = 0; i < nedges; i++) {
        if (u[i] < 0 || u[i] >= nnodes || v[i] < 0 || v[i] >= nnodes || u[i] == v[i])
            FAIL(ORC_BAD_INPUT, "bad edge %lld", (long long)i);
        int32_t e[2] = { u[i], v[i] };
        for (int k = 0; k < 2; k++) {
            if (!seen[e[k]]) { seen[e[k]] = 1; nw[e[k]] = w[i]; }
            else if (w[i] >= nw[e[k]]) nw[e[k]] = w[i];
        }
        deg[u[i] + 1]++; deg[v[i] + 1]++;
    }
    /* adjacency (CSR) standing in for nx.Graph: fast_scaled_scores.py:59 */
    for (int32_t k = 0; k < nnodes; k++) deg[k + 1] += deg[k];
    int64_t *fill = malloc(((size_t)nnodes + 1) * sizeof(int64_t));
    int32_t *adj = malloc((size_t)(2 * nedges + 1) * sizeof(int32_t));
    double *adjw = malloc((size_t)(2 * nedges + 1) * sizeof(double));
    if (!fill || !adj || !adjw) FAIL(ORC_NOMEM, "out of memory");
    memcpy(fill, deg, (size_t)nnodes * sizeof(int64_t));
    for (int64_t i = 0; i < nedges; i++) {
        adj[fill[u[i]]] = v[i]; adjw[fill[u[i]]++] = w[i];
        adj[fill[v[i]]] = u[i]; adjw[fill[v[i]]++] = w[i];
    }
    /* duplicate-edge check on small degree lists only would be O(deg^2); sort-free
     * check: for each node, neighbours must be distinct */
    {
        int32_t *mark = malloc(((size_t)nnodes + 1) * sizeof(int32_t));
        for (int32_t k = 0; k < nnodes; k++) mark[k] = -1;
        for (int32_t k = 0; k < nnodes; k++)
            for (int64_t j = deg[k]; j < deg[k + 1]; j++) {
                if (mark[adj[j]] == k) { free(mark); FAIL(ORC_BAD_INPUT, "duplicate edge at node %d", k); }
                mark[adj[j]] = k;
            }
        free(mark);
    }
    for (int64_t i = 0; i < nedges; i++) {
        double max_u = nw[u[i]], max_v = nw[v[i]];              /* :72-77 */
        double b = max_u > max_v ? max_u : max_v;               /* :79 */
        int is_int = 0;
        if (b == w[i]) {                                        /* :80-82 */
            /* get_max_weight = max(get_max_incident(u,v), get_max_incident(v,u)),
             * fast_scaled_scores.py:13-33 (live branch :22-25, G.neighbors),
             * both starting from the int 0 */
            double m = 0; int m_is_int = 1;
            int32_t ends2[2][2] = { { u[i], v[i] }, { v[i], u[i] } };
            for (int k = 0; k < 2; k++) {
                int32_t s = ends2[k][0], t = ends2[k][1];
                for (int64_t j = deg[s]; j < deg[s + 1]; j++)
                    if (adj[j] != t && adjw[j] > m) { m = adjw[j]; m_is_int = 0; }
            }
            b = m; is_int = m_is_int;
        }
        if (b == 0) { b = 1; is_int = 1; }                      /* :83-84 */
        best_alt[i] = b; best_alt_is_int1[i] = (uint8_t)is_int;
        ratio[i] = w[i] / b;                                    /* :86 */
    }
    free(nw); free(seen); free(deg); free(fill); free(adj); free(adjw);
    return ORC_OK;
}

/* ================================================================== */
/* RE_sites.py                                                          */
/* ================================================================== */
/*
 * oracle_re_sites: restates RE_sites.py:6-88 on an in-memory FASTA.
 * Results are returned in first-appearance order of record names (the
 * reference prints in Python-2 dict order, RE_sites.py:61, which no consumer
 * depends on); a repeated name keeps its first position and its LAST sequence
 * (RE_sites.py:19 re-assigns fa[name] = []).
 * enzyme: the -e string.  "DNASE" -> left = right = len/2 (RE_sites.py:39-43).
 */
typedef struct {
    int64_t n;
    char *names_blob; int64_t *name_off;
    int64_t *left, *right, *seqlen;
} oracle_re;

void oracle_re_free(oracle_re *R)
{
    if (!R) return;
    free(R->names_blob); free(R->name_off); free(R->left); free(R->right); free(R->seqlen);
    memset(R, 0, sizeof *R);
}

/* non-overlapping leftmost literal matches == re.finditer on a literal
 * pattern (RE_sites.py:66) */
static void count_pattern(const char *seq, int64_t n, const char *pat, int64_t m, int64_t half,
                          int64_t *left, int64_t *right)
{
    if (m <= 0 || m > n) return;
    int64_t i = 0;
    while (i + m <= n) {
        const char *q = memmem(seq + i, (size_t)(n - i), pat, (size_t)m);
        if (!q) break;
        int64_t p = q - seq;
        if (p < half) (*left)++; else (*right)++;          /* RE_sites.py:69-74 */
        i = p + m;
    }
}

int oracle_re_sites(const char *fasta, size_t nbytes, const char *enzyme, oracle_re *out)
{
    memset(out, 0, sizeof *out);
    /* ---- parse_fasta, RE_sites.py:6-28 ---- */
    size_t cap_rec = 1024, nrec = 0;
    char **seq = malloc(cap_rec * sizeof(char *));
    int64_t *slen = malloc(cap_rec * sizeof(int64_t));
    int64_t *scap = malloc(cap_rec * sizeof(int64_t));
    const char **nm = malloc(cap_rec * sizeof(char *));
    int64_t *nml = malloc(cap_rec * sizeof(int64_t));
    strmap idx; sm_init(&idx, 1 << 20);
    size_t idx_n = 0;
    int64_t cur = -1;
    size_t p = 0;
    while (p < nbytes) {
        size_t e = line_end(fasta, nbytes, p);
        const char *ln = fasta + p; size_t ll = e - p;
        if (ln[0] == '>') {
            if (memchr(ln, ':', ll)) FAIL(ORC_EXIT1, "Error, your fasta records contain a ':'"); /* :13-15 */
            /* long_name = ln[1:].rstrip(); short = long_name.split()[0]  :18-19 */
            tok_t t[1];
            int k = split_ws(ln + 1, ll - 1, t, 1);
            if (k < 1) FAIL(ORC_INDEX_ERROR, "IndexError: empty fasta header");
            int32_t id = sm_get(&idx, t[0].p, t[0].n);
            if (id < 0) {
                if (nrec == cap_rec) {
                    cap_rec *= 2;
                    seq = realloc(seq, cap_rec * sizeof(char *)); slen = realloc(slen, cap_rec * sizeof(int64_t));
                    scap = realloc(scap, cap_rec * sizeof(int64_t)); nm = realloc(nm, cap_rec * sizeof(char *));
                    nml = realloc(nml, cap_rec * sizeof(int64_t));
                }
                if ((idx_n + 1) * 2 > idx.cap) {   /* grow the index */
                    strmap n2; sm_init(&n2, idx.cap);
                    for (size_t i = 0; i < idx.cap; i++)
                        if (idx.s[i].p) sm_put(&n2, idx.s[i].p, idx.s[i].n, idx.s[i].v);
                    sm_free(&idx); idx = n2;
                }
                id = (int32_t)nrec++;
                seq[id] = NULL; scap[id] = 0; nm[id] = t[0].p; nml[id] = (int64_t)t[0].n;
                sm_put(&idx, t[0].p, t[0].n, id); idx_n++;
            }
            slen[id] = 0;                                   /* fa[name] = [] */
            cur = id;
        } else {
            if (cur < 0) FAIL(ORC_KEY_ERROR, "KeyError: sequence line before any header"); /* :22 */
            /* ln.rstrip(): strip trailing whitespace only */
            size_t r = ll;
            while (r > 0 && is_pyspace((unsigned char)ln[r - 1])) r--;
            if (slen[cur] + (int64_t)r > scap[cur]) {
                int64_t nc = scap[cur] ? scap[cur] * 2 : 1 << 16;
                while (nc < slen[cur] + (int64_t)r) nc *= 2;
                seq[cur] = realloc(seq[cur], (size_t)nc); scap[cur] = nc;
            }
            memcpy(seq[cur] + slen[cur], ln, r);
            slen[cur] += (int64_t)r;
        }
        p = e;
    }
    out->n = (int64_t)nrec;
    out->left = calloc(nrec + 1, sizeof(int64_t)); out->right = calloc(nrec + 1, sizeof(int64_t));
    out->seqlen = calloc(nrec + 1, sizeof(int64_t));
    out->name_off = calloc(nrec + 2, sizeof(int64_t));
    size_t tot = 0;
    for (size_t i = 0; i < nrec; i++) tot += (size_t)nml[i];
    out->names_blob = malloc(tot + 1);
    for (size_t i = 0; i < nrec; i++) {
        memcpy(out->names_blob + out->name_off[i], nm[i], (size_t)nml[i]);
        out->name_off[i + 1] = out->name_off[i] + nml[i];
        out->seqlen[i] = slen[i];
    }
    int rc = ORC_OK;
    if (strcmp(enzyme, "DNASE") == 0) {                     /* RE_sites.py:39-43 */
        for (size_t i = 0; i < nrec; i++) out->left[i] = out->right[i] = slen[i] / 2;
    } else {
        /* RE_sites.py:45-58 */
        char *ez = strdup(enzyme);
        { char *w = ez; for (char *r = ez; *r; r++) if (*r != ' ') *w++ = *r; *w = 0; }  /* replace(' ','') */
        char pats[256][64]; int npat = 0;
        char *s = ez;
        for (;;) {
            char *c = strchr(s, ',');
            size_t L = c ? (size_t)(c - s) : strlen(s);
            int ok = L > 0 && L < 64, hasN = 0;
            for (size_t i = 0; ok && i < L; i++) {
                if (!strchr("ACGTN", s[i])) ok = 0;
                if (s[i] == 'N') hasN = 1;
            }
            if (!ok || npat + 4 > 256) { rc = ORC_EXIT1; snprintf(g_err, sizeof g_err, "Error, enzyme should be restriction site sequence"); break; }
            if (hasN) {
                const char rep[4] = { 'G', 'A', 'T', 'C' };  /* RE_sites.py:53-56 */
                for (int r = 0; r < 4; r++) {
                    for (size_t i = 0; i < L; i++) pats[npat][i] = s[i] == 'N' ? rep[r] : s[i];
                    pats[npat][L] = 0; npat++;
                }
            } else { memcpy(pats[npat], s, L); pats[npat][L] = 0; npat++; }
            if (!c) break;
            s = c + 1;
        }
        if (rc == ORC_OK)
            for (size_t i = 0; i < nrec; i++) {                 /* RE_sites.py:61-88 */
                int64_t half = slen[i] / 2;                     /* py2 int / int */
                for (int k = 0; k < npat; k++)
                    count_pattern(seq[i], slen[i], pats[k], (int64_t)strlen(pats[k]), half,
                                  &out->left[i], &out->right[i]);
            }
        free(ez);
    }
    for (size_t i = 0; i < nrec; i++) free(seq[i]);
    free(seq); free(slen); free(scap); free(nm); free(nml); sm_free(&idx);
    return rc;
}

/* ================================================================== */
/* coverage: break_contigs_start.cpp / break_contigs.cpp                */
/* ================================================================== */
/* istringstream >> long on a token */
static int cov_tokens(const char *ln, size_t ll, tok_t *t, int64_t *start, int64_t *end)
{
    /* iss >> contig >> start >> end >> read  (break_contigs.cpp:212,
     * break_contigs_start.cpp:321): whitespace-separated; a numeric field that
     * does not start with a number fails the extraction. operator>> for long
     * stops at the first non-digit, so "12x" would leave "x" to be read as the
     * read name; generators never produce that, and it is rejected here. */
    int k = split_ws(ln, ll, t, 4);
    if (k < 4) return -1;
    if (parse_long(t[1], start) || parse_long(t[2], end)) return -1;
    return 0;
}

/* Kadane with the reference's tie rules over delta[i] = (cov[i] < cutoff) ? +5 : -5
 * on [lo,hi): break_contigs_start.cpp:432-449, break_contigs.cpp:317-335. */
static void kadane(const int32_t *cov, int64_t lo, int64_t hi, double cutoff,
                   int64_t init_start, int64_t init_end, int64_t *ostart, int64_t *oend)
{
    int max_so_far = -100000, max_ending_here = 0;
    int64_t start = init_start, end = init_end, s = lo;
    for (int64_t i = lo; i < hi; i++) {
        max_ending_here += ((double)cov[i] < cutoff) ? 5 : -5;
        if (max_so_far < max_ending_here) { max_so_far = max_ending_here; start = s; end = i; }
        if (max_ending_here < 0) { max_ending_here = 0; s = i + 1; }
    }
    *ostart = start; *oend = end;
}

/*
 * oracle_cov_start: break_contigs_start.cpp:271-463.
 * Contig table as in oracle_make_links (lengths as long, :286-294).
 * Output: for each contig that prints a line, its index and consensus
 * position (:456-460); order = contig index (reference: unordered_map order).
 */
int oracle_cov_start(const char *bed, size_t nbytes, int32_t ncontig, const char *names_blob,
                     const int64_t *name_off, const int64_t *length,
                     int32_t *out_contig, int64_t *out_pos, int64_t *out_n)
{
    strmap names;
    if (sm_init(&names, (size_t)ncontig)) FAIL(ORC_NOMEM, "out of memory");
    for (int32_t i = 0; i < ncontig; i++)
        sm_put(&names, names_blob + name_off[i], (size_t)(name_off[i + 1] - name_off[i]), i);
    int32_t **cov = calloc((size_t)ncontig + 1, sizeof(int32_t *));
    *out_n = 0;
    size_t p = 0;
    int first = 1;
    tok_t prev_read = { "", 0 };
    int32_t prev_contig = -2;   /* "" */
    int64_t prev_start = -1, prev_end = -1;
    while (p < nbytes) {
        size_t e = line_end(bed, nbytes, p);
        size_t ll = e - p; if (ll && bed[e - 1] == '\n') ll--;   /* getline drops '\n' */
        const char *ln = bed + p;
        p = e;
        if (first) {                                  /* :315-319 (an empty first line keeps it "first") */
            if (ll > 0) first = 0;
            continue;
        }
        tok_t t[4]; int64_t start, end;
        if (cov_tokens(ln, ll, t, &start, &end)) break;          /* :321-322 */
        int32_t c = sm_get(&names, t[0].p, t[0].n);
        int64_t clen = c >= 0 ? length[c] : 0;                   /* operator[] default 0 */
        /* unknown names all behave as length 0; they still need distinct
         * identities for prev_contig == contig, but that test is only reached
         * for contigs >= 1 Mb, which are known. */
        if (clen < 1000000) {                                    /* :325-332 */
            prev_contig = c >= 0 ? c : -1; prev_start = start; prev_end = end; prev_read = t[3];
            continue;
        }
        if (!cov[c]) cov[c] = calloc((size_t)clen + 10, sizeof(int32_t));   /* :333-337 */
        /* read.substr(0,len-2) == prev_read.substr(0,len-2) && prev_contig == contig  :339 */
        tok_t r = t[3], pr = prev_read;
        if (r.n >= 2) r.n -= 2;            /* size_t wrap for n<2 -> whole string */
        if (pr.n >= 2) pr.n -= 2;
        if (tok_eq(r, pr) && prev_contig == c) {
            int64_t ss, se;                                      /* :341-370 */
            if (prev_end <= end) { ss = prev_start; se = end; } else { ss = start; se = prev_end; }
            if (ss >= 0 && se < clen && se - ss <= (int64_t)(int)(clen / 10)) {  /* :371-373 */
                if (ss >= clen + 10 || se + 1 < 0) FAIL(ORC_BAD_INPUT, ".at() out of range");
                cov[c][ss] += 1; cov[c][se + 1] -= 1;            /* :375-376 */
            }
        }
        prev_contig = c; prev_start = start; prev_end = end; prev_read = t[3];
    }
    for (int32_t c = 0; c < ncontig; c++) {
        if (!cov[c]) continue;
        int64_t sz = length[c] + 10;
        int32_t *v = cov[c];
        double sum = v[0];
        for (int64_t i = 1; i < sz; i++) { v[i] += v[i - 1]; sum += v[i]; }   /* :397-401 */
        double average = sum / (double)sz;                       /* :410 */
        double possum = 0; int npos = 0;
        for (int div = 5; div <= 15; div++) {                    /* :412-454 */
            double cutoff = average / div;
            int64_t s0, e0;
            kadane(v, sz / 100, 9 * sz / 10, cutoff, sz / 100, sz / 100, &s0, &e0);
            if (s0 >= sz / 100 + 5000 && e0 <= 9 * sz / 10 - 5000) { possum += (double)((s0 + e0) / 2); npos++; }
        }
        if (npos > 0) {                                          /* :456-460 */
            out_contig[*out_n] = c; out_pos[*out_n] = (int64_t)(possum / npos); (*out_n)++;
        }
        free(v);
    }
    free(cov); sm_free(&names);
    return ORC_OK;
}

/*
 * oracle_cov_joins: break_contigs.cpp:162-358.
 * joins: CSR over scaffolds: join_off[ncontig+1], join_pos[] = every position
 * on that scaffold's breakpoints line (load_breakpoints filters, :84-87);
 * in_breakfile[c] = scaffold has a line in the breakpoints file.
 * Output: for each scaffold that has coverage AND a breakpoints line
 * (:261-267): out_scaf[k]; suspicious positions in CSR out_sus_off/out_sus_pos;
 * totals (:354-355).
 */
int oracle_cov_joins(const char *bed, size_t nbytes, int32_t ncontig, const char *names_blob,
                     const int64_t *name_off, const int64_t *length,
                     const uint8_t *in_breakfile, const int64_t *join_off, const int64_t *join_pos,
                     int32_t *out_scaf, int64_t *out_sus_off, int64_t *out_sus_pos, int64_t *out_nscaf,
                     int64_t *total_joins, int64_t *suspicious_joins)
{
    const int64_t W = 2000000;
    strmap names;
    if (sm_init(&names, (size_t)ncontig)) FAIL(ORC_NOMEM, "out of memory");
    for (int32_t i = 0; i < ncontig; i++)
        sm_put(&names, names_blob + name_off[i], (size_t)(name_off[i + 1] - name_off[i]), i);
    int32_t **cov = calloc((size_t)ncontig + 1, sizeof(int32_t *));
    size_t p = 0;
    int first = 1;
    tok_t prev_read = { "", 0 };
    int32_t prev_contig = -2;
    int64_t prev_start = -1, prev_end = -1;
    while (p < nbytes) {
        size_t e = line_end(bed, nbytes, p);
        size_t ll = e - p; if (ll && bed[e - 1] == '\n') ll--;
        const char *ln = bed + p;
        p = e;
        if (first) { if (ll > 0) first = 0; continue; }          /* :206-210 */
        tok_t t[4]; int64_t start, end;
        if (cov_tokens(ln, ll, t, &start, &end)) break;          /* :212-213 */
        int32_t c = sm_get(&names, t[0].p, t[0].n);
        int64_t clen = c >= 0 ? length[c] : 0;
        if (clen < 1000000) {                                    /* :218-225 */
            prev_contig = c >= 0 ? c : -1; prev_start = start; prev_end = end; prev_read = t[3];
            continue;
        }
        if (!cov[c]) cov[c] = calloc((size_t)clen + 10, sizeof(int32_t));   /* :226-230 */
        tok_t r = t[3], pr = prev_read;
        if (r.n >= 2) r.n -= 2;
        if (pr.n >= 2) pr.n -= 2;
        if (tok_eq(r, pr) && prev_contig == c && prev_end <= start) {        /* :232-235 */
            /* check_if_inside(contig, prev_start, end+1)  :144-160,238 */
            int inside = 0;
            if (in_breakfile[c])
                for (int64_t j = join_off[c]; j < join_off[c + 1]; j++) {
                    int64_t pos = join_pos[j];
                    if (!(pos - W >= 0 && pos + W <= clen)) continue;        /* :84-87 */
                    if (prev_start >= pos - W && end + 1 <= pos + W) { inside = 1; break; }
                }
            if (inside && end - prev_start <= (int64_t)(int)(clen / 10)) {   /* :240 */
                if (prev_start < 0 || prev_start >= clen + 10 || end + 1 < 0 || end + 1 >= clen + 10)
                    FAIL(ORC_BAD_INPUT, ".at() out of range");
                cov[c][prev_start] += 1; cov[c][end + 1] -= 1;               /* :242-243 */
            }
        }
        prev_contig = c; prev_start = start; prev_end = end; prev_read = t[3];
    }
    *out_nscaf = 0; *total_joins = 0; *suspicious_joins = 0;
    out_sus_off[0] = 0;
    int32_t *local = malloc((size_t)(2 * W) * sizeof(int32_t));
    for (int32_t c = 0; c < ncontig; c++) {
        if (!cov[c]) continue;
        if (!in_breakfile[c]) { free(cov[c]); continue; }        /* :264 */
        int64_t k = *out_nscaf;
        out_scaf[k] = c;
        int64_t nsus = out_sus_off[k];
        for (int64_t j = join_off[c]; j < join_off[c + 1]; j++) {
            int64_t pos = join_pos[j];
            if (!(pos - W >= 0 && pos + W <= length[c])) continue;
            int64_t start = pos - W, end = pos + W;
            /* :280-287: local prefix seeded with the RAW diff value at start-1
             * (start == 0 reads tmpcov[-1], UB in the reference; generators avoid
             * joins at exactly 2,000,000 -- here it is read as 0) */
            int64_t run = start > 0 ? cov[c][start - 1] : 0;
            double sum = 0;
            for (int64_t q = start; q < end; q++) { run += cov[c][q]; local[q - start] = (int32_t)run; sum += (double)(int32_t)run; }
            int64_t sz = end - start;
            double average = sum / (double)sz;                   /* :295 */
            int votes = 0;
            for (int div = 5; div <= 15; div++) {                /* :298-342 */
                double cutoff = average / div;
                int64_t s0, e0;
                kadane(local, sz / 3, 2 * sz / 3, cutoff, sz / 3, 2 * sz / 3, &s0, &e0);
                if (pos >= s0 + start && pos <= e0 + start) votes++;   /* :337-340 */
            }
            if (votes >= 8) { (*suspicious_joins)++; out_sus_pos[nsus++] = pos; }   /* :344-348 */
            (*total_joins)++;
        }
        out_sus_off[k + 1] = nsus;
        (*out_nscaf)++;
        free(cov[c]);
    }
    free(local); free(cov); sm_free(&names);
    return ORC_OK;
}

/*
 * oracle_cov_analyse: the per-segment arithmetic shared by both coverage
 * programs, exposed so tests can hammer the tie rules on small arrays:
 * inclusive prefix (seeded), exact total, and the 11 Kadane windows on
 * [lo,hi) -- the same kadane() the two restatements above use
 * (break_contigs_start.cpp:397-449, break_contigs.cpp:283-335).
 */
int oracle_cov_analyse(const int32_t *diff, int64_t len, int32_t seed, int64_t lo, int64_t hi,
                       int64_t *out_sum, int64_t *out_start, int64_t *out_end)
{
    int32_t *cov = malloc((size_t)(len + 1) * sizeof(int32_t));
    if (!cov) FAIL(ORC_NOMEM, "out of memory");
    int32_t run = seed;
    double sum = 0;
    for (int64_t i = 0; i < len; i++) { run += diff[i]; cov[i] = run; sum += (double)run; }
    *out_sum = (int64_t)sum;
    double average = sum / (double)len;
    for (int div = 5; div <= 15; div++)
        kadane(cov, lo, hi, average / div, lo, lo, &out_start[div - 5], &out_end[div - 5]);
    free(cov);
    return ORC_OK;
}
