/* gzb_zip.c -- reference-facing host driver: what parallel_zip() does around the codec.
 *
 * Mirrors `off_t parallel_zip(int pack_level)` (reference src/parallel.c:699-834) and its
 * write_thread (:490-558) with the globals passed in gzb_zip_opts.  The structure is the
 * reference's own -- one reader, N compressors, one in-order writer, joined by bounded queues
 * (src/parallel.c:106-143, 270-340) -- with the GPU engine where the compress threads stood:
 *
 *   reader   (caller's thread)  reads ifd in batches of BATCH bytes into pinned slots, keeps the
 *                               32 KiB that precede a batch right in front of it (:767-774), and
 *                               looks one batch ahead to know which one is final (:764)
 *   workers  (2 per GPU)        each uses a context of the process-wide pool (its own streams): batch i goes to worker
 *                               i mod W, which calls gzb_deflate_host on it.  Two workers per GPU
 *                               put the copies of one batch under the kernels of the other.
 *   writer   (one thread)       header, fragments in batch order, CRCs folded with
 *                               crc32_combine (:544), trailer (:496-507, :538, :555-556)
 *
 * Between two regular files none of this is serial (zip_direct below): every worker reads its own batches and places its
 * fragments at their final offsets (reserved with fallocate, filled through a shared mapping), the copies spread over one
 * I/O thread per CPU.
 *
 * Batches are independent given their dictionary, so with n_gpus > 1 consecutive batches simply
 * go to different GPUs: contiguous shards on chunk boundaries, no collective.
 * With opts->members every batch is closed as a gzip member of its own (no dictionary across batches): the
 * "per-GPU members concatenated" layout, which the reference's own decoder reads only the first member of.
 * Differences by design: errors are reported, empty input yields the valid
 * 20-byte member instead of a deadlock (F4), and FNAME is the caller's choice (F5).
 */
#define _GNU_SOURCE
#include "gzb200.h"
#include "../csrc/gzb_internal.h"
#include <errno.h>
#include <fcntl.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#define MAX_GPUS 16
#define MAX_WORKERS (2 * MAX_GPUS)
#define MAX_SLOTS (MAX_WORKERS + 2)

static double now_s(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return (double)t.tv_sec + 1e-9 * (double)t.tv_nsec; }

static uint64_t batch_bytes(void)               /* multiple of both chunk sizes */
{
    const char *e = getenv("GZB_ZIP_BATCH_MB");
    uint64_t mb = e ? (uint64_t)atoi(e) : 128u;   /* slots are page-locked (~0.3 s per GiB): 4 slots of 2 x 128 MiB */
    return (mb ? mb : 1u) << 20;
}

static ssize_t readn(int fd, uint8_t *buf, size_t len)
{
    size_t got = 0;
    while (got < len) {
        ssize_t r = read(fd, buf + got, len - got);
        if (r < 0) { if (errno == EINTR) continue; return -1; }
        if (r == 0) break;
        got += (size_t)r;
    }
    return (ssize_t)got;
}
static int writen(int fd, const uint8_t *buf, size_t len)
{
    while (len) {
        ssize_t r = write(fd, buf, len);
        if (r < 0) { if (errno == EINTR) continue; return -1; }
        buf += r; len -= (size_t)r;
    }
    return 0;
}

/* A regular file is read by g_readers threads with pread (one read() stream from the page cache is ~5 GB/s, a quarter
 * of what one GPU takes): 8 per GPU, at most 32 and at most the CPUs that are online; anything else -- a pipe, a tty --
 * sequentially, like the reference's readn (:66-84).  A regular output file is written the same way (pwrite). */
#define READERS 32
static int g_readers = 8;
typedef struct { int fd; uint8_t *buf; size_t len; off_t off; ssize_t got; } pread_job;
static void *pread_worker(void *arg)
{
    pread_job *j = (pread_job *)arg;
    size_t got = 0;
    while (got < j->len) {
        ssize_t r = pread(j->fd, j->buf + got, j->len - got, j->off + (off_t)got);
        if (r < 0) { if (errno == EINTR) continue; j->got = -1; return NULL; }
        if (r == 0) break;
        got += (size_t)r;
    }
    j->got = (ssize_t)got;
    return NULL;
}
static ssize_t read_batch(int fd, uint8_t *buf, size_t len, int regular, off_t *pos)
{
    if (!regular) return readn(fd, buf, len);
    pread_job j[READERS]; pthread_t th[READERS]; int started[READERS];
    const int nr = g_readers;
    const size_t part = ((len / (size_t)nr) + 4095u) & ~(size_t)4095u;
    for (int k = 0; k < nr; k++) {
        const size_t lo = (size_t)k * part < len ? (size_t)k * part : len;
        const size_t hi = lo + part < len ? lo + part : len;
        j[k].fd = fd; j[k].buf = buf + lo; j[k].len = hi - lo; j[k].off = *pos + (off_t)lo; j[k].got = 0;
        started[k] = k > 0 && j[k].len && pthread_create(&th[k], NULL, pread_worker, &j[k]) == 0;
        if (k > 0 && !started[k]) pread_worker(&j[k]);
    }
    pread_worker(&j[0]);
    ssize_t total = 0; int shorted = 0;
    for (int k = 0; k < nr; k++) {
        if (started[k]) pthread_join(th[k], NULL);
        if (j[k].got < 0) return -1;
        if (!shorted) total += j[k].got;            /* bytes after a short part belong to a file that grew: ignore */
        if ((size_t)j[k].got < j[k].len) shorted = 1;
    }
    *pos += (off_t)total;
    return total;
}

typedef struct { int fd; const uint8_t *buf; size_t len; off_t off; int rc; } pwrite_job;
static void *pwrite_worker(void *arg)
{
    pwrite_job *j = (pwrite_job *)arg;
    size_t done = 0;
    while (done < j->len) {
        ssize_t r = pwrite(j->fd, j->buf + done, j->len - done, j->off + (off_t)done);
        if (r < 0) { if (errno == EINTR) continue; j->rc = -1; return NULL; }
        done += (size_t)r;
    }
    j->rc = 0;
    return NULL;
}
/* len bytes at *pos of a regular file, in parts of at least 4 MiB on up to g_readers threads; *pos advances */
static int write_batch(int fd, const uint8_t *buf, size_t len, int regular, off_t *pos)
{
    if (!regular) return writen(fd, buf, len);
    int np = (int)(len >> 22); if (np > g_readers) np = g_readers; if (np < 1) np = 1;
    pwrite_job j[READERS]; pthread_t th[READERS]; int started[READERS];
    const size_t part = ((len / (size_t)np) + 4095u) & ~(size_t)4095u;
    for (int k = 0; k < np; k++) {
        const size_t lo = (size_t)k * part < len ? (size_t)k * part : len;
        const size_t hi = (k == np - 1) ? len : (lo + part < len ? lo + part : len);
        j[k].fd = fd; j[k].buf = buf + lo; j[k].len = hi - lo; j[k].off = *pos + (off_t)lo; j[k].rc = 0;
        started[k] = k > 0 && j[k].len && pthread_create(&th[k], NULL, pwrite_worker, &j[k]) == 0;
        if (k > 0 && !started[k]) pwrite_worker(&j[k]);
    }
    pwrite_worker(&j[0]);
    int rc = 0;
    for (int k = 0; k < np; k++) { if (started[k]) pthread_join(th[k], NULL); if (j[k].rc) rc = -1; }
    *pos += (off_t)len;
    return rc;
}

enum { S_FREE = 0, S_READ, S_DONE };

typedef struct {
    uint8_t *in;            /* pinned: GZB_DICT bytes of headroom, then the batch */
    uint8_t *out;           /* pinned: gzb_bound(batch) */
    uint64_t n;             /* bytes of the batch */
    int state, final, have_dict;
    gzb_result res;
} slot;

typedef struct pipe_s {
    pthread_mutex_t mu; pthread_cond_t cv;
    slot s[MAX_SLOTS]; int nslots, nworkers;
    uint64_t n_read;        /* batches published by the reader */
    int eof;                /* the reader is through: batch n_read - 1 is the last one */
    int rc;                 /* first error; everybody stops */
    int level; uint32_t flags; uint64_t batch, dict_sz, out_cap;
    int ofd; gzb_zip_opts *o; uint32_t crc;
    int out_regular; off_t opos;       /* regular output file: written with pwrite at opos (several threads per batch) */
    gzb_ctx *ctx[MAX_WORKERS];
} pipe_t;

typedef struct { pipe_t *p; int w; } worker_arg;

static void fail(pipe_t *p, int rc)
{
    pthread_mutex_lock(&p->mu);
    if (p->rc == GZB_OK) p->rc = rc;
    pthread_cond_broadcast(&p->cv);
    pthread_mutex_unlock(&p->mu);
}

static void *worker(void *arg)
{
    pipe_t *p = ((worker_arg *)arg)->p;
    const int w = ((worker_arg *)arg)->w;
    for (uint64_t i = (uint64_t)w;; i += (uint64_t)p->nworkers) {
        slot *s = &p->s[i % (uint64_t)p->nslots];
        pthread_mutex_lock(&p->mu);
        /* batch i is there, and so is the knowledge whether it is the last one */
        while (p->rc == GZB_OK && !(p->n_read > i + 1 || (p->eof && p->n_read > i)) && !(p->eof && p->n_read <= i))
            pthread_cond_wait(&p->cv, &p->mu);
        const int stop = p->rc != GZB_OK || (p->eof && p->n_read <= i);
        const int final = p->eof && p->n_read == i + 1;
        pthread_mutex_unlock(&p->mu);
        if (stop) return NULL;
        const uint8_t *data = s->in + GZB_DICT;
        int rc = gzb_deflate_host(p->ctx[w], p->level, p->flags | ((final || p->o->members) ? GZB_F_FINAL : 0u), data, s->n,
                                  (p->dict_sz && s->have_dict) ? data - p->dict_sz : NULL, s->out, p->out_cap, &s->res);
        if (rc != GZB_OK) { fail(p, rc); return NULL; }
        pthread_mutex_lock(&p->mu);
        s->final = final; s->state = S_DONE;
        pthread_cond_broadcast(&p->cv);
        pthread_mutex_unlock(&p->mu);
        if (final) return NULL;
    }
}

static void *writer(void *arg)
{
    pipe_t *p = (pipe_t *)arg;
    for (uint64_t i = 0;; i++) {
        slot *s = &p->s[i % (uint64_t)p->nslots];
        pthread_mutex_lock(&p->mu);
        while (p->rc == GZB_OK && s->state != S_DONE) pthread_cond_wait(&p->cv, &p->mu);
        const int stop = p->rc != GZB_OK;
        pthread_mutex_unlock(&p->mu);
        if (stop) return NULL;
        if (p->o->members && i > 0) {             /* a member per batch: its own header (no name) ... */
            uint8_t h[16];
            const int hn = gzb_gzip_header(p->level, p->o->mtime, NULL, h, sizeof(h));
            if (hn < 0 || write_batch(p->ofd, h, (size_t)hn, p->out_regular, &p->opos)) { fail(p, GZB_E_IO); return NULL; }
            p->o->bytes_out += (uint64_t)hn;
        }
        if (write_batch(p->ofd, s->out, s->res.out_bytes, p->out_regular, &p->opos)) { fail(p, GZB_E_IO); return NULL; }
        if (p->o->members) {                      /* ... and its own trailer */
            uint8_t tr[8];
            gzb_gzip_trailer(s->res.crc32, s->n, tr, sizeof(tr));
            if (write_batch(p->ofd, tr, 8, p->out_regular, &p->opos)) { fail(p, GZB_E_IO); return NULL; }
            p->o->bytes_out += 8;
        }
        p->crc = gzb_crc32_combine(p->crc, s->res.crc32, s->n);
        p->o->bytes_in += s->n; p->o->bytes_out += s->res.out_bytes;
        const int final = s->final;
        pthread_mutex_lock(&p->mu);
        s->state = S_FREE;
        pthread_cond_broadcast(&p->cv);
        pthread_mutex_unlock(&p->mu);
        if (final) return NULL;
    }
}


/* ---- what outlives a call ---------------------------------------------------------------------------------------
 * Creating a context (CUDA streams, 2 GiB of device scratch) and page-locking the batch slots (~0.3 s per GiB) cost more
 * than compressing a few GiB does, so both are kept for the next call of the process: the reference's buffer pools live
 * for a whole parallel_zip() (src/parallel.c:327-347), these live until gzb_zip_release().  Calls are serialised. */
static struct {
    pthread_mutex_t mu;
    gzb_ctx *ctx[MAX_WORKERS]; int nctx, ng;
    uint8_t *in[MAX_SLOTS], *out[MAX_SLOTS]; uint64_t batch, out_cap;
} g_pool = { PTHREAD_MUTEX_INITIALIZER, {0}, 0, 0, {0}, {0}, 0, 0 };

static void pool_free_slots(void)
{
    for (int i = 0; i < MAX_SLOTS; i++) {
        if (g_pool.in[i]) gzb_host_free(g_pool.in[i]);
        if (g_pool.out[i]) gzb_host_free(g_pool.out[i]);
        g_pool.in[i] = g_pool.out[i] = NULL;
    }
    g_pool.batch = g_pool.out_cap = 0;
}
void gzb_zip_release(void)
{
    pthread_mutex_lock(&g_pool.mu);
    pool_free_slots();
    for (int w = 0; w < g_pool.nctx; w++) { gzb_destroy(g_pool.ctx[w]); g_pool.ctx[w] = NULL; }
    g_pool.nctx = g_pool.ng = 0;
    pthread_mutex_unlock(&g_pool.mu);
}
/* contexts for wpg workers on each of ng GPUs (worker w on GPU w % ng); the pool's lock is held */
static int pool_contexts(int ng, int wpg)
{
    if (g_pool.ng != ng) {
        for (int w = 0; w < g_pool.nctx; w++) { gzb_destroy(g_pool.ctx[w]); g_pool.ctx[w] = NULL; }
        g_pool.nctx = 0; g_pool.ng = ng;
    }
    while (g_pool.nctx < wpg * ng) {
        int rc = gzb_create(g_pool.nctx % ng, &g_pool.ctx[g_pool.nctx]);
        if (rc != GZB_OK) return rc;
        g_pool.nctx++;
    }
    return GZB_OK;
}

/* ---- file to file ------------------------------------------------------------------------------------------------
 * When both ends are regular files nothing has to pass through ONE reader and ONE writer: batch i is the bytes
 * [base + i * BATCH, ...) of the input, its dictionary the 32 KiB in front of them, and its place in the output is known
 * as soon as the sizes of the batches before it are.  So the batches are dealt round-robin to one worker per GPU, and each
 * worker runs them through its context's own pipeline (gzb_internal_deflate_batches: H2D of batch k+1 || kernels of batch k
 * || D2H of batch k-1 over the context's pinned staging): the `fetch` callback reads a batch (dictionary included, one
 * contiguous range) straight into the staging buffer, the `emit` callback writes the fragment at its offset; the copies
 * between page cache and staging are cut into pieces and done by a pool of I/O threads (one per CPU) that all workers share.
 * The serial reader / in-order writer above remain for pipes, terminals, append-mode output and the member-per-batch mode. */
static size_t IO_PIECE = (size_t)2 << 20;      /* GZB_ZIP_IO_PIECE_KB */
#define IO_QUEUE 8192
#define IO_MAX_THREADS 128
typedef struct { int pending, rc; } io_group;
enum { IO_READ = 0, IO_WRITE = 1, IO_COPY = 2 };     /* pread into buf; pwrite from buf; memcpy(dst, buf) */
typedef struct { int fd, wr; uint8_t *buf, *dst; size_t len; off_t off; io_group *grp; } io_piece;
typedef struct {
    pthread_mutex_t mu; pthread_cond_t work, done;
    /* One ring per kind of piece (head: next to take, tail: next free).  Reads run on as many threads as there are.
     * pwrite()s into the one output file exclude each other in the kernel (the inode's lock), so at most cap[IO_WRITE] = 1
     * thread is in one at a time -- more only queue on that lock and slow each other down (2 GPUs, 24 CPUs, 16 GiB text:
     * 16.3 GB/s with one writer at a time, 13.7 with two, 13.9 with four).  Copies into the output's mapping scale to a
     * handful of threads and fall off beyond (tools/micro_fallocwrite.c: 7.2 GB/s with 4, 6.0 with 32): cap[IO_COPY] = 6. */
    io_piece q[3][IO_QUEUE]; unsigned head[3], tail[3];      /* by kind: IO_READ, IO_WRITE, IO_COPY */
    int busy[3], cap[3];                                      /* threads in a piece of each kind; how many may be */
    int stop, nth; pthread_t th[IO_MAX_THREADS];
} io_pool;

static void *io_thread(void *arg)
{
    io_pool *io = (io_pool *)arg;
    pthread_mutex_lock(&io->mu);
    for (;;) {
        int k;
        for (;;) {
            /* output first (a fragment that waits keeps its pinned buffer), then the reads */
            if (io->head[IO_WRITE] != io->tail[IO_WRITE] && io->busy[IO_WRITE] < io->cap[IO_WRITE]) { k = IO_WRITE; break; }
            if (io->head[IO_COPY] != io->tail[IO_COPY] && io->busy[IO_COPY] < io->cap[IO_COPY]) { k = IO_COPY; break; }
            if (io->head[IO_READ] != io->tail[IO_READ]) { k = IO_READ; break; }
            if (io->stop && io->head[IO_WRITE] == io->tail[IO_WRITE] && io->head[IO_COPY] == io->tail[IO_COPY]) { k = -1; break; }
            pthread_cond_wait(&io->work, &io->mu);
        }
        if (k < 0) break;
        io_piece pc = io->q[k][io->head[k]++ % IO_QUEUE];
        io->busy[k]++;
        pthread_mutex_unlock(&io->mu);
        int rc = 0; size_t done = 0;
        if (pc.wr == IO_COPY) { memcpy(pc.dst, pc.buf, pc.len); done = pc.len; }
        while (done < pc.len) {
            ssize_t r = pc.wr ? pwrite(pc.fd, pc.buf + done, pc.len - done, pc.off + (off_t)done)
                              : pread(pc.fd, pc.buf + done, pc.len - done, pc.off + (off_t)done);
            if (r < 0 && errno == EINTR) continue;
            if (r <= 0) { rc = -1; break; }        /* a short file is an error here: its size was taken at the start */
            done += (size_t)r;
        }
        pthread_mutex_lock(&io->mu);
        if (rc) pc.grp->rc = rc;
        io->busy[k]--;
        if (k != IO_READ && io->head[k] != io->tail[k]) pthread_cond_signal(&io->work);            /* a waiting piece of this kind may go now */
        if (--pc.grp->pending == 0) pthread_cond_broadcast(&io->done);
    }
    pthread_mutex_unlock(&io->mu);
    return NULL;
}

/* queue the copy of [off, off + len) of fd to / from buf in pieces of `piece` bytes; g counts them down */
static void io_submit(io_pool *io, io_group *g, int fd, int wr, uint8_t *buf, size_t len, off_t off, size_t piece, uint8_t *dst)
{
    pthread_mutex_lock(&io->mu);
    for (size_t o = 0; o < len; o += piece) {
        const int k = wr;
        while (io->tail[k] - io->head[k] == IO_QUEUE) pthread_cond_wait(&io->done, &io->mu);     /* full: never in practice */
        io_piece *pc = &io->q[k][io->tail[k]++ % IO_QUEUE];
        pc->fd = fd; pc->wr = wr; pc->buf = buf + o; pc->dst = dst ? dst + o : NULL; pc->len = len - o < piece ? len - o : piece; pc->off = off + (off_t)o; pc->grp = g;
        g->pending++;
    }
    pthread_cond_broadcast(&io->work);
    pthread_mutex_unlock(&io->mu);
}
/* until every piece of g is done; returns the group's status and clears it */
static int io_wait(io_pool *io, io_group *g)
{
    pthread_mutex_lock(&io->mu);
    while (g->pending) pthread_cond_wait(&io->done, &io->mu);
    const int rc = g->rc; g->rc = 0;
    pthread_mutex_unlock(&io->mu);
    return rc;
}

typedef struct {
    pthread_mutex_t mu; pthread_cond_t cv;
    io_pool *io;
    int ifd, ofd, level, rc; uint32_t flags;
    off_t base, obase;                 /* first input byte; first output byte after the header */
    uint64_t n, batch, dict_sz, out_cap, nb;
    int nw;                            /* workers: batch i belongs to worker i % nw */
    int mfd, map_off; long pagesz;     /* output reopened read-write for the shared mapping (-1: pwrite only); set when fallocate is not supported */
    uint8_t *map; size_t map_len; off_t map_lo;   /* ONE mapping of everything the fragments can come to, from file offset map_lo on */
    uint64_t known;                    /* batches [0, known) have their sizes in; pre[known] is final */
    uint64_t *pre;                     /* pre[i] = bytes of the fragments before batch i  (nb + 1 entries) */
    uint64_t *size; uint32_t *crc; uint8_t *have;
    gzb_ctx *ctx[MAX_WORKERS];
} direct_t;
typedef struct { direct_t *d; int w; io_group rg, wg[2]; double t_fetch, t_wait, t_write, t_all; } direct_arg;   /* t_*: seconds, for GZB_ZIP_VERBOSE */

static void direct_fail(direct_t *d, int rc)
{
    pthread_mutex_lock(&d->mu);
    if (d->rc == GZB_OK) d->rc = rc;
    pthread_cond_broadcast(&d->cv);
    pthread_mutex_unlock(&d->mu);
}

static int direct_fetch(void *user, uint64_t k, uint8_t *dst, uint32_t dict_sz, uint64_t *len, int *have_dict, int *final)
{
    direct_t *d = ((direct_arg *)user)->d;
    const uint64_t i = k * (uint64_t)d->nw + (uint64_t)((direct_arg *)user)->w;      /* batches are dealt round-robin to the workers */
    const uint64_t off = i * d->batch, n = d->n - off < d->batch ? d->n - off : d->batch;
    const uint64_t dict = i > 0 ? dict_sz : 0;                            /* the bytes in front of the batch are its dictionary */
    pthread_mutex_lock(&d->mu);
    const int bad = d->rc != GZB_OK;
    pthread_mutex_unlock(&d->mu);
    if (bad) return GZB_E_IO;                                             /* somebody else failed: stop reading */
    const double t0 = now_s();
    io_submit(d->io, &((direct_arg *)user)->rg, d->ifd, IO_READ, dst - dict, (size_t)(dict + n), d->base + (off_t)(off - dict), IO_PIECE, NULL);
    if (io_wait(d->io, &((direct_arg *)user)->rg)) return GZB_E_IO;
    ((direct_arg *)user)->t_fetch += now_s() - t0;
    *len = n; *have_dict = dict != 0; *final = i + 1 == d->nb;
    return GZB_OK;
}

static int direct_emit(void *user, uint64_t k, const uint8_t *src, const gzb_result *r)
{
    direct_t *d = ((direct_arg *)user)->d;
    const uint64_t i = k * (uint64_t)d->nw + (uint64_t)((direct_arg *)user)->w;
    const double t0 = now_s();
    pthread_mutex_lock(&d->mu);
    d->size[i] = r->out_bytes; d->crc[i] = r->crc32; d->have[i] = 1;
    while (d->known < d->nb && d->have[d->known]) { d->pre[d->known + 1] = d->pre[d->known] + d->size[d->known]; d->known++; }
    pthread_cond_broadcast(&d->cv);
    /* every worker takes its batches in rising order and announces a size before it waits: the oldest batch never waits */
    while (d->rc == GZB_OK && d->known < i) pthread_cond_wait(&d->cv, &d->mu);
    const int bad = d->rc != GZB_OK;
    const uint64_t at = d->pre[i];
    pthread_mutex_unlock(&d->mu);
    if (bad) return GZB_E_IO;
    /* The worker does not wait for the fragment's bytes to be in the file: direct_emit_done does, two batches later. */
    direct_arg *a = (direct_arg *)user;
    const off_t lo = d->obase + (off_t)at;
    const size_t len = (size_t)r->out_bytes;
    if (d->map && !d->map_off && len) {
        /* Reserve, then fill: fallocate() gives the fragment's range its pages (the one serial step, ~15 GB/s; a full file
         * system is an error code HERE, never a SIGBUS later), then the bytes go into a shared mapping of that range from
         * all I/O threads at once -- page-table work instead of the inode's lock (tools/micro_fallocwrite.c: 7 GB/s
         * against 3.8 GB/s for pwrite on the B200 hosts). */
        if (fallocate(d->mfd, 0, lo, (off_t)len) != 0) {
            if (errno == EOPNOTSUPP || errno == ENOSYS || errno == EINVAL) { pthread_mutex_lock(&d->mu); d->map_off = 1; pthread_mutex_unlock(&d->mu); }
            else return GZB_E_IO;
        } else {
            io_submit(d->io, &a->wg[k & 1], -1, IO_COPY, (uint8_t *)(uintptr_t)src, len, 0, IO_PIECE, d->map + (lo - d->map_lo));
            a->t_wait += now_s() - t0;
            return GZB_OK;
        }
    }
    /* pwrite()s into one file do not run side by side (the inode's lock: 3.8 GB/s into a fresh tmpfs file however many
     * threads write, tools/micro_pwrite.c): few large pieces, one writer at a time */
    io_submit(d->io, &a->wg[k & 1], d->ofd, IO_WRITE, (uint8_t *)(uintptr_t)src, len, lo, (size_t)16 << 20, NULL);
    a->t_wait += now_s() - t0;
    return GZB_OK;
}

static int direct_emit_done(void *user, uint64_t k)
{
    direct_arg *a = (direct_arg *)user;
    const double t0 = now_s();
    const int rc = io_wait(a->d->io, &a->wg[k & 1]) ? GZB_E_IO : GZB_OK;
    a->t_write += now_s() - t0;
    return rc;
}

static void *direct_worker(void *arg)
{
    direct_arg *a = (direct_arg *)arg;
    direct_t *d = a->d;
    gzb_batch_io io;
    memset(&io, 0, sizeof(io));
    io.user = a; io.max_bytes = d->batch; io.blocking = 1;               /* waiting workers leave their CPUs to the I/O threads */
    io.n_batches = d->nb > (uint64_t)a->w ? (d->nb - (uint64_t)a->w + (uint64_t)d->nw - 1) / (uint64_t)d->nw : 0;
    io.fetch = direct_fetch; io.emit = direct_emit; io.emit_done = direct_emit_done;
    const double t0 = now_s();
    const int rc = gzb_internal_deflate_batches(d->ctx[a->w], d->level, d->flags, &io);
    a->t_all = now_s() - t0;
    if (rc != GZB_OK) direct_fail(d, rc);
    return NULL;
}

/* the body of a member, file to file: n input bytes at base -> fragments from obase on.  The pool's lock is held. */
static int zip_direct(int ifd, int ofd, gzb_zip_opts *o, uint32_t flags, int nw, off_t base, uint64_t n, off_t obase,
                      uint64_t batch, uint64_t out_cap, uint32_t *crc_out, uint64_t *out_bytes)
{
    static io_pool io;                                                    /* one call at a time (g_pool.mu) */
    direct_t D, *d = &D;
    memset(d, 0, sizeof(*d)); memset(&io, 0, sizeof(io));
    d->ifd = ifd; d->ofd = ofd; d->level = o->level; d->flags = flags; d->base = base; d->obase = obase; d->n = n;
    d->batch = batch; d->out_cap = out_cap; d->dict_sz = (flags & GZB_F_RSYNCABLE) ? 0 : GZB_DICT;
    d->nb = (n + batch - 1) / batch; d->rc = GZB_OK; d->io = &io;
    d->nw = nw;
    d->mfd = -1; d->pagesz = sysconf(_SC_PAGESIZE);
    if (!getenv("GZB_ZIP_NO_MMAP") && d->pagesz > 0) {                    /* the caller's descriptor is usually write-only: a mapping needs read-write */
        char path[64];
        snprintf(path, sizeof(path), "/proc/self/fd/%d", ofd);
        d->mfd = open(path, O_RDWR | O_CLOEXEC);
        if (d->mfd >= 0) {
            /* address space only: pages exist where fallocate has put them, and nothing else is ever touched.  One mapping
             * for the whole call -- mmap / munmap per fragment take the process's mmap lock against every page fault of the
             * fillers */
            d->map_lo = obase & ~(off_t)(d->pagesz - 1);
            d->map_len = (size_t)(obase - d->map_lo) + (size_t)gzb_bound(n, flags) + (size_t)out_cap;
            void *m = mmap(NULL, d->map_len, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_NORESERVE, d->mfd, d->map_lo);
            if (m == MAP_FAILED) { close(d->mfd); d->mfd = -1; } else d->map = (uint8_t *)m;
        }
    }
    for (int w = 0; w < nw; w++) d->ctx[w] = g_pool.ctx[w];
    d->pre = (uint64_t *)calloc(d->nb + 1, sizeof(uint64_t)); d->size = (uint64_t *)calloc(d->nb, sizeof(uint64_t));
    d->crc = (uint32_t *)calloc(d->nb, sizeof(uint32_t)); d->have = (uint8_t *)calloc(d->nb, 1);
    int rc = (d->pre && d->size && d->crc && d->have) ? GZB_OK : GZB_E_NOMEM;
    pthread_mutex_init(&d->mu, NULL); pthread_cond_init(&d->cv, NULL);
    pthread_mutex_init(&io.mu, NULL); pthread_cond_init(&io.work, NULL); pthread_cond_init(&io.done, NULL);
    long ncpu = sysconf(_SC_NPROCESSORS_ONLN);
    int want = ncpu > 0 ? (int)ncpu : 8;
    { const char *e = getenv("GZB_ZIP_IO_THREADS"); if (e && atoi(e) > 0) want = atoi(e); }
    io.cap[IO_READ] = IO_MAX_THREADS; io.cap[IO_WRITE] = 1; io.cap[IO_COPY] = 6;
    { const char *e = getenv("GZB_ZIP_WRITERS"); if (e && atoi(e) > 0) io.cap[IO_WRITE] = atoi(e); }
    { const char *e = getenv("GZB_ZIP_FILLERS"); if (e && atoi(e) > 0) io.cap[IO_COPY] = atoi(e); }
    { const char *e = getenv("GZB_ZIP_IO_PIECE_KB"); IO_PIECE = (size_t)(e && atoi(e) >= 64 ? atoi(e) : 2048) << 10; }
    if (want > IO_MAX_THREADS) want = IO_MAX_THREADS;
    pthread_t th[MAX_WORKERS]; direct_arg wa[MAX_WORKERS]; int nth = 0;
    if (rc == GZB_OK) {
        for (; io.nth < want; io.nth++) if (pthread_create(&io.th[io.nth], NULL, io_thread, &io)) break;
        if (io.nth == 0) rc = GZB_E_NOMEM;
    }
    if (rc == GZB_OK) {
        for (int w = 0; w < nw; w++) {
            memset(&wa[w], 0, sizeof(wa[w])); wa[w].d = d; wa[w].w = w;
            if (pthread_create(&th[w], NULL, direct_worker, &wa[w])) { direct_fail(d, GZB_E_NOMEM); break; }
            nth++;
        }
        for (int w = 0; w < nth; w++) pthread_join(th[w], NULL);
        rc = d->rc;
        if (getenv("GZB_ZIP_VERBOSE"))
            for (int w = 0; w < nth; w++)
                fprintf(stderr, "gzb_zip_fd: worker %d: %.3f s in all, of which reading %.3f s, waiting for its offset %.3f s, waiting for its writes %.3f s (%d I/O threads)\n",
                        w, wa[w].t_all, wa[w].t_fetch, wa[w].t_wait, wa[w].t_write, io.nth);
    }
    pthread_mutex_lock(&io.mu); io.stop = 1; pthread_cond_broadcast(&io.work); pthread_mutex_unlock(&io.mu);
    for (int k = 0; k < io.nth; k++) pthread_join(io.th[k], NULL);
    if (rc == GZB_OK) {
        uint32_t crc = 0;
        for (uint64_t i = 0; i < d->nb; i++) {
            const uint64_t off = i * batch, len = n - off < batch ? n - off : batch;
            crc = gzb_crc32_combine(crc, d->crc[i], len);                 /* src/parallel.c:544 */
        }
        *crc_out = crc; *out_bytes = d->pre[d->nb];
    }
    if (d->map) munmap(d->map, d->map_len);
    if (d->mfd >= 0) close(d->mfd);
    free(d->pre); free(d->size); free(d->crc); free(d->have);
    pthread_mutex_destroy(&d->mu); pthread_cond_destroy(&d->cv);
    pthread_mutex_destroy(&io.mu); pthread_cond_destroy(&io.work); pthread_cond_destroy(&io.done);
    return rc;
}

/* Inputs that end within SMALL bytes -- most files a gzip ever sees -- take one call on pageable memory: no pinned
 * slots, no threads. */
#define SMALL ((size_t)32 << 20)
static int zip_small(gzb_ctx *ctx, int ofd, gzb_zip_opts *o, uint32_t flags, const uint8_t *data, uint64_t n)
{
    uint8_t *out = NULL, tr[8];
    gzb_result r;
    int rc = GZB_OK;
    const uint64_t cap = gzb_bound(n, flags);
    if (!(out = (uint8_t *)malloc(cap))) rc = GZB_E_NOMEM;
    if (rc == GZB_OK) rc = gzb_deflate_host(ctx, o->level, flags | GZB_F_FINAL, data, n, NULL, out, cap, &r);
    if (rc == GZB_OK && writen(ofd, out, r.out_bytes)) rc = GZB_E_IO;
    if (rc == GZB_OK) {
        gzb_gzip_trailer(r.crc32, n, tr, sizeof(tr));
        if (writen(ofd, tr, 8)) rc = GZB_E_IO;
        o->bytes_in = n; o->bytes_out += r.out_bytes + 8;
    }
    free(out);
    return rc;
}

int gzb_zip_fd(int ifd, int ofd, gzb_zip_opts *o)
{
    if (!o || o->level < 1 || o->level > 9) return o ? GZB_E_LEVEL : GZB_E_ARG;
    int ng = o->n_gpus > 0 ? o->n_gpus : 1;
    if (ng > MAX_GPUS) ng = MAX_GPUS;
    if (ng > gzb_device_count()) ng = gzb_device_count();
    if (ng < 1) return GZB_E_NO_DEVICE;
    if (o->name && strlen(o->name) > 4000) return GZB_E_ARG;
    const int verbose = getenv("GZB_ZIP_VERBOSE") != NULL;
    const double t0 = now_s();

    pipe_t P, *p = &P;
    memset(p, 0, sizeof(*p));
    struct stat st;
    off_t pos = 0;
    const int regular = fstat(ifd, &st) == 0 && S_ISREG(st.st_mode) && (pos = lseek(ifd, 0, SEEK_CUR)) >= 0;
    const uint64_t in_left = regular && st.st_size > pos ? (uint64_t)(st.st_size - pos) : 0;   /* as of now: a file that grows meanwhile is cut here */
    /* a regular output file that is not in append mode is written with pwrite at explicit offsets */
    const int oflags = fcntl(ofd, F_GETFL);
    p->out_regular = fstat(ofd, &st) == 0 && S_ISREG(st.st_mode) && oflags >= 0 && !(oflags & O_APPEND) && (p->opos = lseek(ofd, 0, SEEK_CUR)) >= 0;
    {
        long ncpu = sysconf(_SC_NPROCESSORS_ONLN);
        int nr = 8 * ng; if (nr > READERS) nr = READERS; if (ncpu > 0 && nr > (int)ncpu) nr = (int)ncpu; if (nr < 1) nr = 1;
        g_readers = nr;
    }
    pthread_mutex_lock(&g_pool.mu);                  /* one call at a time: contexts and pinned slots are the pool's */
#ifdef F_SETPIPE_SZ
    if (!regular) { (void)fcntl(ifd, F_SETPIPE_SZ, 1 << 20); (void)fcntl(ofd, F_SETPIPE_SZ, 1 << 20); }   /* pipes: fewer, larger transfers */
#endif
    pthread_mutex_init(&p->mu, NULL); pthread_cond_init(&p->cv, NULL);
    p->level = o->level; p->flags = o->rsyncable ? GZB_F_RSYNCABLE : 0u;
    p->dict_sz = o->rsyncable ? 0 : GZB_DICT;
    p->batch = batch_bytes(); p->out_cap = gzb_bound(p->batch, p->flags);
    p->nworkers = 2 * ng; p->nslots = p->nworkers + 2;
    p->ofd = ofd; p->o = o; p->rc = GZB_OK;
    pthread_t th[MAX_WORKERS], wth; worker_arg wa[MAX_WORKERS];
    int nth = 0, have_writer = 0, rc = GZB_OK;
    uint8_t hdr[10 + 4096], *peek = NULL;
    double t_setup = t0, t_stream = t0;

    int hn = gzb_gzip_header(o->level, o->mtime, o->name, hdr, sizeof(hdr));
    if (hn < 0) { rc = hn; goto done; }
    if (write_batch(ofd, hdr, (size_t)hn, p->out_regular, &p->opos)) { rc = GZB_E_IO; goto done; }
    o->bytes_in = 0; o->bytes_out = (uint64_t)hn;

    /* file to file, and more than a small input: no serial stage at all (zip_direct) */
    if (regular && p->out_regular && !o->members && in_left > SMALL && !getenv("GZB_ZIP_NO_DIRECT")) {
        /* Workers (thread + context) per GPU.  The batches of one worker overlap inside gzb_internal_deflate_batches, but
         * read + H2D of a batch take about as long as its kernels, so a lone GPU is only kept busy by three of them (16 GiB
         * text, tmpfs to tmpfs: 18.1 / 20.0 / 21.4 / 20.9 GB/s with 1 / 2 / 3 / 4 workers; the GPU itself does 21.7).  With
         * several GPUs the output file is the limit long before (2 GPUs: 30.1 / 25.0 / 24.2 GB/s with 1 / 2 / 3 per GPU). */
        int wpg = ng == 1 ? 3 : 1;
        { const char *e = getenv("GZB_ZIP_WORKERS_PER_GPU"); if (e && atoi(e) > 0) wpg = atoi(e); }
        if (wpg * ng > MAX_WORKERS) wpg = MAX_WORKERS / ng;
        { const uint64_t nb = (in_left + p->batch - 1) / p->batch; while (wpg > 1 && (uint64_t)(wpg - 1) * (uint64_t)ng >= nb) wpg--; }
        const int nw = wpg * ng;
        if ((rc = pool_contexts(ng, wpg)) != GZB_OK) goto done;
        t_setup = now_s();
        uint32_t crc = 0; uint64_t ob = 0; uint8_t tr[8];
        rc = zip_direct(ifd, ofd, o, p->flags, nw, pos, in_left, p->opos, p->batch, p->out_cap, &crc, &ob);
        if (rc == GZB_OK) {
            pos += (off_t)in_left; p->opos += (off_t)ob; o->bytes_in = in_left; o->bytes_out += ob;
            gzb_gzip_trailer(crc, in_left, tr, sizeof(tr));
            if (write_batch(ofd, tr, 8, 1, &p->opos)) rc = GZB_E_IO; else o->bytes_out += 8;
        }
        t_stream = now_s();
        goto done;
    }

    /* look at the first SMALL bytes: is that all there is? */
    const size_t peek_len = SMALL < p->batch ? SMALL : (size_t)p->batch;
    if (!(peek = (uint8_t *)malloc(peek_len))) { rc = GZB_E_NOMEM; goto done; }
    ssize_t pk = read_batch(ifd, peek, peek_len, regular, &pos);
    if (pk < 0) { rc = GZB_E_IO; goto done; }
    if ((rc = pool_contexts(ng, 2)) != GZB_OK) goto done;
    for (int w = 0; w < p->nworkers; w++) p->ctx[w] = g_pool.ctx[w];
    if ((size_t)pk < peek_len) {
        if (p->out_regular) lseek(ofd, p->opos, SEEK_SET);          /* the small path writes sequentially */
        rc = zip_small(p->ctx[0], ofd, o, p->flags, peek, (uint64_t)pk);
        if (p->out_regular) p->opos = lseek(ofd, 0, SEEK_CUR);
        goto done;
    }
    if (g_pool.batch != p->batch || g_pool.out_cap != p->out_cap) { pool_free_slots(); g_pool.batch = p->batch; g_pool.out_cap = p->out_cap; }
    for (int i = 0; i < p->nslots; i++) { p->s[i].in = g_pool.in[i]; p->s[i].out = g_pool.out[i]; }
    for (int w = 0; w < p->nworkers; w++) {
        wa[w].p = p; wa[w].w = w;
        if (pthread_create(&th[w], NULL, worker, &wa[w])) { fail(p, GZB_E_NOMEM); break; }
        nth++;
    }
    if (pthread_create(&wth, NULL, writer, p)) fail(p, GZB_E_NOMEM); else have_writer = 1;
    t_setup = now_s();

    /* the reader */
    for (uint64_t i = 0;; i++) {
        slot *s = &p->s[i % (uint64_t)p->nslots];
        if (!s->in) {                             /* slots are page-locked on first use, under the work of the earlier batches */
            const int si = (int)(i % (uint64_t)p->nslots);
            int arc = gzb_host_alloc(p->ctx[0], GZB_DICT + p->batch, (void **)&s->in);
            g_pool.in[si] = s->in;
            if (arc == GZB_OK) { arc = gzb_host_alloc(p->ctx[0], p->out_cap, (void **)&s->out); g_pool.out[si] = s->out; }
            if (arc != GZB_OK) { fail(p, arc); break; }
        }
        pthread_mutex_lock(&p->mu);
        while (p->rc == GZB_OK && s->state != S_FREE) pthread_cond_wait(&p->cv, &p->mu);
        const int stop = p->rc != GZB_OK;
        pthread_mutex_unlock(&p->mu);
        if (stop) break;
        size_t pre = 0;
        if (i == 0) { memcpy(s->in + GZB_DICT, peek, (size_t)pk); pre = (size_t)pk; }
        ssize_t got = pre < p->batch ? read_batch(ifd, s->in + GZB_DICT + pre, p->batch - pre, regular, &pos) : 0;
        if (got < 0) { fail(p, GZB_E_IO); break; }
        got += (ssize_t)pre;
        if (got == 0 && i > 0) {                  /* nothing after a full batch: the previous one was the last */
            pthread_mutex_lock(&p->mu);
            p->eof = 1; pthread_cond_broadcast(&p->cv);
            pthread_mutex_unlock(&p->mu);
            break;
        }
        s->n = (uint64_t)got; s->have_dict = 0;
        if (i > 0 && p->dict_sz && !o->members) { /* the dictionary of a batch: the tail of the one before (which was full) */
            const slot *prev = &p->s[(i - 1) % (uint64_t)p->nslots];
            memcpy(s->in + GZB_DICT - p->dict_sz, prev->in + GZB_DICT + prev->n - p->dict_sz, p->dict_sz);
            s->have_dict = 1;
        }
        pthread_mutex_lock(&p->mu);
        s->state = S_READ; p->n_read = i + 1;
        if ((uint64_t)got < p->batch) p->eof = 1; /* a short batch is the last */
        pthread_cond_broadcast(&p->cv);
        const int last = p->eof;
        pthread_mutex_unlock(&p->mu);
        if (last) break;
    }
    if (p->rc != GZB_OK) {                        /* wake everybody who waits for a batch that will not come */
        pthread_mutex_lock(&p->mu); p->eof = 1; pthread_cond_broadcast(&p->cv); pthread_mutex_unlock(&p->mu);
    }
    for (int w = 0; w < nth; w++) pthread_join(th[w], NULL);
    if (have_writer) pthread_join(wth, NULL);
    rc = p->rc;
    if (rc == GZB_OK && !o->members) {
        uint8_t tr[8];
        gzb_gzip_trailer(p->crc, o->bytes_in, tr, sizeof(tr));
        if (write_batch(ofd, tr, 8, p->out_regular, &p->opos)) rc = GZB_E_IO; else o->bytes_out += 8;
    }
    t_stream = now_s();
done:
    free(peek);
    pthread_mutex_unlock(&g_pool.mu);                /* contexts and slots stay in the pool (gzb_zip_release) */
    pthread_mutex_destroy(&p->mu); pthread_cond_destroy(&p->cv);
    if (regular) lseek(ifd, pos, SEEK_SET);          /* leave the descriptors where sequential I/O would */
    if (p->out_regular) lseek(ofd, p->opos, SEEK_SET);
    if (verbose)
        fprintf(stderr, "gzb_zip_fd: rc %d, %llu -> %llu bytes; contexts + threads %.3f s, read/compress/write %.3f s (%.2f GB/s), teardown %.3f s\n",
                rc, (unsigned long long)o->bytes_in, (unsigned long long)o->bytes_out, t_setup - t0, t_stream - t_setup,
                t_stream > t_setup ? (double)o->bytes_in / (t_stream - t_setup) / 1e9 : 0.0, now_s() - t_stream);
    return rc;
}
