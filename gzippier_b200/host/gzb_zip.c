/* gzb_zip.c -- reference-facing host driver: what parallel_zip() does around the codec.
 *
 * Mirrors `off_t parallel_zip(int pack_level)` (reference src/parallel.c:699-834) and its
 * write_thread (:490-558) with the globals passed in gzb_zip_opts: read ifd on the fixed chunk
 * grid, look one batch ahead to know which batch is final (:764), hand every batch to the GPU
 * engine together with the 32 KiB that precede it (:767-774), write header / fragments in order /
 * trailer (:496-507, :538, :555-556), fold CRCs with crc32_combine (:544).
 * Differences by design: errors are reported, empty input yields the valid
 * 20-byte member instead of a deadlock (F4), and FNAME is the caller's choice (F5).
 *
 * With n_gpus > 1 every batch is cut into contiguous shards on chunk boundaries, one host thread
 * and one context per GPU, no collective: the host concatenates the shard bodies in order.
 */
#include "gzb200.h"
#include <errno.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#define MAX_GPUS 16
#define SHARD_BYTES ((uint64_t)256 << 20)       /* per GPU per batch; multiple of the chunk size */

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

typedef struct {
    gzb_ctx *ctx; int level; uint32_t flags;
    const uint8_t *in; uint64_t n; const uint8_t *dict;
    uint8_t *out; uint64_t out_cap; gzb_result res; int rc;
} shard_job;

static void *shard_worker(void *arg)
{
    shard_job *j = (shard_job *)arg;
    j->rc = gzb_deflate_host(j->ctx, j->level, j->flags, j->in, j->n, j->dict, j->out, j->out_cap, &j->res);
    return NULL;
}

int gzb_zip_fd(int ifd, int ofd, gzb_zip_opts *o)
{
    if (!o || o->level < 1 || o->level > 9) return o ? GZB_E_LEVEL : GZB_E_ARG;
    int ng = o->n_gpus > 0 ? o->n_gpus : 1;
    if (ng > MAX_GPUS) ng = MAX_GPUS;
    if (ng > gzb_device_count()) ng = gzb_device_count();
    if (ng < 1) return GZB_E_NO_DEVICE;
    const uint32_t flags = o->rsyncable ? GZB_F_RSYNCABLE : 0u;
    const uint64_t dict_sz = o->rsyncable ? 0 : GZB_DICT;
    const uint64_t batch = SHARD_BYTES * (uint64_t)ng;
    gzb_ctx *ctx[MAX_GPUS] = {0};
    uint8_t *buf[2] = {NULL, NULL}, *outb[MAX_GPUS] = {0};
    uint8_t dict_keep[GZB_DICT];
    int rc = GZB_OK, have_dict = 0;
    uint8_t hdr[10 + 4096];

    for (int g = 0; g < ng; g++) if ((rc = gzb_create(g, &ctx[g])) != GZB_OK) goto done;
    for (int i = 0; i < 2; i++) if (!(buf[i] = (uint8_t *)malloc(batch ? batch : 1))) { rc = GZB_E_NOMEM; goto done; }
    for (int g = 0; g < ng; g++) if (!(outb[g] = (uint8_t *)malloc(gzb_bound(SHARD_BYTES, flags)))) { rc = GZB_E_NOMEM; goto done; }

    if (o->name && strlen(o->name) > 4000) { rc = GZB_E_ARG; goto done; }
    int hn = gzb_gzip_header(o->level, o->mtime, o->name, hdr, sizeof(hdr));
    if (hn < 0) { rc = hn; goto done; }
    if (writen(ofd, hdr, (size_t)hn)) { rc = GZB_E_IO; goto done; }
    o->bytes_in = 0; o->bytes_out = (uint64_t)hn;

    uint32_t crc = 0;
    int cur = 0;
    ssize_t got = readn(ifd, buf[cur], batch);
    if (got < 0) { rc = GZB_E_IO; goto done; }
    for (;;) {
        ssize_t next = ((uint64_t)got == batch) ? readn(ifd, buf[cur ^ 1], batch) : 0;   /* look ahead: who is final? */
        if (next < 0) { rc = GZB_E_IO; goto done; }
        const int final_batch = (next == 0);
        /* shards on chunk boundaries */
        shard_job jobs[MAX_GPUS]; pthread_t th[MAX_GPUS]; int nj = 0;
        uint64_t off = 0;
        do {
            uint64_t nb = (uint64_t)got - off < SHARD_BYTES ? (uint64_t)got - off : SHARD_BYTES;
            shard_job *j = &jobs[nj];
            memset(j, 0, sizeof(*j));
            j->ctx = ctx[nj]; j->level = o->level;
            j->in = buf[cur] + off; j->n = nb;
            j->dict = dict_sz ? (off ? buf[cur] + off - dict_sz : (have_dict ? dict_keep : NULL)) : NULL;
            j->flags = flags | ((final_batch && off + nb == (uint64_t)got) ? GZB_F_FINAL : 0u);
            j->out = outb[nj]; j->out_cap = gzb_bound(SHARD_BYTES, flags);
            off += nb; nj++;
        } while (off < (uint64_t)got);
        for (int k = 1; k < nj; k++) if (pthread_create(&th[k], NULL, shard_worker, &jobs[k])) { jobs[k].rc = GZB_E_NOMEM; th[k] = 0; shard_worker(&jobs[k]); }
        shard_worker(&jobs[0]);
        for (int k = 1; k < nj; k++) if (th[k]) pthread_join(th[k], NULL);
        for (int k = 0; k < nj; k++) {
            if (jobs[k].rc != GZB_OK) { rc = jobs[k].rc; goto done; }
            if (writen(ofd, jobs[k].out, jobs[k].res.out_bytes)) { rc = GZB_E_IO; goto done; }
            crc = gzb_crc32_combine(crc, jobs[k].res.crc32, jobs[k].n);
            o->bytes_in += jobs[k].n; o->bytes_out += jobs[k].res.out_bytes;
        }
        if (final_batch) break;
        if (dict_sz) { memcpy(dict_keep, buf[cur] + (uint64_t)got - dict_sz, dict_sz); have_dict = 1; }   /* a non-final batch is full */
        cur ^= 1; got = next;
    }
    uint8_t tr[8];
    gzb_gzip_trailer(crc, o->bytes_in, tr, sizeof(tr));
    if (writen(ofd, tr, 8)) { rc = GZB_E_IO; goto done; }
    o->bytes_out += 8;
done:
    for (int g = 0; g < ng; g++) { if (ctx[g]) gzb_destroy(ctx[g]); free(outb[g]); }
    free(buf[0]); free(buf[1]);
    return rc;
}
