/* gzb200 -- minimal gzip-compatible front end over gzb_zip_fd(), for tests that read like the
 * reference's (tests/parallel: `gzip -j N -c < in | gzip -d | cmp`).  stdin -> stdout only.
 *   gzb200 [-1..-9] [-j NGPUS] [--rsyncable] [--members] [-N name] < in > out.gz
 * Option letters follow the reference's getopt table (src/gzip.c:275-315); -j here counts GPUs. */
#include "gzb200.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

int main(int argc, char **argv)
{
    gzb_zip_opts o; memset(&o, 0, sizeof(o)); o.level = 6; o.n_gpus = 1;
    for (int i = 1; i < argc; i++) {
        const char *a = argv[i];
        if (a[0] == '-' && a[1] >= '1' && a[1] <= '9' && !a[2]) o.level = a[1] - '0';
        else if (!strcmp(a, "-j") && i + 1 < argc) o.n_gpus = atoi(argv[++i]);
        else if (!strcmp(a, "--rsyncable")) o.rsyncable = 1;
        else if (!strcmp(a, "-N") && i + 1 < argc) o.name = argv[++i];
        else if (!strcmp(a, "--members")) o.members = 1;
        else if (!strcmp(a, "-c")) {}
        else { fprintf(stderr, "usage: gzb200 [-1..-9] [-j NGPUS] [--rsyncable] [--members] [-N name] < in > out.gz\n"); return 2; }
    }
    if (o.n_gpus < 1) { fprintf(stderr, "gzb200: invalid number of GPUs\n"); return 2; }
    int rc = gzb_zip_fd(0, 1, &o);
    if (rc != GZB_OK) { fprintf(stderr, "gzb200: %s\n", gzb_strerror(rc)); return 1; }
    return 0;
}
