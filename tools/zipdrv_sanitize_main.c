#define _GNU_SOURCE
#include "gzb200.h"
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>
/* three gzb_zip_fd calls file to file on tmpfs (the second and third find the pool warm); exit code = calls that failed */
int main(int argc, char **argv)
{
    const size_t n = (size_t)(argc > 1 ? atoi(argv[1]) : 40) << 20;
    unsigned char *d = malloc(n); unsigned x = 12345; for (size_t i = 0; i < n; i++) { x = x * 1664525u + 1013904223u; d[i] = (unsigned char)(x >> 24); }
    const char *src = "/dev/shm/gzb_tsan_in.bin", *dst = "/dev/shm/gzb_tsan_out.gz";
    int fd = open(src, O_CREAT | O_TRUNC | O_WRONLY, 0600); if (write(fd, d, n) != (ssize_t)n) return 2; close(fd);
    int bad = 0;
    for (int rep = 0; rep < 3; rep++) {
        int ifd = open(src, O_RDONLY), ofd = open(dst, O_CREAT | O_TRUNC | O_WRONLY, 0600);
        gzb_zip_opts o; memset(&o, 0, sizeof o); o.level = 6; o.n_gpus = argc > 2 ? atoi(argv[2]) : 1;
        int rc = gzb_zip_fd(ifd, ofd, &o);
        printf("rep %d rc %d in %llu out %llu\n", rep, rc, (unsigned long long)o.bytes_in, (unsigned long long)o.bytes_out); fflush(stdout);
        if (rc != 0 || o.bytes_in != n) bad++;
        close(ifd); close(ofd);
    }
    gzb_zip_release();
    unlink(src); unlink(dst);
    free(d); fflush(stdout);
    return bad;
}
