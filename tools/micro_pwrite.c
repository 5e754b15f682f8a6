/* memory -> a FRESH tmpfs file with T threads of pwrite (pages are allocated and zeroed by the write), and the same into a
 * file that already has its pages.  usage: micro_pwrite [MiB] [piece KiB] */
#define _GNU_SOURCE
#include <fcntl.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>
static int fd; static unsigned char *buf; static size_t n, piece; static int T;
static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }
static void *wr(void *a)
{
    const size_t k = (size_t)a;
    for (size_t off = k * piece; off < n; off += (size_t)T * piece) {
        size_t len = n - off < piece ? n - off : piece, done = 0;
        while (done < len) { ssize_t r = pwrite(fd, buf + off + done, len - done, (off_t)(off + done)); if (r <= 0) exit(2); done += (size_t)r; }
    }
    return NULL;
}
int main(int argc, char **argv)
{
    n = (size_t)(argc > 1 ? atoi(argv[1]) : 2048) << 20; piece = (size_t)(argc > 2 ? atoi(argv[2]) : 2048) << 10;
    const char *path = "/dev/shm/micro_pwrite.bin";
    buf = malloc(n); memset(buf, 1, n);
    long nc = sysconf(_SC_NPROCESSORS_ONLN);
    for (int fresh = 1; fresh >= 0; fresh--)
        for (T = 1; T <= 2 * nc; T *= 2) {
            if (fresh) unlink(path);
            fd = open(path, O_CREAT | O_RDWR, 0600);
            pthread_t th[256]; double t = now();
            for (size_t k = 0; k < (size_t)T; k++) pthread_create(&th[k], NULL, wr, (void *)k);
            for (int k = 0; k < T; k++) pthread_join(th[k], NULL);
            t = now() - t; printf("%s file, %3d threads, %zu KiB pieces: %6.2f GB/s\n", fresh ? "fresh   " : "existing", T, piece >> 10, n / t / 1e9);
            close(fd);
        }
    unlink(path); return 0;
}
