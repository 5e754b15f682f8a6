/* tmpfs -> memory with T threads of pread: what the host side of gzb_zip_fd can hope for on a box.
 * usage: micro_pread [MiB] [piece MiB]     (file in /dev/shm, created and removed here) */
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
static void *rd(void *a)
{
    const size_t k = (size_t)a;
    for (size_t off = k * piece; off < n; off += (size_t)T * piece) {
        size_t len = n - off < piece ? n - off : piece, done = 0;
        while (done < len) { ssize_t r = pread(fd, buf + off + done, len - done, (off_t)(off + done)); if (r <= 0) exit(2); done += (size_t)r; }
    }
    return NULL;
}
int main(int argc, char **argv)
{
    n = (size_t)(argc > 1 ? atoi(argv[1]) : 2048) << 20; piece = (size_t)(argc > 2 ? atoi(argv[2]) : 4) << 20;
    const char *path = "/dev/shm/micro_pread.bin";
    buf = malloc(n); memset(buf, 1, n);
    fd = open(path, O_CREAT | O_RDWR | O_TRUNC, 0600);
    for (size_t o = 0; o < n; o += 1 << 26) if (write(fd, buf + o, n - o < (1 << 26) ? n - o : (1 << 26)) < 0) return 1;
    long nc = sysconf(_SC_NPROCESSORS_ONLN);
    for (T = 1; T <= 2 * nc; T *= 2) {
        pthread_t th[256]; double t = now();
        for (size_t k = 0; k < (size_t)T; k++) pthread_create(&th[k], NULL, rd, (void *)k);
        for (int k = 0; k < T; k++) pthread_join(th[k], NULL);
        t = now() - t; printf("%3d threads, %zu MiB pieces: %6.2f GB/s\n", T, piece >> 20, n / t / 1e9);
    }
    { double t = now(); unsigned char *b2 = malloc(n); memcpy(b2, buf, n); t = now() - t; printf("memcpy first touch 1 thread: %.2f GB/s\n", n / t / 1e9);
      t = now(); memcpy(b2, buf, n); t = now() - t; printf("memcpy warm 1 thread: %.2f GB/s\n", n / t / 1e9); }
    close(fd); unlink(path); return 0;
}
