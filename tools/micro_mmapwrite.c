/* memory -> a FRESH tmpfs file through a shared mapping (ftruncate to the bound, memcpy from T threads, ftruncate to the
 * size): page faults allocate the pages, no inode lock.  usage: micro_mmapwrite [MiB] [piece KiB] */
#define _GNU_SOURCE
#include <fcntl.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <time.h>
#include <unistd.h>
static unsigned char *buf, *map; static size_t n, piece; static int T;
static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }
static void *wr(void *a)
{
    const size_t k = (size_t)a;
    for (size_t off = k * piece; off < n; off += (size_t)T * piece) memcpy(map + off, buf + off, n - off < piece ? n - off : piece);
    return NULL;
}
int main(int argc, char **argv)
{
    n = (size_t)(argc > 1 ? atoi(argv[1]) : 2048) << 20; piece = (size_t)(argc > 2 ? atoi(argv[2]) : 2048) << 10;
    const char *path = "/dev/shm/micro_mmapwrite.bin";
    buf = malloc(n); memset(buf, 1, n);
    long nc = sysconf(_SC_NPROCESSORS_ONLN);
    for (T = 1; T <= 2 * nc; T *= 2) {
        unlink(path);
        int fd = open(path, O_CREAT | O_RDWR, 0600);
        double t = now();
        if (ftruncate(fd, (off_t)(4 * n))) return 1;                       /* the bound, well above what gets written */
        map = mmap(NULL, 4 * n, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
        if (map == MAP_FAILED) return 2;
        pthread_t th[256];
        for (size_t k = 0; k < (size_t)T; k++) pthread_create(&th[k], NULL, wr, (void *)k);
        for (int k = 0; k < T; k++) pthread_join(th[k], NULL);
        munmap(map, 4 * n);
        if (ftruncate(fd, (off_t)n)) return 1;
        t = now() - t; printf("fresh file via mmap, %3d threads, %zu KiB pieces: %6.2f GB/s\n", T, piece >> 10, n / t / 1e9);
        close(fd);
    }
    unlink(path); return 0;
}
