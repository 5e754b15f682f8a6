/* memory -> a FRESH tmpfs file, "reserve then fill": fallocate() a fragment's range (serial; a full file system is an error
 * code here, not a SIGBUS later), then memcpy into a shared mapping from T threads (minor faults only).
 * usage: micro_fallocwrite [MiB] [fragment MiB] */
#define _GNU_SOURCE
#include <fcntl.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <time.h>
#include <unistd.h>
static unsigned char *buf, *map; static size_t n, frag; static int T, fd;
static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }
static size_t reserved;             /* bytes fallocate()d so far (one thread reserves, the others fill behind it) */
static pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER; static pthread_cond_t cv = PTHREAD_COND_INITIALIZER;
static void *reserver(void *a)
{
    (void)a;
    for (size_t off = 0; off < n; off += frag) {
        size_t len = n - off < frag ? n - off : frag;
        if (fallocate(fd, 0, (off_t)off, (off_t)len)) { perror("fallocate"); exit(3); }
        pthread_mutex_lock(&mu); reserved = off + len; pthread_cond_broadcast(&cv); pthread_mutex_unlock(&mu);
    }
    return NULL;
}
static void *filler(void *a)
{
    const size_t k = (size_t)a, piece = 2u << 20;
    for (size_t off = k * piece; off < n; off += (size_t)T * piece) {
        size_t len = n - off < piece ? n - off : piece;
        pthread_mutex_lock(&mu); while (reserved < off + len) pthread_cond_wait(&cv, &mu); pthread_mutex_unlock(&mu);
        memcpy(map + off, buf + off, len);
    }
    return NULL;
}
int main(int argc, char **argv)
{
    n = (size_t)(argc > 1 ? atoi(argv[1]) : 2048) << 20; frag = (size_t)(argc > 2 ? atoi(argv[2]) : 27) << 20;
    const char *path = "/dev/shm/micro_fallocwrite.bin";
    buf = malloc(n); memset(buf, 1, n);
    long nc = sysconf(_SC_NPROCESSORS_ONLN);
    { unlink(path); fd = open(path, O_CREAT | O_RDWR, 0600); double t = now();
      for (size_t off = 0; off < n; off += frag) if (fallocate(fd, 0, (off_t)off, (off_t)(n - off < frag ? n - off : frag))) { perror("fallocate"); return 3; }
      t = now() - t; printf("fallocate alone (fresh file, %zu MiB ranges): %6.2f GB/s\n", frag >> 20, n / t / 1e9); close(fd); }
    for (T = 1; T <= 2 * nc; T *= 2) {
        unlink(path);
        fd = open(path, O_CREAT | O_RDWR, 0600);
        reserved = 0;
        double t = now();
        map = mmap(NULL, n, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
        if (map == MAP_FAILED) return 2;
        pthread_t th[256], rt;
        pthread_create(&rt, NULL, reserver, NULL);
        for (size_t k = 0; k < (size_t)T; k++) pthread_create(&th[k], NULL, filler, (void *)k);
        for (int k = 0; k < T; k++) pthread_join(th[k], NULL);
        pthread_join(rt, NULL);
        munmap(map, n);
        t = now() - t; printf("reserve then fill, fresh file, %3d filler threads: %6.2f GB/s\n", T, n / t / 1e9);
        close(fd);
    }
    unlink(path); return 0;
}
