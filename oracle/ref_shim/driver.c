/* refgzip -- a ~100 line stand-in for the parts of the reference's src/gzip.c
 * that its compression hot path needs (the globals of src/gzip.c:179-244 and
 * five trivial helpers of src/util.c), linked against the reference's OWN,
 * UNMODIFIED src/zip.c, src/parallel.c, src/unzip.c and vendored zlib 1.2.11.
 *
 * TEST INFRASTRUCTURE ONLY (oracle/_ref): used by tests/, smoke() and
 * bench.py's cpu_baseline / --impl reference leg.  Never on the product path.
 *
 *   refgzip [-1..-9] [-j N] [--rsyncable] [-d]   < in > out
 *
 * stdin -> stdout only, so parallel mode writes no FNAME field
 * (src/parallel.c:496-507) and MTIME is 0 like `gzip < pipe`
 * (tests/reproducible in the reference).
 */
#include <config.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>
#include <errno.h>
#include "tailor.h"
#include "gzip.h"

/* globals the reference defines in src/gzip.c:179-244 */
uch inbuf[INBUFSIZE + INBUF_EXTRA];
uch outbuf[OUTBUFSIZE + OUTBUF_EXTRA];
uch window[2L * WSIZE];
int method = DEFLATED, level = 6, exit_code = 0, verbose = 0, quiet = 0;
int test = 0, to_stdout = 1, save_orig_name = 0, no_name = 0, pkzip = 0;
int threads = 0, rsync = 0, ifd = 0, ofd = 1;
unsigned insize = 0, inptr = 0, outcnt = 0;
off_t bytes_in = 0, bytes_out = 0, ifile_size = -1;
char ifname[1024] = "stdin", ofname[1024] = "stdout";
char *program_name = "refgzip";
struct timespec time_stamp;

extern int zip(int in, int out);
extern int unzip(int in, int out);

/* helpers the reference defines in src/util.c */
noreturn void abort_gzip(void) { exit(ERROR); }
noreturn void gzip_error(char const *m) { fprintf(stderr, "\n%s: %s: %s\n", program_name, ifname, m); abort_gzip(); }
noreturn void xalloc_die(void) { fprintf(stderr, "\n%s: memory exhausted\n", program_name); abort_gzip(); }
noreturn void read_error(void) { fprintf(stderr, "\n%s: read error: %s\n", program_name, strerror(errno)); abort_gzip(); }
noreturn void write_error(void) { fprintf(stderr, "\n%s: write error: %s\n", program_name, strerror(errno)); abort_gzip(); }
void warning(char const *m) { if (!quiet) fprintf(stderr, "%s: %s: warning: %s\n", program_name, ifname, m); if (exit_code == OK) exit_code = WARNING; }
int read_buffer(int fd, voidp buf, unsigned int cnt) { return (int)read(fd, buf, cnt); }
ulg updcrc(const uch *s, unsigned n) { (void)s; (void)n; return 0; } /* decode-side only */
void write_buf(int fd, voidp buf, unsigned cnt) { if (write(fd, buf, cnt) != (ssize_t)cnt) write_error(); bytes_out += cnt; }
void flush_window(void) {}
int fill_inbuf(int eof_ok, int max_fill) { (void)eof_ok; (void)max_fill; return EOF; }
char *gzip_base_name(char *fname) { char *p = strrchr(fname, '/'); return p ? p + 1 : fname; }   /* src/util.c; used by the GZIP_B200 seam only */

int main(int argc, char **argv)
{
    int decompress = 0;
    for (int i = 1; i < argc; i++) {
        const char *a = argv[i];
        if (a[0] == '-' && a[1] >= '1' && a[1] <= '9' && !a[2]) level = a[1] - '0';
        else if (!strcmp(a, "-j") && i + 1 < argc) threads = atoi(argv[++i]);
        else if (!strcmp(a, "--rsyncable")) rsync = 1;
        else if (!strcmp(a, "-d")) decompress = 1;
        else { fprintf(stderr, "usage: refgzip [-1..-9] [-j N] [--rsyncable] [-d] < in > out\n"); return 2; }
    }
    time_stamp.tv_sec = 0; time_stamp.tv_nsec = -1;   /* "no timestamp": MTIME=0, no range warning */
    no_name = 1;
    if (decompress) { int r = unzip(0, 1); return r == OK ? exit_code : ERROR; }
    zip(0, 1);
    return exit_code;
}
