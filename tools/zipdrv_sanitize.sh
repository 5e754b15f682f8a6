#!/bin/bash
# ThreadSanitizer / AddressSanitizer over the fd driver (gzippier_b200/host/gzb_zip.c) on the CPU, linked with the tests' mock
# engine (tests/mock_engine.c: stored blocks) -- the host side's race / memory check where compute-sanitizer for the GPU side is
# closed on the pool (profiles/r02_sanitizer_closed.txt).  usage: tools/zipdrv_sanitize.sh > profiles/rXX_host_driver_sanitizers.txt
cd "$(dirname "$0")/.." && mkdir -p tools/build
for san in thread address; do
  gcc -O1 -g -fsanitize=$san -I include -o tools/build/drv_$san tools/zipdrv_sanitize_main.c gzippier_b200/host/gzb_zip.c tests/mock_engine.c oracle/crc32_oracle.c -lpthread || exit 1
  for cfg in "" "GZB_ZIP_NO_MMAP=1" "GZB_ZIP_NO_DIRECT=1" "MOCK_GPUS=4" "GZB_ZIP_WORKERS_PER_GPU=3" "MOCK_GPUS=3 GZB_ZIP_NO_DIRECT=1" "MOCK_FAIL_AFTER=7"; do
    ng=1; case "$cfg" in *MOCK_GPUS=4*) ng=4;; *MOCK_GPUS=3*) ng=3;; esac
    out=$(env GZB_ZIP_BATCH_MB=2 $cfg TSAN_OPTIONS="halt_on_error=0" ASAN_OPTIONS="detect_leaks=1" timeout 900 tools/build/drv_$san 40 $ng 2>&1)
    echo "-fsanitize=$san [${cfg:-defaults}] n_gpus=$ng: exit $?, $(echo "$out" | grep -c 'rc 0') of 3 calls ok, sanitizer reports: $(echo "$out" | grep -c -E 'WARNING: ThreadSanitizer|ERROR: AddressSanitizer|ERROR: LeakSanitizer')"
  done
done
