// Stress driver for the host-side input pipeline (csrc/io/record_loader.cpp, crc32c.cpp), built by tools/sanitize_host.sh
// under ThreadSanitizer and AddressSanitizer + UBSan: three reader threads, the caller's dequeue, then the batch ring with
// its own thread, and a destroy while all of them are busy - three rounds.  Usage: loader_stress <file.tfrecords> [<corrupt.tfrecords>]
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
extern "C" {
void* dtm_loader_create(const char* const* files, int n_files, int batch, int image_bytes, int num_threads, uint64_t seed, int shuffle, int min_after);
int dtm_loader_next(void*, uint8_t*, int32_t*);
int dtm_loader_ring_start(void*, uint8_t* const*, int, uint64_t);
int dtm_loader_ring_acquire(void*);
int dtm_loader_ring_release(void*, int);
uint64_t dtm_loader_records_read(void*);
void dtm_loader_destroy(void*);
}
namespace dtm { int dtm_fail(const char* m) { fprintf(stderr, "fail: %s\n", m); return -1; } }
int main(int argc, char** argv) {
  const char* f[1] = {argv[1]};
  for (int round = 0; round < 3; ++round) {
    void* h = dtm_loader_create(f, 1, 64, 784, 3, round, 1, 500);
    std::vector<uint8_t> px(64 * 784); std::vector<int32_t> lab(64);
    for (int i = 0; i < 50; ++i) if (dtm_loader_next(h, px.data(), lab.data())) return 1;
    std::vector<std::vector<uint8_t>> bufs(3, std::vector<uint8_t>(64 * 784 + 64 * 4));
    uint8_t* p[3] = {bufs[0].data(), bufs[1].data(), bufs[2].data()};
    if (dtm_loader_ring_start(h, p, 3, 64 * 784)) return 2;
    for (int i = 0; i < 200; ++i) { int idx = dtm_loader_ring_acquire(h); if (idx < 0) return 3; volatile uint8_t v = p[idx][i]; (void)v; dtm_loader_ring_release(h, idx); }
    printf("round %d read %lu\n", round, (unsigned long)dtm_loader_records_read(h));
    dtm_loader_destroy(h);                       // while producers and the ring thread are busy
  }
  if (argc > 2) {                                // a corrupt file: the error must surface, and tear-down must stay clean
    const char* g[2] = {argv[1], argv[2]};
    void* h = dtm_loader_create(g, 2, 64, 784, 3, 1, 1, 500);
    std::vector<uint8_t> px(64 * 784); std::vector<int32_t> lab(64);
    int failed = 0;
    for (int i = 0; i < 400 && !failed; ++i) failed = dtm_loader_next(h, px.data(), lab.data()) != 0;
    printf("corrupt file: %s\n", failed ? "error reported" : "NOT DETECTED");
    dtm_loader_destroy(h);
    if (!failed) return 4;
  }
  return 0;
}
