// Host-side sanitizer harness for the symmetric-memory runtime (csrc/comm_runtime.cu) and the TMA host utilities
// (csrc/tma_util.cu): SURVEY.md section 5.2 asks for race / memory checking of the native runtime; the device side is
// covered by the emulated-rank stress tests, this covers the host side.  Built twice by `make sanitize`
// (-fsanitize=address,undefined and -fsanitize=thread) and run on any box:
//   * without a GPU every entry point must fail cleanly (error text set, no crash, nothing leaked or raced on);
//   * with a GPU, threads create / export / import / destroy arenas concurrently (the fd hand-over that
//     parallel/symm.py::FdExchange does between processes, here between threads of one process).
#include <stdio.h>
#include <string.h>
#include <unistd.h>

#include <atomic>
#include <thread>
#include <vector>

#include "pdjob_b200.h"

static std::atomic<int> failures{0};
#define EXPECT(c)                                                       \
  do {                                                                  \
    if (!(c)) {                                                         \
      fprintf(stderr, "%s:%d: expectation failed: %s (last error: %s)\n", __FILE__, __LINE__, #c, pdj_last_error()); \
      failures++;                                                       \
    }                                                                   \
  } while (0)

static void worker(int id, bool have_gpu) {
  for (int it = 0; it < 20; ++it) {
    size_t rounded = 0;
    const int rc = pdj_arena_round_size(0, (size_t)(1 + id + it) << 20, 2, &rounded);
    if (!have_gpu) {
      EXPECT(rc != 0);
      EXPECT(strlen(pdj_last_error()) > 0);  // thread-local text, set by THIS thread's failing call
      PdjArena* a = nullptr;
      EXPECT(pdj_arena_create(0, 1 << 20, &a) != 0 && a == nullptr);
      PdjArena* b = nullptr;
      EXPECT(pdj_arena_import_fd(0, -1, 1 << 20, &b) != 0 && b == nullptr);
      continue;
    }
    EXPECT(rc == 0 && rounded >= ((size_t)(1 + id + it) << 20));
    PdjArena* a = nullptr;
    EXPECT(pdj_arena_create(0, rounded, &a) == 0 && a != nullptr);
    if (a == nullptr) continue;
    EXPECT(pdj_arena_bytes(a) == rounded && pdj_arena_ptr(a) != nullptr);
    int fd = -1;
    EXPECT(pdj_arena_export_fd(a, &fd) == 0 && fd >= 0);
    PdjArena* peer = nullptr;  // a second mapping of the same memory, as a peer rank would make
    EXPECT(pdj_arena_import_fd(0, fd, rounded, &peer) == 0 && peer != nullptr);
    if (fd >= 0) close(fd);
    if (peer) EXPECT(pdj_arena_destroy(peer) == 0);
    EXPECT(pdj_arena_destroy(a) == 0);
    PdjArena* bad = nullptr;
    EXPECT(pdj_arena_import_fd(0, -1, rounded, &bad) != 0);
  }
}

int main() {
  const bool have_gpu = pdj_device_count() > 0;
  printf("sanitize_host: %s\n", have_gpu ? "GPU present: concurrent arena create / export / import / destroy"
                                         : "no GPU: every entry point must fail cleanly");
  std::vector<std::thread> th;
  for (int i = 0; i < 8; ++i) th.emplace_back(worker, i, have_gpu);
  for (auto& t : th) t.join();
  printf("sanitize_host: %d failed expectations\n", failures.load());
  return failures.load() ? 1 : 0;
}
