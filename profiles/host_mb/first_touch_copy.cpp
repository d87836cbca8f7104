#include <sys/mman.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
#ifndef MADV_POPULATE_WRITE
#define MADV_POPULATE_WRITE 23
#endif
int main(int argc, char** argv) {
  int mode = atoi(argv[1]), nt = atoi(argv[2]);
  size_t n = (size_t)2 << 30;
  char* src = (char*)malloc(64 << 20); memset(src, 1, 64 << 20);
  for (int rep = 0; rep < 2; ++rep) {
    char* dst = (char*)mmap(0, n, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
    auto t0 = std::chrono::steady_clock::now();
    if (mode == 1) { int r = madvise(dst, n, MADV_HUGEPAGE); if (r) perror("hugepage"); }
    std::vector<std::thread> th;
    size_t per = n / nt;
    for (int t = 0; t < nt; ++t) th.emplace_back([=] {
      char* d = dst + per * t;
      if (mode == 2) { int r = madvise(d, per, MADV_POPULATE_WRITE); if (r) perror("populate"); }
      for (size_t o = 0; o < per; o += 64 << 20) memcpy(d + o, src, std::min<size_t>(64 << 20, per - o));
    });
    for (auto& x : th) x.join();
    double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    printf("mode %d threads %d: %.2f GB/s\n", mode, nt, n / s / 1e9);
    munmap(dst, n);
  }
}
