// tools/ctx_time.cu -- where does opening N devices go: driver initialisation (cuInit, grows with the
// GPUs in the box) or the primary contexts, and do contexts created by separate threads overlap?
//   ctx_time            serial        ctx_time p     one thread per device
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <thread>
#include <vector>
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main(int argc, char** argv) {
  const bool parallel = argc > 1 && argv[1][0] == 'p';
  double t0 = now();
  int n = 0;
  cudaGetDeviceCount(&n);
  double t1 = now();
  std::printf("cudaGetDeviceCount (driver init): %.3f s, %d devices\n", t1 - t0, n);
  if (parallel) {
    std::vector<std::thread> pool;
    for (int d = 0; d < n; d++) pool.emplace_back([d] { cudaSetDevice(d); cudaFree(nullptr); });
    for (auto& t : pool) t.join();
    std::printf("%d primary contexts, one thread each: %.3f s\n", n, now() - t1);
  } else {
    for (int d = 0; d < n; d++) {
      double a = now();
      cudaSetDevice(d);
      cudaFree(nullptr);
      std::printf("primary context of device %d: %.3f s\n", d, now() - a);
    }
    std::printf("%d primary contexts, serial: %.3f s\n", n, now() - t1);
  }
  return 0;
}
