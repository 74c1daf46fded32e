// Micro-benchmark: what does the BOX allow for device -> pinned-host copies when 1, 2, 4 .. all GPUs copy at once?
// (The end-to-end number of bench.py is a batched get into pinned host memory: it can never beat this.)
// Plain cudaMemcpyAsync of 256 MiB per copy, one host thread and one stream per GPU, all started behind a barrier.
// Three kinds of pinned buffer:
//   default   cudaHostAlloc(cudaHostAllocDefault) from the GPU's thread, wherever the OS puts the pages
//   portable  cudaHostAlloc(cudaHostAllocPortable)
//   numa      the thread first pins itself to the CPUs of the GPU's NUMA node (sysfs), then allocates and touches
// Prints one JSON line per (kind, active GPUs): per-GPU GB/s min / mean, and the sum.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o host_d2h_ceiling host_d2h_ceiling.cu -lpthread
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include <sched.h>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

static int numa_node_of(int dev) {
  char bus[32] = "";
  if (cudaDeviceGetPCIBusId(bus, sizeof(bus), dev) != cudaSuccess) return -1;
  for (char* c = bus; *c; ++c) *c = (char)tolower(*c);
  std::string path = std::string("/sys/bus/pci/devices/") + bus + "/numa_node";
  FILE* f = fopen(path.c_str(), "r");
  if (!f) return -1;
  int node = -1;
  if (fscanf(f, "%d", &node) != 1) node = -1;
  fclose(f);
  return node;
}

static bool pin_to_node(int node) {
  if (node < 0) return false;
  char path[128];
  snprintf(path, sizeof(path), "/sys/devices/system/node/node%d/cpulist", node);
  FILE* f = fopen(path, "r");
  if (!f) return false;
  char line[4096] = "";
  if (!fgets(line, sizeof(line), f)) { fclose(f); return false; }
  fclose(f);
  cpu_set_t set;
  CPU_ZERO(&set);
  for (char* tok = strtok(line, ",\n"); tok; tok = strtok(nullptr, ",\n")) {
    int lo = 0, hi = 0;
    if (sscanf(tok, "%d-%d", &lo, &hi) == 2) { for (int c = lo; c <= hi; ++c) CPU_SET(c, &set); }
    else if (sscanf(tok, "%d", &lo) == 1) CPU_SET(lo, &set);
  }
  return sched_setaffinity(0, sizeof(set), &set) == 0;
}

int main() {
  int n = 0;
  CK(cudaGetDeviceCount(&n));
  if (n > 8) n = 8;
  const size_t bytes = 256ull << 20;
  const int reps = 8;
  std::vector<uint8_t*> dbuf(n);
  std::vector<cudaStream_t> st(n);
  std::vector<cudaEvent_t> e0(n), e1(n);
  std::vector<int> node(n);
  for (int g = 0; g < n; ++g) {
    CK(cudaSetDevice(g));
    CK(cudaMalloc(&dbuf[g], bytes));
    CK(cudaMemset(dbuf[g], g + 1, bytes));
    CK(cudaStreamCreateWithFlags(&st[g], cudaStreamNonBlocking));
    CK(cudaEventCreate(&e0[g]));
    CK(cudaEventCreate(&e1[g]));
    node[g] = numa_node_of(g);
  }
  printf("{\"gpus\": %d, \"host_cpus\": %u, \"gpu_numa_nodes\": [", n, std::thread::hardware_concurrency());
  for (int g = 0; g < n; ++g) printf("%d%s", node[g], g + 1 < n ? ", " : "");
  printf("]}\n");
  const char* kinds[] = {"default", "portable", "numa"};
  for (int kind = 0; kind < 3; ++kind) {
    for (int dir = 0; dir < 2; ++dir) {  // 0: device -> host, 1: host -> device
      for (int active = 1; active <= n; active *= 2) {
        std::vector<double> gbps(active, 0.0);
        std::atomic<int> ready{0};
        std::vector<int> pinned_ok(active, 0);
        auto work = [&](int g) {
          CK(cudaSetDevice(g));
          if (kind == 2) pinned_ok[g] = pin_to_node(node[g]) ? 1 : 0;
          uint8_t* h = nullptr;
          CK(cudaHostAlloc(&h, bytes, kind == 1 ? cudaHostAllocPortable : cudaHostAllocDefault));
          memset(h, 1, bytes);
          for (int w = 0; w < 2; ++w)
            CK(cudaMemcpyAsync(dir ? (void*)dbuf[g] : (void*)h, dir ? (void*)h : (void*)dbuf[g], bytes, dir ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost, st[g]));
          CK(cudaStreamSynchronize(st[g]));
          ready.fetch_add(1);
          while (ready.load() < active) {}
          CK(cudaEventRecord(e0[g], st[g]));
          for (int r = 0; r < reps; ++r)
            CK(cudaMemcpyAsync(dir ? (void*)dbuf[g] : (void*)h, dir ? (void*)h : (void*)dbuf[g], bytes, dir ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost, st[g]));
          CK(cudaEventRecord(e1[g], st[g]));
          CK(cudaEventSynchronize(e1[g]));
          float ms = 0;
          CK(cudaEventElapsedTime(&ms, e0[g], e1[g]));
          gbps[g] = (double)reps * bytes / (ms * 1e-3) / 1e9;
          CK(cudaFreeHost(h));
        };
        std::vector<std::thread> th;
        for (int g = 0; g < active; ++g) th.emplace_back(work, g);
        for (auto& t : th) t.join();
        double mn = 1e30, sum = 0;
        for (double v : gbps) { mn = v < mn ? v : mn; sum += v; }
        printf("{\"pinned\": \"%s\", \"dir\": \"%s\", \"active_gpus\": %d, \"per_gpu_GBps_min\": %.2f, \"per_gpu_GBps_mean\": %.2f, \"sum_GBps\": %.1f%s}\n",
               kinds[kind], dir ? "h2d" : "d2h", active, mn, sum / active, sum, kind == 2 ? (pinned_ok[0] ? ", \"affinity_set\": true" : ", \"affinity_set\": false") : "");
        fflush(stdout);
      }
    }
  }
  return 0;
}
