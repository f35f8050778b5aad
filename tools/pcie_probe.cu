// pcie_probe.cu — concurrent pinned H2D / D2H bandwidth of 1, 2, 4, 8 GPUs from ONE process,
// with the host buffers (a) wherever the allocating thread's pages land and (b) bound to the
// NUMA node each GPU hangs off (mbind, no libnuma).  One cudaMemcpyAsync per 64 MB chunk.
//   nvcc -O2 -std=c++17 -o /tmp/pcie_probe tools/pcie_probe.cu && /tmp/pcie_probe [MB per GPU]
#include <cuda_runtime.h>
#include <sys/mman.h>
#include <sys/syscall.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <string>
#include <thread>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

static int gpu_numa_node(int dev) {
    char bus[32];
    if (cudaDeviceGetPCIBusId(bus, sizeof bus, dev) != cudaSuccess) return -1;
    for (char* p = bus; *p; ++p) *p = (char)tolower(*p);
    std::ifstream f(std::string("/sys/bus/pci/devices/") + bus + "/numa_node");
    int node = -1;
    f >> node;
    return node;
}

static long mbind_node(void* p, size_t bytes, int node) {
    unsigned long mask[16] = {0};
    mask[node / 64] = 1ul << (node % 64);
    return syscall(SYS_mbind, p, bytes, 2 /*MPOL_BIND*/, mask, 1024ul, 0u);
}

struct Buf { void* h = nullptr; void* d = nullptr; cudaStream_t up, down; };

int main(int argc, char** argv) {
    const size_t mb = argc > 1 ? (size_t)atol(argv[1]) : 1024;
    const size_t bytes = mb << 20, chunk = 64u << 20;
    int n = 0;
    CK(cudaGetDeviceCount(&n));
    printf("devices %d, %zu MB per GPU per direction, 64 MB chunks\n", n, mb);
    for (int g = 0; g < n; ++g) printf("  gpu %d numa node %d\n", g, gpu_numa_node(g));
    for (int placed = 0; placed < 2; ++placed) {
        std::vector<Buf> up((size_t)n), dn((size_t)n);
        for (int g = 0; g < n; ++g) {
            CK(cudaSetDevice(g));
            for (Buf* b : {&up[(size_t)g], &dn[(size_t)g]}) {
                b->h = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
                if (placed && gpu_numa_node(g) >= 0 && mbind_node(b->h, bytes, gpu_numa_node(g)) != 0) perror("mbind");
                memset(b->h, 1, bytes);
                CK(cudaHostRegister(b->h, bytes, cudaHostRegisterPortable));
                CK(cudaMalloc(&b->d, bytes));
                CK(cudaStreamCreateWithFlags(&b->up, cudaStreamNonBlocking));
                CK(cudaStreamCreateWithFlags(&b->down, cudaStreamNonBlocking));
            }
        }
        printf("host buffers %s\n", placed ? "bound to each GPU's NUMA node (mbind)" : "placed by first touch of the main thread");
        for (int mode = 0; mode < 3; ++mode) {   // 0: H2D, 1: D2H, 2: both
            for (int G = 1; G <= n; G *= 2) {
                auto run = [&](int g) {
                    CK(cudaSetDevice(g));
                    for (size_t off = 0; off < bytes; off += chunk) {
                        if (mode != 1) CK(cudaMemcpyAsync((char*)up[(size_t)g].d + off, (char*)up[(size_t)g].h + off, chunk, cudaMemcpyHostToDevice, up[(size_t)g].up));
                        if (mode != 0) CK(cudaMemcpyAsync((char*)dn[(size_t)g].h + off, (char*)dn[(size_t)g].d + off, chunk, cudaMemcpyDeviceToHost, dn[(size_t)g].down));
                    }
                    CK(cudaStreamSynchronize(up[(size_t)g].up));
                    CK(cudaStreamSynchronize(dn[(size_t)g].down));
                };
                double best = 1e30;
                for (int rep = 0; rep < 3; ++rep) {
                    const auto t0 = std::chrono::steady_clock::now();
                    std::vector<std::thread> th;
                    for (int g = 0; g < G; ++g) th.emplace_back(run, g);
                    for (auto& t : th) t.join();
                    const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
                    if (s < best) best = s;
                }
                const double per_dir = (double)bytes * G / best / 1e9;
                printf("  %-8s G=%d: %7.2f ms  %7.1f GB/s per direction total, %6.1f per GPU\n",
                       mode == 0 ? "H2D" : mode == 1 ? "D2H" : "H2D+D2H", G, best * 1e3, per_dir, per_dir / G);
            }
        }
        for (int g = 0; g < n; ++g)
            for (Buf* b : {&up[(size_t)g], &dn[(size_t)g]}) {
                CK(cudaSetDevice(g));
                CK(cudaHostUnregister(b->h));
                munmap(b->h, bytes);
                CK(cudaFree(b->d));
                CK(cudaStreamDestroy(b->up));
                CK(cudaStreamDestroy(b->down));
            }
    }
    return 0;
}
