// athread_emu.cpp — implementation of the emulator declared in emu/athread.h.
// Test infrastructure only (see the header).
#include "athread.h"

#include <pthread.h>
#include <sched.h>
#include <sys/mman.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace {
constexpr int kCpes = 64, kMesh = 8;
constexpr size_t kStack = 1u << 20;  // per-CPE stack
constexpr size_t kHeap = 4u << 20;   // per-CPE ldm_malloc heap (the real LDM holds 256 KB)
constexpr size_t kArena = kStack + kHeap;

char* g_slab = nullptr;  // kCpes arenas: [stack | heap]
pthread_t g_threads[kCpes];
pthread_barrier_t g_barrier;
bool g_running = false;
void (*g_fn)(void*) = nullptr;
void* g_arg = nullptr;
ptrdiff_t g_marker_off[kCpes];
bool g_failed = false;  // the CPE stacks were not congruent: the kernel was not run

__thread int t_id = -1;
__thread size_t t_heap_top = 0;

char* arena(int id) { return g_slab + (size_t)id * kArena; }

// A local address of the calling CPE -> the same object in CPE `id`.
void* translate(void* p, int id) {
    char* base = arena(t_id);
    char* c = static_cast<char*>(p);
    if (c < base || c >= base + kArena) {
        std::fprintf(stderr, "athread emulator: address %p is not local to CPE %d\n", p, t_id);
        std::abort();
    }
    return arena(id) + (c - base);
}

void strided(char* dense, char* mem, size_t len, size_t bsize, size_t stride, bool to_dense) {
    if (bsize == 0 || bsize > len) bsize = len;
    for (size_t done = 0; done < len; done += bsize, dense += bsize, mem += bsize + stride) {
        const size_t n = len - done < bsize ? len - done : bsize;
        if (to_dense) std::memcpy(dense, mem, n);
        else std::memcpy(mem, dense, n);
    }
}

void bump(athread_rply_t* r) { __atomic_fetch_add(const_cast<int*>(r), 1, __ATOMIC_RELEASE); }

void wait_value(athread_rply_t* r, int value) {
    while (__atomic_load_n(const_cast<int*>(r), __ATOMIC_ACQUIRE) < value) sched_yield();
}

void* trampoline(void* idp) {
    volatile char marker = 0;
    t_id = (int)(intptr_t)idp;
    t_heap_top = 0;
    sb200_emu_row = t_id / kMesh;
    sb200_emu_col = t_id % kMesh;
    g_marker_off[t_id] = const_cast<char*>(&marker) - arena(t_id);
    pthread_barrier_wait(&g_barrier);
    // identical stack layout in every CPE is what makes stack addresses translatable; every
    // CPE takes the same decision, so nobody is left waiting at a barrier
    bool congruent = true;
    for (int i = 1; i < kCpes; ++i) congruent = congruent && g_marker_off[i] == g_marker_off[0];
    if (!congruent) {
        g_failed = true;
        return nullptr;
    }
    g_fn(g_arg);
    return nullptr;
}
}  // namespace

extern "C" {
__thread int sb200_emu_row = 0, sb200_emu_col = 0;

int athread_init(void) { return 0; }
// 1 if the last spawn could not run its kernel (reported by athread_join as -2 too); the
// reference's Stencil::run ignores athread_join's result, so the driver asks here
static bool g_last_failed = false;
int sb200_emu_last_spawn_failed(void) { return g_last_failed ? 1 : 0; }

int athread_spawn(void* fn, void* arg) {
    if (g_running) return -1;
    if (!g_slab) {
        void* p = mmap(nullptr, kArena * kCpes, PROT_READ | PROT_WRITE,
                       MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        if (p == MAP_FAILED) return -1;
        g_slab = static_cast<char*>(p);
    }
    g_fn = reinterpret_cast<void (*)(void*)>(fn);
    g_arg = arg;
    pthread_barrier_init(&g_barrier, nullptr, kCpes);
    for (int i = 0; i < kCpes; ++i) {
        pthread_attr_t a;
        pthread_attr_init(&a);
        pthread_attr_setstack(&a, arena(i), kStack);
        if (pthread_create(&g_threads[i], &a, trampoline, (void*)(intptr_t)i) != 0) {
            std::fprintf(stderr, "athread emulator: pthread_create failed\n");
            std::abort();
        }
        pthread_attr_destroy(&a);
    }
    g_running = true;
    return 0;
}

int athread_join(void) {
    if (!g_running) return -1;
    for (int i = 0; i < kCpes; ++i) pthread_join(g_threads[i], nullptr);
    pthread_barrier_destroy(&g_barrier);
    g_running = false;
    g_last_failed = g_failed;
    if (g_failed) {
        g_failed = false;
        return -2;
    }
    return 0;
}

void* ldm_malloc(size_t bytes) {
    const size_t need = (bytes + 63) & ~(size_t)63;
    if (t_heap_top + need > kHeap) return nullptr;
    void* p = arena(t_id) + kStack + t_heap_top;
    t_heap_top += need;
    return p;
}
// The kernels allocate once up front and free at exit: a bump allocator is enough.
void ldm_free(void*, size_t) {}

void athread_dma_get(void* d, void* s, size_t len) { std::memcpy(d, s, len); }
void athread_dma_put(void* d, void* s, size_t len) { std::memcpy(d, s, len); }
void athread_dma_get_stride(void* d, void* s, size_t len, size_t bsize, size_t stride) {
    strided(static_cast<char*>(d), static_cast<char*>(s), len, bsize, stride, true);
}
void athread_dma_put_stride(void* d, void* s, size_t len, size_t bsize, size_t stride) {
    strided(static_cast<char*>(s), static_cast<char*>(d), len, bsize, stride, false);
}
void athread_dma_iget(void* d, void* s, size_t len, athread_rply_t* reply) {
    std::memcpy(d, s, len);
    bump(reply);
}
void athread_dma_iput(void* d, void* s, size_t len, athread_rply_t* reply) {
    std::memcpy(d, s, len);
    bump(reply);
}
void athread_dma_iget_stride(void* d, void* s, size_t len, size_t bsize, size_t stride, athread_rply_t* reply) {
    athread_dma_get_stride(d, s, len, bsize, stride);
    bump(reply);
}
void athread_dma_iput_stride(void* d, void* s, size_t len, size_t bsize, size_t stride, athread_rply_t* reply) {
    athread_dma_put_stride(d, s, len, bsize, stride);
    bump(reply);
}
void athread_dma_wait_value(athread_rply_t* reply, int value) { wait_value(reply, value); }

void athread_rma_iput(void* src, athread_rply_t* local_reply, size_t len, int remote_id, void* remote_dst,
                      athread_rply_t* remote_reply) {
    std::memcpy(translate(remote_dst, remote_id), src, len);
    bump(static_cast<athread_rply_t*>(translate(const_cast<int*>(remote_reply), remote_id)));
    bump(local_reply);
}
void athread_rma_wait_value(athread_rply_t* reply, int value) { wait_value(reply, value); }

void athread_ssync_array(void) { pthread_barrier_wait(&g_barrier); }
}
