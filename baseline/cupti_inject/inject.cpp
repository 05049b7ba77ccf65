// Device-time probe for the UNMODIFIED reference binary (baseline tooling, not product code).
//
// The reference has no timers and links the CUDA runtime statically, so the only
// way to learn how long its kernels ran without touching its source is the driver's own injection
// hook: when CUDA_INJECTION64_PATH names this library, the driver loads it at cuInit and calls
// InitializeInjection(). We switch on CUPTI's activity records for kernels and memcpys (start/end
// timestamps taken by the GPU, no kernel replay, no serialisation) and, at process exit, write one
// JSON line per process to $REF_CUPTI_OUT.<pid>:
//   {"pid":..,"kernels":N,"kernel_ns":sum of (end-start),"first_ns":..,"last_ns":..,
//    "h2d_bytes":..,"d2h_bytes":..,"memcpy_ns":..}
// run_ref.py takes kernel_ns per rank and the max over ranks as the reference's device time.
#include <cupti.h>

#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <unistd.h>

namespace {

std::atomic<uint64_t> g_kernels{0}, g_kernel_ns{0}, g_memcpy_ns{0}, g_h2d{0}, g_d2h{0};
std::atomic<uint64_t> g_first{~0ull}, g_last{0};
std::atomic<int> g_done{0};

void CUPTIAPI buffer_requested(uint8_t** buffer, size_t* size, size_t* max_records) {
    *size = 1 << 20;
    *buffer = static_cast<uint8_t*>(aligned_alloc(8, *size));
    *max_records = 0;
}

void atomic_min(std::atomic<uint64_t>& a, uint64_t v) {
    uint64_t cur = a.load();
    while (v < cur && !a.compare_exchange_weak(cur, v)) {}
}
void atomic_max(std::atomic<uint64_t>& a, uint64_t v) {
    uint64_t cur = a.load();
    while (v > cur && !a.compare_exchange_weak(cur, v)) {}
}

void CUPTIAPI buffer_completed(CUcontext, uint32_t, uint8_t* buffer, size_t, size_t valid) {
    CUpti_Activity* rec = nullptr;
    while (valid > 0 && cuptiActivityGetNextRecord(buffer, valid, &rec) == CUPTI_SUCCESS) {
        if (rec->kind == CUPTI_ACTIVITY_KIND_CONCURRENT_KERNEL || rec->kind == CUPTI_ACTIVITY_KIND_KERNEL) {
            const auto* k = reinterpret_cast<const CUpti_ActivityKernel9*>(rec);
            g_kernels.fetch_add(1);
            g_kernel_ns.fetch_add(k->end - k->start);
            atomic_min(g_first, k->start);
            atomic_max(g_last, k->end);
        } else if (rec->kind == CUPTI_ACTIVITY_KIND_MEMCPY) {
            const auto* m = reinterpret_cast<const CUpti_ActivityMemcpy6*>(rec);
            g_memcpy_ns.fetch_add(m->end - m->start);
            if (m->copyKind == CUPTI_ACTIVITY_MEMCPY_KIND_HTOD) g_h2d.fetch_add(m->bytes);
            if (m->copyKind == CUPTI_ACTIVITY_MEMCPY_KIND_DTOH) g_d2h.fetch_add(m->bytes);
        }
    }
    free(buffer);
}

void finish() {
    if (g_done.exchange(1)) return;
    cuptiActivityFlushAll(1);
    const char* base = getenv("REF_CUPTI_OUT");
    if (!base) return;
    char path[512];
    snprintf(path, sizeof(path), "%s.%d", base, (int)getpid());
    if (FILE* f = fopen(path, "w")) {
        fprintf(f,
                "{\"pid\":%d,\"kernels\":%llu,\"kernel_ns\":%llu,\"first_ns\":%llu,\"last_ns\":%llu,"
                "\"h2d_bytes\":%llu,\"d2h_bytes\":%llu,\"memcpy_ns\":%llu}\n",
                (int)getpid(), (unsigned long long)g_kernels.load(), (unsigned long long)g_kernel_ns.load(),
                (unsigned long long)(g_kernels.load() ? g_first.load() : 0), (unsigned long long)g_last.load(),
                (unsigned long long)g_h2d.load(), (unsigned long long)g_d2h.load(), (unsigned long long)g_memcpy_ns.load());
        fclose(f);
    }
}

}  // namespace

extern "C" int InitializeInjection(void) {
    if (cuptiActivityRegisterCallbacks(buffer_requested, buffer_completed) != CUPTI_SUCCESS) return 0;
    cuptiActivityEnable(CUPTI_ACTIVITY_KIND_CONCURRENT_KERNEL);
    cuptiActivityEnable(CUPTI_ACTIVITY_KIND_MEMCPY);
    atexit(finish);
    return 1;
}
