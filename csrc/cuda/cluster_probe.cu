// Thread-block clusters for the TILE kernel's start-up? A measured experiment.
//
// A TILE item's CTA starts by fetching what it needs into shared memory with TMA bulk copies: the 4 KB
// pair-value table (the same for every CTA), a Seq1 window (the same for the CTAs that work on the same
// offsets) and its own Seq2 letters. In a cluster the common part could be fetched ONCE and delivered to
// all CTAs of the cluster by one multicast bulk copy (cp.async.bulk ... .multicast::cluster) instead of once
// per CTA. This program measures what that buys on a cold L2, which is the regime of the 18 us solve:
//   variant A: plain launch, every CTA copies common + private bytes itself (what align_kernel.cu does)
//   variant B: cluster of C CTAs, CTA 0 multicasts the common bytes to all C, every CTA copies its private bytes
//   variant C: cluster of C CTAs, no multicast (the cost of the cluster launch + cluster barrier alone)
// per CTA: nanoseconds from kernel entry until all its bytes have landed (%globaltimer); per launch: CUDA events.
// TMA multicast loads here read only this GPU's memory (one cudaMalloc'ed buffer).
//
// Build: make cluster_probe      Run (on a B200): build/cluster_probe > profiles/cluster_probe.json
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "ptx_helpers.cuh"

namespace cg = cooperative_groups;
using namespace sa_ptx;

#define CHECK(x)                                                             \
    do {                                                                     \
        cudaError_t e__ = (x);                                               \
        if (e__ != cudaSuccess) {                                            \
            fprintf(stderr, "%s failed: %s\n", #x, cudaGetErrorString(e__)); \
            exit(1);                                                         \
        }                                                                    \
    } while (0)

constexpr uint32_t kCommon = 4096 + 2048;  // table + Seq1 window
constexpr uint32_t kPrivate = 1536;        // this CTA's Seq2 letters
constexpr int kThreads = 512;

__device__ __forceinline__ void bulk_g2s_multicast(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar, uint16_t mask) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;" ::"r"(dst),
        "l"(src), "r"(bytes), "r"(bar), "h"(mask)
        : "memory");
}

template <int kCluster, bool kMulticast>
__global__ void __launch_bounds__(kThreads, 1) probe_kernel(const unsigned char* common, const unsigned char* priv, unsigned long long* landed_ns,
                                                            uint32_t* checksum) {
    __shared__ __align__(128) unsigned char s_common[kCommon];
    __shared__ __align__(128) unsigned char s_priv[kPrivate];
    __shared__ __align__(8) unsigned long long s_bar;
    const unsigned long long t0 = global_ns();
    const uint32_t bar = smem_u32(&s_bar);
    uint32_t rank = 0;
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    if (kCluster > 1) {
        cg::cluster_group cl = cg::this_cluster();
        rank = cl.block_rank();
        cl.sync();  // every CTA's barrier exists before a peer's copy can signal it
    } else {
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        mbar_expect_tx(bar, kCommon + kPrivate);
        bulk_g2s(smem_u32(s_priv), priv + (size_t)blockIdx.x * kPrivate, kPrivate, bar);
        if (!kMulticast) bulk_g2s(smem_u32(s_common), common, kCommon, bar);
        else if (rank == 0) bulk_g2s_multicast(smem_u32(s_common), common, kCommon, bar, (uint16_t)((1u << kCluster) - 1u));
    }
    mbar_wait(bar, 0);
    const unsigned long long t1 = global_ns();
    uint32_t acc = 0;
    for (uint32_t i = threadIdx.x; i < kCommon; i += kThreads) acc += s_common[i];
    for (uint32_t i = threadIdx.x; i < kPrivate; i += kThreads) acc += s_priv[i] * 3u;
    atomicAdd(&checksum[blockIdx.x], acc);
    if (threadIdx.x == 0) landed_ns[blockIdx.x] = t1 - t0;
    if (kCluster > 1) cg::this_cluster().sync();  // nobody leaves while a peer may still be written to
}

__global__ void flush_kernel(uint4* buf, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) buf[i] = make_uint4(i, 1, 2, 3);
}

template <int kCluster, bool kMulticast>
void run(const char* name, int ctas, const unsigned char* d_common, const unsigned char* d_priv, unsigned long long* d_ns, uint32_t* d_sum,
         uint32_t want_common, const std::vector<uint32_t>& want_priv, uint4* d_flush, size_t flush_n, bool last) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(ctas);
    cfg.blockDim = dim3(kThreads);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kCluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = kCluster > 1 ? 1 : 0;
    cudaEvent_t e0, e1;
    CHECK(cudaEventCreate(&e0));
    CHECK(cudaEventCreate(&e1));
    std::vector<float> ev;
    std::vector<unsigned long long> med, mx;
    bool ok = true;
    for (int it = 0; it < 25; ++it) {
        CHECK(cudaMemset(d_sum, 0, sizeof(uint32_t) * ctas));
        flush_kernel<<<592, 256>>>(d_flush, flush_n);  // cold L2, as in a fresh solve
        CHECK(cudaDeviceSynchronize());
        CHECK(cudaEventRecord(e0));
        CHECK(cudaLaunchKernelEx(&cfg, probe_kernel<kCluster, kMulticast>, d_common, d_priv, d_ns, d_sum));
        CHECK(cudaEventRecord(e1));
        CHECK(cudaDeviceSynchronize());
        float ms = 0;
        CHECK(cudaEventElapsedTime(&ms, e0, e1));
        std::vector<unsigned long long> ns(ctas);
        std::vector<uint32_t> sum(ctas);
        CHECK(cudaMemcpy(ns.data(), d_ns, sizeof(unsigned long long) * ctas, cudaMemcpyDeviceToHost));
        CHECK(cudaMemcpy(sum.data(), d_sum, sizeof(uint32_t) * ctas, cudaMemcpyDeviceToHost));
        for (int c = 0; c < ctas; ++c) ok = ok && sum[c] == want_common + 3u * want_priv[c];
        if (it < 5) continue;  // warm-up
        std::sort(ns.begin(), ns.end());
        ev.push_back(ms * 1000.f);
        med.push_back(ns[ctas / 2]);
        mx.push_back(ns.back());
    }
    std::sort(ev.begin(), ev.end());
    std::sort(med.begin(), med.end());
    std::sort(mx.begin(), mx.end());
    printf("  \"%s\": {\"cluster\": %d, \"multicast\": %s, \"ctas\": %d, \"bytes_landed_ns_median_cta\": %llu, \"bytes_landed_ns_slowest_cta\": %llu, "
           "\"launch_us_events\": %.2f, \"contents_ok\": %s}%s\n",
           name, kCluster, kMulticast ? "true" : "false", ctas, med[med.size() / 2], mx[mx.size() / 2], ev[ev.size() / 2], ok ? "true" : "false",
           last ? "" : ",");
}

int main() {
    const int ctas = 144;  // divisible by every cluster size tried; one CTA per SM
    std::vector<unsigned char> h_common(kCommon), h_priv((size_t)ctas * kPrivate);
    uint32_t want_common = 0;
    for (uint32_t i = 0; i < kCommon; ++i) want_common += (h_common[i] = (unsigned char)(i * 7 + 3));
    std::vector<uint32_t> want_priv(ctas, 0);
    for (int c = 0; c < ctas; ++c)
        for (uint32_t i = 0; i < kPrivate; ++i) want_priv[c] += (h_priv[(size_t)c * kPrivate + i] = (unsigned char)(i * 13 + c));
    unsigned char *d_common, *d_priv;
    unsigned long long* d_ns;
    uint32_t* d_sum;
    uint4* d_flush;
    const size_t flush_n = (256u << 20) / sizeof(uint4);
    CHECK(cudaMalloc(&d_common, kCommon));
    CHECK(cudaMalloc(&d_priv, h_priv.size()));
    CHECK(cudaMalloc(&d_ns, sizeof(unsigned long long) * ctas));
    CHECK(cudaMalloc(&d_sum, sizeof(uint32_t) * ctas));
    CHECK(cudaMalloc(&d_flush, flush_n * sizeof(uint4)));
    CHECK(cudaMemcpy(d_common, h_common.data(), kCommon, cudaMemcpyHostToDevice));
    CHECK(cudaMemcpy(d_priv, h_priv.data(), h_priv.size(), cudaMemcpyHostToDevice));
    printf("{\n  \"what\": \"TILE start-up: %u common + %u private bytes per CTA, cold L2, 144 CTAs of 512 threads\",\n", kCommon, kPrivate);
#define RUN(name, C, M, last) run<C, M>(name, ctas, d_common, d_priv, d_ns, d_sum, want_common, want_priv, d_flush, flush_n, last)
    RUN("plain_launch_every_cta_copies", 1, false, false);
    RUN("cluster2_no_multicast", 2, false, false);
    RUN("cluster2_multicast", 2, true, false);
    RUN("cluster4_no_multicast", 4, false, false);
    RUN("cluster4_multicast", 4, true, false);
    RUN("cluster8_no_multicast", 8, false, false);
    RUN("cluster8_multicast", 8, true, true);
    printf("}\n");
    return 0;
}
