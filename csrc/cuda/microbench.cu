// Pipe-rate micro-benchmarks for the roofline of the scoring kernels (sm_100a).
//
// The scan kernels are integer add / max over shared-memory lookups, so their "speed of light" is
// set by the SM's issue slots, the ALU pipe (IADD3, LOP3, PRMT, VIMNMX), the FMA pipe (IMAD) and the
// shared-memory pipe (LDS wavefronts) — none of which is in MEASURED_PEAKS.json. This program
// measures them on the box, plus the TMEM read/write rate (tcgen05.ld / tcgen05.st) that decides
// whether a tensor-core formulation that returns per-cell values can pay off.
//
// Method: one 512-thread CTA per SM (16 warps = 4 per SM sub-partition), every thread runs ITER
// trips of a fully unrolled body of independent chains, cycles taken with clock64() around the
// loop inside the kernel (per CTA; the maximum over CTAs is reported). Output: one JSON object.
//
// Build: make microbench   Run (on a B200): build/microbench > profiles/MEASURED_PIPES.json
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CHECK(x)                                                                          \
    do {                                                                                  \
        cudaError_t e__ = (x);                                                            \
        if (e__ != cudaSuccess) {                                                         \
            fprintf(stderr, "%s failed: %s\n", #x, cudaGetErrorString(e__));              \
            exit(1);                                                                      \
        }                                                                                 \
    } while (0)

constexpr int kThreads = 512;
constexpr int kIter = 512;
constexpr int kChains = 8;

enum Op { OP_IADD3, OP_IMAD, OP_VMAX2, OP_VIADDMAX, OP_LOP3, OP_PRMT, OP_MIX_ALU3, OP_MIX_IMAD, OP_COUNT };

template <int kOp>
__global__ void __launch_bounds__(kThreads, 1) alu_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, long long* cycles) {
    uint32_t a[kChains], b[kChains], c[kChains];
#pragma unroll
    for (int i = 0; i < kChains; ++i) {
        a[i] = in[threadIdx.x + i * 32];
        b[i] = in[threadIdx.x + 1024 + i];
        c[i] = in[threadIdx.x + 2048 + i];
    }
    const uint32_t one = in[4095];  // runtime 1: keeps mad.lo a real IMAD
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < kIter; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int i = 0; i < kChains; ++i) {
                if (kOp == OP_IADD3) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b[i]));
                if (kOp == OP_IMAD) asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b[i]), "r"(one));
                if (kOp == OP_VMAX2) asm volatile("max.u16x2 %0, %0, %1;" : "+r"(a[i]) : "r"(b[i]));
                if (kOp == OP_VIADDMAX) a[i] = (uint32_t)__viaddmax_s32((int)a[i], (int)b[i], (int)c[i]);
                if (kOp == OP_LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b[i]), "r"(c[i]));
                if (kOp == OP_PRMT) asm volatile("prmt.b32 %0, %0, %1, 0x4440;" : "+r"(a[i]) : "r"(b[i]));
                if (kOp == OP_MIX_ALU3) {  // the scan's arithmetic as compiled today: 3 ALU-pipe instructions per packed cell
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b[i]));
                    asm volatile("{.reg .b32 t; add.u32 t, %1, 0x7fff7fff; sub.u32 t, t, %2; max.u16x2 %0, %0, t;}" : "+r"(c[i]) : "r"(a[i]), "r"(a[(i + 1) % kChains]));
                }
                if (kOp == OP_MIX_IMAD) {  // same arithmetic with the prefix add moved to the FMA pipe (IMAD)
                    asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b[i]), "r"(one));
                    asm volatile("{.reg .b32 t; add.u32 t, %1, 0x7fff7fff; sub.u32 t, t, %2; max.u16x2 %0, %0, t;}" : "+r"(c[i]) : "r"(a[i]), "r"(a[(i + 1) % kChains]));
                }
            }
        }
    }
    const long long t1 = clock64();
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < kChains; ++i) acc ^= a[i] ^ c[i];
    out[blockIdx.x * kThreads + threadIdx.x] = acc;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

// Conflict-free LDS (lane i reads bank i); kWidth = 1, 2 or 4 words per lane (LDS.32/.64/.128); kGather:
// each lane reads a different bank of a different row pattern (a permutation, still conflict free).
// The row offset advances by a runtime stride every trip so that ptxas cannot hoist the loads.
template <int kWidth, bool kGather>
__global__ void __launch_bounds__(kThreads, 1) lds_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, long long* cycles) {
    __shared__ __align__(16) uint32_t sm[8192 + 1024];
    for (int i = threadIdx.x; i < 8192 + 1024; i += kThreads) sm[i] = i * 2654435761u;
    __syncthreads();
    const uint32_t stride = in[4094];  // runtime 128
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t bank = kGather ? ((lane * 5 + 3) & 31) : lane;
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(sm) + bank * 4 * kWidth;
    uint32_t acc[kChains] = {0};
    uint32_t off = 0;
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < kIter; ++it) {
        off = (off + stride) & 0xfffu;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int i = 0; i < kChains; ++i) {
                const uint32_t addr = base + off + ((u * kChains + i) * 128 * kWidth) % (8192 * 4 - 4096 - 512);
                if (kWidth == 1) {
                    uint32_t v;
                    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
                    acc[i] ^= v;
                } else if (kWidth == 2) {
                    uint32_t v0, v1;
                    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v0), "=r"(v1) : "r"(addr));
                    acc[i] ^= v0 ^ v1;
                } else {
                    uint32_t v0, v1, v2, v3;
                    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v0), "=r"(v1), "=r"(v2), "=r"(v3) : "r"(addr));
                    acc[i] ^= v0 ^ v1 ^ v2 ^ v3;
                }
            }
        }
    }
    const long long t1 = clock64();
    uint32_t a = 0;
#pragma unroll
    for (int i = 0; i < kChains; ++i) a ^= acc[i];
    out[blockIdx.x * kThreads + threadIdx.x] = a;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

// The STREAM inner loop in isolation on a synthetic 32-row profile block: per row 9 LDS + 9 prefix
// adds + 8 (IADD3 + VIMNMX.U16x2). kImad moves the 9 prefix adds to the FMA pipe.
template <bool kImad>
__global__ void __launch_bounds__(kThreads, 1) scan_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, long long* cycles) {
    extern __shared__ __align__(16) uint32_t prof[];  // 16 x 4 KB: one profile block per warp
    for (int i = threadIdx.x; i < 16 * 32 * 32; i += kThreads) prof[i] = (i * 2654435761u >> 20) & 0x000F000Fu;
    __syncthreads();
    const uint32_t one = in[4095];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned char* pb = reinterpret_cast<const unsigned char*>(prof + warp * 1024);
    uint32_t L[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) L[k] = ((lane + k * 7) & 31) * 4;  // letter byte offsets (conflict free: distinct banks not required, rows are)
    uint32_t P[9], best[8];
#pragma unroll
    for (int r = 0; r <= 8; ++r) P[r] = 0;
#pragma unroll
    for (int r = 0; r < 8; ++r) best[r] = 0;
    constexpr uint32_t BIG2 = 0x7fff7fffu;
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < kIter; ++it) {
#pragma unroll 1
        for (int row = 0; row < 32; row += 8) {
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                uint32_t v[9];
#pragma unroll
                for (int r = 0; r <= 8; ++r) v[r] = *reinterpret_cast<const uint32_t*>(pb + L[u + r] + (row + u) * 128);
#pragma unroll
                for (int r = 0; r <= 8; ++r) {
                    if (kImad) asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(P[r]) : "r"(v[r]), "r"(one));
                    else P[r] += v[r];
                }
#pragma unroll
                for (int r = 0; r < 8; ++r) best[r] = __vmaxu2(best[r], P[r] + BIG2 - P[r + 1]);
            }
        }
#pragma unroll
        for (int r = 0; r <= 8; ++r) P[r] &= 0x00ff00ffu;  // keep the halves from overflowing across trips
    }
    const long long t1 = clock64();
    uint32_t a = 0;
#pragma unroll
    for (int r = 0; r < 8; ++r) a ^= best[r] ^ P[r];
    out[blockIdx.x * kThreads + threadIdx.x] = a;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

// The same scan in "difference form": per lane only the eight neighbour differences D[r] = P[r] - P[r+1]
// (one 3-input IADD3 each: D += v[r] - v[r+1]), their running maxima two rows at a time (VIMNMX3.U16x2)
// and ONE running total (the ninth diagonal); the other eight totals follow from the final differences.
__global__ void __launch_bounds__(kThreads, 1) scan_dform_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, long long* cycles) {
    extern __shared__ __align__(16) uint32_t prof[];
    for (int i = threadIdx.x; i < 16 * 32 * 32; i += kThreads) prof[i] = (i * 2654435761u >> 20) & 0x000F000Fu;
    __syncthreads();
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned char* pb = reinterpret_cast<const unsigned char*>(prof + warp * 1024);
    uint32_t L[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) L[k] = ((lane + k * 7) & 31) * 4;
    constexpr uint32_t BIG2 = 0x7fff7fffu;
    uint32_t D[8], best[8], P8 = in[threadIdx.x] & 0xffu;
#pragma unroll
    for (int r = 0; r < 8; ++r) D[r] = BIG2, best[r] = 0;
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < kIter; ++it) {
#pragma unroll 1
        for (int row = 0; row < 32; row += 8) {
#pragma unroll
            for (int u = 0; u < 8; u += 2) {
                uint32_t v[9], w[9];
#pragma unroll
                for (int r = 0; r <= 8; ++r) v[r] = *reinterpret_cast<const uint32_t*>(pb + L[u + r] + (row + u) * 128);
#pragma unroll
                for (int r = 0; r <= 8; ++r) w[r] = *reinterpret_cast<const uint32_t*>(pb + L[u + 1 + r] + (row + u + 1) * 128);
                uint32_t d0[8];
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    d0[r] = D[r] + v[r] - v[r + 1];
                    D[r] = d0[r] + w[r] - w[r + 1];
                    best[r] = __vimax3_u16x2(best[r], d0[r], D[r]);
                }
                P8 += v[8] + w[8];
            }
        }
#pragma unroll
        for (int r = 0; r < 8; ++r) D[r] = (D[r] & 0x00ff00ffu) | 0x7f007f00u;
    }
    const long long t1 = clock64();
    uint32_t a = P8;
#pragma unroll
    for (int r = 0; r < 8; ++r) a ^= best[r] ^ D[r];
    out[blockIdx.x * kThreads + threadIdx.x] = a;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

// TMEM read / write rate: `nwarps` warps, each in its own 32-lane quarter (warp % 4), sweep all
// 512 columns with tcgen05.ld.32x32b.x32 (32 registers per thread = 4 KB per warp instruction).
template <int kShape>  // 0: ld 32x32b.x32, 1: ld 32x32b.x8, 2: ld 16x256b.x8 (32 regs), 3: st 32x32b.x32
__global__ void __launch_bounds__(kThreads, 1) tmem_kernel(uint32_t* __restrict__ out, long long* cycles, int nwarps, int iters) {
    __shared__ uint32_t tmem_base_s;
    const uint32_t warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&tmem_base_s)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = tmem_base_s;
    const uint32_t lane_base = (warp & 3u) * 32u;
    uint32_t acc = 0;
    // initialise every column this warp will read (also the STTM path)
    if (warp < 4) {
        for (uint32_t col = 0; col < 512; col += 32) {
            const uint32_t ta = tbase + (lane_base << 16) + col;
            asm volatile(
                "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(ta),
                "r"(col + threadIdx.x)
                : "memory");
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const long long t0 = clock64();
    if ((int)warp < nwarps) {
#pragma unroll 1
        for (int it = 0; it < iters; ++it) {
#pragma unroll 1
            for (uint32_t col = 0; col < 512; col += 128) {
#pragma unroll
                for (uint32_t c4 = 0; c4 < 128; c4 += 32) {
                    const uint32_t ta = tbase + (lane_base << 16) + col + c4;
                    uint32_t r[32];
                    if (kShape == 0) {
                        asm volatile(
                            "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                            : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                              "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
                              "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
                              "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                            : "r"(ta)
                            : "memory");
                        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                        for (int k = 0; k < 32; ++k) acc ^= r[k];
                    } else if (kShape == 1) {
#pragma unroll
                        for (uint32_t c8 = 0; c8 < 32; c8 += 8) {
                            asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                         : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                                         : "r"(ta + c8)
                                         : "memory");
                            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                            for (int k = 0; k < 8; ++k) acc ^= r[k];
                        }
                    } else if (kShape == 2) {
                        // 16x256b.x8: 16 lanes x (8 x 256 bit) = 32 registers per thread, two issues cover 32 lanes
#pragma unroll
                        for (uint32_t half = 0; half < 2; ++half) {
                            asm volatile(
                                "tcgen05.ld.sync.aligned.16x256b.x4.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                                : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                                  "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                                : "r"(ta + ((half * 16u) << 16))
                                : "memory");
                            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                            for (int k = 0; k < 16; ++k) acc ^= r[k];
                        }
                    } else {
                        asm volatile(
                            "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(ta),
                            "r"(acc + it)
                            : "memory");
                    }
                }
            }
        }
        if (kShape == 3) asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    const long long t1 = clock64();
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    out[blockIdx.x * kThreads + threadIdx.x] = acc;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tbase) : "memory");
}

static long long max_cycles(const long long* d, int n) {
    std::vector<long long> h(n);
    CHECK(cudaMemcpy(h.data(), d, n * sizeof(long long), cudaMemcpyDeviceToHost));
    long long m = 0;
    for (long long v : h) m = v > m ? v : m;
    return m;
}

int main() {
    int dev = 0, sms = 0, clk_khz = 0;
    CHECK(cudaSetDevice(dev));
    CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    CHECK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, dev));
    uint32_t *in, *out;
    long long* cyc;
    std::vector<uint32_t> hin(4096);
    for (int i = 0; i < 4096; ++i) hin[i] = (uint32_t)(i * 2654435761u) & 0x00ff00ffu;
    hin[4095] = 1u;
    hin[4094] = 128u;
    CHECK(cudaMalloc(&in, 4096 * 4));
    CHECK(cudaMalloc(&out, (size_t)sms * kThreads * 4));
    CHECK(cudaMalloc(&cyc, sms * sizeof(long long)));
    CHECK(cudaMemcpy(in, hin.data(), 4096 * 4, cudaMemcpyHostToDevice));

    printf("{\n \"device_sms\": %d, \"sm_clock_khz\": %d, \"threads_per_cta\": %d, \"ctas\": %d,\n", sms, clk_khz, kThreads, sms);
    printf(" \"unit\": \"warp-instructions per clock per SM sub-partition (4 per SM); lanes/clk/SM = x128\",\n");
    const double warps_per_smsp = kThreads / 32 / 4.0;
    auto report_alu = [&](const char* name, long long c, double instr_per_trip) {
        const double per_warp = (double)kIter * 4 * kChains * instr_per_trip;
        const double rate = per_warp * warps_per_smsp / (double)c;
        printf(" \"%s\": {\"cycles\": %lld, \"warp_instr_per_clk_per_smsp\": %.4f},\n", name, c, rate);
    };
#define RUN_ALU(OP, NAME, IPT)                                      \
    for (int rep = 0; rep < 2; ++rep) {                             \
        alu_kernel<OP><<<sms, kThreads>>>(in, out, cyc);            \
        CHECK(cudaDeviceSynchronize());                             \
    }                                                               \
    report_alu(NAME, max_cycles(cyc, sms), IPT);
    RUN_ALU(OP_IADD3, "iadd3 (ptxas folds two adds into one 3-input IADD3)", 0.5)
    RUN_ALU(OP_IMAD, "imad", 1.0)
    RUN_ALU(OP_VMAX2, "vimnmx3_u16x2 (ptxas folds two max into one VIMNMX3)", 0.5)
    RUN_ALU(OP_VIADDMAX, "viaddmnmx_s32", 1.0)
    RUN_ALU(OP_LOP3, "lop3", 1.0)
    RUN_ALU(OP_PRMT, "prmt", 1.0)
    RUN_ALU(OP_MIX_ALU3, "scan_arith_3alu (packed cells/clk/smsp)", 1.0)
    RUN_ALU(OP_MIX_IMAD, "scan_arith_imad (packed cells/clk/smsp)", 1.0)

    auto report_lds = [&](const char* name, long long c, int width) {
        const double per_warp = (double)kIter * 4 * kChains;
        printf(" \"%s\": {\"cycles\": %lld, \"warp_instr_per_clk_per_sm\": %.4f, \"bytes_per_clk_per_sm\": %.1f},\n", name, c,
               per_warp * (kThreads / 32) / (double)c, per_warp * (kThreads / 32) * 128.0 * width / (double)c);
    };
    for (int rep = 0; rep < 2; ++rep) { lds_kernel<1, false><<<sms, kThreads>>>(in, out, cyc); CHECK(cudaDeviceSynchronize()); }
    report_lds("lds32", max_cycles(cyc, sms), 1);
    for (int rep = 0; rep < 2; ++rep) { lds_kernel<1, true><<<sms, kThreads>>>(in, out, cyc); CHECK(cudaDeviceSynchronize()); }
    report_lds("lds32_gather", max_cycles(cyc, sms), 1);
    for (int rep = 0; rep < 2; ++rep) { lds_kernel<2, false><<<sms, kThreads>>>(in, out, cyc); CHECK(cudaDeviceSynchronize()); }
    report_lds("lds64", max_cycles(cyc, sms), 2);
    for (int rep = 0; rep < 2; ++rep) { lds_kernel<4, false><<<sms, kThreads>>>(in, out, cyc); CHECK(cudaDeviceSynchronize()); }
    report_lds("lds128", max_cycles(cyc, sms), 4);
    CHECK(cudaFuncSetAttribute(scan_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    CHECK(cudaFuncSetAttribute(scan_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    auto report_scan = [&](const char* name, long long c) {
        const double packed_cells = (double)kIter * 32 * 8 * 32 * (kThreads / 32);  // trips x rows x offsets/lane x lanes x warps
        printf(" \"%s\": {\"cycles\": %lld, \"packed_cells_per_clk_per_sm\": %.3f, \"cells_per_clk_per_sm\": %.3f},\n", name, c,
               packed_cells / (double)c, 2.0 * packed_cells / (double)c);
    };
    for (int rep = 0; rep < 2; ++rep) { scan_kernel<false><<<sms, kThreads, 65536>>>(in, out, cyc); CHECK(cudaDeviceSynchronize()); }
    report_scan("stream_inner_loop_iadd", max_cycles(cyc, sms));
    for (int rep = 0; rep < 2; ++rep) { scan_kernel<true><<<sms, kThreads, 65536>>>(in, out, cyc); CHECK(cudaDeviceSynchronize()); }
    report_scan("stream_inner_loop_imad", max_cycles(cyc, sms));
    CHECK(cudaFuncSetAttribute(scan_dform_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    for (int rep = 0; rep < 2; ++rep) { scan_dform_kernel<<<sms, kThreads, 65536>>>(in, out, cyc); CHECK(cudaDeviceSynchronize()); }
    report_scan("stream_inner_loop_difference_form", max_cycles(cyc, sms));

    const int iters = 64;
    auto report_tmem = [&](const char* name, long long c, int nwarps) {
        const double bytes = (double)iters * 512 * 32 * 4 * nwarps;  // columns x lanes x 4 B per warp sweep
        printf(" \"%s_w%d\": {\"cycles\": %lld, \"bytes_per_clk_per_sm\": %.1f},\n", name, nwarps, c, bytes / (double)c);
    };
    for (int nw : {1, 4, 8, 16}) {
        tmem_kernel<0><<<sms, kThreads>>>(out, cyc, nw, iters); CHECK(cudaDeviceSynchronize());
        tmem_kernel<0><<<sms, kThreads>>>(out, cyc, nw, iters); CHECK(cudaDeviceSynchronize());
        report_tmem("tmem_ld_32x32b_x32", max_cycles(cyc, sms), nw);
    }
    for (int nw : {4, 16}) {
        tmem_kernel<1><<<sms, kThreads>>>(out, cyc, nw, iters); CHECK(cudaDeviceSynchronize());
        report_tmem("tmem_ld_32x32b_x8", max_cycles(cyc, sms), nw);
        tmem_kernel<2><<<sms, kThreads>>>(out, cyc, nw, iters); CHECK(cudaDeviceSynchronize());
        report_tmem("tmem_ld_16x256b_x4", max_cycles(cyc, sms), nw);
        tmem_kernel<3><<<sms, kThreads>>>(out, cyc, nw, iters); CHECK(cudaDeviceSynchronize());
        report_tmem("tmem_st_32x32b_x32", max_cycles(cyc, sms), nw);
    }
    printf(" \"done\": true\n}\n");
    return 0;
}
