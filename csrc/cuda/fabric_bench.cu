// Collective micro-kernels for the fabric sweep (BASELINE.json configs[4]): the two data-path
// primitives of the solver in isolation, so they can be timed from 1 KB to 1 GB against NCCL and
// against the reference's MPI path (main.c:149-152 Bcast, :195-197 Gather).
//   * broadcast: rank 0 pushes a byte range into every peer's symmetric buffer (NVLS multimem.st
//     or per-peer P2P stores), chunk by chunk, each chunk followed by a system-scope release
//     flag; the other ranks' kernels return when every chunk flag has arrived.
//   * pull: every rank other than 0 copies a byte range OUT OF rank 0's buffer into its own (LDG.128 over
//     NVLink, STG.128 local) — the data movement the solver uses since round 2; no flags, the range is
//     resident at rank 0 before anybody launches.
//   * argmax reduce: every rank folds its packed 64-bit (score, ~index) keys into rank 0's array
//     with system-scope atomicMax over NVLink, then bumps rank 0's arrival counter; rank 0's
//     kernel returns when all ranks have arrived.
#include <cuda_runtime.h>

#include <cstdint>
#include <string>

#include "ptx_helpers.cuh"
#include "seqalign.h"

namespace {

constexpr int kThreads = 512;
constexpr uint32_t kCtlCtasDone = 1, kCtlRanksDone = 2, kCtlError = 3, kCtlFlags = 64;

struct BParams {
    unsigned char* peer[SA_MAX_WORLD];
    unsigned char* mc;
    uint32_t rank, world, epoch, use_mc;
    uint32_t off_ctrl, off_blob;
    unsigned long long bytes, chunk;
    uint32_t nchunks;
    unsigned long long timeout_ns;
    const unsigned long long* keys;  // argmax: this rank's local keys
    unsigned long long nkeys;
};

using namespace sa_ptx;  // csrc/cuda/ptx_helpers.cuh

__global__ void __launch_bounds__(kThreads) bcast_kernel(const BParams p) {
    const uint32_t tid = threadIdx.x;
    uint32_t* const my_ctl = reinterpret_cast<uint32_t*>(p.peer[p.rank] + p.off_ctrl);
    if (p.rank == 0) {
        for (uint32_t c = blockIdx.x; c < p.nchunks; c += gridDim.x) {
            const unsigned long long base = (unsigned long long)c * p.chunk;
            const unsigned long long left = p.bytes - base;
            const uint32_t n16 = (uint32_t)(((left < p.chunk ? left : p.chunk) + 15ull) >> 4);
            const int4* src = reinterpret_cast<const int4*>(p.peer[0] + p.off_blob + base);
            if (p.use_mc) {
                int4* dst = reinterpret_cast<int4*>(p.mc + p.off_blob + base);
                for (uint32_t i = tid; i < n16; i += kThreads) multimem_st_v4(dst + i, src[i]);
            } else {
                for (uint32_t i = tid; i < n16; i += kThreads) {
                    const int4 v = src[i];
                    for (uint32_t r = 1; r < p.world; ++r)
                        reinterpret_cast<int4*>(p.peer[r] + p.off_blob + base)[i] = v;
                }
            }
            __syncthreads();
            if (tid == 0) {
                __threadfence_system();  // one fence, then relaxed flag stores (not world-1 releases)
                for (uint32_t r = 1; r < p.world; ++r)
                    st_relaxed_sys(reinterpret_cast<uint32_t*>(p.peer[r] + p.off_ctrl) + kCtlFlags + c, p.epoch);
            }
        }
    } else if (tid == 0) {
        const unsigned long long t0 = global_ns();
        for (uint32_t c = blockIdx.x; c < p.nchunks; c += gridDim.x) {
            while ((int)(ld_acquire_sys(my_ctl + kCtlFlags + c) - p.epoch) < 0) {
                __nanosleep(64);
                if (global_ns() - t0 > p.timeout_ns) {
                    atomicExch(my_ctl + kCtlError, 3u);
                    return;
                }
            }
        }
    }
}

__global__ void __launch_bounds__(kThreads) pull_kernel(const BParams p) {
    const int4* const src = reinterpret_cast<const int4*>(p.peer[0] + p.off_blob);
    int4* const dst = reinterpret_cast<int4*>(p.peer[p.rank] + p.off_blob);
    const unsigned long long n16 = (p.bytes + 15ull) >> 4, stride = (unsigned long long)gridDim.x * kThreads;
    unsigned long long i = (unsigned long long)blockIdx.x * kThreads + threadIdx.x;
    for (; i + 3 * stride < n16; i += 4 * stride) {  // four loads in flight per thread
        const int4 a = src[i], b = src[i + stride], c = src[i + 2 * stride], d = src[i + 3 * stride];
        dst[i] = a, dst[i + stride] = b, dst[i + 2 * stride] = c, dst[i + 3 * stride] = d;
    }
    for (; i < n16; i += stride) dst[i] = src[i];
}

__global__ void __launch_bounds__(kThreads) argmax_kernel(const BParams p) {
    const uint32_t tid = threadIdx.x;
    unsigned long long* const dst = reinterpret_cast<unsigned long long*>(p.peer[0] + p.off_blob);
    const unsigned long long stride = (unsigned long long)gridDim.x * kThreads;
    for (unsigned long long i = (unsigned long long)blockIdx.x * kThreads + tid; i < p.nkeys; i += stride)
        atomicMax_system(dst + i, p.keys[i]);  // RED.MAX.64 over NVLink when rank != 0
    __syncthreads();
    if (tid == 0) {
        uint32_t* const my_ctl = reinterpret_cast<uint32_t*>(p.peer[p.rank] + p.off_ctrl);
        __threadfence_system();
        const unsigned int done = atomicAdd(&my_ctl[kCtlCtasDone], 1u);
        if (done + 1u == gridDim.x) {
            my_ctl[kCtlCtasDone] = 0u;
            uint32_t* const arrived = reinterpret_cast<uint32_t*>(p.peer[0] + p.off_ctrl) + kCtlRanksDone;
            __threadfence_system();
            atomicAdd_system(arrived, 1u);
            if (p.rank == 0) {
                const unsigned long long t0 = global_ns();
                while (ld_acquire_sys(arrived) < p.world) {
                    __nanosleep(128);
                    if (global_ns() - t0 > p.timeout_ns) {
                        atomicExch(my_ctl + kCtlError, 4u);
                        break;
                    }
                }
                *arrived = 0u;
            }
        }
    }
}

int fill(const sa_fabric_desc* f, BParams* p) {
    if (!f || f->world < 1 || f->world > SA_MAX_WORLD || f->rank >= f->world) {
        sa_set_last_error("fabric bench: bad descriptor");
        return SA_ERR_ARG;
    }
    for (uint32_t r = 0; r < f->world; ++r) p->peer[r] = static_cast<unsigned char*>(f->peer_base[r]);
    p->mc = static_cast<unsigned char*>(f->mc_base);
    p->use_mc = (f->use_multicast && f->mc_base) ? 1u : 0u;
    p->rank = f->rank;
    p->world = f->world;
    p->epoch = f->epoch;
    p->off_ctrl = f->layout.off_ctrl;
    p->off_blob = f->layout.off_blob;
    p->timeout_ns = (unsigned long long)(f->timeout_ms ? f->timeout_ms : 10000u) * 1000000ull;
    return SA_OK;
}

int check_launch(const char* what) {
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        const std::string msg = std::string(what) + ": " + cudaGetErrorString(e);
        sa_set_last_error(msg.c_str());
        return SA_ERR_CUDA;
    }
    return SA_OK;
}

}  // namespace

extern "C" {

int sa_cuda_fabric_bcast(const sa_fabric_desc* f, uint64_t bytes, int ctas, void* stream) {
    BParams p{};
    int rc = fill(f, &p);
    if (rc != SA_OK) return rc;
    if (bytes == 0 || bytes > f->layout.max_blob_bytes) {
        sa_set_last_error("sa_cuda_fabric_bcast: size outside the fabric's blob region");
        return SA_ERR_TOO_LARGE;
    }
    p.bytes = bytes;
    const unsigned long long per = (bytes + SA_FABRIC_MAX_CHUNKS - 1) / SA_FABRIC_MAX_CHUNKS;
    p.chunk = per < 8192 ? 8192 : (per + 8191) / 8192 * 8192;
    p.nchunks = (uint32_t)((bytes + p.chunk - 1) / p.chunk);
    uint32_t grid = ctas > 0 ? (uint32_t)ctas : 296u;
    if (grid > p.nchunks) grid = p.nchunks;
    bcast_kernel<<<grid, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
    return check_launch("sa_cuda_fabric_bcast");
}

int sa_cuda_fabric_pull(const sa_fabric_desc* f, uint64_t bytes, int ctas, void* stream) {
    BParams p{};
    int rc = fill(f, &p);
    if (rc != SA_OK) return rc;
    if (bytes == 0 || bytes > f->layout.max_blob_bytes) {
        sa_set_last_error("sa_cuda_fabric_pull: size outside the fabric's blob region");
        return SA_ERR_TOO_LARGE;
    }
    if (f->rank == 0) return SA_OK;  // the source has nothing to do
    p.bytes = bytes;
    uint32_t grid = ctas > 0 ? (uint32_t)ctas : 296u;
    const uint64_t need = ((bytes + 15) / 16 + kThreads - 1) / kThreads;
    if (grid > need) grid = (uint32_t)need;
    pull_kernel<<<grid, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
    return check_launch("sa_cuda_fabric_pull");
}

int sa_cuda_fabric_argmax(const sa_fabric_desc* f, const void* d_keys, uint64_t nkeys, int ctas, void* stream) {
    BParams p{};
    int rc = fill(f, &p);
    if (rc != SA_OK) return rc;
    if (!d_keys || nkeys == 0 || nkeys * 8 > f->layout.max_blob_bytes) {
        sa_set_last_error("sa_cuda_fabric_argmax: size outside the fabric's blob region");
        return SA_ERR_TOO_LARGE;
    }
    p.keys = static_cast<const unsigned long long*>(d_keys);
    p.nkeys = nkeys;
    uint32_t grid = ctas > 0 ? (uint32_t)ctas : 296u;
    const uint64_t need = (nkeys + kThreads - 1) / kThreads;
    if (grid > need) grid = (uint32_t)need;
    argmax_kernel<<<grid, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
    return check_launch("sa_cuda_fabric_argmax");
}

}  // extern "C"
