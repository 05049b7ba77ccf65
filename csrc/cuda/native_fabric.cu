// Native bring-up of the multi-GPU fabric: one PROCESS per GPU with no Python, torch or NCCL.
//
// The reference's multi-rank entry is `mpiexec -np N ./final` (makefile:11) + MPI_Init (main.c:62-64):
// a native launcher plus a communicator. Here `final --gpus N` forks N workers (csrc/host/
// native_launch.cpp); each worker calls sa_native_worker() below, which
//   1. allocates its symmetric buffer with the CUDA virtual-memory API (cuMemCreate),
//   2. exports it as a POSIX file descriptor and trades descriptors with the other workers through
//      the launcher (a Unix socket switchboard, SCM_RIGHTS) — the analogue of MPI_Init's wire-up,
//   3. maps every peer's buffer (cuMemImportFromShareableHandle + cuMemMap + cuMemSetAccess) and, when
//      the GPUs sit behind an NVSwitch, binds all buffers to one multicast object (cuMulticastCreate),
//   4. launches sa_cuda_fabric_solve — the same single kernel per GPU the Python path launches.
// The driver entry points are fetched with cudaGetDriverEntryPoint, so the library still loads on
// a machine without a driver (the CPU-only test box).
#include <cuda.h>
#include <cuda_runtime.h>
#include <sys/socket.h>
#include <sys/uio.h>
#include <unistd.h>

#include <cerrno>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../host/native_wire.h"
#include "seqalign.h"



namespace {

struct DriverApi {
    CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
    CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
    CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
    CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
    CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
    CUresult (*MemUnmap)(CUdeviceptr, size_t);
    CUresult (*MemRelease)(CUmemGenericAllocationHandle);
    CUresult (*MemAddressFree)(CUdeviceptr, size_t);
    CUresult (*DeviceGet)(CUdevice*, int);
    CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
    CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
    CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
    CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
    CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
    CUresult (*GetErrorString)(CUresult, const char**);
};

bool load_driver(DriverApi& d, std::string& why) {
    struct Item {
        const char* name;
        void** slot;
    } items[] = {
        {"cuMemGetAllocationGranularity", (void**)&d.MemGetAllocationGranularity},
        {"cuMemCreate", (void**)&d.MemCreate},
        {"cuMemExportToShareableHandle", (void**)&d.MemExportToShareableHandle},
        {"cuMemImportFromShareableHandle", (void**)&d.MemImportFromShareableHandle},
        {"cuMemAddressReserve", (void**)&d.MemAddressReserve},
        {"cuMemMap", (void**)&d.MemMap},
        {"cuMemSetAccess", (void**)&d.MemSetAccess},
        {"cuMemUnmap", (void**)&d.MemUnmap},
        {"cuMemRelease", (void**)&d.MemRelease},
        {"cuMemAddressFree", (void**)&d.MemAddressFree},
        {"cuDeviceGet", (void**)&d.DeviceGet},
        {"cuDeviceGetAttribute", (void**)&d.DeviceGetAttribute},
        {"cuMulticastCreate", (void**)&d.MulticastCreate},
        {"cuMulticastAddDevice", (void**)&d.MulticastAddDevice},
        {"cuMulticastBindMem", (void**)&d.MulticastBindMem},
        {"cuMulticastGetGranularity", (void**)&d.MulticastGetGranularity},
        {"cuGetErrorString", (void**)&d.GetErrorString},
    };
    for (const Item& it : items) {
        cudaDriverEntryPointQueryResult q;
        *it.slot = nullptr;
        const cudaError_t e = cudaGetDriverEntryPoint(it.name, it.slot, cudaEnableDefault, &q);
        if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !*it.slot) {
            why = std::string("driver entry point ") + it.name + " not available";
            return false;
        }
    }
    return true;
}

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

}  // namespace

// --------------------------------------------------------------------------- collective helpers
namespace {

struct Wire {
    int sock, rank, world;
    bool all_fds(int mine, int* all) {  // all-gather one descriptor per rank
        sa_wire_msg m = {SA_WIRE_FDS, (uint32_t)rank, 0, 1, 0};
        if (sa_wire_send(sock, &m, &mine) != 0) return false;
        int fds[SA_WIRE_MAX_FDS];
        if (sa_wire_recv(sock, &m, fds) != 0 || m.type != SA_WIRE_FDS || (int)m.nfds != world) return false;
        memcpy(all, fds, sizeof(int) * world);
        return true;
    }
    bool root_fd(int mine, int* got) {  // rank 0's descriptor (or -1) to everyone
        sa_wire_msg m = {SA_WIRE_ROOT_FD, (uint32_t)rank, 0, (rank == 0 && mine >= 0) ? 1u : 0u, 0};
        if (sa_wire_send(sock, &m, &mine) != 0) return false;
        int fds[SA_WIRE_MAX_FDS];
        if (sa_wire_recv(sock, &m, fds) != 0 || m.type != SA_WIRE_ROOT_FD) return false;
        *got = m.nfds ? fds[0] : -1;
        return true;
    }
    bool min(int64_t v, int64_t* out) {  // all-reduce(min); with v = 0 a plain barrier
        sa_wire_msg m = {SA_WIRE_MIN, (uint32_t)rank, v, 0, 0};
        if (sa_wire_send(sock, &m, nullptr) != 0) return false;
        int fds[SA_WIRE_MAX_FDS];
        if (sa_wire_recv(sock, &m, fds) != 0 || m.type != SA_WIRE_MIN) return false;
        if (out) *out = m.value;
        return true;
    }
};

size_t round_up(size_t v, size_t g) { return (v + g - 1) / g * g; }

}  // namespace

#define NF_FAIL(msg)                                                             \
    do {                                                                         \
        std::snprintf(err, sizeof(err), "rank %d: %s", a->rank, (msg));          \
        sa_set_last_error(err);                                                  \
        return SA_ERR_CUDA;                                                      \
    } while (0)
#define NF_CU(call)                                                                              \
    do {                                                                                         \
        const CUresult r__ = (call);                                                             \
        if (r__ != CUDA_SUCCESS) {                                                               \
            const char* s__ = nullptr;                                                           \
            d.GetErrorString(r__, &s__);                                                         \
            std::snprintf(err, sizeof(err), "rank %d: %s: %s", a->rank, #call, s__ ? s__ : "?"); \
            sa_set_last_error(err);                                                              \
            return SA_ERR_CUDA;                                                                  \
        }                                                                                        \
    } while (0)
#define NF_RT(call)                                                                                   \
    do {                                                                                              \
        const cudaError_t r__ = (call);                                                               \
        if (r__ != cudaSuccess) {                                                                     \
            std::snprintf(err, sizeof(err), "rank %d: %s: %s", a->rank, #call, cudaGetErrorString(r__)); \
            sa_set_last_error(err);                                                                   \
            return SA_ERR_CUDA;                                                                       \
        }                                                                                             \
    } while (0)

extern "C" int sa_native_worker(sa_native_args* a) {
    char err[512];
    if (!a || a->world < 1 || a->world > SA_MAX_WORLD || a->rank < 0 || a->rank >= a->world) {
        sa_set_last_error("sa_native_worker: bad argument");
        return SA_ERR_ARG;
    }
    const double t_begin = now_s();
    Wire w = {a->sock, a->rank, a->world};
    NF_RT(cudaSetDevice(a->device));
    NF_RT(cudaFree(nullptr));  // primary context
    DriverApi d;
    std::string why;
    if (!load_driver(d, why)) NF_FAIL(why.c_str());
    CUdevice dev;
    NF_CU(d.DeviceGet(&dev, a->device));
    a->times[0] = now_s() - t_begin;  // CUDA start-up

    sa_fabric_layout_t lay;
    if (sa_fabric_layout(a->max_seq, a->max_blob_bytes, &lay) != SA_OK) return SA_ERR_ARG;

    // ---- decide about multicast together (every rank must agree: it changes who sends the header how)
    int mc_supported = 0;
    d.DeviceGetAttribute(&mc_supported, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
    int64_t mc_vote = (a->world > 1 && a->use_multicast != 0 && mc_supported) ? 1 : 0;
    if (!w.min(mc_vote, &mc_vote)) NF_FAIL("launcher went away (multicast vote)");

    CUmemAllocationProp prop;
    memset(&prop, 0, sizeof(prop));
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = dev;
    prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t gran = 0;
    NF_CU(d.MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    CUmulticastObjectProp mprop;
    memset(&mprop, 0, sizeof(mprop));
    mprop.numDevices = (unsigned)a->world;
    mprop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    if (mc_vote) {
        size_t mgran = 0;
        mprop.size = round_up(lay.total_bytes, gran);
        if (d.MulticastGetGranularity(&mgran, &mprop, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mgran > gran) gran = mgran;
    }
    const size_t bytes = round_up(lay.total_bytes, gran);
    mprop.size = bytes;

    // ---- own buffer, exported; everyone's buffers imported and mapped
    CUmemGenericAllocationHandle own;
    NF_CU(d.MemCreate(&own, bytes, &prop, 0));
    int own_fd = -1;
    NF_CU(d.MemExportToShareableHandle(&own_fd, own, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    int fds[SA_WIRE_MAX_FDS];
    if (!w.all_fds(own_fd, fds)) NF_FAIL("launcher went away (handle exchange)");
    close(own_fd);
    CUmemAccessDesc acc;
    memset(&acc, 0, sizeof(acc));
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = dev;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    CUdeviceptr base[SA_MAX_WORLD] = {0};
    for (int r = 0; r < a->world; ++r) {
        CUmemGenericAllocationHandle h = own;
        if (r != a->rank) NF_CU(d.MemImportFromShareableHandle(&h, (void*)(uintptr_t)fds[r], CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
        close(fds[r]);
        NF_CU(d.MemAddressReserve(&base[r], bytes, gran, 0, 0));
        NF_CU(d.MemMap(base[r], bytes, 0, h, 0));
        NF_CU(d.MemSetAccess(base[r], bytes, &acc, 1));
        if (r != a->rank) NF_CU(d.MemRelease(h));  // the mapping keeps the memory alive
    }
    NF_RT(cudaMemset((void*)base[a->rank], 0, bytes));
    NF_RT(cudaDeviceSynchronize());
    a->times[1] = now_s() - t_begin;  // buffers mapped

    // ---- one multicast object over all buffers (NVSwitch): optional, all ranks or none
    CUdeviceptr mc_va = 0;
    CUmemGenericAllocationHandle mc = 0;
    if (mc_vote) {
        bool ok = true;
        int mc_fd = -1, got = -1;
        if (a->rank == 0) {
            ok = d.MulticastCreate(&mc, &mprop) == CUDA_SUCCESS &&
                 d.MemExportToShareableHandle(&mc_fd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) == CUDA_SUCCESS;
            if (!ok) mc_fd = -1;
        }
        if (!w.root_fd(mc_fd, &got)) NF_FAIL("launcher went away (multicast handle)");
        if (mc_fd >= 0) close(mc_fd);
        if (a->rank != 0)
            ok = got >= 0 && d.MemImportFromShareableHandle(&mc, (void*)(uintptr_t)got, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) == CUDA_SUCCESS;
        if (got >= 0) close(got);
        ok = ok && d.MulticastAddDevice(mc, dev) == CUDA_SUCCESS;
        int64_t all = ok ? 1 : 0;  // every device must have been added before anyone binds memory
        if (!w.min(all, &all)) NF_FAIL("launcher went away (multicast add)");
        ok = all != 0;
        if (ok) {
            ok = d.MulticastBindMem(mc, 0, own, 0, bytes, 0) == CUDA_SUCCESS && d.MemAddressReserve(&mc_va, bytes, gran, 0, 0) == CUDA_SUCCESS &&
                 d.MemMap(mc_va, bytes, 0, mc, 0) == CUDA_SUCCESS && d.MemSetAccess(mc_va, bytes, &acc, 1) == CUDA_SUCCESS;
        }
        all = ok ? 1 : 0;
        if (!w.min(all, &all)) NF_FAIL("launcher went away (multicast bind)");
        if (!all) mc_va = 0;
    }
    a->used_multicast = mc_va != 0;

    // ---- the solve: rank 0 stages the blob in its own buffer, every rank launches ONE kernel
    sa_fabric_desc f;
    memset(&f, 0, sizeof(f));
    f.rank = (uint32_t)a->rank;
    f.world = (uint32_t)a->world;
    f.use_multicast = mc_va ? 1u : 0u;
    f.timeout_ms = a->timeout_ms;
    for (int r = 0; r < a->world; ++r) f.peer_base[r] = (void*)base[r];
    f.mc_base = (void*)mc_va;
    f.layout = lay;
    if (a->rank == 0) NF_RT(cudaMemcpy((char*)base[0] + lay.off_blob, a->blob, a->header->total_bytes, cudaMemcpyHostToDevice));
    if (!w.min(0, nullptr)) NF_FAIL("launcher went away (before the launch)");  // every buffer zeroed and mapped
    a->times[2] = now_s() - t_begin;  // ready to launch
    const int reps = a->repeat > 0 ? a->repeat : 1;
    cudaEvent_t e0, e1;
    NF_RT(cudaEventCreate(&e0));
    NF_RT(cudaEventCreate(&e1));
    float best_ms = 1e30f;
    int rc = SA_OK;
    for (int it = 0; it < reps && rc == SA_OK; ++it) {
        f.epoch = (uint32_t)(it + 1);
        NF_RT(cudaEventRecord(e0, nullptr));
        rc = sa_cuda_fabric_solve(&f, a->rank == 0 ? a->header : nullptr, nullptr);
        NF_RT(cudaEventRecord(e1, nullptr));
        NF_RT(cudaStreamSynchronize(nullptr));
        float ms = 0;
        NF_RT(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best_ms) best_ms = ms;
        if (reps > 1 && !w.min(0, nullptr)) NF_FAIL("launcher went away (between solves)");
    }
    a->kernel_ms = best_ms;
    uint32_t errword = 0;
    NF_RT(cudaMemcpy(&errword, (char*)base[a->rank] + lay.off_ctrl + 12, 4, cudaMemcpyDeviceToHost));
    if (rc == SA_OK && errword) {
        std::snprintf(err, sizeof(err), "rank %d: a wait on another GPU timed out (code %u)", a->rank, errword);
        sa_set_last_error(err);
        rc = SA_ERR_CUDA;
    }
    if (rc == SA_OK && a->rank == 0 && a->header->nseq)
        NF_RT(cudaMemcpy(a->results, (char*)base[0] + lay.off_out, (size_t)a->header->nseq * 16, cudaMemcpyDeviceToHost));
    a->times[3] = now_s() - t_begin;  // results on the host
    w.min(0, nullptr);                // nobody unmaps while a peer's kernel may still touch its memory

    for (int r = 0; r < a->world; ++r) {
        d.MemUnmap(base[r], bytes);
        d.MemAddressFree(base[r], bytes);
    }
    if (mc_va) {
        d.MemUnmap(mc_va, bytes);
        d.MemAddressFree(mc_va, bytes);
    }
    if (mc) d.MemRelease(mc);
    d.MemRelease(own);
    return rc;
}
