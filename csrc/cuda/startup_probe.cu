// Where does a cold CUDA process spend its first second? Times the steps every one-shot CLI run pays
// before its first kernel can start (the reference pays the same in every `./final` run, plus PTX JIT):
// driver initialisation, primary context, module load of a trivial kernel, first launch + sync.
// Build: make startup_probe     Run: build/startup_probe [device]     (one JSON object on stdout)
#include <cuda.h>
#include <cuda_runtime.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>

__global__ void noop(int* p) {
    if (p) *p = 1;
}

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main(int argc, char** argv) {
    const int dev = argc > 1 ? std::atoi(argv[1]) : 0;
    const double t0 = now();
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);  // cuInit happens here
    const double t1 = now();
    if (e != cudaSuccess) {
        std::fprintf(stderr, "cudaGetDeviceCount: %s\n", cudaGetErrorString(e));
        return 1;
    }
    cudaSetDevice(dev);
    cudaFree(nullptr);  // primary context
    const double t2 = now();
    int* d = nullptr;
    cudaMalloc(&d, 4);
    const double t3 = now();
    noop<<<1, 32>>>(d);  // module load + first launch
    e = cudaDeviceSynchronize();
    const double t4 = now();
    noop<<<1, 32>>>(d);
    cudaDeviceSynchronize();
    const double t5 = now();
    const char* vis = std::getenv("CUDA_VISIBLE_DEVICES");
    std::printf(
        "{\"visible_devices\": %d, \"CUDA_VISIBLE_DEVICES\": \"%s\", \"driver_init_s\": %.4f, \"primary_context_s\": %.4f, \"first_malloc_s\": %.4f, "
        "\"module_load_first_launch_s\": %.4f, \"second_launch_s\": %.6f, \"total_s\": %.4f, \"status\": \"%s\"}\n",
        n, vis ? vis : "", t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4, t5 - t0, cudaGetErrorString(e));
    return 0;
}
