// launch-floor microbenchmark: event-timed empty kernels, regular vs cooperative
#include <cstdio>
#include <cuda_runtime.h>
#include <algorithm>
#include <vector>
__global__ void empty(int* p) { if (p && threadIdx.x == 9999) *p = 1; }
int main() {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    int* d; cudaMalloc(&d, 4);
    auto run = [&](const char* name, int grid, int block, bool coop) {
        std::vector<float> v;
        for (int i = 0; i < 300; ++i) {
            cudaEventRecord(a);
            if (coop) { void* args[] = {&d}; cudaLaunchCooperativeKernel((void*)empty, dim3(grid), dim3(block), args, 0, 0); }
            else empty<<<grid, block>>>(d);
            cudaEventRecord(b);
            cudaEventSynchronize(b);
            float ms; cudaEventElapsedTime(&ms, a, b); v.push_back(ms * 1000);
        }
        std::sort(v.begin(), v.end());
        printf("%-28s grid=%4d block=%4d  min=%.2f us  median=%.2f us\n", name, grid, block, v[0], v[150]);
    };
    run("regular", 1, 32, false);
    run("regular", 1, 1024, false);
    run("regular", 148, 1024, false);
    run("cooperative", 1, 1024, true);
    run("cooperative", 148, 1024, true);
    std::vector<float> v;
    for (int i = 0; i < 300; ++i) { cudaEventRecord(a); cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms, a, b); v.push_back(ms * 1000); }
    std::sort(v.begin(), v.end());
    printf("event pair, no kernel: min=%.2f median=%.2f us\n", v[0], v[150]);
    return 0;
}
