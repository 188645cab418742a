// What does clock64() tick at?  Spin for ~200 us in one thread and compare clock64 with %globaltimer (ns).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/clock_probe tools/clock_probe.cu
#include <cstdio>
__global__ void k(unsigned long long* out) {
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    long long c0 = clock64();
    do {
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    } while (t1 - t0 < 200000);
    long long c1 = clock64();
    out[0] = t1 - t0;
    out[1] = (unsigned long long)(c1 - c0);
}
int main() {
    unsigned long long* d;
    cudaMalloc(&d, 16);
    for (int i = 0; i < 3; i++) {
        k<<<1, 1>>>(d);
        unsigned long long h[2];
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        printf("globaltimer %llu ns, clock64 %llu ticks -> %.3f GHz\n", h[0], h[1], (double)h[1] / (double)h[0]);
    }
    return 0;
}
