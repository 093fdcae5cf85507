// How much of a kernel launch's host cost is the size of its parameter block?  (GPU box)
//   nvcc -arch=sm_100a -O3 scripts/launch_cost.cu -o /tmp/launch_cost && /tmp/launch_cost
#include <cstdio>
#include <cstdint>
#include <chrono>
#include <cuda_runtime.h>
template <int N> struct Blk { uint8_t a[N]; };
template <int N> __global__ void k(const __grid_constant__ Blk<N> b, int *out) {
    if (b.a[threadIdx.x % N] == 255 && out) out[0] = 1;
}
template <int N> double run(cudaStream_t *st, int ns, int iters) {
    static Blk<N> b;
    for (int i = 0; i < 6000; i++) { k<N><<<1024, 128, 0, st[i % ns]>>>(b, nullptr); if (i % 4 == 3) cudaStreamSynchronize(st[(i + 1) % ns]); }
    cudaDeviceSynchronize();
    auto t0 = std::chrono::steady_clock::now();
    double in_launch = 0;
    for (int i = 0; i < iters; i++) {
        cudaStreamSynchronize(st[i % ns]);
        auto a = std::chrono::steady_clock::now();
        k<N><<<1024, 128, 0, st[i % ns]>>>(b, nullptr);
        in_launch += std::chrono::duration<double>(std::chrono::steady_clock::now() - a).count();
    }
    cudaDeviceSynchronize();
    double tot = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    printf("params %5d B: launch call %.2f us, loop %.2f us per iteration\n", N, in_launch / iters * 1e6, tot / iters * 1e6);
    return tot;
}
int main() {
    cudaStream_t st[4];
    for (int i = 0; i < 4; i++) cudaStreamCreateWithFlags(&st[i], cudaStreamNonBlocking);
    run<64>(st, 4, 20000);
    run<1024>(st, 4, 20000);
    run<2048>(st, 4, 20000);
    run<4096>(st, 4, 20000);
    run<64>(st, 4, 20000);
    return 0;
}
