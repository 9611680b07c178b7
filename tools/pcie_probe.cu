// pcie_probe.cu — what the box's PCIe link gives: H2D alone, D2H alone, both at once, for frame-sized
// and large pinned copies. Sets the ceiling the end-to-end numbers are read against (debug tool).
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <chrono>
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main()
{
    const size_t IN = 8294400, OUT = 3110400;
    const int N = 64;
    char *h_in, *h_out, *d_in, *d_out;
    cudaMallocHost(&h_in, IN * N); cudaMallocHost(&h_out, OUT * N);
    cudaMalloc(&d_in, IN * N); cudaMalloc(&d_out, OUT * N);
    cudaStream_t s1, s2; cudaStreamCreate(&s1); cudaStreamCreate(&s2);
    for (int rep = 0; rep < 2; rep++) {
        double t0 = now();
        for (int k = 0; k < 4; k++) cudaMemcpyAsync(d_in, h_in, IN * N, cudaMemcpyHostToDevice, s1);
        cudaStreamSynchronize(s1);
        double t1 = now();
        if (rep) printf("H2D large (531 MB x4): %.2f GB/s\n", 4.0 * IN * N / (t1 - t0) / 1e9);
        t0 = now();
        for (int k = 0; k < 4; k++) cudaMemcpyAsync(h_out, d_out, OUT * N, cudaMemcpyDeviceToHost, s2);
        cudaStreamSynchronize(s2);
        t1 = now();
        if (rep) printf("D2H large (199 MB x4): %.2f GB/s\n", 4.0 * OUT * N / (t1 - t0) / 1e9);
        t0 = now();
        for (int k = 0; k < 4 * N; k++) cudaMemcpyAsync(d_in + (k % N) * IN, h_in + (k % N) * IN, IN, cudaMemcpyHostToDevice, s1);
        cudaStreamSynchronize(s1);
        t1 = now();
        if (rep) printf("H2D frame-sized (8.3 MB x256): %.2f GB/s = %.0f frames/s\n", 4.0 * IN * N / (t1 - t0) / 1e9, 4.0 * N / (t1 - t0));
        t0 = now();
        for (int k = 0; k < 4 * N; k++) {
            cudaMemcpyAsync(d_in + (k % N) * IN, h_in + (k % N) * IN, IN, cudaMemcpyHostToDevice, s1);
            cudaMemcpyAsync(h_out + (k % N) * OUT, d_out + (k % N) * OUT, OUT, cudaMemcpyDeviceToHost, s2);
        }
        cudaStreamSynchronize(s1); cudaStreamSynchronize(s2);
        t1 = now();
        if (rep) printf("H2D+D2H frame-sized concurrent: H2D %.2f GB/s, D2H %.2f GB/s = %.0f frames/s\n",
                        4.0 * IN * N / (t1 - t0) / 1e9, 4.0 * OUT * N / (t1 - t0) / 1e9, 4.0 * N / (t1 - t0));
        // 2D copies with device pitch 7680 -> same as 1D here; with 3 plane copies for D2H
        t0 = now();
        for (int k = 0; k < 4 * N; k++) {
            cudaMemcpy2DAsync(d_in + (k % N) * IN, 7680, h_in + (k % N) * IN, 7680, 7680, 1080, cudaMemcpyHostToDevice, s1);
            char *ho = h_out + (k % N) * OUT, *dd = d_out + (k % N) * OUT;
            cudaMemcpy2DAsync(ho, 1920, dd, 2048, 1920, 1080, cudaMemcpyDeviceToHost, s2);
            cudaMemcpy2DAsync(ho + 2073600, 960, dd + 2211840, 1024, 960, 540, cudaMemcpyDeviceToHost, s2);
            cudaMemcpy2DAsync(ho + 2592000, 960, dd + 2211840 + 552960, 1024, 960, 540, cudaMemcpyDeviceToHost, s2);
        }
        cudaStreamSynchronize(s1); cudaStreamSynchronize(s2);
        t1 = now();
        if (rep) printf("same with cudaMemcpy2DAsync, re-pitched D2H (2048->1920): %.0f frames/s\n", 4.0 * N / (t1 - t0));
    }
    return 0;
}
