// Random pointer chase: dependent ld.cg latency over buffers of several sizes (after a warm pass).
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <numeric>
#include <algorithm>
#include <random>
#include <cuda_runtime.h>
__global__ void chase(const int *p, int start, int n, long long *out, int *sink)
{
    int idx = start; long long t0 = clock64();
    for (int i = 0; i < n; ++i) idx = __ldcg(p + (size_t)idx * 32);
    long long t1 = clock64(); out[0] = (t1 - t0) / n; *sink = idx;
}
__global__ void touch(int *p, size_t nlines) { size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; if (i < nlines) p[i * 32 + 1] = 1; }
__global__ void readall(const int *p, size_t nlines, int *sink) { size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; if (i < nlines && __ldcg(p + i * 32) == -7) *sink = 1; }
int main()
{
    long long *out; int *sink; cudaMallocManaged(&out, 64); cudaMalloc(&sink, 64);
    for (size_t mb : {1, 8, 32, 48, 64, 100, 200, 1000}) {
        size_t nlines = mb * 1024 * 1024 / 128;
        std::vector<int> perm(nlines); std::iota(perm.begin(), perm.end(), 0);
        std::mt19937 g(1); std::shuffle(perm.begin(), perm.end(), g);
        std::vector<int> buf(nlines * 32, 0);
        for (size_t i = 0; i < nlines; ++i) buf[(size_t)perm[i] * 32] = perm[(i + 1) % nlines];
        int *d; cudaMalloc(&d, nlines * 128); cudaMemcpy(d, buf.data(), nlines * 128, cudaMemcpyHostToDevice);
        for (int rep = 0; rep < 3; ++rep) readall<<<(nlines + 255) / 256, 256>>>(d, nlines, sink);
        chase<<<1, 1>>>(d, perm[0], 2000, out, sink); cudaDeviceSynchronize();
        long long a = out[0];
        touch<<<(nlines + 255) / 256, 256>>>(d, nlines);      // lines now dirty, written by all SMs
        chase<<<1, 1>>>(d, perm[5], 2000, out, sink); cudaDeviceSynchronize();
        printf("%5zu MB: chase after read-warm %lld cyc, after write-warm %lld cyc\n", mb, a, out[0]);
        cudaFree(d);
    }
    return 0;
}
