// Latency microbenchmarks of the primitives the tree kernels are made of (B200).
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;
#define N 64
__global__ void k_ldcg(const int *p, long long *out, int *sink)
{
    int idx = 0; long long t0 = clock64();
    for (int i = 0; i < N; ++i) idx = __ldcg(p + idx);
    long long t1 = clock64(); out[0] = (t1 - t0) / N; *sink = idx;
}
__global__ void k_ldca(const int *p, long long *out, int *sink)
{
    int idx = 0; long long t0 = clock64();
    for (int i = 0; i < N; ++i) idx = p[idx];
    long long t1 = clock64(); out[0] = (t1 - t0) / N; *sink = idx;
}
__global__ void k_atom(int *p, long long *out, int *sink)
{
    int idx = 0; long long t0 = clock64();
    for (int i = 0; i < N; ++i) idx = atomicOr(p + idx * 64 + 64, 0) ;
    long long t1 = clock64(); out[0] = (t1 - t0) / N; *sink = idx;
}
__global__ void k_dadd(double x, long long *out, double *sink)
{
    double a = x; long long t0 = clock64();
#pragma unroll
    for (int i = 0; i < 256; ++i) a = a + x;
    long long t1 = clock64(); out[0] = (t1 - t0); *sink = a;
}
__global__ void k_dsetp(double x, long long *out, double *sink)
{
    double a = x; long long t0 = clock64();
#pragma unroll
    for (int i = 0; i < 256; ++i) a = (a < x + i) ? a + 1.0 : a;
    long long t1 = clock64(); out[0] = (t1 - t0); *sink = a;
}
__global__ void k_shfl(double x, long long *out, double *sink)
{
    double a = x + threadIdx.x; long long t0 = clock64();
#pragma unroll
    for (int i = 0; i < 256; ++i) a = __shfl_sync(0xffffffffu, a, (threadIdx.x + 1) & 31);
    long long t1 = clock64(); out[0] = (t1 - t0); *sink = a;
}
__global__ void k_sync(long long *out)
{
    long long t0 = clock64();
    for (int i = 0; i < N; ++i) __syncthreads();
    long long t1 = clock64(); if (threadIdx.x == 0) out[0] = (t1 - t0) / N;
}
__global__ void k_csync(long long *out, int *g)
{
    cg::cluster_group c = cg::this_cluster();
    long long t0 = clock64();
    for (int i = 0; i < N; ++i) { g[threadIdx.x + blockIdx.x * blockDim.x] = i; c.sync(); }
    long long t1 = clock64(); if (c.thread_rank() == 0) out[0] = (t1 - t0) / N;
}
__global__ void k_fence(long long *out, int *g)
{
    long long t0 = clock64();
    for (int i = 0; i < N; ++i) { __stcg(g + threadIdx.x, i); __threadfence(); }
    long long t1 = clock64(); if (threadIdx.x == 0) out[0] = (t1 - t0) / N;
}
__global__ void k_pow(double x, long long *out, double *sink)
{
    double a = x + threadIdx.x * 1e-3; long long t0 = clock64();
    for (int i = 0; i < 16; ++i) a = pow(a + 1e-6, 0.6);
    long long t1 = clock64(); out[0] = (t1 - t0) / 16; *sink = a;
}
int main()
{
    int *p, *sink; long long *out; double *ds;
    cudaMalloc(&p, 1 << 26); cudaMemset(p, 0, 1 << 26);
    cudaMalloc(&sink, 64); cudaMalloc(&ds, 64); cudaMallocManaged(&out, 64);
    auto rep = [&](const char *n) { cudaDeviceSynchronize(); printf("%-28s %lld cycles  (%s)\n", n, out[0], cudaGetErrorString(cudaGetLastError())); };
    for (int w = 0; w < 2; ++w) {
        k_ldcg<<<1, 1>>>(p, out, sink); rep("ldcg dependent (L2)");
        k_ldca<<<1, 1>>>(p, out, sink); rep("ld dependent (L1)");
        k_atom<<<1, 1>>>(p, out, sink); rep("atomicOr returning");
        k_dadd<<<1, 32>>>(1.0, out, ds); rep("256 dependent DADD");
        k_dsetp<<<1, 32>>>(1.0, out, ds); rep("256 dependent DSETP+sel+DADD");
        k_shfl<<<1, 32>>>(1.0, out, ds); rep("256 dependent shfl(double)");
        k_pow<<<1, 32>>>(1.0, out, ds); rep("pow(double) dependent");
        k_sync<<<1, 512>>>(out); rep("__syncthreads 512 thr");
        k_fence<<<1, 512>>>(out, p); rep("stcg+__threadfence 512 thr");
        for (int cs : {1, 2, 8, 16}) {
            cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3(cs); cfg.blockDim = dim3(512);
            cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            if (cs > 8) cudaFuncSetAttribute(k_csync, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
            cudaLaunchKernelEx(&cfg, k_csync, out, p);
            char nm[64]; snprintf(nm, 64, "st+cluster.sync x%d", cs); rep(nm);
        }
    }
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0); printf("clock attr %d kHz\n", clk);
    return 0;
}
