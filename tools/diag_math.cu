// Device self-check of the exact-division shortcuts and the f32x2 forms against the scalar ops.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../faster_voxelpose_b200/csrc/fvp_geom.cuh"
namespace fvp { char *error_buffer() { static char b[512]; return b; } }
using namespace fvp;
__device__ unsigned long long bad[8];
__device__ __forceinline__ uint32_t rng(uint32_t &s) { s ^= s << 13; s ^= s >> 17; s ^= s << 5; return s; }
__device__ __forceinline__ float rnd_float(uint32_t &s, int emin, int emax) {
    uint32_t m = rng(s) & 0x807fffffu; int e = emin + (int)(rng(s) % (uint32_t)(emax - emin + 1));
    return __uint_as_float(m | ((uint32_t)(e + 127) << 23));
}
__global__ void k(int iters, DivConst dc) {
    uint32_t s = 1234567u + blockIdx.x * 7919u + threadIdx.x * 104729u;
    for (int it = 0; it < iters; ++it) {
        float a0 = rnd_float(s, -12, 14), a1 = rnd_float(s, -12, 14), b = rnd_float(s, -3, 14);
        float2 q = div2_same_den(a0, a1, b);
        if (q.x != __fdiv_rn(a0, b) || q.y != __fdiv_rn(a1, b)) atomicAdd(&bad[0], 1ull);
        float c = div_const(a0, dc);
        if (c != __fdiv_rn(a0, dc.c)) atomicAdd(&bad[1], 1ull);
        float2 m = __fmul2_rn(make_float2(a0, a1), make_float2(b, a0));
        if (m.x != __fmul_rn(a0, b) || m.y != __fmul_rn(a1, a0)) atomicAdd(&bad[2], 1ull);
        float2 f = __ffma2_rn(make_float2(a0, a1), make_float2(b, a0), make_float2(a1, b));
        if (f.x != __fmaf_rn(a0, b, a1) || f.y != __fmaf_rn(a1, a0, b)) atomicAdd(&bad[3], 1ull);
        float2 ad = __fadd2_rn(make_float2(a0, a1), make_float2(b, a0));
        if (ad.x != __fadd_rn(a0, b) || ad.y != __fadd_rn(a1, a0)) atomicAdd(&bad[4], 1ull);
        float cl = clamp_keep_nan(a0, -1.0f, 1920.0f);
        float cr = a0 < -1.0f ? -1.0f : (a0 > 1920.0f ? 1920.0f : a0);
        if (cl != cr) atomicAdd(&bad[5], 1ull);
    }
}
int main() {
    float cs[] = {5.f, 960.f, 512.f, 239.f, 127.f, 3.f};
    for (float c : cs) {
        unsigned long long z[8] = {0}; cudaMemcpyToSymbol(bad, z, sizeof(z));
        DivConst dc; dc.c = c; dc.r = (float)(1.0L / (long double)c);
        k<<<592, 256>>>(2000, dc);
        cudaError_t e = cudaDeviceSynchronize();
        cudaMemcpyFromSymbol(z, bad, sizeof(z));
        printf("c=%g err=%s div2=%llu divc=%llu mul2=%llu fma2=%llu add2=%llu clamp=%llu\n", c, cudaGetErrorString(e), z[0], z[1], z[2], z[3], z[4], z[5]);
    }
    return 0;
}
