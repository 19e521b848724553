// tcgen05 probe: the properties the fused Res2DBlock kernel (csrc/fvp_res2d.cu) builds on, checked on the device
// against a host loop before any of it is trusted:
//   * kind::tf32 MMA, M = 128, N = 32, K = 8 per instruction, both operands K-major in the NO-SWIZZLE canonical
//     layout: a "plane" per group of four channels, [plane][row][4 floats]; rows are 16 bytes apart everywhere
//     (SBO = 128 bytes), planes LBO bytes apart;
//   * a shared-memory descriptor may start at ANY row (start address in 16-byte units): an implicit-GEMM tap is a
//     shifted start address over one staged pixel raster;
//   * accumulation over taps / k-steps / M-tiles in TMEM, tcgen05.commit -> mbarrier, tcgen05.ld 32x32b.x32 (thread t of
//     warp w reads row 32w + t, 32 consecutive columns), tcgen05.st to pre-load an accumulator.
//     D[m][n] = init[m][n] + sum_t sum_c A[m + shift[t]][c] * W[t][n][c]
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o umma_probe tools/umma_probe.cu && ./umma_probe
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

constexpr int C = 32, NOUT = 32, TAPS = 9, MT = 3, ROWS = 128 * MT, NPX = ROWS + 160 + 1;  // NPX odd multiple: planes shift banks
constexpr int PLANES = C / 4;

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(unsigned saddr, unsigned lbo_bytes, unsigned sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fffu);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32;
    d |= (uint64_t)1 << 46;  // descriptor version of sm_100
    return d;                // base offset 0, layout type 0 = no swizzle
}

// c_format F32 (1 << 4), a/b format TF32 (2 << 7, 2 << 10), K-major both, N >> 3 at bit 17, M >> 4 at bit 24
constexpr unsigned IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(NOUT >> 3) << 17) | ((128u >> 4) << 24);

__global__ void __launch_bounds__(128, 1)
probe_kernel(const float *__restrict__ a, const float *__restrict__ w, const float *__restrict__ init, const int *__restrict__ shift,
             float *__restrict__ d) {
    extern __shared__ __align__(128) unsigned char smem[];
    float4 *sA = reinterpret_cast<float4 *>(smem);                 // [PLANES][NPX]
    float4 *sW = sA + PLANES * NPX;                               // [TAPS][PLANES][NOUT]
    __shared__ __align__(8) unsigned long long bar;
    __shared__ unsigned tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;

    for (int e = tid; e < PLANES * NPX; e += 128) {
        const int pl = e / NPX, px = e - pl * NPX;
        sA[e] = *reinterpret_cast<const float4 *>(a + (size_t)px * C + pl * 4);
    }
    for (int e = tid; e < TAPS * PLANES * NOUT; e += 128) {
        const int t = e / (PLANES * NOUT), r = e - t * PLANES * NOUT, pl = r / NOUT, n = r - pl * NOUT;
        sW[e] = *reinterpret_cast<const float4 *>(w + ((size_t)t * NOUT + n) * C + pl * 4);
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 128;" ::"r"(smem_u32(&tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy smem writes -> visible to the MMA unit
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const unsigned tm = tmem_base;

    // pre-load the accumulators: thread t of warp w owns TMEM lane 32w + t
    for (int mt = 0; mt < MT; ++mt) {
        unsigned v[32];
        const float *src = init + ((size_t)mt * 128 + tid) * NOUT;
        for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(src[j]);
        const unsigned addr = tm + ((unsigned)(warp * 32) << 16) + mt * NOUT;
        asm volatile(
            "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,"
            "%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(addr),
            "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
            "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]),
            "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]),
            "r"(v[31]));
    }
    asm volatile("tcgen05.wait::st.sync.aligned;");
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");

    if (tid == 0) {
        const unsigned a0 = smem_u32(sA), w0 = smem_u32(sW);
        for (int mt = 0; mt < MT; ++mt)
            for (int t = 0; t < TAPS; ++t) {
                const int sh = shift[t];
                for (int ks = 0; ks < C / 8; ++ks) {
                    const uint64_t da = make_desc(a0 + (unsigned)((2 * ks * NPX + mt * 128 + sh) * 16), NPX * 16, 128);
                    const uint64_t db = make_desc(w0 + (unsigned)(((t * PLANES + 2 * ks) * NOUT) * 16), NOUT * 16, 128);
                    const unsigned dcol = tm + mt * NOUT;
                    asm volatile(
                        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(dcol),
                        "l"(da), "l"(db), "r"(IDESC), "r"(1u)
                        : "memory");
                }
            }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    {   // everyone waits for the MMAs
        unsigned done = 0;
        while (!done)
            asm volatile(
                "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                : "=r"(done)
                : "r"(smem_u32(&bar)), "r"(0u)
                : "memory");
    }
    asm volatile("tcgen05.fence::after_thread_sync;");
    for (int mt = 0; mt < MT; ++mt) {
        unsigned v[32];
        const unsigned addr = tm + ((unsigned)(warp * 32) << 16) + mt * NOUT;
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,"
            "%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
              "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
              "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
              "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
            : "r"(addr));
        asm volatile("tcgen05.wait::ld.sync.aligned;");
        float *dst = d + ((size_t)mt * 128 + tid) * NOUT;
        for (int j = 0; j < 32; ++j) dst[j] = __uint_as_float(v[j]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 128;" ::"r"(tm));
}


// Issue rate of the M = 128, K = 8 TF32 MMA from shared-memory operands as a function of N (the A operand, 4 KB per
// instruction, is re-read for every instruction): cycles per MMA over a long dependent-accumulator stream.
template <int N>
__global__ void __launch_bounds__(128, 1) rate_kernel(long long *out, int iters) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) unsigned long long bar;
    __shared__ unsigned tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int e = tid; e < (PLANES * NPX + 2 * 256) * 4; e += 128) reinterpret_cast<float *>(smem)[e] = 0.0f;
    if (tid == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const unsigned tm = tmem_base;
    constexpr unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(N >> 3) << 17) | ((128u >> 4) << 24);
    long long t0 = 0, t1 = 0;
    if (tid == 0) {
        const unsigned a0 = smem_u32(smem), w0 = a0 + PLANES * NPX * 16;
        t0 = clock64();
        for (int i = 0; i < iters; ++i) {
            const int mt = i & 3, sh = (i * 7) & 63, ks = i & 3;
            const uint64_t da = make_desc(a0 + (unsigned)((2 * ks * NPX + sh) * 16), NPX * 16, 128);
            const uint64_t db = make_desc(w0, N * 16, 128);
            asm volatile(
                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tm + (unsigned)(mt * (N <= 128 ? N : 0))),
                "l"(da), "l"(db), "r"(idesc), "r"(1u)
                : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    unsigned done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
    if (tid == 0) { t1 = clock64(); out[0] = t1 - t0; }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm));
}

template <int N>
static int run_rate() {
    long long *dout, h = 0;
    if (cudaMalloc(&dout, 8) != cudaSuccess) return 2;
    const size_t smem = 16 * ((size_t)PLANES * NPX + 2 * 256);
    cudaFuncSetAttribute(rate_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int iters = 20000;
    rate_kernel<N><<<1, 128, smem>>>(dout, iters);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("rate<%d>: %s\n", N, cudaGetErrorString(cudaGetLastError())); return 2; }
    cudaMemcpy(&h, dout, 8, cudaMemcpyDeviceToHost);
    printf("rate: M=128 N=%-3d K=8 tf32: %.1f cycles per MMA (%d issued back to back, one CTA)\n", N, (double)h / iters, iters);
    cudaFree(dout);
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
// Weight-stationary form (tcgen05.mma.ws): A = the filters, M = 32 output channels; B = 256 pixels of the staged raster
// (K-major planes, shifted start address per tap); D[cout][pixel] expected at TMEM lane cout + 32 * (pixel / 64),
// column pixel % 64 (the "1 x 4 datapath" layout of M = 32).  Also its issue rate.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int WS_N = 256, WS_TILES = 2;
template <int WM> __host__ __device__ constexpr unsigned idesc_ws() { return (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(WS_N >> 3) << 17) | ((unsigned)(WM >> 4) << 24); }

template <int WM>
__global__ void __launch_bounds__(128, 1)
ws_kernel(const float *__restrict__ a, const float *__restrict__ w, const int *__restrict__ shift, float *__restrict__ dump,
          long long *cycles, int rate_iters) {
    extern __shared__ __align__(128) unsigned char smem[];
    float4 *sA = reinterpret_cast<float4 *>(smem);   // pixels [PLANES][NPXW]
    constexpr int NPXW = WS_TILES * WS_N + 160 + 1;
    float4 *sW = sA + PLANES * NPXW;                 // filters [TAPS][PLANES][WM]
    constexpr int COLS = WS_N * WM / 128;            // TMEM columns of one accumulator tile
    constexpr unsigned IDESC_WS = idesc_ws<WM>();
    __shared__ __align__(8) unsigned long long bar[2];
    __shared__ unsigned tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int e = tid; e < PLANES * NPXW; e += 128) {
        const int pl = e / NPXW, px = e - pl * NPXW;
        sA[e] = *reinterpret_cast<const float4 *>(a + (size_t)px * C + pl * 4);
    }
    for (int e = tid; e < TAPS * PLANES * WM; e += 128) {
        const int t = e / (PLANES * WM), r = e - t * PLANES * WM, pl = r / WM, n = r - pl * WM;
        sW[e] = *reinterpret_cast<const float4 *>(w + ((size_t)t * WM + n) * C + pl * 4);
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[0])));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[1])));
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const unsigned tm = tmem_base;
    const unsigned a0 = smem_u32(sA), w0 = smem_u32(sW);
    if (tid == 0) {
        for (int mt = 0; mt < WS_TILES; ++mt)
            for (int t = 0; t < TAPS; ++t)
                for (int ks = 0; ks < C / 8; ++ks) {
                    const uint64_t dw = make_desc(w0 + (unsigned)(((t * PLANES + 2 * ks) * WM) * 16), WM * 16, 128);
                    const uint64_t dp = make_desc(a0 + (unsigned)((2 * ks * NPXW + mt * WS_N + shift[t]) * 16), NPXW * 16, 128);
                    asm volatile(
                        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                        "tcgen05.mma.ws.cta_group::1.kind::tf32.collector::b0::discard [%0], %1, %2, %3, p;\n\t}" ::"r"(tm + mt * COLS),
                        "l"(dw), "l"(dp), "r"(IDESC_WS), "r"((unsigned)((t | ks) != 0))
                        : "memory");
                }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar[0])) : "memory");
    }
    unsigned done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(smem_u32(&bar[0])), "r"(0u) : "memory");
    asm volatile("tcgen05.fence::after_thread_sync;");
    for (int blk = 0; blk < WS_TILES * COLS / 32; ++blk) {   // 32 columns at a time: dump[lane][col]
        unsigned v[32];
        const unsigned addr = tm + ((unsigned)(warp * 32) << 16) + blk * 32;
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,"
            "%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
              "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
              "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
              "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
            : "r"(addr));
        asm volatile("tcgen05.wait::ld.sync.aligned;");
        for (int j = 0; j < 32; ++j) dump[(size_t)tid * (WS_TILES * COLS) + blk * 32 + j] = __uint_as_float(v[j]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    // issue rate: the same instruction back to back (8 accumulator tiles in rotation)
    if (tid == 0) {
        const long long t0 = clock64();
        for (int i = 0; i < rate_iters; ++i) {
            const int ks = i & 3, t = i % 9;
            const uint64_t dw = make_desc(w0 + (unsigned)(((t * PLANES + 2 * ks) * WM) * 16), WM * 16, 128);
            const uint64_t dp = make_desc(a0 + (unsigned)((2 * ks * NPXW + ((i * 7) & 63)) * 16), NPXW * 16, 128);
            asm volatile(
                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                "tcgen05.mma.ws.cta_group::1.kind::tf32.collector::b0::discard [%0], %1, %2, %3, p;\n\t}" ::"r"(tm + (unsigned)((i % (512 / COLS)) * COLS)),
                "l"(dw), "l"(dp), "r"(IDESC_WS), "r"(1u)
                : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar[1])) : "memory");
        unsigned ok = 0;
        while (!ok)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok) : "r"(smem_u32(&bar[1])), "r"(0u) : "memory");
        cycles[0] = clock64() - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm));
}

template <int WM>
static int run_ws(const std::vector<float> &a, const std::vector<float> &w0, const std::vector<int> &shift) {
    constexpr int COLS = WS_N * WM / 128;
    std::vector<float> w((size_t)TAPS * WM * C);
    srand(11 + WM);
    for (auto &x : w) x = (float)(rand() % 33 - 16) / 8.0f;
    (void)w0;
    constexpr int NPXW = WS_TILES * WS_N + 160 + 1;
    float *da, *dw, *dd;
    int *ds;
    long long *dc, cyc = 0;
    const size_t nd = (size_t)128 * WS_TILES * COLS;
    std::vector<float> dump(nd, -1.0f);
    if (cudaMalloc(&da, (size_t)NPXW * C * 4) || cudaMalloc(&dw, w.size() * 4) || cudaMalloc(&dd, nd * 4) || cudaMalloc(&ds, 9 * 4) ||
        cudaMalloc(&dc, 8))
        return 2;
    cudaMemcpy(da, a.data(), (size_t)NPXW * C * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dw, w.data(), w.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(ds, shift.data(), 9 * 4, cudaMemcpyHostToDevice);
    const size_t smem = 16 * ((size_t)PLANES * NPXW + (size_t)TAPS * PLANES * WM);
    cudaFuncSetAttribute(ws_kernel<WM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int iters = 20000;
    ws_kernel<WM><<<1, 128, smem>>>(da, dw, ds, dd, dc, iters);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("ws_kernel: %s\n", cudaGetErrorString(cudaGetLastError())); return 2; }
    cudaMemcpy(dump.data(), dd, nd * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(&cyc, dc, 8, cudaMemcpyDeviceToHost);
    long bad = 0;
    for (int mt = 0; mt < WS_TILES; ++mt)
        for (int m = 0; m < WM; ++m)
            for (int n = 0; n < WS_N; ++n) {
                double ref = 0;
                for (int t = 0; t < TAPS; ++t)
                    for (int c = 0; c < C; ++c)
                        ref += (double)w[((size_t)t * WM + m) * C + c] * a[(size_t)(mt * WS_N + n + shift[t]) * C + c];
                constexpr int G = 128 / WM;   // groups of WS_N / G positions side by side in the lanes
                const int lane = m + WM * (n / (WS_N / G)), col = mt * COLS + n % (WS_N / G);
                const float got = dump[(size_t)lane * (WS_TILES * COLS) + col];
                if (got != (float)ref && bad++ < 6) printf("ws mismatch tile %d cout %d px %d: lane %d col %d got %f want %f\n", mt, m, n, lane, col, got, ref);
            }
    printf("umma_probe.ws: M=%d (filters) x N=256 (pixels), %d tiles, D[cout][px] at lane cout + %d*(px/%d), column px%%%d: %ld mismatches -> %s\n",
           WM, WS_TILES, WM, WS_N * WM / 128, WS_N * WM / 128, bad, bad ? "FAIL" : "PASS (exact)");
    printf("rate: .ws M=%d N=256 K=8 tf32: %.1f cycles per MMA (%d back to back)\n", WM, (double)cyc / iters, iters);
    return bad ? 1 : 0;
}

#define CK(x)                                                                                 \
    do {                                                                                      \
        cudaError_t e_ = (x);                                                                 \
        if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 2; } \
    } while (0)

int main() {
    // values exactly representable in TF32 with exactly representable sums: multiples of 1/8 in [-2, 2]
    std::vector<float> a((size_t)NPX * C), w((size_t)TAPS * NOUT * C), init((size_t)ROWS * NOUT), d((size_t)ROWS * NOUT, -1.0f);
    std::vector<int> shift = {0, 1, 2, 68, 69, 70, 136, 137, 138};   // the nine taps of a 3x3 stencil over a 68-wide raster
    srand(7);
    for (auto &x : a) x = (float)(rand() % 33 - 16) / 8.0f;
    for (auto &x : w) x = (float)(rand() % 33 - 16) / 8.0f;
    for (auto &x : init) x = (float)(rand() % 65 - 32) / 4.0f;
    float *da, *dw, *di, *dd;
    int *ds;
    CK(cudaMalloc(&da, a.size() * 4)); CK(cudaMalloc(&dw, w.size() * 4)); CK(cudaMalloc(&di, init.size() * 4));
    CK(cudaMalloc(&dd, d.size() * 4)); CK(cudaMalloc(&ds, shift.size() * 4));
    CK(cudaMemcpy(da, a.data(), a.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dw, w.data(), w.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(di, init.data(), init.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ds, shift.data(), shift.size() * 4, cudaMemcpyHostToDevice));
    const size_t smem = sizeof(float4) * ((size_t)PLANES * NPX + (size_t)TAPS * PLANES * NOUT);
    CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    probe_kernel<<<1, 128, smem>>>(da, dw, di, ds, dd);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(d.data(), dd, d.size() * 4, cudaMemcpyDeviceToHost));
    long bad = 0;
    double worst = 0;
    for (int m = 0; m < ROWS; ++m)
        for (int n = 0; n < NOUT; ++n) {
            double ref = init[(size_t)m * NOUT + n];
            for (int t = 0; t < TAPS; ++t)
                for (int c = 0; c < C; ++c) ref += (double)a[(size_t)(m + shift[t]) * C + c] * w[((size_t)t * NOUT + n) * C + c];
            const double err = fabs(ref - d[(size_t)m * NOUT + n]);
            if (err > worst) worst = err;
            if (err != 0.0 && bad++ < 5) printf("mismatch m=%d n=%d got %f want %f\n", m, n, d[(size_t)m * NOUT + n], ref);
        }
    printf("umma_probe: %d x %d outputs, %d taps x %d channels, smem %zu B: %ld mismatches, worst |err| %g -> %s\n", ROWS, NOUT, TAPS,
           C, smem, bad, worst, bad ? "FAIL" : "PASS (exact)");
    for (int ws_rc : {run_ws<32>(a, w, shift), run_ws<64>(a, w, shift)}) {   // M = 128: the nine taps of 128 filters do not fit beside the raster
        if (ws_rc == 2) return 2;
        bad += ws_rc;
    }
    if (run_rate<32>() || run_rate<64>() || run_rate<128>() || run_rate<256>()) return 2;
    return bad ? 1 : 0;
}
