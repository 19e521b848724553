// Projection + gather kernels and their C entry points.
//
// Data layout in HBM
//   heatmaps   channels-last [B][V][H][W][Jp], Jp = 4*ceil(J/4): one bilinear tap of all joints is
//              Jp*4 contiguous bytes, read as Jp/4 128-bit loads by Jp/4 adjacent lanes.
//   cubes      [C][J][nx][ny][nz] (the reference's layout, z fastest).
//   cameras    [R][V][24] floats, axis tables 3 x 1-D floats, windows [C][12] ints.
#include "fvp_geom.cuh"

namespace fvp {

// ------------------------------------------------------------------------------------------------
// a1-a4: the per-sequence sample grid of the reference (setup kernel / cached-grid A/B mode)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
sample_grid_kernel(const float *__restrict__ cams, const __grid_constant__ fvp_geom g,
                   const float *__restrict__ ax, const float *__restrict__ ay, const float *__restrict__ az,
                   int X, int Y, int Z, float2 *__restrict__ out) {
    const long nb = (long)X * Y * Z;
    const int v = blockIdx.y;
    const Camera cam = load_camera(cams + (size_t)v * FVP_CAM_STRIDE);
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nb; i += (long)gridDim.x * blockDim.x) {
        const int z = (int)(i % Z), y = (int)((i / Z) % Y), x = (int)(i / ((long)Z * Y));
        float gx, gy;
        sample_coords(cam, g, __ldg(ax + x), __ldg(ay + y), __ldg(az + z), gx, gy);
        out[(size_t)v * nb + i] = make_float2(gx, gy);
    }
}

// ------------------------------------------------------------------------------------------------
// layout pre-pass: planar [BV][J][H][W] -> channels-last [BV][H][W][Jp]
// ------------------------------------------------------------------------------------------------
constexpr int CL_PIX = 64;  // pixels per block

__global__ void __launch_bounds__(256)
channels_last_kernel(const float *__restrict__ in, int J, int HW, int Jp, float *__restrict__ out) {
    extern __shared__ float tile[];  // [Jp][CL_PIX + 1]
    const int bv = blockIdx.y;
    const int pix0 = blockIdx.x * CL_PIX;
    const int npix = min(CL_PIX, HW - pix0);
    const float *src = in + (size_t)bv * J * HW + pix0;
    for (int e = threadIdx.x; e < J * CL_PIX; e += blockDim.x) {
        const int j = e / CL_PIX, p = e % CL_PIX;
        if (p < npix) tile[j * (CL_PIX + 1) + p] = __ldg(src + (size_t)j * HW + p);
    }
    __syncthreads();
    float *dst = out + ((size_t)bv * HW + pix0) * Jp;
    for (int e = threadIdx.x; e < npix * Jp; e += blockDim.x) {
        const int p = e / Jp, j = e % Jp;
        dst[e] = j < J ? tile[j * (CL_PIX + 1) + p] : 0.0f;
    }
}

// ------------------------------------------------------------------------------------------------
// a11-a12: per-proposal window arithmetic (project_individual.py:255-269,274)
// ------------------------------------------------------------------------------------------------
struct WindowConsts {
    float scale[3], bias[3], whole[3], ind[3];
    int fine[3], voxels[3];
};

__global__ void individual_windows_kernel(const float *__restrict__ props, const uint8_t *__restrict__ valid,
                                          const int *__restrict__ frame, int N,
                                          const __grid_constant__ WindowConsts c, int *__restrict__ win,
                                          float *__restrict__ offset) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    int bad = (valid != nullptr && valid[n] == 0) ? 1 : 0;
    int *w = win + (size_t)n * FVP_WIN_STRIDE;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float ctr = __fadd_rn(__fmul_rn(props[7 * n + a], c.scale[a]), c.bias[a]);
        const int t = __float2int_rn(ctr);  // torch.round: half to even
        const float off = __fadd_rn(
            __fsub_rn(__fmul_rn(__fdiv_rn((float)t, (float)(c.fine[a] - 1)), c.whole[a]), __fdiv_rn(c.whole[a], 2.0f)),
            __fdiv_rn(c.ind[a], 2.0f));
        int m = 0;
        if (a < 2) {
            const float mf =
                __fmul_rn(__fdiv_rn(__fsub_rn(1.0f, props[7 * n + 5 + a]), 2.0f), (float)(c.voxels[a] - 1));
            m = (int)mf;  // .int() truncates toward zero
            if (m < 0) m = 0;
        }
        const int s = t + m >= 0 ? t + m : 0;
        const int e = t + c.voxels[a] - m <= c.fine[a] ? t + c.voxels[a] - m : c.fine[a];
        w[a] = t;
        w[3 + a] = s;
        w[6 + a] = e;
        offset[3 * n + a] = off;
        bad |= (s >= e);
    }
    w[9] = bad;
    w[10] = frame[n];
    w[11] = 0;
}

// ------------------------------------------------------------------------------------------------
// a5 / a12: fused projection + bilinear gather + view mean + clamp, "direct" mode
// ------------------------------------------------------------------------------------------------
// One CTA = one brick of BX*BY*BZ = 256 voxels of one cube.
//   phase 1  thread = voxel: project into each view (exact fp32 order), set up the four taps,
//            hand (tap coordinates, weights) over through shared memory;
//   phase 2  Q = Jp/4 adjacent lanes = one voxel: each lane gathers 4 channels of every tap with one
//            128-bit read-only load and runs the reference's fma chain on them.
// The brick is 4x8x8 (x,y,z): 8 consecutive z per warp instruction keep cube stores in whole 32-byte
// sectors and the four-lane groups of a warp land on neighbouring pixels.
constexpr int BX = 4, BY = 8, BZ = 8, BRICK = BX * BY * BZ;
constexpr int VCHUNK = 8;  // views staged per pass: 8 * 256 * 24 B = 48 KB of dynamic shared memory

struct ProjectParams {
    const float *hm;         // [B][V][H][W][Jp]
    const float *cams;       // [R][V][24]
    const int *frame_rig;    // [B]
    const float *ax, *ay, *az;
    const float2 *grid;      // cached sample grid or null
    const int *win;          // [C][12] or null
    float *cubes;            // [C][J][nx][ny][nz]
    fvp_geom g;
    int Jp, n_ax, n_ay, n_az, nx, ny, nz, nbx, nby, nbz;
    float half_wm1, half_hm1, fV;
    int vchunk;              // views staged per pass (<= VCHUNK)
};

__device__ __forceinline__ float finish_voxel(float acc, float fV) {
    // mean over ALL views (F4), clamp(0,1) keeping NaN, and canonicalise -0 to +0
    float v = __fdiv_rn(acc, fV);
    v = clamp_keep_nan(v, 0.0f, 1.0f);
    return __fadd_rn(v, 0.0f);
}

template <int Q>
__global__ void __launch_bounds__(BRICK)
project_direct_kernel(const __grid_constant__ ProjectParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4 *s_w = reinterpret_cast<float4 *>(smem_raw);                             // [vchunk][BRICK]
    uint2 *s_xy = reinterpret_cast<uint2 *>(smem_raw + sizeof(float4) * p.vchunk * BRICK);  // [vchunk][BRICK]

    const int tid = threadIdx.x;
    const int c = blockIdx.y;
    int brick = blockIdx.x;
    const int bz = brick % p.nbz; brick /= p.nbz;
    const int by = brick % p.nby;
    const int bx = brick / p.nby;

    // window of this cube
    int tl[3] = {0, 0, 0}, st[3] = {0, 0, 0}, en[3] = {p.n_ax, p.n_ay, p.n_az}, skip = 0, frame = c;
    if (p.win != nullptr) {
        const int *w = p.win + (size_t)c * FVP_WIN_STRIDE;
#pragma unroll
        for (int a = 0; a < 3; ++a) { tl[a] = __ldg(w + a); st[a] = __ldg(w + 3 + a); en[a] = __ldg(w + 6 + a); }
        skip = __ldg(w + 9);
        frame = __ldg(w + 10);
    }
    const int V = p.g.V, J = p.g.J, H = p.g.H, W = p.g.W;

    // phase-1 voxel of this thread
    const int i = bx * BX + (tid >> 6), j = by * BY + ((tid >> 3) & 7), k = bz * BZ + (tid & 7);
    const int gi = tl[0] + i, gj = tl[1] + j, gk = tl[2] + k;
    const bool in_cube = i < p.nx && j < p.ny && k < p.nz;
    const bool in_win = in_cube && !skip && gi >= st[0] && gi < en[0] && gj >= st[1] && gj < en[1] &&
                        gk >= st[2] && gk < en[2];
    const bool any = __syncthreads_or(in_win);

    float acc[Q][4];
#pragma unroll
    for (int q = 0; q < Q; ++q) { acc[q][0] = acc[q][1] = acc[q][2] = acc[q][3] = 0.0f; }

    if (any) {
        const int rig = __ldg(p.frame_rig + frame);
        float wx = 0.f, wy = 0.f, wz = 0.f;
        if (in_win) { wx = __ldg(p.ax + gi); wy = __ldg(p.ay + gj); wz = __ldg(p.az + gk); }
        const size_t ntab = (size_t)p.n_ax * p.n_ay * p.n_az;
        const size_t tab = in_win ? ((size_t)gi * p.n_ay + gj) * p.n_az + gk : 0;

        for (int v0 = 0; v0 < V; v0 += p.vchunk) {
            const int nv = min(p.vchunk, V - v0);
            if (v0) __syncthreads();  // previous chunk fully consumed
            for (int vv = 0; vv < nv; ++vv) {
                float4 wgt = make_float4(0.f, 0.f, 0.f, 0.f);
                uint2 xy = make_uint2(0u, 0u);
                if (in_win) {
                    float gx, gy;
                    if (p.grid != nullptr) {
                        const float2 gg = __ldg(p.grid + ((size_t)rig * V + v0 + vv) * ntab + tab);
                        gx = gg.x; gy = gg.y;
                    } else {
                        const Camera cam = load_camera(p.cams + ((size_t)rig * V + v0 + vv) * FVP_CAM_STRIDE);
                        sample_coords(cam, p.g, wx, wy, wz, gx, gy);
                    }
                    const Taps t = tap_setup(gx, gy, W, H, p.half_wm1, p.half_hm1);
                    wgt = make_float4(t.nw, t.ne, t.sw, t.se);
                    xy = make_uint2((unsigned)t.x0 | ((unsigned)t.x1 << 16), (unsigned)t.y0 | ((unsigned)t.y1 << 16));
                }
                s_w[vv * BRICK + tid] = wgt;
                s_xy[vv * BRICK + tid] = xy;
            }
            __syncthreads();

            // phase 2
#pragma unroll
            for (int q = 0; q < Q; ++q) {
                const int item = tid + q * BRICK;
                const int lv = item / Q, quad = item % Q;
                for (int vv = 0; vv < nv; ++vv) {
                    const float4 wgt = s_w[vv * BRICK + lv];
                    if (wgt.x == 0.0f && wgt.y == 0.0f && wgt.z == 0.0f && wgt.w == 0.0f) continue;
                    const uint2 xy = s_xy[vv * BRICK + lv];
                    const int x0 = xy.x & 0xffff, x1 = xy.x >> 16, y0 = xy.y & 0xffff, y1 = xy.y >> 16;
                    const float4 *img = reinterpret_cast<const float4 *>(
                                            p.hm + ((size_t)frame * V + v0 + vv) * ((size_t)H * W * p.Jp)) + quad;
                    const int qs = p.Jp >> 2;  // float4 per pixel
                    const float4 a = __ldg(img + (size_t)(y0 * W + x0) * qs);
                    const float4 b = __ldg(img + (size_t)(y0 * W + x1) * qs);
                    const float4 cc = __ldg(img + (size_t)(y1 * W + x0) * qs);
                    const float4 d = __ldg(img + (size_t)(y1 * W + x1) * qs);
                    // out = fma(se_val,se, fma(sw_val,sw, fma(ne_val,ne, nw_val*nw))); sum views in order
                    float t0 = __fmaf_rn(d.x, wgt.w, __fmaf_rn(cc.x, wgt.z, __fmaf_rn(b.x, wgt.y, __fmul_rn(a.x, wgt.x))));
                    float t1 = __fmaf_rn(d.y, wgt.w, __fmaf_rn(cc.y, wgt.z, __fmaf_rn(b.y, wgt.y, __fmul_rn(a.y, wgt.x))));
                    float t2 = __fmaf_rn(d.z, wgt.w, __fmaf_rn(cc.z, wgt.z, __fmaf_rn(b.z, wgt.y, __fmul_rn(a.z, wgt.x))));
                    float t3 = __fmaf_rn(d.w, wgt.w, __fmaf_rn(cc.w, wgt.z, __fmaf_rn(b.w, wgt.y, __fmul_rn(a.w, wgt.x))));
                    acc[q][0] = __fadd_rn(acc[q][0], t0);
                    acc[q][1] = __fadd_rn(acc[q][1], t1);
                    acc[q][2] = __fadd_rn(acc[q][2], t2);
                    acc[q][3] = __fadd_rn(acc[q][3], t3);
                }
            }
        }
    }

    // store: every voxel of the brick that lies inside the cube is written (0 outside the window)
    const size_t vox = (size_t)p.nx * p.ny * p.nz;
#pragma unroll
    for (int q = 0; q < Q; ++q) {
        const int item = tid + q * BRICK;
        const int lv = item / Q, quad = item % Q;
        const int oi = bx * BX + (lv >> 6), oj = by * BY + ((lv >> 3) & 7), ok = bz * BZ + (lv & 7);
        if (oi >= p.nx || oj >= p.ny || ok >= p.nz) continue;
        float *dst = p.cubes + ((size_t)c * J + quad * 4) * vox + ((size_t)oi * p.ny + oj) * p.nz + ok;
#pragma unroll
        for (int e = 0; e < 4; ++e)
            if (quad * 4 + e < J) dst[(size_t)e * vox] = finish_voxel(acc[q][e], p.fV);
    }
}

// ------------------------------------------------------------------------------------------------
// a6 / a13: orthographic max-planes of a stack of cubes (values >= 0 or NaN)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float max_keep_nan(float a, float b) {  // torch.max propagates NaN
    return (a != a || a > b) ? a : b;
}

__global__ void __launch_bounds__(256)
planes3_kernel(const float *__restrict__ cubes, int nx, int ny, int nz, float *__restrict__ xy,
               float *__restrict__ xz, float *__restrict__ yz) {
    extern __shared__ float sm[];
    const int pitch = nz + 1;
    float *slab = sm;                    // [ny][nz+1]
    float *run = sm + (size_t)ny * pitch;  // [ny][nz] running max over x
    const int cj = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    const float *src = cubes + (size_t)cj * nx * ny * nz;
    for (int x = 0; x < nx; ++x) {
        const float *sl = src + (size_t)x * ny * nz;
        for (int y = warp; y < ny; y += nwarp)
            for (int z = lane; z < nz; z += 32) {
                const float v = __ldg(sl + y * nz + z);
                slab[y * pitch + z] = v;
                if (yz != nullptr) run[y * nz + z] = x ? max_keep_nan(run[y * nz + z], v) : v;
            }
        __syncthreads();
        for (int r = threadIdx.x; r < ny + nz; r += blockDim.x) {
            if (r < ny) {
                if (xy != nullptr) {
                    float m = slab[r * pitch];
                    for (int z = 1; z < nz; ++z) m = max_keep_nan(m, slab[r * pitch + z]);
                    xy[((size_t)cj * nx + x) * ny + r] = m;
                }
            } else if (xz != nullptr) {
                const int z = r - ny;
                float m = slab[z];
                for (int y = 1; y < ny; ++y) m = max_keep_nan(m, slab[y * pitch + z]);
                xz[((size_t)cj * nx + x) * nz + z] = m;
            }
        }
        __syncthreads();
    }
    if (yz != nullptr)
        for (int y = warp; y < ny; y += nwarp)
            for (int z = lane; z < nz; z += 32) yz[((size_t)cj * ny + y) * nz + z] = run[y * nz + z];
}

}  // namespace fvp

// ================================================================================================
// C entry points
// ================================================================================================
using namespace fvp;

static int check_geom(const fvp_geom *g) {
    FVP_REQUIRE(g != nullptr, "geom is null");
    FVP_REQUIRE(g->V > 0 && g->J > 0 && g->H > 1 && g->W > 1, "geom: bad V/J/H/W (%d,%d,%d,%d)", g->V, g->J, g->H, g->W);
    FVP_REQUIRE(g->W <= 65535 && g->H <= 65535, "geom: heatmap larger than 65535 px per side");
    return FVP_OK;
}

extern "C" int fvp_sample_grid(const float *cams, const fvp_geom *g, const float *ax, const float *ay,
                               const float *az, int X, int Y, int Z, float *out, void *stream) {
    if (int rc = check_geom(g)) return rc;
    FVP_REQUIRE(cams && ax && ay && az && out, "fvp_sample_grid: null pointer");
    FVP_REQUIRE(X > 0 && Y > 0 && Z > 0, "fvp_sample_grid: bad dims");
    const long nb = (long)X * Y * Z;
    dim3 grid((unsigned)std::min<long>((nb + 255) / 256, 148L * 16), (unsigned)g->V);
    sample_grid_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(cams, *g, ax, ay, az, X, Y, Z,
                                                                 reinterpret_cast<float2 *>(out));
    return check_launch("sample_grid_kernel");
}

extern "C" int fvp_heatmaps_channels_last(const float *planar, int BV, int J, int H, int W, int Jp, float *out,
                                          void *stream) {
    FVP_REQUIRE(planar && out, "fvp_heatmaps_channels_last: null pointer");
    FVP_REQUIRE(BV > 0 && J > 0 && H > 0 && W > 0, "fvp_heatmaps_channels_last: bad dims");
    FVP_REQUIRE(Jp >= J && Jp % 4 == 0 && Jp <= 64, "fvp_heatmaps_channels_last: Jp must be a multiple of 4 in [J,64]");
    FVP_REQUIRE(BV <= 65535, "fvp_heatmaps_channels_last: B*V > 65535");
    const int HW = H * W;
    dim3 grid((unsigned)ceil_div(HW, CL_PIX), (unsigned)BV);
    const size_t smem = sizeof(float) * (size_t)Jp * (CL_PIX + 1);
    channels_last_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(planar, J, HW, Jp, out);
    return check_launch("channels_last_kernel");
}

extern "C" int fvp_individual_windows(const float *props, const uint8_t *valid, const int32_t *frame, int N,
                                      const float scale[3], const float bias[3], const int32_t fine[3],
                                      const float whole_size[3], const float ind_size[3], const int32_t voxels[3],
                                      int32_t *win, float *offset, void *stream) {
    FVP_REQUIRE(props && frame && win && offset, "fvp_individual_windows: null pointer");
    FVP_REQUIRE(scale && bias && fine && whole_size && ind_size && voxels, "fvp_individual_windows: null constants");
    if (N == 0) return FVP_OK;
    FVP_REQUIRE(N > 0, "fvp_individual_windows: N < 0");
    WindowConsts c;
    for (int a = 0; a < 3; ++a) {
        c.scale[a] = scale[a]; c.bias[a] = bias[a]; c.whole[a] = whole_size[a]; c.ind[a] = ind_size[a];
        c.fine[a] = fine[a]; c.voxels[a] = voxels[a];
        FVP_REQUIRE(fine[a] > 1 && voxels[a] > 0, "fvp_individual_windows: bad fine/voxels");
    }
    individual_windows_kernel<<<ceil_div(N, 128), 128, 0, (cudaStream_t)stream>>>(props, valid, frame, N, c, win, offset);
    return check_launch("individual_windows_kernel");
}

template <int Q>
static int launch_direct(const ProjectParams &p, dim3 grid, size_t smem, cudaStream_t s) {
    project_direct_kernel<Q><<<grid, BRICK, smem, s>>>(p);
    return check_launch("project_direct_kernel");
}

extern "C" int fvp_project_cubes(const float *hm_cl, int B, int Jp, const float *cams, const int32_t *frame_rig,
                                 const fvp_geom *g, const float *ax, const float *ay, const float *az, int n_ax,
                                 int n_ay, int n_az, const float *cached_grid, int C, const int32_t *win, int nx,
                                 int ny, int nz, float *cubes, unsigned flags, void *stream) {
    if (int rc = check_geom(g)) return rc;
    FVP_REQUIRE(hm_cl && cams && frame_rig && ax && ay && az && cubes, "fvp_project_cubes: null pointer");
    FVP_REQUIRE(B > 0 && n_ax > 0 && n_ay > 0 && n_az > 0 && nx > 0 && ny > 0 && nz > 0, "fvp_project_cubes: bad dims");
    FVP_REQUIRE(Jp >= g->J && Jp % 4 == 0 && Jp <= 32, "fvp_project_cubes: Jp must be a multiple of 4 in [J,32]");
    if (C == 0) return FVP_OK;
    FVP_REQUIRE(C > 0 && C <= 65535, "fvp_project_cubes: C out of range (1..65535)");
    FVP_REQUIRE(win != nullptr || (C <= B && nx <= n_ax && ny <= n_ay && nz <= n_az),
                "fvp_project_cubes: without windows cube c samples frame c over the whole table");
    const bool cached = (flags & FVP_FLAG_CACHED_GRID) != 0;
    FVP_REQUIRE(!cached || cached_grid != nullptr, "fvp_project_cubes: FVP_FLAG_CACHED_GRID needs cached_grid");
    const unsigned mode = flags & 0xfu;
    FVP_REQUIRE(mode == FVP_MODE_DIRECT, "fvp_project_cubes: mode %u not available in this build", mode);

    ProjectParams p;
    p.hm = hm_cl; p.cams = cams; p.frame_rig = frame_rig; p.ax = ax; p.ay = ay; p.az = az;
    p.grid = cached ? reinterpret_cast<const float2 *>(cached_grid) : nullptr;
    p.win = win; p.cubes = cubes; p.g = *g; p.Jp = Jp;
    p.n_ax = n_ax; p.n_ay = n_ay; p.n_az = n_az; p.nx = nx; p.ny = ny; p.nz = nz;
    p.nbx = ceil_div(nx, BX); p.nby = ceil_div(ny, BY); p.nbz = ceil_div(nz, BZ);
    p.half_wm1 = g->wm1 * 0.5f; p.half_hm1 = g->hm1 * 0.5f; p.fV = (float)g->V;
    p.vchunk = std::min(g->V, VCHUNK);
    const size_t smem = (size_t)p.vchunk * BRICK * (sizeof(float4) + sizeof(uint2));
    dim3 grid((unsigned)(p.nbx * p.nby * p.nbz), (unsigned)C);
    cudaStream_t s = (cudaStream_t)stream;
    switch (Jp / 4) {
        case 1: return launch_direct<1>(p, grid, smem, s);
        case 2: return launch_direct<2>(p, grid, smem, s);
        case 3: return launch_direct<3>(p, grid, smem, s);
        case 4: return launch_direct<4>(p, grid, smem, s);
        case 5: return launch_direct<5>(p, grid, smem, s);
        case 6: return launch_direct<6>(p, grid, smem, s);
        case 7: return launch_direct<7>(p, grid, smem, s);
        case 8: return launch_direct<8>(p, grid, smem, s);
    }
    return fail(FVP_E_ARG, "fvp_project_cubes: unsupported Jp %d", Jp);
}

extern "C" int fvp_planes3(const float *cubes, int CJ, int nx, int ny, int nz, float *xy, float *xz, float *yz,
                           void *stream) {
    FVP_REQUIRE(cubes != nullptr, "fvp_planes3: null cubes");
    FVP_REQUIRE(nx > 0 && ny > 0 && nz > 0, "fvp_planes3: bad dims");
    if (CJ == 0 || (!xy && !xz && !yz)) return FVP_OK;
    FVP_REQUIRE(CJ > 0, "fvp_planes3: CJ < 0");
    const size_t smem = sizeof(float) * ((size_t)ny * (nz + 1) + (size_t)ny * nz);
    FVP_REQUIRE(smem <= 200 * 1024, "fvp_planes3: ny*nz too large for shared memory (%zu B)", smem);
    if (smem > 48 * 1024)
        FVP_CUDA(cudaFuncSetAttribute(planes3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    planes3_kernel<<<CJ, 256, smem, (cudaStream_t)stream>>>(cubes, nx, ny, nz, xy, xz, yz);
    return check_launch("planes3_kernel");
}
