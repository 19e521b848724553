// Projection + gather kernels and their C entry points.
//
// Data layout in HBM
//   heatmaps   row-paired [B][V][2][H/2+1][W][Q][2][4], Q = ceil(J/4) (fvp_heatmaps_rowpair): the north and
//              south rows of a bilinear footprint of 4 joints are 32 contiguous bytes, one 256-bit load.
//   cubes      [C][J][nx][ny][nz] (the reference's layout, z fastest).
//   cameras    [R][V][24] floats, axis tables 3 x 1-D floats, windows [C][12] ints.
#include <cstring>

#include "fvp_project.cuh"

namespace fvp {

// ------------------------------------------------------------------------------------------------
// a1-a4: the per-sequence sample grid of the reference (setup kernel / cached-grid A/B mode)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
sample_grid_kernel(const float *__restrict__ cams, const __grid_constant__ GeomDev g,
                   const float *__restrict__ ax, const float *__restrict__ ay, const float *__restrict__ az,
                   int X, int Y, int Z, float2 *__restrict__ out) {
    const long nb = (long)X * Y * Z;
    const int v = blockIdx.y;
    const Camera cam = load_camera(cams + (size_t)v * FVP_CAM_STRIDE);
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nb; i += (long)gridDim.x * blockDim.x) {
        const int z = (int)(i % Z), y = (int)((i / Z) % Y), x = (int)(i / ((long)Z * Y));
        out[(size_t)v * nb + i] = sample_coords(cam, g, __ldg(ax + x), __ldg(ay + y), __ldg(az + z));
    }
}

// a1-a4 + the tap set-up of grid_sample, per (view, voxel): what a rig's sample grid becomes once the heatmap
// size is fixed.  Per view: npad float4 weights (nw, ne, sw, se) followed by npad packed records whose pixel
// index already carries the view's offset inside a frame (tap_setup in fvp_geom.cuh).
__global__ void __launch_bounds__(256)
tap_table_kernel(const float *__restrict__ cams, const __grid_constant__ GeomDev g,
                 const float *__restrict__ ax, const float *__restrict__ ay, const float *__restrict__ az,
                 int X, int Y, int Z, long npad, unsigned char *__restrict__ out) {
    const long nb = (long)X * Y * Z;
    const int v = blockIdx.y;
    const Camera cam = load_camera(cams + (size_t)v * FVP_CAM_STRIDE);
    float4 *wgt_out = reinterpret_cast<float4 *>(out + (size_t)v * npad * TAP_ENTRY_BYTES);
    unsigned *rec_out = reinterpret_cast<unsigned *>(out + (size_t)v * npad * TAP_ENTRY_BYTES + (size_t)npad * 16);
    const unsigned view_px = 2u * (unsigned)(g.HP * g.W);
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nb; i += (long)gridDim.x * blockDim.x) {
        const int z = (int)(i % Z), y = (int)((i / Z) % Y), x = (int)(i / ((long)Z * Y));
        float4 wgt;
        unsigned rec = tap_setup(sample_coords(cam, g, __ldg(ax + x), __ldg(ay + y), __ldg(az + z)), g, wgt);
        if (rec & TAP_LIVE) rec += (unsigned)v * view_px;
        wgt_out[i] = wgt;
        rec_out[i] = rec;
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
// layout pre-pass for the projection kernels: planar [BV][J][H][W] -> row-paired
// [BV][2][HP][W][Q][2][4] (see tap_setup in fvp_geom.cuh).  One block = one (bv, copy, pair) row and
// a chunk of RP_PIX columns: the two source rows of all joints are read along x (coalesced) into
// shared memory and written back as whole 8*Jp-byte pixel records.
// ------------------------------------------------------------------------------------------------
constexpr int RP_PIX = 128;

// One CTA: 128 pixels of rows 2m-1, 2m, 2m+1 of every joint of one (frame, view) -> pair m of both copies
// (copy 0 = rows 2m, 2m+1; copy 1 = rows 2m-1, 2m).  QT > 0 fixes Q at compile time (index arithmetic is
// what bounds this kernel, not HBM); QT == 0 is the generic form.  VEC: W % 4 == 0, 16-byte global loads.
// SPLIT (17 joints, QT == 5): channels 0..15 go out as 128-byte records (Q = 4) and channel 16 as one float4 per
// record position, {north x, south x, north x+1, south x+1}, into `ext` -- the layout of the row-walking kernel
// for 17-joint rigs (a 160-byte record would straddle two 128-byte lines).  Occupancy covers all 17 channels.
template <int QT, bool VEC, bool SPLIT = false>
__global__ void __launch_bounds__(256)
rowpair_kernel(const float *__restrict__ in, int J, int H, int W, int Qrt, int HP, float *__restrict__ out,
               unsigned *__restrict__ occ, int occ_stride, int V, float4 *__restrict__ ext) {
    extern __shared__ float tile[];  // [3][4Q][RP_PIX + 1]
    constexpr int pitch = RP_PIX + 1;
    static_assert(!SPLIT || QT == 5, "the split layout is 16 + 1 channels");
    const int Q = QT > 0 ? QT : Qrt, Jp = 4 * Q;
    const int bv = blockIdx.z, m = blockIdx.y;
    const int x0 = blockIdx.x * RP_PIX, npix = min(RP_PIX, W - x0);
    const float *src = in + (size_t)bv * J * H * W;
    if (VEC) {
        for (int e = threadIdx.x; e < 3 * Jp * (RP_PIX / 4); e += 256) {
            const int x = (e & (RP_PIX / 4 - 1)) * 4, rj = e / (RP_PIX / 4), r = rj / Jp, j = rj - r * Jp;
            const int row = 2 * m - 1 + r;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (x < npix && j < J && row >= 0 && row < H)
                v = __ldg(reinterpret_cast<const float4 *>(src + ((size_t)j * H + row) * W + x0 + x));
            float *t = tile + rj * pitch + x;
            t[0] = v.x; t[1] = v.y; t[2] = v.z; t[3] = v.w;
        }
    } else {
        for (int e = threadIdx.x; e < 3 * Jp * RP_PIX; e += 256) {
            const int x = e & (RP_PIX - 1), rj = e / RP_PIX, r = rj / Jp, j = rj - r * Jp;
            const int row = 2 * m - 1 + r;
            float v = 0.0f;
            if (x < npix && j < J && row >= 0 && row < H) v = __ldg(src + ((size_t)j * H + row) * W + x0 + x);
            tile[rj * pitch + x] = v;
        }
    }
    if (SPLIT && threadIdx.x < 3) {   // channel 16 of the pixel right of this chunk, into the padding column
        const int row = 2 * m - 1 + (int)threadIdx.x;
        float v = 0.0f;
        if (x0 + RP_PIX < W && row >= 0 && row < H) v = __ldg(src + ((size_t)16 * H + row) * W + x0 + RP_PIX);
        tile[(threadIdx.x * Jp + 16) * pitch + RP_PIX] = v;
    }
    __syncthreads();
    // Occupancy of the pixel records (optional): per frame, word 0 counts the records with any non-zero bit, then one
    // bit per strip of 8 consecutive records, set when any of the records [8s, 8s+8] is non-zero -- the west AND the
    // east record of every sample whose west record lies in the strip.  The projection kernels drop samples over
    // clear strips: all their taps are +0 and add exactly nothing (input heatmaps rendered from 2-D poses are mostly 0).
    if (occ != nullptr) {
        const int copy = threadIdx.x >> 7, x = threadIdx.x & (RP_PIX - 1), lane = threadIdx.x & 31;
        bool nz = false;
        if (x < npix) {
            const float *t = tile + (size_t)(1 - copy) * Jp * pitch + x;   // copy 0: tile rows 1, 2; copy 1: rows 0, 1
            for (int rj = 0; rj < 2 * Jp; ++rj) nz |= __float_as_uint(t[rj * pitch]) != 0u;
        }
        const unsigned b = __ballot_sync(0xffffffffu, nz);
        if (b) {
            const int frame = bv / V, view = bv - frame * V;
            unsigned *occf = occ + (size_t)frame * occ_stride;
            const long rec0 = ((long)(view * 2 + copy) * HP + m) * W + x0 + (x & ~31);   // record of lane 0, frame relative
            if (lane == 0) atomicAdd(occf, (unsigned)__popc(b));
            if (lane < 6) {   // the 32 records of this warp touch strips (rec0 >> 3) - 1 ... (rec0 >> 3) + 4
                const long sidx = (rec0 >> 3) - 1 + lane;
                const long lo = 8 * sidx - rec0, hi = lo + 8;   // lanes whose records lie in [8s, 8s+8]
                const int a = (int)max(lo, 0L), e = (int)min(hi, 31L);
                if (sidx >= 0 && a <= e) {
                    const unsigned mask = (e - a == 31 ? 0xffffffffu : ((1u << (e - a + 1)) - 1u) << a);
                    if (b & mask) atomicOr(occf + 1 + (sidx >> 5), 1u << (sidx & 31));
                }
            }
        }
    }
    // record of pixel x: [q][r][4 channels] = 2Q float4, written contiguously (QO quads per record go out)
    const int QO = SPLIT ? 4 : Q;
#pragma unroll
    for (int copy = 0; copy < 2; ++copy) {
        float4 *dst = reinterpret_cast<float4 *>(out + ((((size_t)bv * 2 + copy) * HP + m) * W + x0) * (8 * QO));
        for (int e = threadIdx.x; e < npix * 2 * QO; e += 256) {
            const int x = e / (2 * QO), w = e - x * (2 * QO), q = w >> 1, r = (w & 1) + 1 - copy;
            const float *t = tile + (r * Jp + q * 4) * pitch + x;
            dst[e] = make_float4(t[0], t[pitch], t[2 * pitch], t[3 * pitch]);
        }
    }
    if (SPLIT) {
        const int copy = threadIdx.x >> 7, x = threadIdx.x & (RP_PIX - 1);
        if (x < npix) {
            const float *tn = tile + ((1 - copy) * Jp + 16) * pitch + x, *ts = tn + Jp * pitch;   // north / south row of the pair
            ext[(((size_t)bv * 2 + copy) * HP + m) * W + x0 + x] = make_float4(tn[0], ts[0], tn[1], ts[1]);
        }
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
// a5 / a12 (+ a6 / a13): fused projection + bilinear gather + view mean + clamp, with the cube and /
// or its three orthographic max-planes as output
// ------------------------------------------------------------------------------------------------
// One CTA = one brick of BX*BY*BZ = 256 voxels of one cube.
//   phase 1  thread = voxel: project into each view (exact fp32 order), set up the four taps, hand a
//            20-byte record (packed tap position + four weights) over through shared memory;
//   phase 2  Q = Jp/4 adjacent lanes = one voxel: each lane gathers 4 channels of every tap with one
//            128-bit read-only load and runs the reference's fma chain on them two channels at a
//            time (f32x2);
//   epilogue mean over ALL views (exact division by the constant V), clamp, then either the cube
//            store, or block-level max-planes in shared memory flushed with atomicMax on the
//            unsigned view of the non-negative floats (NaN > every finite value, like torch.max).
// The brick is 4x8x8 (x,y,z): 8 consecutive z per warp instruction keep cube stores in whole 32-byte
// sectors and the Q-lane groups of a warp land on neighbouring pixels.
// asynchronous global -> shared copies (LDGSTS): the tap-table rows of a brick go straight into the staging
// buffers phase 2 reads, without passing through registers
__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void cp_async4(void *smem, const void *gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// TAB: phase 1 copies precomputed tap records (fvp_tap_table) instead of projecting.
// Resident CTAs per SM.  Projecting on the fly is latency-bound and wants 5 (the bound below caps it at 48
// registers); with tap tables the bound only asks for 3 (up to 80 registers) so that ptxas does not spill: the
// kernel needs 62, which makes 4 CTAs resident -- 4.05 ms against 4.23 ms when forced to 5 and 4.71 ms at 6
// (Panoptic, 320 cubes).  This generic kernel serves the cube interface, positions projected in the kernel or read
// from a cached grid, and channel counts other than 16; planes-only launches with tap tables and 16 channels go to
// the row-walking kernel (fvp_project_walk.cu).
template <int Q, bool CUBE, bool PLANES, bool TAB>
__global__ void __launch_bounds__(BRICK, TAB ? (Q <= 4 ? 4 : 3) : 5)
project_fused_kernel(const __grid_constant__ ProjectParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4 *s_w = reinterpret_cast<float4 *>(smem_raw);                                          // [vchunk][BRICK]
    unsigned *s_rec = reinterpret_cast<unsigned *>(smem_raw + sizeof(float4) * p.vchunk * BRICK);  // [vchunk][BRICK]

    const int tid = threadIdx.x;
    const int c = blockIdx.y;
    const unsigned brick = blockIdx.x;
    const unsigned byx = p.m_nbz ? __umulhi(brick, p.m_nbz) : brick / (unsigned)p.nbz;
    const int bz = (int)(brick - byx * (unsigned)p.nbz);
    const int bx = (int)(p.m_nby ? __umulhi(byx, p.m_nby) : byx / (unsigned)p.nby);
    const int by = (int)byx - bx * p.nby;

    // window of this cube
    int tl[3] = {0, 0, 0}, st[3] = {0, 0, 0}, en[3] = {p.n_ax, p.n_ay, p.n_az}, skip = 0, frame = c;
    if (p.win != nullptr) {
        const int *w = p.win + (size_t)c * FVP_WIN_STRIDE;
#pragma unroll
        for (int a = 0; a < 3; ++a) { tl[a] = __ldg(w + a); st[a] = __ldg(w + 3 + a); en[a] = __ldg(w + 6 + a); }
        skip = __ldg(w + 9);
        frame = __ldg(w + 10);
    }
    const int V = p.g.V, J = p.g.J;

    // phase-1 voxel of this thread
    const int i = bx * BX + (tid >> 6), j = by * BY + ((tid >> 3) & 7), k = p.z0 + bz * BZ + (tid & 7);
    const int gi = tl[0] + i, gj = tl[1] + j, gk = tl[2] + k;
    const bool in_cube = i < p.nx && j < p.ny && k < p.z1;
    const bool in_win = in_cube && !skip && gi >= st[0] && gi < en[0] && gj >= st[1] && gj < en[1] &&
                        gk >= st[2] && gk < en[2];
    const bool any = __syncthreads_or(in_win);
    if (!CUBE && !any) return;  // planes start from zero: a brick outside the window has nothing to add

    float2 acc[Q][2];
#pragma unroll
    for (int q = 0; q < Q; ++q) acc[q][0] = acc[q][1] = make_float2(0.0f, 0.0f);

    if (any) {
        const int rig = __ldg(p.frame_rig + frame);
        float wx = 0.f, wy = 0.f, wz = 0.f;
        if (in_win) { wx = __ldg(p.ax + gi); wy = __ldg(p.ay + gj); wz = __ldg(p.az + gk); }
        const size_t ntab = (size_t)p.n_ax * p.n_ay * p.n_az;
        const size_t tab = in_win ? ((size_t)gi * p.n_ay + gj) * p.n_az + gk : 0;
        FVP_ASSERT(tab < ntab && (!in_win || (gi >= 0 && gj >= 0 && gk >= 0 && gi < p.n_ax && gj < p.n_ay && gk < p.n_az)));
        const unsigned view_px = 2u * (unsigned)(p.g.HP * p.g.W);  // pixel records per view (both copies)
        const float4 *img0 = p.hm + (size_t)frame * V * view_px * (2 * Q);
        // occupancy of this frame's pixel records (fvp_heatmaps_rowpair_occ); consulted when at least a quarter of
        // them is all zero: samples over clear strips would add exactly +0 and are dropped before the gather
        const unsigned *const occf = p.occ ? p.occ + (size_t)frame * p.occ_stride : nullptr;
        const bool sparse = occf != nullptr && 4ull * __ldg(occf) < 3ull * V * view_px;

        for (int v0 = 0; v0 < V; v0 += p.vchunk) {
            const int nv = min(p.vchunk, V - v0);
            if (v0) __syncthreads();  // previous chunk fully consumed
            unsigned my_live = 0u;    // does any sample of this thread survive in this chunk (1 = unknown: records not in registers)
            if constexpr (TAB) {
                const size_t npad = (ntab + 3) & ~(size_t)3;
                const unsigned char *tv0 = reinterpret_cast<const unsigned char *>(p.grid) + ((size_t)rig * V + v0) * npad * TAP_ENTRY_BYTES;
                if (sparse) {
                    // occupancy shortcut: the packed records come into registers first (all views of the chunk in
                    // flight), then their strip bits; only samples that survive copy their weights
#pragma unroll 1
                    for (int vb = 0; vb < nv; vb += 4) {   // four views at a time: their loads overlap, the registers stay few
                        unsigned r[4], ow[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u)
                            r[u] = (vb + u < nv && in_win)
                                       ? __ldg(reinterpret_cast<const unsigned *>(tv0 + (size_t)(vb + u) * npad * TAP_ENTRY_BYTES + npad * 16) + tab)
                                       : 0u;
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            FVP_ASSERT(1 + ((r[u] & TAP_INDEX) >> 8) < (unsigned)p.occ_stride);
                            ow[u] = __ldg(occf + 1 + ((r[u] & TAP_INDEX) >> 8));
                        }
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int vv = vb + u;
                            if (vv < nv) {
                                const unsigned bit = (ow[u] >> (((r[u] & TAP_INDEX) >> 3) & 31u)) & 1u;
                                const unsigned rec = (bit | (r[u] & TAP_NOSKIP)) ? r[u] : 0u;
                                s_rec[vv * BRICK + tid] = rec;
                                my_live |= rec;
                                if (rec & TAP_LIVE) cp_async16(s_w + vv * BRICK + tid, tv0 + (size_t)vv * npad * TAP_ENTRY_BYTES + tab * 16);
                            }
                        }
                    }
                } else {
                    my_live = TAP_LIVE;
                    for (int vv = 0; vv < nv; ++vv) {
                        if (in_win) {
                            const unsigned char *tv = tv0 + (size_t)vv * npad * TAP_ENTRY_BYTES;
                            asm("" : "+l"(tv));  // opaque: keeps the row pointers out of the loop-carried state
                            cp_async16(s_w + vv * BRICK + tid, tv + tab * 16);
                            cp_async4(s_rec + vv * BRICK + tid, tv + npad * 16 + tab * 4);
                        } else {
                            s_rec[vv * BRICK + tid] = 0u;
                        }
                    }
                }
                cp_async_wait_all();
            } else
            for (int vv = 0; vv < nv; ++vv) {
                float4 wgt = make_float4(0.f, 0.f, 0.f, 0.f);
                unsigned rec = 0u;
                if (in_win) {
                    float2 gxy;
                    if (p.grid != nullptr) {
                        size_t view = (size_t)rig * V + v0 + vv;
                        asm("" : "+l"(view));  // opaque: keeps this pointer out of the loop-carried state
                        gxy = __ldg(p.grid + view * ntab + tab);
                    } else {
                        const Camera cam = load_camera(p.cams + ((size_t)rig * V + v0 + vv) * FVP_CAM_STRIDE);
                        gxy = sample_coords(cam, p.g, wx, wy, wz);
                    }
                    rec = tap_setup(gxy, p.g, wgt);
                    if (rec & TAP_LIVE) rec += (unsigned)(v0 + vv) * view_px;  // index within the frame
                    if (sparse && (rec & TAP_LIVE) && !(rec & TAP_NOSKIP)) {   // nothing but zeros under this sample
                        const unsigned idx = rec & TAP_INDEX;
                        if (!((__ldg(occf + 1 + (idx >> 8)) >> ((idx >> 3) & 31u)) & 1u)) rec = 0u;
                    }
                }
                s_w[vv * BRICK + tid] = wgt;
                s_rec[vv * BRICK + tid] = rec;
                my_live |= rec;
            }
            // (barrier) a brick none of whose samples survives in any view adds nothing to planes that start from zero
            const bool chunk_live = __syncthreads_or((int)(my_live >> 31));
            if (!CUBE && !chunk_live && nv == V) return;
            if (!chunk_live) continue;

            // phase 2.  Views outermost: the Q passes of one view walk adjacent x-slabs of the brick, so the pixel
            // records a pass pulled into L1 are still resident for the next pass (with the passes outermost every
            // pass came back to a view after the four other views of all resident CTAs had gone through L1).
            // Each voxel still sums its views in order.
            const unsigned *pr = s_rec;
            const float4 *pw = s_w;
            for (int vv = 0; vv < nv; ++vv, pr += BRICK, pw += BRICK) {
#pragma unroll
                for (int q = 0; q < Q; ++q) {
                    const int item = tid + q * BRICK;
                    const int lv = item / Q, quad = item - lv * Q;
                    const unsigned rec = pr[lv];
                    if (!(rec & TAP_LIVE)) continue;
                    const float4 wgt = pw[lv];
                    // one 256-bit load = this quad's north and south taps of a column; east = next pixel
                    FVP_ASSERT((rec & TAP_INDEX) + ((rec >> 30) & 1u) < (unsigned)V * view_px);
                    const float4 *t0 = img0 + 2 * quad + (size_t)(rec & TAP_INDEX) * (2 * Q);
                    const Tap2 west = ldg256(t0), east = ldg256(t0 + ((rec & TAP_DX) ? 2 * Q : 0));
                    const float4 a = west.n, cc = west.s, b = east.n, d = east.s;
                    // out = fma(se_val,se, fma(sw_val,sw, fma(ne_val,ne, nw_val*nw))); views summed in order
                    const float2 nw = make_float2(wgt.x, wgt.x), ne = make_float2(wgt.y, wgt.y);
                    const float2 sw = make_float2(wgt.z, wgt.z), se = make_float2(wgt.w, wgt.w);
                    float2 lo = __fmul2_rn(make_float2(a.x, a.y), nw), hi = __fmul2_rn(make_float2(a.z, a.w), nw);
                    lo = __ffma2_rn(make_float2(b.x, b.y), ne, lo);   hi = __ffma2_rn(make_float2(b.z, b.w), ne, hi);
                    lo = __ffma2_rn(make_float2(cc.x, cc.y), sw, lo); hi = __ffma2_rn(make_float2(cc.z, cc.w), sw, hi);
                    lo = __ffma2_rn(make_float2(d.x, d.y), se, lo);   hi = __ffma2_rn(make_float2(d.z, d.w), se, hi);
                    acc[q][0] = __fadd2_rn(acc[q][0], lo);
                    acc[q][1] = __fadd2_rn(acc[q][1], hi);
                }
            }
        }
    }

    // epilogue
    const size_t vox = (size_t)p.nx * p.ny * p.nz;
    constexpr int CH = 4 * Q;                 // channels per cell in the shared planes
    constexpr int NXY = BX * BY, NXZ = BX * BZ, NYZ = BY * BZ;
    // PLANES: the tap records are dead after the last phase 2; their shared memory becomes
    //   s_xy[NXY][CH] | s_yz[NYZ][CH] | s_xz[NXZ][CH] | (Q == 4) s_tmp[BX][BY][BZ][CH]
    unsigned *s_xy = reinterpret_cast<unsigned *>(smem_raw);
    unsigned *s_yz = s_xy + NXY * CH, *s_xz = s_yz + NYZ * CH, *s_tmp = s_xz + NXZ * CH;
    if (PLANES) {
        __syncthreads();
        if (Q != 4) {
            for (int e = tid; e < (NXY + NYZ + NXZ) * CH; e += BRICK) s_xy[e] = 0u;
            __syncthreads();
        }
    }
    unsigned yz_run[4] = {0u, 0u, 0u, 0u};
#pragma unroll
    for (int q = 0; q < Q; ++q) {
        const int item = tid + q * BRICK;
        const int lv = item / Q, quad = item - lv * Q;
        const int li = lv >> 6, lj = (lv >> 3) & 7, lk = lv & 7;
        const int oi = bx * BX + li, oj = by * BY + lj, ok = p.z0 + bz * BZ + lk;
        const bool inside = oi < p.nx && oj < p.ny && ok < p.z1;
        // mean over ALL views (F4), clamp(0,1) keeping NaN, and canonicalise -0 to +0
        float val[4] = {acc[q][0].x, acc[q][0].y, acc[q][1].x, acc[q][1].y};
        div_const4(val, p.g.fV);
#pragma unroll
        for (int e = 0; e < 4; ++e) val[e] = __fadd_rn(clamp_keep_nan(val[e], 0.0f, 1.0f), 0.0f);
        if (CUBE && inside) {
            // every voxel of the brick that lies inside the cube is written (0 outside the window)
            FVP_ASSERT(oi >= 0 && oj >= 0 && ok >= 0 && ok < p.nz);
            float *dst = p.cubes + ((size_t)c * J + quad * 4) * vox + ((size_t)oi * p.ny + oj) * p.nz + ok;
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if (quad * 4 + e < J) __stcs(dst + (size_t)e * vox, val[e]);   // written once, read by another kernel: keep L2 for the heatmaps
        }
        if (PLANES) {
            uint4 u;
            u.x = inside ? __float_as_uint(val[0]) : 0u;
            u.y = inside ? __float_as_uint(val[1]) : 0u;
            u.z = inside ? __float_as_uint(val[2]) : 0u;
            u.w = inside ? __float_as_uint(val[3]) : 0u;
            if (Q == 4) {
                // pass q holds x = q; a warp holds the 8 z of one (x,y) column x 4 channel quads
                // yz (max over x): the same thread sees every x of its (y,z,quad) -> registers
                yz_run[0] = max(yz_run[0], u.x); yz_run[1] = max(yz_run[1], u.y);
                yz_run[2] = max(yz_run[2], u.z); yz_run[3] = max(yz_run[3], u.w);
                // xy (max over z) and xz (max over y): park the values, reduce after the barrier
                *reinterpret_cast<uint4 *>(s_tmp + ((li * BY + lj) * BZ + lk) * CH + quad * 4) = u;
            } else {
                const unsigned uu[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    if (uu[e]) {
                        atomicMax(s_xy + (li * BY + lj) * CH + quad * 4 + e, uu[e]);
                        atomicMax(s_xz + (li * BZ + lk) * CH + quad * 4 + e, uu[e]);
                        atomicMax(s_yz + (lj * BZ + lk) * CH + quad * 4 + e, uu[e]);
                    }
            }
        }
    }
    if (PLANES) {
        if (Q == 4) {
            const int lv = tid >> 2, quad = tid & 3;  // the (y,z,quad) this thread owns in every pass
            *reinterpret_cast<uint4 *>(s_yz + (((lv >> 3) & 7) * BZ + (lv & 7)) * CH + quad * 4) =
                make_uint4(yz_run[0], yz_run[1], yz_run[2], yz_run[3]);
        }
        __syncthreads();
        if constexpr (Q == 4) {
            // one pass over the parked brick: threads 0..127 reduce xy[x][y] over z, threads 128..255 xz[x][z] over y,
            // four channels (one 16-byte shared load) at a time
            static_assert(NXY * Q == BRICK / 2 && NXZ * Q == BRICK / 2, "reduction mapping assumes 4x8x8 bricks, Q = 4");
            const int half = tid >> 7, cell = (tid & 127) >> 2, quad = tid & 3;
            const int li = cell >> 3, l2 = cell & 7;             // (x, y) for xy, (x, z) for xz
            const unsigned *src = s_tmp + (half ? (li * BY * BZ + l2) : (li * BY + l2) * BZ) * CH + quad * 4;
            const int step = half ? BZ * CH : CH;                // walk y (xz) or z (xy)
            uint4 m = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const uint4 v = *reinterpret_cast<const uint4 *>(src + k * step);
                m.x = max(m.x, v.x); m.y = max(m.y, v.y); m.z = max(m.z, v.z); m.w = max(m.w, v.w);
            }
            *reinterpret_cast<uint4 *>((half ? s_xz : s_xy) + cell * CH + quad * 4) = m;
            __syncthreads();
        }
        // flush the block planes with atomicMax; channel-fastest order = coalesced for the NHWC layout
        const int a_xy = bx * BX, b_xy = by * BY, zs = bz * BZ;  // zs: slab-local z of the brick
        const int jc = p.nhwc ? CH : J;                          // channels stored per cell
        unsigned *const base_xy = p.pxy ? p.pxy + (size_t)c * jc * p.nx * p.ny : nullptr;
        unsigned *const base_yz = p.pyz ? p.pyz + (size_t)c * jc * p.ny * p.nzs : nullptr;
        unsigned *const base_xz = p.pxz ? p.pxz + (size_t)c * jc * p.nx * p.nzs : nullptr;
        // one plane at a time: cell shapes and trip counts are compile-time constants
        auto flush = [&](const unsigned *sp, unsigned *plane, const int ncell, const int div, const int a0, const int b0,
                         const int na, const int nb) {
            if (plane == nullptr) return;
#pragma unroll
            for (int e = tid; e < ncell * CH; e += BRICK) {
                const unsigned m = sp[e];
                if (!m) continue;                                // planes start from zero; pad channels stay zero
                const int ch = e % CH, cell = e / CH;
                const int a = a0 + cell / div, b = b0 + cell % div;
                if (a >= na || b >= nb || ch >= jc) continue;
                const unsigned idx = p.nhwc ? (unsigned)((a * nb + b) * CH + ch) : (unsigned)((ch * na + a) * nb + b);
                atomicMax(plane + idx, m);
            }
        };
        flush(s_xy, base_xy, NXY, BY, a_xy, b_xy, p.nx, p.ny);
        flush(s_yz, base_yz, NYZ, BZ, b_xy, zs, p.ny, p.nzs);
        flush(s_xz, base_xz, NXZ, BZ, a_xy, zs, p.nx, p.nzs);
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

// The same for 64 x 64 (y, z) slices -- the per-person cubes -- without the serial shared-memory scans: one CTA per (cube,
// joint) walks x; a warp owns 8 rows of the slice, a lane 4 rows x 4 z (float4 loads, two whole rows per warp instruction).
// yz (max over x) stays in 16 registers for the whole walk; xy (max over z) is 3 local maxima + 4 shuffles per row; xz (max
// over y) 3 local maxima + 1 shuffle per z, then 8 warps meet through 2 KB of shared memory.  The next slice is in flight
// while this one is reduced.
__global__ void __launch_bounds__(256)
planes3_64_kernel(const float *__restrict__ cubes, int nx, float *__restrict__ xy, float *__restrict__ xz, float *__restrict__ yz) {
    constexpr int N = 64;
    __shared__ float s_xz[2][8][N];
    const int cj = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int r2 = lane >> 4, zg = lane & 15;
    const float4 *src = reinterpret_cast<const float4 *>(cubes + (size_t)cj * nx * N * N) + (size_t)(warp * 8 + r2) * (N / 4) + zg;
    float4 cur[4], nxt[4], run[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) cur[i] = __ldcs(src + (size_t)(2 * i) * (N / 4));
    for (int x = 0; x < nx; ++x) {
        if (x + 1 < nx) {
            const float4 *sn = src + (size_t)(x + 1) * (N * N / 4);
#pragma unroll
            for (int i = 0; i < 4; ++i) nxt[i] = __ldcs(sn + (size_t)(2 * i) * (N / 4));
        }
        float4 colmax = cur[0];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float4 v = cur[i];
            if (x == 0) run[i] = v;
            else {
                run[i].x = max_keep_nan(run[i].x, v.x); run[i].y = max_keep_nan(run[i].y, v.y);
                run[i].z = max_keep_nan(run[i].z, v.z); run[i].w = max_keep_nan(run[i].w, v.w);
            }
            if (i) {
                colmax.x = max_keep_nan(colmax.x, v.x); colmax.y = max_keep_nan(colmax.y, v.y);
                colmax.z = max_keep_nan(colmax.z, v.z); colmax.w = max_keep_nan(colmax.w, v.w);
            }
            if (xy != nullptr) {     // row warp * 8 + r2 + 2 i: over this lane's 4 z, then over the 16 lanes of the row
                float m = max_keep_nan(max_keep_nan(v.x, v.y), max_keep_nan(v.z, v.w));
#pragma unroll
                for (int o = 8; o >= 1; o >>= 1) m = max_keep_nan(m, __shfl_xor_sync(0xffffffffu, m, o));
                if (zg == 0) xy[((size_t)cj * nx + x) * N + warp * 8 + r2 + 2 * i] = m;
            }
        }
        if (xz != nullptr) {         // over this lane's 4 rows (done), the other half-warp's 4 rows, then the 8 warps
            colmax.x = max_keep_nan(colmax.x, __shfl_xor_sync(0xffffffffu, colmax.x, 16));
            colmax.y = max_keep_nan(colmax.y, __shfl_xor_sync(0xffffffffu, colmax.y, 16));
            colmax.z = max_keep_nan(colmax.z, __shfl_xor_sync(0xffffffffu, colmax.z, 16));
            colmax.w = max_keep_nan(colmax.w, __shfl_xor_sync(0xffffffffu, colmax.w, 16));
            if (r2 == 0) *reinterpret_cast<float4 *>(&s_xz[x & 1][warp][4 * zg]) = colmax;
            __syncthreads();         // one barrier per slice: the buffer written now is read before slice x + 2 writes it again
            if (tid < N) {
                float m = s_xz[x & 1][0][tid];
#pragma unroll
                for (int k = 1; k < 8; ++k) m = max_keep_nan(m, s_xz[x & 1][k][tid]);
                xz[((size_t)cj * nx + x) * N + tid] = m;
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) cur[i] = nxt[i];
    }
    if (yz != nullptr) {
        float4 *dst = reinterpret_cast<float4 *>(yz + (size_t)cj * N * N) + (size_t)(warp * 8 + r2) * (N / 4) + zg;
#pragma unroll
        for (int i = 0; i < 4; ++i) dst[(size_t)(2 * i) * (N / 4)] = run[i];
    }
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

// r = RN(1/c) when the three-operation exact division is known to hold for c (positive, normal,
// moderate exponent, mantissa not all ones -- Markstein's condition), else 0 = generic division.
static DivConst make_div_const(float c) {
    DivConst d;
    d.c = c;
    d.r = 0.0f;
    uint32_t bits;
    memcpy(&bits, &c, 4);
    const int expo = (int)((bits >> 23) & 0xff) - 127;
    if (c > 0.0f && expo >= -20 && expo <= 20 && (bits & 0x7fffffu) != 0x7fffffu) d.r = (float)(1.0L / (long double)c);
    return d;
}

static GeomDev make_geom_dev(const fvp_geom *g) {
    GeomDev d;
    d.V = g->V; d.J = g->J; d.H = g->H; d.W = g->W;
    d.clamp_max = g->clamp_max;
    for (int i = 0; i < 6; ++i) d.t[i] = g->t[i];
    d.hm_w = g->hm_w; d.hm_h = g->hm_h;
    d.img_w = make_div_const(g->img_w); d.img_h = make_div_const(g->img_h);
    d.wm1 = make_div_const(g->wm1); d.hm1 = make_div_const(g->hm1);
    d.fV = make_div_const((float)g->V);
    d.half_wm1 = g->wm1 * 0.5f; d.half_hm1 = g->hm1 * 0.5f;
    d.HP = g->H / 2 + 1;
    return d;
}

extern "C" int fvp_sample_grid(const float *cams, const fvp_geom *g, const float *ax, const float *ay,
                               const float *az, int X, int Y, int Z, float *out, void *stream) {
    if (int rc = check_geom(g)) return rc;
    FVP_REQUIRE(cams && ax && ay && az && out, "fvp_sample_grid: null pointer");
    FVP_REQUIRE((reinterpret_cast<uintptr_t>(cams) & 31) == 0, "fvp_sample_grid: camera table must be 32-byte aligned");
    FVP_REQUIRE(X > 0 && Y > 0 && Z > 0, "fvp_sample_grid: bad dims");
    const long nb = (long)X * Y * Z;
    dim3 grid((unsigned)std::min<long>((nb + 255) / 256, 148L * 16), (unsigned)g->V);
    sample_grid_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(cams, make_geom_dev(g), ax, ay, az, X, Y, Z,
                                                                 reinterpret_cast<float2 *>(out));
    return check_launch("sample_grid_kernel");
}

extern "C" int fvp_tap_table(const float *cams, const fvp_geom *g, const float *ax, const float *ay, const float *az,
                             int X, int Y, int Z, void *out, void *stream) {
    if (int rc = check_geom(g)) return rc;
    FVP_REQUIRE(cams && ax && ay && az && out, "fvp_tap_table: null pointer");
    FVP_REQUIRE((reinterpret_cast<uintptr_t>(cams) & 31) == 0, "fvp_tap_table: camera table must be 32-byte aligned");
    FVP_REQUIRE((reinterpret_cast<uintptr_t>(out) & 15) == 0, "fvp_tap_table: out must be 16-byte aligned");
    FVP_REQUIRE(X > 0 && Y > 0 && Z > 0, "fvp_tap_table: bad dims");
    FVP_REQUIRE(2L * g->V * (g->H / 2 + 1) * g->W < (1L << 29), "fvp_tap_table: V * H * W too large for the packed tap record");
    const long nb = (long)X * Y * Z, npad = (nb + 3) & ~3L;
    dim3 grid((unsigned)std::min<long>((nb + 255) / 256, 148L * 16), (unsigned)g->V);
    tap_table_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(cams, make_geom_dev(g), ax, ay, az, X, Y, Z, npad,
                                                               reinterpret_cast<unsigned char *>(out));
    return check_launch("tap_table_kernel");
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

static int rowpair_impl(const char *who, const float *planar, int B, int V, int J, int H, int W, int Jp, float *out,
                        uint32_t *occ, void *stream, bool split = false) {
    FVP_REQUIRE(planar && out, "%s: null pointer", who);
    const long BV = (long)B * V;
    FVP_REQUIRE(B > 0 && V > 0 && J > 0 && H > 1 && W > 1, "%s: bad dims", who);
    FVP_REQUIRE(Jp >= J && Jp % 4 == 0 && Jp <= 32, "%s: Jp must be a multiple of 4 in [J,32]", who);
    const int HP = H / 2 + 1;
    FVP_REQUIRE(BV <= 65535 && HP <= 65535, "%s: B*V or H too large for the launch grid", who);
    dim3 grid((unsigned)ceil_div(W, RP_PIX), (unsigned)HP, (unsigned)BV);
    const int Q = Jp / 4;
    const size_t smem = sizeof(float) * 3 * (size_t)Jp * (RP_PIX + 1);
    const bool vec = W % 4 == 0 && (reinterpret_cast<uintptr_t>(planar) & 15) == 0;
    cudaStream_t s = (cudaStream_t)stream;
    int occ_stride = 0;
    if (occ != nullptr) {
        occ_stride = (int)fvp_occupancy_words(V, H, W);
        FVP_CUDA(cudaMemsetAsync(occ, 0, sizeof(uint32_t) * (size_t)B * occ_stride, s));
    }
    auto launch = [&](auto kern) {
        if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        kern<<<grid, 256, smem, s>>>(planar, J, H, W, Q, HP, out, occ, occ_stride, V,
                                     split ? reinterpret_cast<float4 *>(out) + (size_t)BV * 2 * HP * W * 8 : nullptr);
    };
    if (split) {
        FVP_REQUIRE(J == 17 && Jp == 20, "%s: the split layout is defined for 17 joints", who);
        FVP_REQUIRE((reinterpret_cast<uintptr_t>(out) & 15) == 0, "%s: out must be 16-byte aligned", who);
        if (vec) launch(rowpair_kernel<5, true, true>);
        else launch(rowpair_kernel<5, false, true>);
    } else if (vec && Q == 4) launch(rowpair_kernel<4, true>);
    else if (vec && Q == 5) launch(rowpair_kernel<5, true>);
    else if (vec) launch(rowpair_kernel<0, true>);
    else launch(rowpair_kernel<0, false>);
    return check_launch("rowpair_kernel");
}

extern "C" long fvp_occupancy_words(int V, int H, int W) {
    const long records = 2L * V * (H / 2 + 1) * W, strips = (records + 7) / 8;
    return 1 + (strips + 31) / 32;
}

extern "C" int fvp_heatmaps_rowpair(const float *planar, int BV, int J, int H, int W, int Jp, float *out, void *stream) {
    return rowpair_impl("fvp_heatmaps_rowpair", planar, BV, 1, J, H, W, Jp, out, nullptr, stream);
}

extern "C" int fvp_heatmaps_rowpair_occ(const float *planar, int B, int V, int J, int H, int W, int Jp, float *out,
                                        uint32_t *occ, void *stream) {
    FVP_REQUIRE(occ != nullptr, "fvp_heatmaps_rowpair_occ: null occ");
    return rowpair_impl("fvp_heatmaps_rowpair_occ", planar, B, V, J, H, W, Jp, out, occ, stream);
}

extern "C" int fvp_heatmaps_rowpair_split17(const float *planar, int B, int V, int H, int W, float *out, uint32_t *occ,
                                            void *stream) {
    return rowpair_impl("fvp_heatmaps_rowpair_split17", planar, B, V, 17, H, W, 20, out, occ, stream, true);
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

template <int Q, bool CUBE, bool PLANES, bool TAB>
static int launch_fused(const ProjectParams &p, dim3 grid, size_t smem, cudaStream_t s) {
    auto kern = project_fused_kernel<Q, CUBE, PLANES, TAB>;
    if (smem > 48 * 1024) FVP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, BRICK, smem, s>>>(p);
    return check_launch("project_fused_kernel");
}

template <int Q, bool TAB>
static int launch_fused_q(const ProjectParams &p, dim3 grid, size_t smem, cudaStream_t s) {
    const bool cube = p.cubes != nullptr, planes = p.pxy || p.pxz || p.pyz;
    if (cube && planes) return launch_fused<Q, true, true, TAB>(p, grid, smem, s);
    if (cube) return launch_fused<Q, true, false, TAB>(p, grid, smem, s);
    return launch_fused<Q, false, true, TAB>(p, grid, smem, s);
}

// Shared implementation of fvp_project_cubes / fvp_project_planes.
static int project_impl(const char *who, const float *hm_cl, int B, int Jp, const float *cams, const int32_t *frame_rig,
                        const fvp_geom *g, const float *ax, const float *ay, const float *az, int n_ax, int n_ay,
                        int n_az, const float *cached_grid, int C, const int32_t *win, int nx, int ny, int nz, int z0,
                        int z1, float *cubes, float *xy, float *xz, float *yz, unsigned flags, const uint32_t *occ,
                        cudaStream_t s) {
    if (int rc = check_geom(g)) return rc;
    FVP_REQUIRE(hm_cl && cams && frame_rig && ax && ay && az, "%s: null pointer", who);
    FVP_REQUIRE((reinterpret_cast<uintptr_t>(cams) & 31) == 0 && (reinterpret_cast<uintptr_t>(hm_cl) & 31) == 0,
                "%s: camera table and row-paired heatmaps must be 32-byte aligned (256-bit loads)", who);
    FVP_REQUIRE(cubes || xy || xz || yz, "%s: no output requested", who);
    FVP_REQUIRE(B > 0 && n_ax > 0 && n_ay > 0 && n_az > 0 && nx > 0 && ny > 0 && nz > 0, "%s: bad dims", who);
    FVP_REQUIRE(Jp >= g->J && Jp % 4 == 0 && Jp <= 32, "%s: Jp must be a multiple of 4 in [J,32]", who);
    FVP_REQUIRE(0 <= z0 && z0 < z1 && z1 <= nz, "%s: z slab [%d,%d) outside [0,%d)", who, z0, z1, nz);
    FVP_REQUIRE(2L * g->V * (g->H / 2 + 1) * g->W < (1L << 29), "%s: V * H * W too large for the packed tap record", who);
    if (C == 0) return FVP_OK;
    FVP_REQUIRE(C > 0 && C <= 65535, "%s: C out of range (1..65535)", who);
    FVP_REQUIRE(win != nullptr || (C <= B && nx <= n_ax && ny <= n_ay && nz <= n_az),
                "%s: without windows cube c samples frame c over the whole table", who);
    const bool taps = (flags & FVP_FLAG_TAP_TABLE) != 0;
    const bool cached = (flags & FVP_FLAG_CACHED_GRID) != 0 || taps;
    FVP_REQUIRE(!cached || cached_grid != nullptr, "%s: FVP_FLAG_CACHED_GRID / FVP_FLAG_TAP_TABLE need cached_grid", who);
    FVP_REQUIRE(!taps || !(flags & FVP_FLAG_CACHED_GRID), "%s: FVP_FLAG_CACHED_GRID and FVP_FLAG_TAP_TABLE exclude each other", who);
    FVP_REQUIRE(!taps || (reinterpret_cast<uintptr_t>(cached_grid) & 15) == 0, "%s: tap table must be 16-byte aligned", who);
    const unsigned mode = flags & 0xfu;
    FVP_REQUIRE(mode == FVP_MODE_DIRECT, "%s: mode %u not available in this build", who, mode);
    const bool split = (flags & FVP_FLAG_SPLIT17) != 0;
    FVP_REQUIRE(!split || (g->J == 17 && Jp == 20), "%s: FVP_FLAG_SPLIT17 is the layout of 17-joint heatmaps (Jp 20)", who);
    FVP_REQUIRE(!split || (taps && cubes == nullptr && option(FVP_OPT_WALK_VARIANT) >= 0),
                "%s: the split 16 + 1 layout is read by the row-walking planes kernel only (tap table, planes, no cubes)", who);

    const int nzs = z1 - z0;
    const bool nhwc = (flags & FVP_FLAG_PLANES_NHWC) != 0;
    const int Jc = nhwc ? Jp : g->J;
    // planes are combined with atomicMax, so they start from zero (all values are >= 0)
    if (xy) FVP_CUDA(cudaMemsetAsync(xy, 0, sizeof(float) * (size_t)C * Jc * nx * ny, s));
    if (xz) FVP_CUDA(cudaMemsetAsync(xz, 0, sizeof(float) * (size_t)C * Jc * nx * nzs, s));
    if (yz) FVP_CUDA(cudaMemsetAsync(yz, 0, sizeof(float) * (size_t)C * Jc * ny * nzs, s));

    ProjectParams p;
    p.hm = reinterpret_cast<const float4 *>(hm_cl); p.cams = cams; p.frame_rig = frame_rig;
    p.hm_ext = split ? p.hm + (size_t)B * g->V * 2 * (g->H / 2 + 1) * g->W * 8 : nullptr;
    p.ax = ax; p.ay = ay; p.az = az;
    p.grid = cached ? reinterpret_cast<const float2 *>(cached_grid) : nullptr;
    p.win = win; p.cubes = cubes;
    p.xseg = 0;
    p.occ = occ; p.occ_stride = occ ? (int)fvp_occupancy_words(g->V, g->H, g->W) : 0;
    p.pxy = reinterpret_cast<unsigned *>(xy); p.pxz = reinterpret_cast<unsigned *>(xz);
    p.pyz = reinterpret_cast<unsigned *>(yz);
    p.g = make_geom_dev(g);
    p.n_ax = n_ax; p.n_ay = n_ay; p.n_az = n_az; p.nx = nx; p.ny = ny; p.nz = nz;
    p.nby = ceil_div(ny, BY); p.nbz = ceil_div(nzs, BZ);
    p.z0 = z0; p.z1 = z1; p.nzs = nzs;
    p.vchunk = std::min(g->V, VCHUNK);
    const bool planes = xy || xz || yz;
    p.nhwc = nhwc ? 1 : 0;
    // the block planes alias the tap records: xy | yz | xz (+ the parked brick for the Jp == 16 path)
    const size_t smem_planes = sizeof(unsigned) * Jp * ((BX * BY + BX * BZ + BY * BZ) + (Jp == 16 ? BRICK : 0));
    const size_t smem = std::max((size_t)p.vchunk * BRICK * (sizeof(float4) + sizeof(unsigned)), planes ? smem_planes : 0);
    dim3 grid((unsigned)(ceil_div(nx, BX) * p.nby * p.nbz), (unsigned)C);
    // q = umulhi(x, ceil(2^32 / d)) == x / d while x * d < 2^32
    auto magic = [&](int d) -> unsigned {
        return (d > 1 && (unsigned long long)grid.x * (unsigned)d < (1ull << 32)) ? (unsigned)(((1ull << 32) + d - 1) / d) : 0u;
    };
    p.m_nby = magic(p.nby);
    p.m_nbz = magic(p.nbz);
    // planes only, tap-table positions, 16 padded channels: the row-walking kernel (fvp_project_walk.cu)
    const int walk = option(FVP_OPT_WALK_VARIANT);
    if (walk >= 0 && walk_supported(p, Jp, taps)) return launch_project_walk(p, C, walk, s);
    if (walk >= 0 && walk_cube_supported(p, Jp, taps)) return launch_project_walk_cube(p, C, s);
    if (split) return fail(FVP_E_ARG, "%s: the split 16 + 1 layout cannot be served by the generic kernel", who);
#define FVP_LAUNCH_Q(q) \
    case q: return taps ? launch_fused_q<q, true>(p, grid, smem, s) : launch_fused_q<q, false>(p, grid, smem, s);
    switch (Jp / 4) {
        FVP_LAUNCH_Q(1) FVP_LAUNCH_Q(2) FVP_LAUNCH_Q(3) FVP_LAUNCH_Q(4)
        FVP_LAUNCH_Q(5) FVP_LAUNCH_Q(6) FVP_LAUNCH_Q(7) FVP_LAUNCH_Q(8)
    }
#undef FVP_LAUNCH_Q
    return fail(FVP_E_ARG, "%s: unsupported Jp %d", who, Jp);
}

extern "C" int fvp_project_cubes_occ(const float *hm_cl, int B, int Jp, const float *cams, const int32_t *frame_rig,
                                     const fvp_geom *g, const float *ax, const float *ay, const float *az, int n_ax,
                                     int n_ay, int n_az, const float *cached_grid, int C, const int32_t *win, int nx,
                                     int ny, int nz, float *cubes, unsigned flags, const uint32_t *occ, void *stream) {
    FVP_REQUIRE(cubes != nullptr, "fvp_project_cubes: null cubes");
    return project_impl("fvp_project_cubes", hm_cl, B, Jp, cams, frame_rig, g, ax, ay, az, n_ax, n_ay, n_az, cached_grid,
                        C, win, nx, ny, nz, 0, nz, cubes, nullptr, nullptr, nullptr, flags, occ, (cudaStream_t)stream);
}

extern "C" int fvp_project_cubes(const float *hm_cl, int B, int Jp, const float *cams, const int32_t *frame_rig,
                                 const fvp_geom *g, const float *ax, const float *ay, const float *az, int n_ax,
                                 int n_ay, int n_az, const float *cached_grid, int C, const int32_t *win, int nx,
                                 int ny, int nz, float *cubes, unsigned flags, void *stream) {
    return fvp_project_cubes_occ(hm_cl, B, Jp, cams, frame_rig, g, ax, ay, az, n_ax, n_ay, n_az, cached_grid, C, win, nx,
                                 ny, nz, cubes, flags, nullptr, stream);
}

extern "C" int fvp_project_planes_occ(const float *hm_cl, int B, int Jp, const float *cams, const int32_t *frame_rig,
                                      const fvp_geom *g, const float *ax, const float *ay, const float *az, int n_ax,
                                      int n_ay, int n_az, const float *cached_grid, int C, const int32_t *win, int nx,
                                      int ny, int nz, int z0, int z1, float *cubes, float *xy, float *xz, float *yz,
                                      unsigned flags, const uint32_t *occ, void *stream) {
    FVP_REQUIRE(xy || xz || yz, "fvp_project_planes: no plane requested");
    return project_impl("fvp_project_planes", hm_cl, B, Jp, cams, frame_rig, g, ax, ay, az, n_ax, n_ay, n_az, cached_grid,
                        C, win, nx, ny, nz, z0, z1, cubes, xy, xz, yz, flags, occ, (cudaStream_t)stream);
}

extern "C" int fvp_project_planes(const float *hm_cl, int B, int Jp, const float *cams, const int32_t *frame_rig,
                                  const fvp_geom *g, const float *ax, const float *ay, const float *az, int n_ax,
                                  int n_ay, int n_az, const float *cached_grid, int C, const int32_t *win, int nx,
                                  int ny, int nz, int z0, int z1, float *cubes, float *xy, float *xz, float *yz,
                                  unsigned flags, void *stream) {
    return fvp_project_planes_occ(hm_cl, B, Jp, cams, frame_rig, g, ax, ay, az, n_ax, n_ay, n_az, cached_grid, C, win, nx,
                                  ny, nz, z0, z1, cubes, xy, xz, yz, flags, nullptr, stream);
}

// a8 without the cube: one 1x1xZ window per top-k cell (human_detection_net.py:199-200), so that
// fvp_project_cubes re-projects just those columns into cols[B*P][J][1][1][Z] = [B][P][J][Z].
__global__ void topk_windows_kernel(const long long *__restrict__ flat, int BP, int P, int Y, int Z,
                                    int *__restrict__ win) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= BP) return;
    const int cell = (int)flat[n];
    const int x = cell / Y, y = cell - x * Y;  // flat index of the flattened [X][Y] map
    int *w = win + (size_t)n * FVP_WIN_STRIDE;
    w[0] = x; w[1] = y; w[2] = 0;
    w[3] = x; w[4] = y; w[5] = 0;
    w[6] = x + 1; w[7] = y + 1; w[8] = Z;
    w[9] = 0; w[10] = n / P; w[11] = 0;
}

extern "C" int fvp_topk_windows(const int64_t *flat, int B, int P, int X, int Y, int Z, int32_t *win, void *stream) {
    FVP_REQUIRE(flat && win, "fvp_topk_windows: null pointer");
    FVP_REQUIRE(B > 0 && P > 0 && X > 0 && Y > 0 && Z > 0, "fvp_topk_windows: bad dims");
    topk_windows_kernel<<<ceil_div(B * P, 128), 128, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const long long *>(flat), B * P, P, Y, Z, win);
    return check_launch("topk_windows_kernel");
}

extern "C" int fvp_planes3(const float *cubes, int CJ, int nx, int ny, int nz, float *xy, float *xz, float *yz,
                           void *stream) {
    FVP_REQUIRE(cubes != nullptr, "fvp_planes3: null cubes");
    FVP_REQUIRE(nx > 0 && ny > 0 && nz > 0, "fvp_planes3: bad dims");
    if (CJ == 0 || (!xy && !xz && !yz)) return FVP_OK;
    FVP_REQUIRE(CJ > 0, "fvp_planes3: CJ < 0");
    if (ny == 64 && nz == 64 && (reinterpret_cast<uintptr_t>(cubes) & 15) == 0 && (!yz || (reinterpret_cast<uintptr_t>(yz) & 15) == 0)) {
        planes3_64_kernel<<<CJ, 256, 0, (cudaStream_t)stream>>>(cubes, nx, xy, xz, yz);
        return check_launch("planes3_64_kernel");
    }
    const size_t smem = sizeof(float) * ((size_t)ny * (nz + 1) + (size_t)ny * nz);
    FVP_REQUIRE(smem <= 200 * 1024, "fvp_planes3: ny*nz too large for shared memory (%zu B)", smem);
    if (smem > 48 * 1024)
        FVP_CUDA(cudaFuncSetAttribute(planes3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    planes3_kernel<<<CJ, 256, smem, (cudaStream_t)stream>>>(cubes, nx, ny, nz, xy, xz, yz);
    return check_launch("planes3_kernel");
}
