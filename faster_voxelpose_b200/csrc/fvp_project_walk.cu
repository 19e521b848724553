// a12 + a13 (and a5 + a6): fused projection -> orthographic max-planes, "row-walking" form.
//   reference: lib/models/project_individual.py:222-293 + lib/models/joint_localization_net.py:158-160
//              lib/models/project_whole.py:119-168 + lib/models/cnns_2d.py:291
//
// Same arithmetic as project_fused_kernel (fvp_project.cu) -- the tap records of fvp_tap_table, the reference's
// fma chain per tap quartet, views summed in order, exact division by V, clamp -- reorganised around what bounds
// it on B200: the L1 data path serves a multi-line gather at ~2 cycles per 128-byte line, every sample (voxel-view)
// needs two lines, so the work that can be taken away is (a) samples whose lines hold nothing but zeros and
// (b) everything that is not the gather:
//   * one CTA owns a (y, z) brick row of a cube and WALKS it along x, one brick of 4 x NW x 8 voxels per (brick,
//     view) step: the window / rig / pointer set-up is paid once per row, bricks outside the window's x range are
//     never visited, and the yz plane (max over x) stays in registers for the whole row;
//   * Q = 4 adjacent lanes are one voxel, a warp is 8 voxels consecutive in z; the tap records come straight from
//     the table (no CTA-wide staging, no barrier per view) and are spread over the lanes, see the kernel comment;
//   * per brick the values are parked once in shared memory; half of the CTA reduces xy (over z), the other half
//     xz (over y), and the owners combine their maxima into the planes with RED.MAX straight from registers.
// What was measured on the way (320 person cubes of 64^3, Panoptic, B200; tools/bench_walk.py):
// every lane group loading its own 4 records + weights per view: 4.18-4.48 ms (generic brick kernel: 4.03 ms);
// the same with constant weights: 3.40 ms = 2.09 cycles per line touched, the L1 rate for multi-line requests;
// L1 prefetch of the next step's lines (prefetch.global.L1): no gain; two samples in flight per thread: no gain.
#include <algorithm>

#include "fvp_project.cuh"

namespace fvp {

namespace {

struct Item {  // the four taps of one voxel-view for one channel quad
    float4 wn, ws, en, es;
};

// Loads are UNCONDITIONAL (a predicated load is a partial definition: ptxas then keeps the destination's old value
// alive around the whole view loop and spills): dead lanes carry record 0, i.e. they read pixel record 0 of their
// frame and the weights of a valid table entry -- one shared wavefront -- and the result is discarded.
typedef unsigned long long u64;

__device__ __forceinline__ u64 mad_wide(unsigned a, unsigned b, u64 c) {  // one IMAD.WIDE.U32
    u64 d;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(d) : "r"(a), "r"(b), "l"(c));
    return d;
}

__device__ __forceinline__ Tap2 ldg256v(u64 p) {
    Tap2 r;
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r.n.x), "=f"(r.n.y), "=f"(r.n.z), "=f"(r.n.w), "=f"(r.s.x), "=f"(r.s.y), "=f"(r.s.z), "=f"(r.s.w)
                 : "l"(p));
    return r;
}

// imgq: byte address of this lane's quad in pixel record 0 of the frame; records are 32*Q bytes
template <int Q>
__device__ __forceinline__ void gather_issue(Item &it, u64 imgq, unsigned rec) {
    const unsigned idx = rec & TAP_INDEX, idxe = idx + ((rec >> 30) & 1u);  // east record: one further when TAP_DX
    const Tap2 west = ldg256v(mad_wide(idx, 32 * Q, imgq)), east = ldg256v(mad_wide(idxe, 32 * Q, imgq));
    it.wn = west.n; it.ws = west.s; it.en = east.n; it.es = east.s;
}

// streaming table reads: one use per launch, keep them out of L1 (the pixel records live there)
__device__ __forceinline__ unsigned ld_record(u64 p, bool pred) {
    unsigned r = 0u;
    asm volatile(
        "{\n\t.reg .pred pp;\n\tsetp.ne.u32 pp, %2, 0;\n\t"
        "@pp ld.global.nc.L1::no_allocate.u32 %0, [%1];\n\t}"
        : "+r"(r)
        : "l"(p), "r"((unsigned)pred));
    return r;
}

// out = fma(se_v,se, fma(sw_v,sw, fma(ne_v,ne, nw_v*nw))), two channels at a time; views summed in order
__device__ __forceinline__ void consume(float2 &a0, float2 &a1, const Item &it, const float4 wgt, unsigned live) {
    const float2 nw = make_float2(wgt.x, wgt.x), ne = make_float2(wgt.y, wgt.y);
    const float2 sw = make_float2(wgt.z, wgt.z), se = make_float2(wgt.w, wgt.w);
    float2 lo = __fmul2_rn(make_float2(it.wn.x, it.wn.y), nw), hi = __fmul2_rn(make_float2(it.wn.z, it.wn.w), nw);
    lo = __ffma2_rn(make_float2(it.en.x, it.en.y), ne, lo); hi = __ffma2_rn(make_float2(it.en.z, it.en.w), ne, hi);
    lo = __ffma2_rn(make_float2(it.ws.x, it.ws.y), sw, lo); hi = __ffma2_rn(make_float2(it.ws.z, it.ws.w), sw, hi);
    lo = __ffma2_rn(make_float2(it.es.x, it.es.y), se, lo); hi = __ffma2_rn(make_float2(it.es.z, it.es.w), se, hi);
    if (live) {
        a0 = __fadd2_rn(a0, lo);
        a1 = __fadd2_rn(a1, hi);
    }
    // pins the sums between the surrounding (volatile) loads: without it the compiler sinks the fma chains of a
    // whole view below its last gather and spills the in-flight pixel records
    asm volatile("" : "+f"(a0.x), "+f"(a0.y), "+f"(a1.x), "+f"(a1.y));
}

__device__ __forceinline__ void red_max4(unsigned *plane, int a, int b, int na, int nb, int quad, const uint4 m,
                                         int nhwc, int J, int CH) {
    if (a >= na || b >= nb) return;
    FVP_ASSERT(a >= 0 && b >= 0 && quad >= 0 && quad * 4 < CH);
    const unsigned mm[4] = {m.x, m.y, m.z, m.w};
    if (nhwc) {
        unsigned *dst = plane + ((size_t)a * nb + b) * CH + quad * 4;
#pragma unroll
        for (int e = 0; e < 4; ++e)
            if (mm[e]) atomicMax(dst + e, mm[e]);
    } else {
#pragma unroll
        for (int e = 0; e < 4; ++e)
            if (mm[e] && quad * 4 + e < J) atomicMax(plane + ((size_t)(quad * 4 + e) * na + a) * nb + b, mm[e]);
    }
}

// ------------------------------------------------------------------------------------------------------------
// Third form: the per-step bookkeeping is spread over the lanes.  Lane (z, quad) of a warp owns ONE sample of the
// step -- x-slab `quad` of voxel z -- for everything that happens before the gather: it loads that sample's packed
// record three steps ahead, fetches its occupancy word two steps ahead, decides one step ahead whether the sample
// survives (some tap inside the image AND the pixel records under it not all zero) and, if so, copies its weights
// into the warp's shared-memory ring.  A step then starts from a warp ballot of the surviving samples: a step with
// none costs ~40 instructions and no gather; otherwise x-slab q is gathered by the four lanes of each voxel after
// a shuffle hands them the record of lane (z, q).  Bricks in which no warp gathered anything skip the epilogue.
// ------------------------------------------------------------------------------------------------------------
//
// X17 (17-joint rigs, Shelf / Campus): channels 0..15 are served exactly as above from 128-byte pixel records; the
// 17th channel lives in its own array of 16-byte entries {north x, south x, north x+1, south x+1} (a 160-byte record
// of 20 channels would straddle two lines for every tap).  The lane that owns a sample gathers that one entry itself
// with the weights it staged, so channel 16 accumulates per (voxel x = quad, z) lane; its maxima go out through
// shuffles (xy: over z, yz: over x) and a small parked array (xz: over the warps = y).  Planes have 20 channels.
//
// CUBE (the cube interface of ProjectLayer.forward, 16-channel layout): the same walk, but the voxels themselves go to HBM
// instead of their maxima.  A warp is then on its own -- no parked brick, no block barriers (the CTA's warps take four
// consecutive z bricks of ONE y row, whose 32-byte runs share 128-byte lines of the cube, z fastest); every voxel of the
// cube is written, zeros outside the window and where nothing was gathered.  Tried and dropped: all-zero bricks written as
// whole 128-byte lines after a block-wide vote -- the barrier per brick ties four warps with different z, hence different
// work, together: 4.0 ms against 2.97.
template <int NW, int MINB, bool X17, bool CUBE = false>
__global__ void __launch_bounds__(NW * 32, MINB)
project_walk3_kernel(const __grid_constant__ ProjectParams p) {
    static_assert(!(CUBE && X17), "the cube variant reads the 16-channel layout");
    constexpr int Q = 4, CH = 16, CHS = X17 ? 20 : 16;   // channels gathered by the lane quads; channel stride of nhwc planes
    constexpr unsigned FULL = 0xffffffffu;
    extern __shared__ __align__(16) unsigned s_park[];  // [4][NW][8][CH], then float4 s_wgt[NW][2][4][8], then X17: [4][NW][8]
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31, zl = lane >> 2, quad = lane & 3;
    const int c = blockIdx.y;
    // planes: CTA = NW y rows of one z brick (p.nbz bricks); cubes: CTA = NW z bricks of one y row (p.nbz groups of NW bricks)
    const int by = (int)blockIdx.x / p.nbz, bz = CUBE ? ((int)blockIdx.x - by * p.nbz) * NW + w : (int)blockIdx.x - by * p.nbz;

    int tl0 = 0, tl1 = 0, tl2 = 0, st0 = 0, st1 = 0, st2 = 0, en0 = p.n_ax, en1 = p.n_ay, en2 = p.n_az, skip = 0, frame = c;
    if (p.win != nullptr) {
        const int *wn = p.win + (size_t)c * FVP_WIN_STRIDE;
        tl0 = __ldg(wn + 0); tl1 = __ldg(wn + 1); tl2 = __ldg(wn + 2);
        st0 = __ldg(wn + 3); st1 = __ldg(wn + 4); st2 = __ldg(wn + 5);
        en0 = __ldg(wn + 6); en1 = __ldg(wn + 7); en2 = __ldg(wn + 8);
        skip = __ldg(wn + 9);
        frame = __ldg(wn + 10);
    }
    // live x range of the row inside the cube (uniform): voxels outside are 0 and add nothing.  Launches with few
    // rows (space grids, z slabs) split every row into gridDim.z segments of p.xseg bricks so that the GPU fills up;
    // the yz maxima of the segments meet in the plane through RED.MAX like everything else.
    int xlo = max(0, st0 - tl0), xhi = min(p.nx, en0 - tl0);
    if (gridDim.z > 1) {
        xlo += (int)blockIdx.z * p.xseg * 4;
        xhi = min(xhi, xlo + p.xseg * 4);
    }
    const int j = CUBE ? by : by * NW + w, k = p.z0 + bz * 8 + zl;
    // CUBE: this lane's voxels are (x, j, k) for every x, channels 4 quad .. 4 quad + 3; x in [xa, xb) <- 0
    const size_t cvox = (size_t)p.nx * p.ny * p.nz, cxs = (size_t)p.ny * p.nz;
    float *const cube = CUBE ? p.cubes + ((size_t)c * p.g.J + quad * 4) * cvox + (size_t)j * p.nz + k : nullptr;
    const bool cube_mine = CUBE && j < p.ny && k < p.nz;
    auto zero_x = [&](int xa, int xb) {
        if (!cube_mine) return;
        for (int x = xa; x < xb; ++x)
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if (quad * 4 + e < p.g.J) __stcs(cube + (size_t)e * cvox + (size_t)x * cxs, 0.0f);
    };
    if (skip || xlo >= xhi) { zero_x(0, p.nx); return; }
    const int gj = tl1 + j, gk = tl2 + k;
    const bool yz_ok = j < p.ny && k < p.z1 && gj >= st1 && gj < en1 && gk >= st2 && gk < en2;
    if (CUBE) {
        if (!__any_sync(0xffffffffu, yz_ok)) { zero_x(0, p.nx); return; }
    } else if (!__syncthreads_or((int)yz_ok)) return;

    const int V = p.g.V;
    const int rig = __ldg(p.frame_rig + frame);
    const size_t ntab = (size_t)p.n_ax * p.n_ay * p.n_az, npad = (ntab + 3) & ~(size_t)3;
    const u64 vstride = npad * TAP_ENTRY_BYTES;
    const u64 tab0 = reinterpret_cast<u64>(p.grid) + (u64)rig * V * vstride;   // weights of view 0; its records npad*16 further
    const unsigned xs = (unsigned)(p.n_ay * p.n_az);
    // table entry of this lane's sample in brick 0: (xlo + quad, y, z); lanes outside the window never load
    const unsigned ent0 = yz_ok ? (unsigned)(((tl0 + xlo + quad) * p.n_ay + gj) * p.n_az + gk) : 0u;
    const unsigned view_px = 2u * (unsigned)(p.g.HP * p.g.W);
    const u64 imgq = reinterpret_cast<u64>(p.hm + (size_t)frame * V * view_px * (2 * Q)) + 32u * quad;
    const u64 img17 = X17 ? reinterpret_cast<u64>(p.hm_ext + (size_t)frame * V * view_px) : 0ull;
    const bool want_xy = p.pxy != nullptr, want_xz = p.pxz != nullptr, want_yz = p.pyz != nullptr;
    const int nb = (xhi - xlo + 3) >> 2, span = xhi - xlo;

    // occupancy of this frame's pixel records; worth consulting when at least a quarter of them is all zero
    const unsigned *const occf = p.occ ? p.occ + (size_t)frame * p.occ_stride : nullptr;
    const bool sparse = occf != nullptr && 4ull * __ldg(occf) < 3ull * V * view_px;

    float4 *const s_wgt = reinterpret_cast<float4 *>(s_park + 4 * NW * 8 * CH) + w * 64;
    unsigned *const s_x17 = s_park + 4 * NW * 8 * CH + NW * 64 * 4;   // [4][NW][8]
    const unsigned s_wgt_mine = (unsigned)__cvta_generic_to_shared(s_wgt + quad * 8 + zl);

    // a stream of (brick, view) steps: uniform view base + this lane's entry
    struct Stream {
        int b, v;
        u64 base;       // tab0 + v * vstride
        unsigned ent;   // ent0 + 4 * b * xs
    };
    auto advance = [&](Stream &t) {
        t.base += vstride;
        if (++t.v == V) { t.v = 0; t.base = tab0; ++t.b; t.ent += 4u * xs; }
    };
    auto load_record = [&](const Stream &t) -> unsigned {
        FVP_ASSERT(!(yz_ok && 4 * t.b + quad < span) || (t.ent < ntab && t.v < V));
        return ld_record(mad_wide(t.ent, 4, t.base + npad * 16), yz_ok && 4 * t.b + quad < span);
    };
    // dead records carry index 0: they read word 1 of the bitmap, valid memory
    auto occ_fetch = [&](unsigned r) -> unsigned {
        FVP_ASSERT(1 + ((r & TAP_INDEX) >> 8) < (unsigned)p.occ_stride);
        return __ldg(occf + 1 + ((r & TAP_INDEX) >> 8));
    };
    auto occ_apply = [&](unsigned r, unsigned ow) -> unsigned {
        const unsigned bit = (ow >> (((r & TAP_INDEX) >> 3) & 31u)) & 1u;
        return (bit | (r & TAP_NOSKIP)) ? r : 0u;
    };
    auto stage_weights = [&](const Stream &t, unsigned r, unsigned st) {
        FVP_ASSERT(!(r >> 31) || (t.ent < ntab && t.v < V && st < 2u));
        if (r >> 31) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s_wgt_mine + st * 512u), "l"(mad_wide(t.ent, 16, t.base)));
        asm volatile("cp.async.commit_group;");
    };

    Stream sl = {0, 0, tab0, ent0};   // records are loaded three steps ahead
    unsigned r0 = load_record(sl); advance(sl);
    Stream sw = sl;                   // weights are staged one step ahead
    unsigned r1 = load_record(sl); advance(sl);
    unsigned r2 = load_record(sl); advance(sl);
    unsigned r3, ow1 = 0u, ow2;
    {
        Stream s0 = {0, 0, tab0, ent0};
        if (sparse) {
            r0 = occ_apply(r0, occ_fetch(r0));
            ow1 = occ_fetch(r1);
        }
        stage_weights(s0, r0, 0u);
    }
    zero_x(0, xlo);
    unsigned stage = 0, brick_live = 0;
    unsigned yz_run[4] = {0u, 0u, 0u, 0u};
    unsigned yz17 = 0u;
    float acc17 = 0.0f;
    float2 acc[4][2];
#pragma unroll
    for (int q = 0; q < 4; ++q) acc[q][0] = acc[q][1] = make_float2(0.0f, 0.0f);

    for (int b = 0, v = 0; b < nb;) {
        asm volatile("cp.async.wait_group 0;" ::: "memory");   // this step's weights have landed ...
        __syncwarp();                                          // ... for every lane, and the other stage is free
        r3 = load_record(sl);
        advance(sl);
        if (sparse) {
            ow2 = occ_fetch(r2);
            r1 = occ_apply(r1, ow1);
        }
        stage_weights(sw, r1, stage ^ 1u);
        advance(sw);

        const unsigned m0 = __ballot_sync(FULL, r0 >> 31);
        brick_live |= m0;
        if (m0) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (m0 & (0x11111111u << q)) {   // warp-uniform: x-slab q has a surviving sample in this warp
                    const unsigned rq = __shfl_sync(FULL, r0, (lane & ~3) | q);
                    Item it;
                    const float4 wg = s_wgt[stage * 32 + q * 8 + zl];
                    FVP_ASSERT((rq & TAP_INDEX) + ((rq >> 30) & 1u) < (unsigned)V * view_px);
                    gather_issue<Q>(it, imgq, rq);
                    consume(acc[q][0], acc[q][1], it, wg, rq >> 31);
                }
            }
            if (X17) {   // channel 16 of this lane's own sample: one 16-byte entry, the weights this lane staged
                const unsigned idx = r0 & TAP_INDEX;
                FVP_ASSERT(idx < (unsigned)V * view_px);
                float4 e;
                asm volatile("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];"
                             : "=f"(e.x), "=f"(e.y), "=f"(e.z), "=f"(e.w) : "l"(mad_wide(idx, 16, img17)));
                const float4 wg = s_wgt[stage * 32 + quad * 8 + zl];
                const bool dx = (r0 >> 30) & 1u;
                const float en = dx ? e.z : e.x, es = dx ? e.w : e.y;
                float t = __fmul_rn(e.x, wg.x);
                t = __fmaf_rn(en, wg.y, t);
                t = __fmaf_rn(e.y, wg.z, t);
                t = __fmaf_rn(es, wg.w, t);
                if (r0 >> 31) acc17 = __fadd_rn(acc17, t);
            }
        }
        if (v == V - 1) {
            // ---- brick epilogue: mean over ALL views (F4), clamp(0,1) keeping NaN, -0 -> +0.  The barrier also orders
            // the previous brick's reduction reads before this brick's parking stores.  A brick in which nothing was
            // gathered is all zero: it adds nothing to planes that start from zero.
            const int x0 = xlo + 4 * b;
            if (CUBE) {
                // the brick's voxels, mean over ALL views, clamped: 32-byte runs along z per (channel, x).  Voxels beyond the
                // window inside the brick, lanes outside it and bricks in which nothing was gathered accumulated nothing:
                // they store +0.
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float val[4] = {acc[q][0].x, acc[q][0].y, acc[q][1].x, acc[q][1].y};
                    acc[q][0] = acc[q][1] = make_float2(0.0f, 0.0f);
                    if (brick_live) {
                        div_const4(val, p.g.fV);
#pragma unroll
                        for (int e = 0; e < 4; ++e) val[e] = __fadd_rn(clamp_keep_nan(val[e], 0.0f, 1.0f), 0.0f);
                    }
                    if (cube_mine && x0 + q < p.nx) {
                        FVP_ASSERT(x0 + q >= 0 && c >= 0 && k >= 0);
#pragma unroll
                        for (int e = 0; e < 4; ++e)
                            if (quad * 4 + e < p.g.J) __stcs(cube + (size_t)e * cvox + (size_t)(x0 + q) * cxs, val[e]);
                    }
                }
            } else if (__syncthreads_or((int)(brick_live != 0u))) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float val[4] = {acc[q][0].x, acc[q][0].y, acc[q][1].x, acc[q][1].y};
                    acc[q][0] = acc[q][1] = make_float2(0.0f, 0.0f);
                    div_const4(val, p.g.fV);
                    uint4 u;
                    u.x = __float_as_uint(__fadd_rn(clamp_keep_nan(val[0], 0.0f, 1.0f), 0.0f));
                    u.y = __float_as_uint(__fadd_rn(clamp_keep_nan(val[1], 0.0f, 1.0f), 0.0f));
                    u.z = __float_as_uint(__fadd_rn(clamp_keep_nan(val[2], 0.0f, 1.0f), 0.0f));
                    u.w = __float_as_uint(__fadd_rn(clamp_keep_nan(val[3], 0.0f, 1.0f), 0.0f));
                    // dead voxels (outside the window / cube, or nothing but zeros under them) accumulated nothing: u == 0
                    yz_run[0] = max(yz_run[0], u.x); yz_run[1] = max(yz_run[1], u.y);
                    yz_run[2] = max(yz_run[2], u.z); yz_run[3] = max(yz_run[3], u.w);
                    if (want_xy || want_xz)
                        *reinterpret_cast<uint4 *>(s_park + ((q * NW + w) * 8 + zl) * CH + quad * 4) = u;
                }
                if (X17) {   // channel 16: this lane is voxel (x0 + quad, j, k)
                    const unsigned u17 = __float_as_uint(__fadd_rn(clamp_keep_nan(div_const(acc17, p.g.fV), 0.0f, 1.0f), 0.0f));
                    acc17 = 0.0f;
                    yz17 = max(yz17, u17);
                    if (want_xy) {
                        unsigned m = max(u17, __shfl_xor_sync(FULL, u17, 4));
                        m = max(m, __shfl_xor_sync(FULL, m, 8));
                        m = max(m, __shfl_xor_sync(FULL, m, 16));
                        if (zl == 0 && m && x0 + quad < p.nx && j < p.ny) {
                            unsigned *pl = p.pxy + (size_t)c * (p.nhwc ? CHS : p.g.J) * p.nx * p.ny;
                            atomicMax(p.nhwc ? pl + ((size_t)(x0 + quad) * p.ny + j) * CHS + 16
                                             : pl + ((size_t)16 * p.nx + x0 + quad) * p.ny + j, m);
                        }
                    }
                    if (want_xz) s_x17[(quad * NW + w) * 8 + zl] = u17;
                }
                if (want_xy || want_xz) {
                    __syncthreads();
                    if (X17 && want_xz && tid < 32) {
                        const int x = tid >> 3, z = tid & 7;
                        unsigned m = 0u;
#pragma unroll
                        for (int y = 0; y < NW; ++y) m = max(m, s_x17[(x * NW + y) * 8 + z]);
                        if (m && x0 + x < p.nx && bz * 8 + z < p.nzs) {
                            unsigned *pl = p.pxz + (size_t)c * (p.nhwc ? CHS : p.g.J) * p.nx * p.nzs;
                            atomicMax(p.nhwc ? pl + ((size_t)(x0 + x) * p.nzs + bz * 8 + z) * CHS + 16
                                             : pl + ((size_t)16 * p.nx + x0 + x) * p.nzs + bz * 8 + z, m);
                        }
                    }
                    constexpr int NXY = 4 * NW * Q;  // (x, y, quad) owners of the xy cells of a brick, then 32 * Q owners of xz cells
                    for (int item = tid; item < NXY + 32 * Q; item += NW * 32) {
                        const int iq = item & 3;
                        if (item < NXY) {
                            if (!want_xy) continue;
                            const int cell = item >> 2, x = cell / NW, y = cell - x * NW;
                            const unsigned *src = s_park + (size_t)cell * 8 * CH + iq * 4;
                            uint4 m = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
                            for (int kk = 0; kk < 8; ++kk) {  // rotated start: the two cells of a quarter-warp hit different banks
                                const uint4 t = *reinterpret_cast<const uint4 *>(src + ((kk + cell) & 7) * CH);
                                m.x = max(m.x, t.x); m.y = max(m.y, t.y); m.z = max(m.z, t.z); m.w = max(m.w, t.w);
                            }
                            red_max4(p.pxy + (size_t)c * (p.nhwc ? CHS : p.g.J) * p.nx * p.ny, x0 + x, by * NW + y, p.nx, p.ny,
                                     iq, m, p.nhwc, p.g.J, CHS);
                        } else {
                            if (!want_xz) continue;
                            const int cell = (item - NXY) >> 2, x = cell >> 3, z = cell & 7;
                            const unsigned *src = s_park + ((size_t)x * NW * 8 + z) * CH + iq * 4;
                            uint4 m = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
                            for (int y = 0; y < NW; ++y) {
                                const uint4 t = *reinterpret_cast<const uint4 *>(src + y * 8 * CH);
                                m.x = max(m.x, t.x); m.y = max(m.y, t.y); m.z = max(m.z, t.z); m.w = max(m.w, t.w);
                            }
                            red_max4(p.pxz + (size_t)c * (p.nhwc ? CHS : p.g.J) * p.nx * p.nzs, x0 + x, bz * 8 + z, p.nx, p.nzs,
                                     iq, m, p.nhwc, p.g.J, CHS);
                        }
                    }
                }
            }
            brick_live = 0u;
            v = 0;
            ++b;
        } else {
            ++v;
        }
        r0 = r1; r1 = r2; r2 = r3; ow1 = ow2;
        stage ^= 1u;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    if (CUBE) {
        zero_x(xlo + 4 * nb, p.nx);
        return;
    }
    // yz: this thread saw every x of its (y, z, quad)
    if (want_yz)
        red_max4(p.pyz + (size_t)c * (p.nhwc ? CHS : p.g.J) * p.ny * p.nzs, j, bz * 8 + zl, p.ny, p.nzs, quad,
                 make_uint4(yz_run[0], yz_run[1], yz_run[2], yz_run[3]), p.nhwc, p.g.J, CHS);
    if (X17 && want_yz) {   // channel 16: max over the four x lanes of (j, k)
        unsigned m = max(yz17, __shfl_xor_sync(FULL, yz17, 1));
        m = max(m, __shfl_xor_sync(FULL, m, 2));
        if (quad == 0 && m && j < p.ny && bz * 8 + zl < p.nzs) {
            unsigned *pl = p.pyz + (size_t)c * (p.nhwc ? CHS : p.g.J) * p.ny * p.nzs;
            atomicMax(p.nhwc ? pl + ((size_t)j * p.nzs + bz * 8 + zl) * CHS + 16
                             : pl + ((size_t)16 * p.ny + j) * p.nzs + bz * 8 + zl, m);
        }
    }
}

template <int NW, int MINB, bool X17 = false>
int launch_walk3(const ProjectParams &p0, int C, cudaStream_t s) {
    ProjectParams p = p0;
    p.nby = ceil_div(p.ny, NW);
    p.nbz = ceil_div(p.nzs, 8);
    auto kern = project_walk3_kernel<NW, MINB, X17>;
    const size_t smem = sizeof(unsigned) * 4 * NW * 8 * 16 + sizeof(float4) * NW * 64 + (X17 ? sizeof(unsigned) * 4 * NW * 8 : 0);
    if (smem > 48 * 1024) FVP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // enough CTAs to fill 148 SMs several times over, at least 4 bricks per segment
    const long rows = (long)p.nby * p.nbz * C, bricks = ceil_div(p.nx, 4);
    long nseg = (2L * 148 * MINB + rows - 1) / rows;
    nseg = std::max(1L, std::min(nseg, (long)std::max(1, (int)bricks / 4)));
    p.xseg = (int)((bricks + nseg - 1) / nseg);
    nseg = (bricks + p.xseg - 1) / p.xseg;
    dim3 grid((unsigned)(p.nby * p.nbz), (unsigned)C, (unsigned)nseg);
    kern<<<grid, NW * 32, smem, s>>>(p);
    return check_launch("project_walk3_kernel");
}

// the cube interface: grid = (y rows x groups of NW z bricks, cubes)
template <int NW, int MINB>
int launch_walk3_cube(const ProjectParams &p0, int C, cudaStream_t s) {
    ProjectParams p = p0;
    p.nby = p.ny;
    p.nbz = ceil_div(ceil_div(p.nzs, 8), NW);
    p.xseg = 0;
    auto kern = project_walk3_kernel<NW, MINB, false, true>;
    const size_t smem = sizeof(unsigned) * 4 * NW * 8 * 16 + sizeof(float4) * NW * 64;
    dim3 grid((unsigned)(p.nby * p.nbz), (unsigned)C, 1u);
    kern<<<grid, NW * 32, smem, s>>>(p);
    return check_launch("project_walk3_kernel<cube>");
}

}  // namespace

bool walk_cube_supported(const ProjectParams &p, int Jp, bool taps) {
    return taps && Jp == 16 && p.hm_ext == nullptr && p.cubes != nullptr && !(p.pxy || p.pxz || p.pyz) && p.z0 == 0 && p.z1 == p.nz;
}

int launch_project_walk_cube(const ProjectParams &p, int C, cudaStream_t s) { return launch_walk3_cube<4, 8>(p, C, s); }

bool walk_supported(const ProjectParams &p, int Jp, bool taps) {
    const bool layout = p.hm_ext != nullptr ? (Jp == 20 && p.g.J == 17) : Jp == 16;
    return taps && layout && p.cubes == nullptr && (p.pxy || p.pxz || p.pyz);
}

int launch_project_walk(const ProjectParams &p, int C, int variant, cudaStream_t s) {
    if (p.hm_ext != nullptr) {   // 17-joint rigs, split 16 + 1 layout
        switch (variant) {
            default: return launch_walk3<4, 8, true>(p, C, s);
            case 6: return launch_walk3<4, 6, true>(p, C, s);
            case 2: return launch_walk3<8, 4, true>(p, C, s);
            case 3: return launch_walk3<8, 3, true>(p, C, s);
        }
    }
    switch (variant) {
        default:   // 128-thread CTAs, up to 8 per SM: 2.06 ms (Gaussian heatmaps) / 3.97 ms (noise) on 320 person cubes
        case 1: return launch_walk3<4, 8>(p, C, s);
        case 2: return launch_walk3<8, 4>(p, C, s);
        case 3: return launch_walk3<8, 3>(p, C, s);
        case 4: return launch_walk3<8, 5>(p, C, s);
        case 5: return launch_walk3<16, 2>(p, C, s);
        case 6: return launch_walk3<4, 6>(p, C, s);
        case 7: return launch_walk3<4, 10>(p, C, s);
        case 8: return launch_walk3<2, 16>(p, C, s);
    }
}

}  // namespace fvp
