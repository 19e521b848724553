// Declarations shared by the projection translation units (fvp_project.cu: the generic brick kernel and
// the C entry points; fvp_project_walk.cu: the row-walking planes kernel).
#pragma once
#include "fvp_geom.cuh"

namespace fvp {

// One brick = BX*BY*BZ = 256 voxels of one cube (generic kernel); the walking kernel uses BX x NW x BZ.
constexpr int BX = 4, BY = 8, BZ = 8, BRICK = BX * BY * BZ;
constexpr int VCHUNK = 8;  // views staged per pass of the generic kernel: 8 * 256 * 20 B = 40 KB of dynamic shared memory

struct ProjectParams {
    const float4 *hm;        // row-paired heatmaps [B][V][2][HP][W][Q][2] float4 (fvp_heatmaps_rowpair)
    const float4 *hm_ext;    // split 16 + 1 layout (fvp_heatmaps_rowpair_split17): channel 16 as [B][V][2][HP][W] float4
                             // {north x, south x, north x+1, south x+1}, `hm` then holds channels 0..15 (Q = 4); else null
    const float *cams;       // [R][V][24]
    const int *frame_rig;    // [B]
    const float *ax, *ay, *az;
    const float2 *grid;      // cached sample grid, tap table (TAB kernels) or null
    const int *win;          // [C][12] or null
    float *cubes;            // [C][J][nx][ny][nz] or null
    unsigned *pxy, *pxz, *pyz;  // [C][J][nx][ny], [C][J][nx][nzs], [C][J][ny][nzs] or null (pre-zeroed)
    GeomDev g;
    int n_ax, n_ay, n_az, nx, ny, nz, nby, nbz;
    unsigned m_nby, m_nbz;   // ceil(2^32 / nby), ceil(2^32 / nbz): brick index decomposition by __umulhi (0 = divide)
    int z0, z1, nzs;         // z slab [z0,z1) of the cube processed by this launch; nzs = z1 - z0
    int vchunk;              // views staged per pass (<= VCHUNK)
    int nhwc;                // planes stored [C][n1][n2][4Q] instead of [C][J][n1][n2]
    const unsigned *occ;     // occupancy of the pixel records, [B][occ_stride] (fvp_heatmaps_rowpair_occ) or null
    int occ_stride;
    int xseg;                // row-walking kernel: bricks per x segment (gridDim.z segments per row)
};

// 256-bit read-only load (sm_100 LDG.E.256): north and south quad of one pixel column
struct __align__(32) Tap2 {
    float4 n, s;
};
__device__ __forceinline__ Tap2 ldg256(const float4 *p) {
    Tap2 r;
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r.n.x), "=f"(r.n.y), "=f"(r.n.z), "=f"(r.n.w), "=f"(r.s.x), "=f"(r.s.y), "=f"(r.s.z), "=f"(r.s.w)
                 : "l"(p));
    return r;
}

// Row-walking planes kernel (fvp_project_walk.cu): planes only, tap-table positions, Jp == 16 or the split 16 + 1 layout.
//   variant  0 = library default, else a tuning build (FVP_OPT_WALK_VARIANT)
bool walk_supported(const ProjectParams &p, int Jp, bool taps);
int launch_project_walk(const ProjectParams &p, int C, int variant, cudaStream_t s);
// the same walk writing the cubes themselves (cube interface, 16-channel layout, tap table, no planes, whole z range)
bool walk_cube_supported(const ProjectParams &p, int Jp, bool taps);
int launch_project_walk_cube(const ProjectParams &p, int C, cudaStream_t s);

int option(int key);  // fvp_abi.cu: values set through fvp_set_option

}  // namespace fvp
