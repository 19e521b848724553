/*
 * fvp_b200.h -- C ABI of libfvp_b200.so: the B200 (sm_100a) voxel-projection hot path of
 * Faster-VoxelPose.
 *
 * The reference (ME495/Faster-VoxelPose) is pure Python/PyTorch and has no FFI; its boundary for
 * this path is the nn.Module interface in lib/models.  The entry points below are
 * what a binding for that path needs: every pointer is a plain device pointer (unless marked
 * HOST), every size an int, `stream` is a cudaStream_t passed as void* (0 = default stream).
 * No torch / ATen type crosses this boundary.  The Python side (faster_voxelpose_b200/_lib.py)
 * binds them with ctypes on tensor.data_ptr().
 *
 * Conventions
 *   - return value: 0 on success, negative FVP_E_* on error; fvp_last_error() gives the text
 *     (thread local).  Launches are asynchronous on `stream`; no call synchronises.
 *   - the caller owns every buffer.  Outputs are fully written (no pre-zeroing needed) unless
 *     stated otherwise.
 *   - all floating point is fp32 and follows the reference's operation order bit for bit
 *     (DESIGN.md "fp32 op order"); indices are int32 on the way in, int64 where the reference
 *     returns int64.
 *   - heatmap values are expected to be FINITE.  A bilinear tap that falls outside the image is not skipped as
 *     F.grid_sample skips it: it is read from the nearest row / column inside the image and weighted with +0, so an
 *     Inf / NaN pixel on the image border would turn a sample that torch computes as finite into NaN.  NaN sample
 *     COORDINATES (degenerate cameras) propagate exactly as in torch.
 *   - heatmaps arrive in the reference's [B,V,J,H,W] ("planar") layout; the gather kernels read the
 *     "row-paired" channel-contiguous layout of fvp_heatmaps_rowpair (joints padded to
 *     Jp = 4*ceil(J/4), north|south rows adjacent) with 256-bit loads.
 *     fvp_heatmaps_channels_last ([B,V,H,W,Jp]) is a plain layout utility.
 */
#ifndef FVP_B200_H
#define FVP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FVP_ABI_VERSION 13
#define FVP_CAM_STRIDE 24 /* floats per camera: R[9] T[3] fx fy cx cy k[3] p[2] pad[3]; tables 32-byte aligned */
#define FVP_WIN_STRIDE 12 /* ints per cube window: tl[3] start[3] end[3] skip frame pad */

#define FVP_OK 0
#define FVP_E_ARG (-1)    /* bad argument (null pointer, non-positive size, unsupported shape) */
#define FVP_E_CUDA (-2)   /* a CUDA call or launch failed; text carries cudaGetErrorString   */

/* Per-layer projection constants.  Replaces the attributes ProjectLayer.__init__ keeps
 * (lib/models/project_whole.py:29-32, project_individual.py:37-40) plus the resize matrix that
 * every forward receives (project_whole.py:119). */
typedef struct fvp_geom {
    int32_t V, J, H, W;   /* views, joints, heatmap rows, heatmap cols                       */
    float clamp_max;      /* max(ORI_IMAGE_SIZE)            project_whole.py:100             */
    float t[6];           /* resize_transform, row major    transforms.py:59-63              */
    float hm_w, hm_h;     /* HEATMAP_SIZE                   project_whole.py:106             */
    float img_w, img_h;   /* IMAGE_SIZE                     project_whole.py:107-108         */
    float wm1, hm1;       /* w-1, h-1                       project_whole.py:111-113         */
} fvp_geom;

/* Projection modes (flags argument of fvp_project_cubes). */
#define FVP_MODE_DIRECT 0u      /* taps gathered straight from global memory through L1       */
#define FVP_FLAG_PLANES_NHWC 32u /* fvp_project_planes writes xy/xz/yz channels-last,
                                   [C][n1][n2][Jp] with channels J..Jp-1 = 0 (what the conv nets read) */
#define FVP_FLAG_CACHED_GRID 16u /* read sample coordinates from `cached_grid` instead of
                                   projecting on the fly (A/B evidence)          */

#define FVP_FLAG_SPLIT17 128u    /* fvp_project_planes: `hm_cl` is the split 16 + 1 layout of
                                   fvp_heatmaps_rowpair_split17 (17 joints, Jp = 20); needs FVP_FLAG_TAP_TABLE,
                                   planes only -- served by the row-walking kernel */
#define FVP_FLAG_TAP_TABLE 64u   /* `cached_grid` is the tap table of fvp_tap_table: the kernel copies
                                   precomputed tap records instead of projecting (the reference caches
                                   its sample grid per sequence the same way, project_whole.py:138-141) */

int fvp_abi_version(void);
const char *fvp_last_error(void);

/* Process-wide tuning switches (A/B evidence for DESIGN.md; results never depend on them).
 *   FVP_OPT_WALK_VARIANT  kernel behind fvp_project_planes when the cube is not requested, Jp == 16 (or the split
 *                         16 + 1 layout of 17-joint heatmaps, which has no other kernel) and the
 *                         positions come from a tap table: 0 = the row-walking kernel with the library's
 *                         default shape, -1 = the generic brick kernel (round-1 form), n > 0 = tuning build n
 *                         of the row-walking kernel (fvp_project_walk.cu). */
#define FVP_OPT_WALK_VARIANT 0
#define FVP_OPT_COUNT 4
int fvp_set_option(int key, int value);
int fvp_get_option(int key, int *value /*HOST*/);

/* Number of SMs of the current device (grid sizing / bench reporting). */
int fvp_device_sm_count(int *out /*HOST*/);

/* a1-a4  lib/utils/cameras.py:30-56 (project_point), lib/utils/transforms.py:59-63,
 *        lib/models/project_whole.py:81-117 (project_grid).
 * Materialises the grid the reference caches per sequence: out[V][X*Y*Z][2] normalised (x,y),
 * voxel order x slowest / z fastest.  ax/ay/az are the per-axis voxel-centre tables
 * (torch.linspace + centre, project_whole.py:62-70). */
int fvp_sample_grid(const float *cams /*[V][24]*/, const fvp_geom *g /*HOST*/, const float *ax,
                    const float *ay, const float *az, int X, int Y, int Z, float *out, void *stream);

/* a1-a4 + the tap set-up of F.grid_sample (project_whole.py:162, project_individual.py:283), cached per
 * rig like the reference's sample grid (project_whole.py:138-141, project_individual.py:236-240).
 * For n = X*Y*Z voxels and npad = 4*ceil(n/4), view v owns bytes [v*20*npad, (v+1)*20*npad) of `out`:
 * npad float4 bilinear weights (nw, ne, sw, se; out-of-image taps forced to +0) followed by npad uint32
 * records (bit 31: some tap inside the image, bit 30: east taps one pixel to the right, bits 0-29: index
 * of the west/north pixel record within a frame of the row-paired heatmaps).  A table is valid for one
 * (rig, fvp_geom, axis tables) triple; rigs are stacked [R][V][20*npad] for FVP_FLAG_TAP_TABLE. */
int fvp_tap_table(const float *cams /*[V][24]*/, const fvp_geom *g /*HOST*/, const float *ax, const float *ay,
                  const float *az, int X, int Y, int Z, void *out, void *stream);

/* Layout pre-pass: planar [BV][J][H][W] -> channels-last [BV][H][W][Jp], channels J..Jp-1 = 0. */
int fvp_heatmaps_channels_last(const float *planar, int BV, int J, int H, int W, int Jp,
                               float *out, void *stream);

/* Layout pre-pass of the projection kernels: planar [BV][J][H][W] -> "row-paired"
 * [BV][2][HP][W][Jp/4][2][4] with HP = H/2 + 1.  For every pixel column the two rows of a bilinear
 * footprint are adjacent (north quad | south quad = 32 bytes, one 256-bit load per channel quad); copy 0
 * pairs rows (2m, 2m+1), copy 1 rows (2m-1, 2m), rows outside the image and channels J..Jp-1 are 0.
 * This is the `hm_cl` argument of fvp_project_cubes / fvp_project_planes. */
int fvp_heatmaps_rowpair(const float *planar, int BV, int J, int H, int W, int Jp, float *out, void *stream);

/* The same pre-pass (planar [B][V][J][H][W]) that also records which pixel records hold anything but +0:
 * occ[B][fvp_occupancy_words(V,H,W)] uint32; per frame, word 0 = number of records with a non-zero bit, then one
 * bit per strip of 8 consecutive records of the frame (record r = ((v*2 + copy)*HP + m)*W + x), bit s set when any
 * of the records [8s, 8s+8] is non-zero.  fvp_project_planes_occ uses it to drop samples whose four taps are all +0
 * -- exact, and most of the work when the heatmaps are rendered from 2-D poses (the reference's INPUT_HM path,
 * lib/dataset/JointsDataset.py:368-447) rather than produced by a backbone.  occ is cleared here. */
long fvp_occupancy_words(int V, int H, int W);
int fvp_heatmaps_rowpair_occ(const float *planar, int B, int V, int J, int H, int W, int Jp, float *out,
                             uint32_t *occ, void *stream);

/* The pre-pass for 17-joint heatmaps (Shelf, Campus: planar [B][V][17][H][W]) in the split 16 + 1 layout: a pixel
 * record of 20 padded channels is 160 bytes and straddles two 128-byte lines for every tap, so channels 0..15 go out
 * as 128-byte records exactly like a 16-channel row-paired array, out[B][V][2][HP][W][4][2][4] (R = B*V*2*HP*W
 * records), followed in the same buffer by channel 16 as one float4 per record position,
 * ((float4 *)out + 8*R)[r] = {north x, south x, north x+1, south x+1} (x+1 entries past the row end are 0 and never
 * weighted).  out holds 36*R floats.  occ as in fvp_heatmaps_rowpair_occ (all 17 channels; NULL = not recorded).
 * Read by fvp_project_planes[_occ] under FVP_FLAG_SPLIT17 (Jp = 20). */
int fvp_heatmaps_rowpair_split17(const float *planar, int B, int V, int H, int W, float *out, uint32_t *occ,
                                 void *stream);

/* a11-a12 window arithmetic  lib/models/project_individual.py:255-269,274.
 * props[N][7] proposal records, valid[N] (may be NULL = all valid), frame[N] frame index of each
 * proposal.  Writes win[N][FVP_WIN_STRIDE] = {tl xyz, start xyz, end xyz, skip, frame, 0} and
 * offset[N][3].  skip = 1 for an empty range (project_individual.py:274) or an invalid slot. */
int fvp_individual_windows(const float *props, const uint8_t *valid, const int32_t *frame, int N,
                           const float scale[3] /*HOST*/, const float bias[3] /*HOST*/,
                           const int32_t fine[3] /*HOST*/, const float whole_size[3] /*HOST*/,
                           const float ind_size[3] /*HOST*/, const int32_t voxels[3] /*HOST*/,
                           int32_t *win, float *offset, void *stream);

/* a5   lib/models/project_whole.py:119-168   (win == NULL: cube c = frame c, whole table)
 * a12  lib/models/project_individual.py:222-293 (win from fvp_individual_windows)
 * Fused projection + bilinear gather + mean over ALL views + clamp(0,1) for C cubes of
 * nx*ny*nz voxels.  Voxel (i,j,k) of cube c sits at table index tl+(i,j,k); voxels outside
 * [start,end) or in a skipped cube are 0.
 *   hm_cl        row-paired heatmaps of fvp_heatmaps_rowpair, [B][V][2][H/2+1][W][Jp/4][2][4]
 *   cams         [R][V][24] rig table; frame_rig[B] picks the rig of each frame
 *   cached_grid  [R][V][n_ax*n_ay*n_az][2] or NULL (required with FVP_FLAG_CACHED_GRID); with
 *                FVP_FLAG_TAP_TABLE the stacked tables of fvp_tap_table, [R][V][20*npad] bytes
 *   cubes        [C][J][nx][ny][nz] */
int fvp_project_cubes(const float *hm_cl, int B, int Jp, const float *cams, const int32_t *frame_rig,
                      const fvp_geom *g /*HOST*/, const float *ax, const float *ay, const float *az,
                      int n_ax, int n_ay, int n_az, const float *cached_grid, int C,
                      const int32_t *win, int nx, int ny, int nz, float *cubes, unsigned flags,
                      void *stream);

/* a5 + a6 / a12 + a13 fused  (the same arguments as fvp_project_cubes, plus the planes)
 *   lib/models/project_whole.py:119-168 + lib/models/cnns_2d.py:291 (z-max consumed by CenterNet),
 *   lib/models/project_individual.py:222-293 + lib/models/joint_localization_net.py:158-160.
 * Projects only the z slab [z0,z1) of every cube (z0 = 0, z1 = nz for the whole cube; the slab is
 * what one rank of a z-sharded dense grid owns) and reduces it onto the orthographic max-planes
 * inside the kernel, so the cube never goes to HBM:
 *   xy[C][J][nx][ny]  max over the slab (partial when the slab is not the whole cube)
 *   xz[C][J][nx][z1-z0], yz[C][J][ny][z1-z0]  slab-local
 * Any plane may be NULL (at least one is required); `cubes` may be NULL, else the slab of the cube
 * is written as well ([C][J][nx][ny][nz], voxels outside the slab untouched).  With
 * FVP_FLAG_PLANES_NHWC the planes are [C][nx][ny][Jp], [C][nx][z1-z0][Jp], [C][ny][z1-z0][Jp]. */
int fvp_project_planes(const float *hm_cl, int B, int Jp, const float *cams, const int32_t *frame_rig,
                       const fvp_geom *g /*HOST*/, const float *ax, const float *ay, const float *az,
                       int n_ax, int n_ay, int n_az, const float *cached_grid, int C,
                       const int32_t *win, int nx, int ny, int nz, int z0, int z1, float *cubes,
                       float *xy, float *xz, float *yz, unsigned flags, void *stream);

/* fvp_project_cubes with the occupancy words of fvp_heatmaps_rowpair_occ for the same heatmaps (NULL = none):
 * identical cubes; samples over all-zero pixel records are not gathered. */
int fvp_project_cubes_occ(const float *hm_cl, int B, int Jp, const float *cams, const int32_t *frame_rig,
                          const fvp_geom *g /*HOST*/, const float *ax, const float *ay, const float *az,
                          int n_ax, int n_ay, int n_az, const float *cached_grid, int C,
                          const int32_t *win, int nx, int ny, int nz, float *cubes, unsigned flags,
                          const uint32_t *occ, void *stream);

/* fvp_project_planes with the occupancy words of fvp_heatmaps_rowpair_occ for the same heatmaps (NULL = none):
 * identical results; samples over all-zero pixel records are not gathered. */
int fvp_project_planes_occ(const float *hm_cl, int B, int Jp, const float *cams, const int32_t *frame_rig,
                           const fvp_geom *g /*HOST*/, const float *ax, const float *ay, const float *az,
                           int n_ax, int n_ay, int n_az, const float *cached_grid, int C,
                           const int32_t *win, int nx, int ny, int nz, int z0, int z1, float *cubes,
                           float *xy, float *xz, float *yz, unsigned flags, const uint32_t *occ, void *stream);

/* a8 without the cube  lib/models/human_detection_net.py:199-200.
 * flat[B][P] top-k cells of the [X][Y] map -> win[B*P][FVP_WIN_STRIDE] = one 1x1xZ column window
 * per proposal (frame = b), so that fvp_project_cubes(..., C = B*P, win, 1, 1, Z, cols) re-projects
 * just those columns: cols[B*P][J][1][1][Z] is feature_1d [B][P][J][Z]. */
int fvp_topk_windows(const int64_t *flat, int B, int P, int X, int Y, int Z, int32_t *win, void *stream);

/* a6 / a13  lib/models/cnns_2d.py:291, lib/models/joint_localization_net.py:158-160.
 * Orthographic max-planes of cubes[CJ][nx][ny][nz] (values must be >= 0 or NaN, which is what
 * fvp_project_cubes emits).  Any of xy[CJ][nx][ny], xz[CJ][nx][nz], yz[CJ][ny][nz] may be NULL. */
int fvp_planes3(const float *cubes, int CJ, int nx, int ny, int nz, float *xy, float *xz, float *yz,
                void *stream);

/* (multi-GPU)  z-slab plane combine over peer memory: the reduce-max of the xy plane and
 * the all-gather of the xz / yz slabs in ONE kernel that loads every rank's partial planes through peer
 * pointers (NVLink / NVSwitch P2P; pointers of symmetric-memory buffers mapped into this process), instead of
 * an NCCL all-reduce plus two all-gathers.  Rank r owns z in [zstart[r], zstart[r+1]), zstart[R] = nz.
 *   xy_peers[R]  HOST array of device pointers: partial xy plane of every rank, n_xy floats each
 *   xz_peers[R], yz_peers[R]  HOST arrays (or NULL): slabs [rows][zstart[r+1]-zstart[r]] of every rank
 *   out_xy[n_xy] = max over ranks (values >= 0 or NaN, NaN wins as in fvp_project_planes),
 *   out_xz[rows_xz][nz], out_yz[rows_yz][nz] = slabs concatenated along z.
 * The caller orders the launch after every rank's producer (a barrier on the symmetric buffers). */
int fvp_planes_combine_peers(const void *const *xy_peers /*HOST*/, const void *const *xz_peers /*HOST*/,
                             const void *const *yz_peers /*HOST*/, int R, const int32_t *zstart /*HOST [R+1]*/,
                             long n_xy, long rows_xz, long rows_yz, float *out_xy, float *out_xz, float *out_yz,
                             void *stream);

/* The same combine reduced inside the NVSwitch (NVLS).  mc_* are MULTICAST addresses of symmetric buffers
 * (one mapping that addresses every rank's copy): rank `rank` takes multimem.ld_reduce.max of its 1/R of the
 * partial xy planes and multimem.st's the result into every rank's full xy plane, and multimem.st's its own
 * xz / yz slab (plain local pointers, [rows][zstart[rank+1]-zstart[rank]]) into place in every rank's full
 * xz / yz planes ([rows][nz]).  Needs a barrier before (all partial planes written) and after (all ranks'
 * shares stored) the launch; the full planes are then complete in every rank's local copy. */
int fvp_planes_combine_multimem(const void *mc_xy_partial, void *mc_xy_full, const float *xz_slab, void *mc_xz_full,
                                const float *yz_slab, void *mc_yz_full, int R, int rank,
                                const int32_t *zstart /*HOST [R+1]*/, long n_xy, long rows_xz, long rows_yz,
                                void *stream);

/* a7  lib/core/proposal.py:34-76 (max_pool2D, nms2D, get_index2D).
 * hm[B][X][Y] -> vals[B][P] (descending), idx2[B][P][2] (flat/X, flat%X -- the reference's
 * divisor), flat[B][P].  Tie rule: larger value, then smaller flat index.
 * scratch[B][X*Y] floats is workspace (receives the NMS'd map, then is consumed) for maps of more than 10 240 cells;
 * smaller maps are kept in shared memory and scratch is left untouched. */
int fvp_nms_topk(const float *hm, int B, int X, int Y, int P, float *scratch, float *vals,
                 int64_t *idx2, int64_t *flat, void *stream);

/* fvp_nms_topk without the scratch argument, for callers that cannot allocate device memory (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same inputs, outputs, tie rule and results as fvp_nms_topk; the
 * workspace is a static device array inside the library (FVP_NMS_STATIC_MAX_CELLS floats = 4 MiB, resident on
 * every GPU that loads the library until the process exits).  B*X*Y must not exceed FVP_NMS_STATIC_MAX_CELLS:
 * larger problems return FVP_E_ARG (fvp_last_error names the limit) -- use fvp_nms_topk for those.  The call
 * performs no cudaMalloc / cudaFree and calls no allocating library routine; the first call on each device looks
 * the array's device address up (cudaGetSymbolAddress, with the caller's current device) and caches it.
 * All calls on a device share the one workspace, which keeps the last call's contents (nothing needs clearing:
 * the kernel overwrites every cell before reading it): two calls that may overlap on the GPU -- different
 * streams, or concurrent graph branches -- RACE; keep them ordered on one stream or use fvp_nms_topk. */
#define FVP_NMS_STATIC_MAX_CELLS (1 << 20)
int fvp_nms_topk_static(const float *hm, int B, int X, int Y, int P, float *vals, int64_t *idx2,
                        int64_t *flat, void *stream);

/* a8 lib/models/human_detection_net.py:191-192,199-200.
 * bbox[B][2][X][Y] -> match_bbox[B][P][2]; cubes[B][J][X][Y][Z] -> cols[B][P][J][Z], both at the
 * top-k cells flat[B][P].  Either output pair may be NULL. */
int fvp_gather_topk(const float *bbox, const float *cubes, const int64_t *flat, int B, int P, int J,
                    int X, int Y, int Z, float *match_bbox, float *cols, void *stream);

/* a9-a10  lib/models/human_detection_net.py:208-218 and ProposalLayer :23-37,82-125 (eval).
 * argmax over z of hm1d[B*P][Z], conf = conf2d*conf1d, record [x,y,z mm, match, conf, bw, bh]. */
int fvp_proposal_records(const int64_t *idx2, const float *conf2d, const float *hm1d,
                         const float *match_bbox, int BP, int Z, const float scale[3] /*HOST*/,
                         const float bias[3] /*HOST*/, float min_score, float *rec, void *stream);

/* Res2DBlock with 32 output channels  lib/models/cnns_2d.py:39-76 (eval mode, BatchNorm folded by the caller):
 *   out = ReLU(conv3x3(ReLU(conv3x3(x, w1) + b1), w2) + b2 + skip(x)),  skip = conv1x1(x, wskip) or x (wskip NULL, cin 32)
 * as one kernel on the tcgen05 tensor cores (TF32 operands, fp32 accumulation in tensor memory; the intermediate
 * activation stays in shared memory).  x[N][H][W][cin] and out[N][H][W][32] are channels-last, cin is 16 or 32.
 * Filters are packed K-major per tap: w1[9][cin/4][32][4] with w1[t][p][o][i] = W1[o][4p+i][t/3][t%3], w2[9][8][32][4]
 * likewise, wskip[cin/4][32][4]; b2 carries the skip convolution's bias as well.  fvp_res2d_supported says whether an
 * image size fits a tile (else the caller keeps its library path).  Meant for use where TF32 convolutions are allowed
 * (torch.backends.cudnn.allow_tf32, torch's default). */
int fvp_res2d_supported(int cin, int H, int W);
/* Diagnostic: counters[8] (DEVICE, int64, caller-zeroed; NULL switches it off) receive the cycles the first CTA of every
 * following fvp_res2d_block_tf32 launch spends staging the tile, issuing / waiting for the MMAs of either convolution
 * and in the two epilogues, and the number of tiles it processed (tools/res2d_phases.py). */
int fvp_res2d_profile(long long *counters);
int fvp_res2d_block_tf32(const float *x, int N, int H, int W, int cin, const float *w1, const float *b1,
                         const float *w2, const float *b2, const float *wskip, float *out, void *stream);

/* a14  lib/models/joint_localization_net.py:32-56 (SoftArgmaxLayer).
 * feat[3][N][J][M], grid[3][M][2] -> pose[3][N][J][2], maxprob[3][N][J]. */
int fvp_softargmax(const float *feat, const float *grid, int N, int J, int M, float beta,
                   float *pose, float *maxprob, void *stream);

/* a14 (confidence mean) + a15  lib/models/joint_localization_net.py:50-51,168-171,83-120.
 * Adds offset[N][3] to pose in place, fuses with weights[3][N][J] -> fused[N][J][3],
 * conf[N] = mean over planes and joints of maxprob. */
int fvp_fuse(float *pose, const float *maxprob, const float *offset, const float *weights, int N,
             int J, float *fused, float *conf, void *stream);

/* (next row)  lib/models/weight_net.py:36-88, front end of WeightNet:
 * conv3x3(1->F) + eval BatchNorm (folded into w,b by the caller) + MaxPool2d(2) + ReLU + global average
 * pool, one kernel: maps[M][n][n] -> out[M][F].  w[F][9] and b[F] are HOST arrays (they travel as
 * kernel parameters); F is 16, 32 or 64. */
int fvp_weightnet_features(const float *maps, int M, int n, const float *w /*HOST*/, const float *b /*HOST*/,
                           int F, float *out, void *stream);

/* The same front end with the convolution on TF32 tensor cores (mma.m16n8k8 over taps 0..7, the ninth tap in
 * fp32): for callers that run their convolutions in TF32 anyway (torch's default; what the reference's
 * Conv2d does on a GPU).  n must be a multiple of 8, maps 16-byte aligned. */
int fvp_weightnet_features_tf32(const float *maps, int M, int n, const float *w /*HOST*/, const float *b /*HOST*/,
                                int F, float *out, void *stream);

/* (next row)  memory-bound glue of the 2-D conv nets' inference path; activations are
 * channels-last fp32 [N][H][W][C], C % 4 == 0.
 * MaxPool2d(2) of lib/models/cnns_2d.py:113,116: x[N][H][W][C] -> out[N][H/2][W/2][C] (NaN-propagating). */
int fvp_maxpool2_nhwc(const float *x, int N, int H, int W, int C, float *out, void *stream);

/* ConvTranspose2d(k=2,s=2) + skip add of lib/models/cnns_2d.py:122-126, after the caller ran the transposed
 * conv as a 1x1 conv to 4*C channels ordered (a, b, c):
 * out[n][2h+a][2w+b][c] = y[n][h][w][(2a+b)*C + c] + s[n][2h+a][2w+b][c]. */
int fvp_upshuffle_add(const float *y, const float *s, int N, int H, int W, int C, float *out, void *stream);

/* P2PNet output layer lib/models/cnns_2d.py:238-240 (1x1 conv, 32 -> Cout <= 20) writing planar maps:
 * out[n][co][y][x] = b[co] + sum_c w[co][c] * (a[n][y][x][c] + up[n][y/2][x/2][2(y&1)+(x&1)][c]).
 * `up` ([N][H/2][W/2][4][C], the un-shuffled 1x1 form of the last up-sampling) may be NULL.
 * w[Cout][32], b[Cout] are HOST arrays (kernel parameters). */
int fvp_pointwise_planar(const float *a, const float *up, int N, int H, int W, int C, const float *w /*HOST*/,
                         const float *b /*HOST*/, int Cout, float *out, void *stream);
/* The same with the input formed on the way in as the output of a residual block,  relu(a + res_bias + res)
 * (cnns_2d.py:75-76): `a` is the block's second convolution without bias, residual add or ReLU, `res` [N][H][W][32] the
 * block's skip input, res_bias[32] (HOST) that convolution's bias. */
int fvp_pointwise_planar_res(const float *a, const float *res, const float *res_bias /*HOST*/, const float *up, int N, int H,
                             int W, int C, const float *w /*HOST*/, const float *b /*HOST*/, int Cout, float *out,
                             void *stream);

/* (next row)  lib/dataset/JointsDataset.py:368-447 generate_input_heatmap (+ :265-279
 * compute_human_scale), data augmentation off: the network's input heatmaps from 2-D poses, on the device.
 * joints[M][N][J][2]: network-input image pixels (after the resize affine), M = frames x views map sets;
 * vis[M][N][J] (0 = skip the joint) or NULL; count[M] = people per map set or NULL (N everywhere).
 * out[M][J][hm_h][hm_w] is zero-filled here, then every Gaussian is max-combined into it.
 * float64 arithmetic up to the exponential, as numpy evaluates the reference; values rounded to fp32. */
int fvp_splat_heatmaps(const float *joints, const unsigned char *vis, const int32_t *count, int M, int N, int J,
                       float sigma, int img_w, int img_h, int hm_w, int hm_h, float *out, void *stream);

/* (next row)  lib/models/cnns_1d.py:165-241 C2CNet (eval mode) in one kernel:
 * x[n][J][L] z-columns -> out[n][L], L = 20.  `weights` is one flat DEVICE buffer of BatchNorm-folded filters,
 * `ops` a HOST table [nops][10] = {type, in, out, add, cin, log2(cout), k, relu, w_off, b_off} over twelve
 * shared-memory activation buffers (ids 0-3: 32 rows x 20, 4-7: 64 x 10, 8-11: 128 x 5); type 0 = conv
 * (filters [cin][k][cout]), 1 = max-pool 2, 2 = transposed conv k2/s2 + ReLU + skip add (filters
 * [cin][2][cout]), 3 = final 1x1 conv to one channel.  Input columns are loaded into buffer 0. */
int fvp_c2c_forward(const float *x, int n, int J, int L, const float *weights, const int32_t *ops /*HOST*/, int nops,
                    float *out, void *stream);

/* (next row)  front layer of P2PNet / CenterNet, lib/models/cnns_2d.py:190-192
 * (conv 7x7, 16 -> 16 channels, pad 3, + folded BatchNorm + optional ReLU) on TF32 tensor cores.
 * in/out: channels-last [N][H][W][16] fp32.  wfrag: DEVICE buffer of 49*2*32*4 floats, the TF32-rounded
 * filters arranged as mma.m16n8k8 B fragments [tap = ky*7+kx][k-step][lane = g*4+t][{(cin ks*8+t, cout g),
 * (ks*8+t+4, g), (ks*8+t, 8+g), (ks*8+t+4, 8+g)}].  bias: DEVICE [16].  Inputs are rounded to TF32 (rna),
 * accumulation is fp32 -- the arithmetic of cuDNN's TF32 engines, which torch selects by default. */
int fvp_conv7x7_c16_tf32(const float *in, const float *wfrag, const float *bias, int N, int H, int W, int relu,
                         float *out, void *stream);
/* The same layer for 20-channel inputs (17-joint rigs: planes padded to 20 channels): wfrag carries 49*32*2 more floats,
 * [tap][lane][n8 tile] = W[8*tile + lane/4][16 + lane%4][tap], consumed as one k4 step per tap. */
int fvp_conv7x7_c20_tf32(const float *in, const float *wfrag, const float *bias, int N, int H, int W, int relu,
                         float *out, void *stream);

/* The front 7x7 layer (fvp_conv7x7_c16_tf32) on the tcgen05 tensor cores.  wpairs[28][cin/4][32][4]: pair-tap pt = 4*ky + pr holds
 * in rows 0..15 the filters of tap (ky, 2pr), in rows 16..31 those of tap (ky, 2pr+1) (zeros for pr = 3):
 * wpairs[pt][p][row][i] = W[row % 16][4p + i][ky][2pr + row / 16], TF32-rounded, p < cin/4; for cin = 20 (17-joint rigs, planes
 * padded to 20 channels) 128 zeros follow (the empty half of the last k-step).  in[N][H][W][cin], cin 16 or 20, out[N][H][W][16],
 * channels-last.  fvp_conv7x7_tc_supported says whether an image size fits (W >= 58). */
int fvp_conv7x7_tc_supported(int cin, int H, int W);
int fvp_conv7x7_tc_tf32(const float *in, const float *wpairs, const float *bias, int N, int H, int W, int cin, int relu,
                        float *out, void *stream);

/* The tail of P2PNet in one kernel on the tcgen05 tensor cores: second convolution + skip addition + ReLU of skip_res1
 * (cnns_2d.py:72-76), `x = decoder_upsample1(x) + skip_x1` (:179-180) and output_layer (:212, :221):
 *   out[n][co][y][x] = bout[co] + sum_c wout[co][c] * ( relu(conv3x3(h, w2)[c] + b2[c] + xres[c]) + shuffle(up)[c] ).
 * h [N][H][W][32]: the block's first convolution, ReLU done; xres [N][H][W][32]: the block's input; up [N][H/2][W/2][4][32]:
 * the last up-sampling as a 1x1 convolution to 4 x 32 channels, (y & 1, x & 1) major (fvp_upshuffle_add's operand);
 * w2packed[9][8][32][4] = W2[cout][4p + i][ky][kx] at [3 ky + kx][p][cout][i], TF32-rounded (DEVICE); b2[32], wout[cout][32],
 * bout[cout] on the HOST (they travel as kernel parameters).  out [N][cout][H][W] planar.  TF32 products with fp32
 * accumulation for the 3x3 convolution (torch's default for convolutions), fp32 FMAs for the 1x1 layer.
 * fvp_res_tail_supported: even H and W, W >= 32, cout <= 20. */
int fvp_res_tail_supported(int H, int W, int cout);
/* Diagnostics: device array of 12 int64 counters (or NULL to stop) that CTA 0 of later launches adds its per-phase cycles to
 * (tools/res_tail_time.py). */
int fvp_res_tail_profile(long long *counters);
int fvp_res_tail_tf32(const float *h, const float *xres, const float *up, int N, int H, int W, const float *w2packed,
                      const float *b2 /*HOST*/, const float *wout /*HOST*/, const float *bout /*HOST*/, int cout, float *out,
                      void *stream);

/* Evaluation, the numeric part  lib/dataset/panoptic.py:224-246 (next-row widening).
 * pred[F][P][J][5] fused poses (x, y, z mm, match flag, score), gt[F][G][J][3] ground-truth joints in DOUBLE (mm, padded to G
 * persons per frame), vis[F][G][J] visibility, ngt[F] persons per frame.  For every pose with match flag >= 0 in a frame
 * with ground truth: err[F][P] = mean distance over the visible (> 0.1) joints to its closest person, best[F][P] = that
 * person's index (first minimum); otherwise err = +inf, best = -1.  Double precision, as numpy computes it. */
int fvp_pose_match(const float *pred, const double *gt, const float *vis, const int32_t *ngt, int F, int P, int J, int G,
                   double *err, int32_t *best, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* FVP_B200_H */
