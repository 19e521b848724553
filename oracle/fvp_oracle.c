/*
 * TEST INFRASTRUCTURE -- plain-C restatement of the Faster-VoxelPose projection hot path.
 *
 * This file is the parity ORACLE.  It is never linked into, imported by or executed from the
 * product (faster_voxelpose_b200/); only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg load it (through oracle/oracle.py).
 *
 * The reference is pure Python/PyTorch; its arithmetic lives in torch 2.11.0 (the version in
 * this image; the reference pins none -- requirements.txt lists no torch, README.md:28-31 names
 * 1.12.0).  The fp32 operation ORDER below was established by probing torch-CPU bit patterns in
 * the authoring container (see DESIGN.md "fp32 op order"):
 *   - torch.mm with K=3 evaluates fma(a2,b2, fma(a1,b1, a0*b0));
 *   - every other elementwise op rounds separately (no contraction);
 *   - grid_sample(bilinear, zeros, align_corners=True) on CPU evaluates
 *       ix = (gx+1)*((W-1)/2);  w = ix-floor(ix); e = 1-w; ...; nw = s*e, ne = s*w, sw = n*e, se = n*w
 *       out = fma(v_se,se, fma(v_sw,sw, fma(v_ne,ne, v_nw*nw)))   (out-of-image taps read 0)
 *   - mean over views = sequential sum, then one division by V.
 * Parity status: PINNED -- checked against tests/golden/*.npz, which were produced by the
 * unmodified reference modules (tests/golden/make_golden.py).
 *
 * Build: gcc -O2 -fopenmp -ffp-contract=off -shared -fPIC (see oracle/build_oracle.py).
 * -ffp-contract=off matters: every fused multiply-add below is an explicit fmaf().
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define CAM_STRIDE 24 /* R[9] T[3] fx fy cx cy k[3] p[2] pad[3] */

typedef struct {
    int V, J, H, W;      /* views, joints, heatmap rows / cols                         */
    float clamp_max;     /* max(ORI_IMAGE_SIZE) -- project_whole.py:100                */
    float t[6];          /* resize_transform 2x3, row major (fp32)                     */
    float hm_w, hm_h;    /* HEATMAP_SIZE as floats                                     */
    float img_w, img_h;  /* IMAGE_SIZE as floats                                       */
    float wm1, hm1;      /* w-1, h-1                                                   */
} orc_geom;

static inline float clamp_nan(float x, float lo, float hi) {
    /* torch.clamp propagates NaN */
    if (x != x) return x;
    return x < lo ? lo : (x > hi ? hi : x);
}

/* cameras.py:30-56 + project_whole.py:96-116 + transforms.py:59-63, one point, one view.
 * Returns the normalised grid_sample coordinate pair. */
static inline void orc_coords(const float *cam, const orc_geom *g, float X, float Y, float Z,
                              float *gx, float *gy) {
    const float *R = cam, *T = cam + 9, *k = cam + 16, *p = cam + 19;
    const float fx = cam[12], fy = cam[13], cx = cam[14], cy = cam[15];
    float d0 = X - T[0], d1 = Y - T[1], d2 = Z - T[2];                     /* x.T - T       */
    float xc = fmaf(R[2], d2, fmaf(R[1], d1, R[0] * d0));                 /* torch.mm, K=3 */
    float yc = fmaf(R[5], d2, fmaf(R[4], d1, R[3] * d0));
    float zc = fmaf(R[8], d2, fmaf(R[7], d1, R[6] * d0));
    float den = zc + 1e-5f;                                               /* no z>0 test   */
    float y0 = xc / den, y1 = yc / den;
    float r = y0 * y0 + y1 * y1;
    float d = ((1.0f + k[0] * r) + (k[1] * r) * r) + ((k[2] * r) * r) * r;
    float u = (y0 * d + ((2.0f * p[0]) * y0) * y1) + p[1] * (r + (2.0f * y0) * y0);
    float v = (y1 * d + ((2.0f * p[1]) * y0) * y1) + p[0] * (r + (2.0f * y1) * y1);
    float px = fx * u + cx, py = fy * v + cy;
    px = clamp_nan(px, -1.0f, g->clamp_max);
    py = clamp_nan(py, -1.0f, g->clamp_max);
    float qx = fmaf(g->t[2], 1.0f, fmaf(g->t[1], py, g->t[0] * px));      /* mm(t,[x;y;1]) */
    float qy = fmaf(g->t[5], 1.0f, fmaf(g->t[4], py, g->t[3] * px));
    qx = (qx * g->hm_w) / g->img_w;
    qy = (qy * g->hm_h) / g->img_h;
    float sx = (qx / g->wm1) * 2.0f - 1.0f;
    float sy = (qy / g->hm1) * 2.0f - 1.0f;
    *gx = clamp_nan(sx, -1.1f, 1.1f);
    *gy = clamp_nan(sy, -1.1f, 1.1f);
}

/* One bilinear sample of all J channels of one view (ATen GridSamplerKernel.cpp, CPU path). */
static inline void orc_tap4(const float *hm /*[J][H][W]*/, const orc_geom *g, float gx, float gy,
                            float *out /*[J]*/) {
    const int W = g->W, H = g->H, J = g->J;
    float ix = (gx + 1.0f) * ((float)(W - 1) / 2.0f);
    float iy = (gy + 1.0f) * ((float)(H - 1) / 2.0f);
    float xw = floorf(ix), yn = floorf(iy);
    float w = ix - xw, e = 1.0f - w, n = iy - yn, s = 1.0f - n;
    float nw = s * e, ne = s * w, sw = n * e, se = n * w;
    int x0ok = (xw >= 0.0f) && (xw <= (float)(W - 1));
    int x1ok = (xw + 1.0f >= 0.0f) && (xw + 1.0f <= (float)(W - 1));
    int y0ok = (yn >= 0.0f) && (yn <= (float)(H - 1));
    int y1ok = (yn + 1.0f >= 0.0f) && (yn + 1.0f <= (float)(H - 1));
    long x0 = (x0ok || x1ok) ? (long)xw : 0, y0 = (y0ok || y1ok) ? (long)yn : 0;
    for (int j = 0; j < J; ++j) {
        const float *pl = hm + (size_t)j * H * W;
        float a = (x0ok && y0ok) ? pl[y0 * W + x0] : 0.0f;
        float b = (x1ok && y0ok) ? pl[y0 * W + x0 + 1] : 0.0f;
        float c = (x0ok && y1ok) ? pl[(y0 + 1) * W + x0] : 0.0f;
        float dd = (x1ok && y1ok) ? pl[(y0 + 1) * W + x0 + 1] : 0.0f;
        out[j] = fmaf(dd, se, fmaf(c, sw, fmaf(b, ne, a * nw)));
    }
}

/* a1-a4: the grid the reference caches per sequence.  out[V][X*Y*Z][2]. */
void orc_sample_grid(const float *cams, const orc_geom *g, const float *ax, const float *ay,
                     const float *az, int X, int Y, int Z, float *out) {
    const long nb = (long)X * Y * Z;
    for (int v = 0; v < g->V; ++v) {
#pragma omp parallel for schedule(static)
        for (long i = 0; i < nb; ++i) {
            int z = (int)(i % Z), y = (int)((i / Z) % Y), x = (int)(i / ((long)Z * Y));
            orc_coords(cams + v * CAM_STRIDE, g, ax[x], ay[y], az[z], out + ((long)v * nb + i) * 2,
                       out + ((long)v * nb + i) * 2 + 1);
        }
    }
}

/* Generic cube projection.  Cube c samples frame frame[c]; its voxel (i,j,k) sits at table
 * index tl[c]+(i,j,k); only voxels with table index in [start[c], end[c]) are sampled, all
 * others stay 0 (project_individual.py:272-290); skip[c] != 0 leaves the cube 0.
 * Space cube (project_whole.py:119-168): tl = 0, start = 0, end = (X,Y,Z).
 * out[C][J][nx][ny][nz]; planes (optional) xy[C][J][nx][ny], xz[C][J][nx][nz], yz[C][J][ny][nz]. */
void orc_project_cubes(const float *heatmaps /*[B][V][J][H][W]*/, const float *cams /*[R][V][24]*/,
                       const int *frame_rig /*[B]*/, const orc_geom *g, const float *ax,
                       const float *ay, const float *az, int C, const int *frame, const int *tl,
                       const int *start, const int *end, const int *skip, int nx, int ny, int nz,
                       float *out) {
    const int V = g->V, J = g->J;
    const size_t plane = (size_t)g->H * g->W, vox = (size_t)nx * ny * nz;
    memset(out, 0, sizeof(float) * (size_t)C * J * vox);
#pragma omp parallel
    {
        float *acc = (float *)malloc(sizeof(float) * 2 * J);
        float *tap = acc + J;
#pragma omp for collapse(2) schedule(dynamic, 8)
        for (int c = 0; c < C; ++c) {
            for (int i = 0; i < nx; ++i) {
                if (skip && skip[c]) continue;
                const int b = frame[c];
                const float *rig = cams + (size_t)frame_rig[b] * V * CAM_STRIDE;
                int gi = tl[3 * c] + i;
                if (gi < start[3 * c] || gi >= end[3 * c]) continue;
                for (int j = 0; j < ny; ++j) {
                    int gj = tl[3 * c + 1] + j;
                    if (gj < start[3 * c + 1] || gj >= end[3 * c + 1]) continue;
                    for (int k = 0; k < nz; ++k) {
                        int gk = tl[3 * c + 2] + k;
                        if (gk < start[3 * c + 2] || gk >= end[3 * c + 2]) continue;
                        for (int v = 0; v < V; ++v) {
                            float gx, gy;
                            orc_coords(rig + v * CAM_STRIDE, g, ax[gi], ay[gj], az[gk], &gx, &gy);
                            orc_tap4(heatmaps + ((size_t)b * V + v) * J * plane, g, gx, gy, tap);
                            for (int q = 0; q < J; ++q) acc[q] = v ? acc[q] + tap[q] : tap[q];
                        }
                        for (int q = 0; q < J; ++q)
                            out[((size_t)c * J + q) * vox + ((size_t)i * ny + j) * nz + k] =
                                clamp_nan(acc[q] / (float)V, 0.0f, 1.0f);
                    }
                }
            }
        }
        free(acc);
    }
}

/* a13 / a6: the three orthographic max-planes of cubes[C][J][nx][ny][nz]
 * (joint_localization_net.py:158-160, cnns_2d.py:291).  NaN-propagating like torch.max. */
static inline float max_nan(float a, float b) { return (a != a || a > b) ? a : ((b != b) ? b : (a > b ? a : b)); }

void orc_planes3(const float *cubes, int CJ, int nx, int ny, int nz, float *xy, float *xz, float *yz) {
#pragma omp parallel for schedule(static)
    for (int c = 0; c < CJ; ++c) {
        const float *q = cubes + (size_t)c * nx * ny * nz;
        float *pxy = xy ? xy + (size_t)c * nx * ny : NULL;
        float *pxz = xz ? xz + (size_t)c * nx * nz : NULL;
        float *pyz = yz ? yz + (size_t)c * ny * nz : NULL;
        for (int i = 0; i < nx; ++i)
            for (int j = 0; j < ny; ++j)
                for (int k = 0; k < nz; ++k) {
                    float v = q[((size_t)i * ny + j) * nz + k];
                    if (pxy) pxy[i * ny + j] = k ? max_nan(pxy[i * ny + j], v) : v;
                    if (pxz) pxz[i * nz + k] = j ? max_nan(pxz[i * nz + k], v) : v;
                    if (pyz) pyz[j * nz + k] = i ? max_nan(pyz[j * nz + k], v) : v;
                }
    }
}

/* a11-a12 window arithmetic (project_individual.py:255-269,274).  All per-axis constants are
 * passed in as the fp32/int values the reference constructor produces (:43-85). */
void orc_individual_windows(const float *props /*[N][7]*/, int N, const float *scale, const float *bias,
                            const int *fine, const float *whole_size, const float *ind_size,
                            const int *voxels, int *tl, float *offset, int *start, int *end, int *skip) {
    for (int n = 0; n < N; ++n) {
        int bad = 0;
        for (int a = 0; a < 3; ++a) {
            float c = props[7 * n + a] * scale[a] + bias[a];
            int t = (int)rintf(c);                                 /* torch.round: half to even */
            tl[3 * n + a] = t;
            offset[3 * n + a] =
                (((float)t / (float)(fine[a] - 1)) * whole_size[a] - whole_size[a] / 2.0f) + ind_size[a] / 2.0f;
            int m = 0;
            if (a < 2) {
                m = (int)(((1.0f - props[7 * n + 5 + a]) / 2.0f) * (float)(voxels[a] - 1)); /* trunc */
                if (m < 0) m = 0;
            }
            int s = t + m >= 0 ? t + m : 0;
            int e = t + voxels[a] - m <= fine[a] ? t + voxels[a] - m : fine[a];
            start[3 * n + a] = s;
            end[3 * n + a] = e;
            bad |= (s >= e);
        }
        skip[n] = bad;
    }
}

/* a7: 3x3 max-equality NMS + top-k (proposal.py:34-76).  Tie rule: larger
 * value first, then smaller flat index.  idx2 uses the reference's divisor X (F8). */
void orc_nms_topk(const float *hm /*[B][X][Y]*/, int B, int X, int Y, int P, float *vals, int64_t *idx2,
                  int64_t *flat) {
    const int n = X * Y;
    float *kept = (float *)malloc(sizeof(float) * n);
    for (int b = 0; b < B; ++b) {
        const float *m = hm + (size_t)b * n;
        for (int x = 0; x < X; ++x)
            for (int y = 0; y < Y; ++y) {
                float mx = -INFINITY;
                for (int dx = -1; dx <= 1; ++dx)
                    for (int dy = -1; dy <= 1; ++dy) {
                        int xx = x + dx, yy = y + dy;
                        if (xx < 0 || yy < 0 || xx >= X || yy >= Y) continue;
                        float v = m[xx * Y + yy];
                        if (v > mx || v != v) mx = v;
                    }
                kept[x * Y + y] = (m[x * Y + y] == mx ? 1.0f : 0.0f) * m[x * Y + y];
            }
        for (int p = 0; p < P; ++p) {
            int best = -1;
            for (int i = 0; i < n; ++i) {
                int taken = 0;
                for (int q = 0; q < p; ++q) taken |= (flat[(size_t)b * P + q] == i);
                if (taken) continue;
                if (best < 0 || kept[i] > kept[best]) best = i;
            }
            vals[(size_t)b * P + p] = kept[best];
            flat[(size_t)b * P + p] = best;
            idx2[((size_t)b * P + p) * 2] = best / X;
            idx2[((size_t)b * P + p) * 2 + 1] = best % X;
        }
    }
    free(kept);
}

/* a9-a10: argmax over z of the 1-D maps, confidence product, 7-float proposal record in eval
 * mode (human_detection_net.py:208-218, 23-37, 82-125). */
void orc_proposal_records(const int64_t *idx2 /*[BP][2]*/, const float *conf2d /*[BP]*/,
                          const float *hm1d /*[BP][Z]*/, const float *bbox /*[BP][2]*/, int BP, int Z,
                          const float *scale, const float *bias, float min_score, float *rec /*[BP][7]*/) {
    for (int i = 0; i < BP; ++i) {
        int best = 0;
        for (int z = 1; z < Z; ++z)
            if (hm1d[(size_t)i * Z + z] > hm1d[(size_t)i * Z + best]) best = z;
        float conf = conf2d[i] * hm1d[(size_t)i * Z + best];
        rec[7 * i + 0] = (float)idx2[2 * i] * scale[0] + bias[0];
        rec[7 * i + 1] = (float)idx2[2 * i + 1] * scale[1] + bias[1];
        rec[7 * i + 2] = (float)best * scale[2] + bias[2];
        rec[7 * i + 3] = (conf > min_score ? 1.0f : 0.0f) - 1.0f;
        rec[7 * i + 4] = conf;
        rec[7 * i + 5] = bbox[2 * i];
        rec[7 * i + 6] = bbox[2 * i + 1];
    }
}

/* a14: soft-argmax over the n*n cells of each (plane, person, joint) map
 * (joint_localization_net.py:32-56).  pose[3][N][J][2]; maxprob[3][N][J]; conf[N] = mean. */
void orc_softargmax(const float *feat /*[3][N][J][M]*/, const float *grid /*[3][M][2]*/, int N, int J, int M,
                    float beta, float *pose, float *conf) {
    double *cs = (double *)calloc(N, sizeof(double));
    for (int a = 0; a < 3; ++a)
        for (int n = 0; n < N; ++n)
            for (int j = 0; j < J; ++j) {
                const float *x = feat + (((size_t)a * N + n) * J + j) * M;
                float mx = -INFINITY;
                for (int i = 0; i < M; ++i) mx = fmaxf(mx, beta * x[i]);
                double sum = 0.0, sx = 0.0, sy = 0.0;
                for (int i = 0; i < M; ++i) {
                    double e = exp((double)(beta * x[i] - mx));
                    sum += e;
                    sx += e * grid[((size_t)a * M + i) * 2];
                    sy += e * grid[((size_t)a * M + i) * 2 + 1];
                }
                pose[((((size_t)a * N + n) * J) + j) * 2] = (float)(sx / sum);
                pose[((((size_t)a * N + n) * J) + j) * 2 + 1] = (float)(sy / sum);
                cs[n] += 1.0 / sum;
            }
    for (int n = 0; n < N; ++n) conf[n] = (float)(cs[n] / (3.0 * J));
    free(cs);
}

/* a15: offset add + weighted plane fusion (joint_localization_net.py:168-171,83-120).
 * pose[3][N][J][2] is updated in place (offsets added); weights[3][N][J]; fused[N][J][3]. */
void orc_fuse(float *pose, const float *offset /*[N][3]*/, const float *weights, int N, int J, float *fused) {
    const size_t S = (size_t)N * J;
    for (int n = 0; n < N; ++n)
        for (int j = 0; j < J; ++j) {
            size_t i = (size_t)n * J + j;
            const float *o = offset + 3 * n;
            float *xy = pose + i * 2, *xz = pose + (S + i) * 2, *yz = pose + (2 * S + i) * 2;
            xy[0] += o[0]; xy[1] += o[1];
            xz[0] += o[0]; xz[1] += o[2];
            yz[0] += o[1]; yz[1] += o[2];
            float wxy = weights[i], wxz = weights[S + i], wyz = weights[2 * S + i];
            float sx = wxy + wxz, sy = wxy + wyz, sz = wxz + wyz;
            fused[3 * i + 0] = (wxy / sx) * xy[0] + (wxz / sx) * xz[0];
            fused[3 * i + 1] = (wxy / sy) * xy[1] + (wyz / sy) * yz[0];
            fused[3 * i + 2] = (wxz / sz) * xz[1] + (wyz / sz) * yz[1];
        }
}
