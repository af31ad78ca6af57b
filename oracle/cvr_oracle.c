/*
 * cvr_oracle.c -- CPU ORACLE (test infrastructure, NOT the product).
 *
 * Scalar restatement of the arithmetic the reference's dense stereo hot path
 * performs.  The reference itself (asad-ahmed-pk/master-thesis, cv_reconstruct)
 * only holds thin glue; the arithmetic lives in two un-vendored third-party
 * libraries:
 *   - OpenCV >= 4.1 (unpinned; cv_reconstruct/CMakeLists.txt:14), calib3d/imgproc:
 *     cv::cvtColor, cv::StereoSGBM::compute, cv::medianBlur, cv::filterSpeckles,
 *     cv::normalize, cv::meanStdDev, cv::threshold, cv::reprojectImageTo3D.
 *   - PCL >= 1.8 (unpinned; cv_reconstruct/CMakeLists.txt:24): pcl::VoxelGrid,
 *     pcl::StatisticalOutlierRemoval.
 *
 * Pinning: the OpenCV half of this file is pinned against the real library
 * (python cv2 4.13.0 in the build container) by tests/golden/make_golden.py ->
 * tests/golden/ (npz files) and tests/test_oracle_golden.py.  The PCL half is a
 * restatement of PCL's published algorithm (filters/impl/voxel_grid.hpp,
 * filters/impl/statistical_outlier_removal.hpp); PCL is absent from the image
 * and the reference's tests hold no golden vectors => PCL-half "parity unpinned".
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library.  The product (libcvrecon_b200.so)
 * never links or calls it.
 *
 * Call sites restated (reference file:line):
 *   gray        cv_reconstruct/src/reconstruct/Reconstruct3D.cpp:78-79
 *   sgbm        cv_reconstruct/src/reconstruct/Reconstruct3D.cpp:39-69, :81
 *   prune       cv_reconstruct/src/system/TrackingFrame.cpp:25-44, 47-67
 *   reproject   cv_reconstruct/src/reconstruct/Reconstruct3D.cpp:26-36, 106-153
 *   triangulate cv_reconstruct/src/reconstruct/Reconstruct3D.cpp:156-216
 *   voxel       cv_reconstruct/src/system/MappingSystem.cpp:297-300, 72-75
 *   sor         cv_reconstruct/src/point_cloud/PointCloudPostProcessor.cpp:20-24, 38-41
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off: the reference's CMake
 * sets no -march/-mfma, so float expressions are evaluated without contraction).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <float.h>

#define ORC_MAX_COST 32767
#define ORC_DISP_SHIFT 4
#define ORC_DISP_SCALE 16

typedef struct orc_sgbm_params {
    int minDisparity;
    int numDisparities;
    int blockSize;
    int P1;
    int P2;
    int disp12MaxDiff;
    int preFilterCap;
    int uniquenessRatio;
    int speckleWindowSize;
    int speckleRange;
    int mode;              /* 0 = MODE_SGBM (5 paths, one pass), 1 = MODE_HH (8 paths, two passes) */
} orc_sgbm_params;

static inline int16_t sat16(int v) { return (int16_t)(v > 32767 ? 32767 : (v < -32768 ? -32768 : v)); }
static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }

/* ------------------------------------------------------------------------- */
/* cv::cvtColor(BGR2GRAY), 8-bit: (3735 B + 19235 G + 9798 R + 2^14) >> 15.   */
/* Reference call: Reconstruct3D.cpp:78-79.                                   */
void orc_gray_bgr(const uint8_t* bgr, size_t step, int W, int H, int channels, uint8_t* gray, size_t gstep)
{
    for (int y = 0; y < H; y++) {
        const uint8_t* s = bgr + (size_t)y * step;
        uint8_t* g = gray + (size_t)y * gstep;
        for (int x = 0; x < W; x++, s += channels)
            g[x] = (uint8_t)((3735 * s[0] + 19235 * s[1] + 9798 * s[2] + 16384) >> 15);
    }
}

/* ------------------------------------------------------------------------- */
/* 3x3 median on int16, BORDER_REPLICATE (cv::medianBlur(disp, disp, 3) inside */
/* StereoSGBM::compute).                                                      */
void orc_median3_s16(const int16_t* src, int16_t* dst, int W, int H)
{
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            int16_t v[9]; int n = 0;
            for (int dy = -1; dy <= 1; dy++)
                for (int dx = -1; dx <= 1; dx++) {
                    int yy = imin(imax(y + dy, 0), H - 1), xx = imin(imax(x + dx, 0), W - 1);
                    v[n++] = src[(size_t)yy * W + xx];
                }
            for (int i = 1; i < 9; i++) { int16_t t = v[i]; int j = i - 1; while (j >= 0 && v[j] > t) { v[j + 1] = v[j]; j--; } v[j + 1] = t; }
            dst[(size_t)y * W + x] = v[4];
        }
}

/* ------------------------------------------------------------------------- */
/* cv::filterSpeckles(img, newVal, maxSpeckleSize, maxDiff): 4-connected       */
/* components over pixels != newVal, edge iff |v(p)-v(q)| <= maxDiff; every    */
/* component with size <= maxSpeckleSize is set to newVal.                    */
void orc_filter_speckles(int16_t* img, int W, int H, int newVal, int maxSpeckleSize, int maxDiff)
{
    size_t N = (size_t)W * H;
    int32_t* label = (int32_t*)calloc(N, sizeof(int32_t));
    int32_t* stack = (int32_t*)malloc(N * sizeof(int32_t));
    int32_t* members = (int32_t*)malloc(N * sizeof(int32_t));
    int cur = 0;
    for (size_t s = 0; s < N; s++) {
        if (img[s] == newVal || label[s]) continue;
        cur++;
        int sp = 0, cnt = 0;
        stack[sp++] = (int32_t)s; label[s] = cur;
        while (sp) {
            int32_t p = stack[--sp];
            members[cnt++] = p;
            int y = p / W, x = p % W; int v = img[p];
            const int nx[4] = { x - 1, x + 1, x, x }, ny[4] = { y, y, y - 1, y + 1 };
            for (int k = 0; k < 4; k++) {
                if (nx[k] < 0 || nx[k] >= W || ny[k] < 0 || ny[k] >= H) continue;
                int32_t q = ny[k] * W + nx[k];
                if (label[q] || img[q] == newVal) continue;
                if (abs(v - img[q]) > maxDiff) continue;
                label[q] = cur; stack[sp++] = q;
            }
        }
        if (cnt <= maxSpeckleSize)
            for (int i = 0; i < cnt; i++) label[members[i]] = -1;  /* erase after the scan so values stay intact */
    }
    for (size_t s = 0; s < N; s++) if (label[s] == -1) img[s] = (int16_t)newVal;
    free(label); free(stack); free(members);
}

/* ------------------------------------------------------------------------- */
/* StereoSGBM::compute (MODE_SGBM / MODE_HH), SURVEY.md Appendix A.0-A.7.      */
/* left/right: 8-bit single channel.  disp: int16 W*H (x16 fixed point).       */
/* Optional stage outputs (may be NULL): C_out, S_out volumes [H][W1][D] int16 */
/* and raw_out (disparity before median/speckle).  stats[0] receives the       */
/* minimum path cost L seen, stats[1] the minimum C (order-independence guard).*/
int orc_sgbm_compute(const uint8_t* left, size_t stepL, const uint8_t* right, size_t stepR,
                     int W, int H, const orc_sgbm_params* prm,
                     int16_t* disp, int16_t* raw_out, int16_t* C_out, int16_t* S_out, int* stats)
{
    const int minD = prm->minDisparity, D = prm->numDisparities, maxD = minD + D;
    const int bs = prm->blockSize > 0 ? prm->blockSize : 5;
    const int SW2 = bs / 2, SH2 = bs / 2;
    const int uniq = prm->uniquenessRatio >= 0 ? prm->uniquenessRatio : 10;
    const int d12 = prm->disp12MaxDiff > 0 ? prm->disp12MaxDiff : 1;
    const int P1 = prm->P1 > 0 ? prm->P1 : 2;
    const int P2 = imax(prm->P2 > 0 ? prm->P2 : 5, P1 + 1);
    const int ftzero = imax(prm->preFilterCap, 15) | 1;
    const int minX1 = imax(maxD, 0), maxX1 = W + imin(minD, 0), W1 = maxX1 - minX1;
    const int INVALID = (minD - 1) * ORC_DISP_SCALE;
    const int npasses = prm->mode == 1 ? 2 : 1;
    int16_t* raw = (int16_t*)malloc((size_t)W * H * sizeof(int16_t));
    for (size_t i = 0; i < (size_t)W * H; i++) raw[i] = (int16_t)INVALID;
    int minLseen = 32767, minCseen = 32767;

    if (W1 > 0 && D > 0) {
        const size_t rowCells = (size_t)W1 * D, cells = rowCells * H;
        int16_t* pix = (int16_t*)malloc(cells * sizeof(int16_t));
        int16_t* hsum = (int16_t*)malloc(cells * sizeof(int16_t));
        int16_t* C = (int16_t*)malloc(cells * sizeof(int16_t));
        int16_t* S = (int16_t*)calloc(cells, sizeof(int16_t));

        /* A.2 per-row pixel cost: BT on clipped x-Sobel + BT on raw intensity >> 2 */
        int* gL = (int*)malloc(sizeof(int) * W * 4); int* gR = gL + W; int* rL = gR + W; int* rR = rL + W;
        for (int y = 0; y < H; y++) {
            int yn = imax(y - 1, 0), ys = imin(y + 1, H - 1);
            const uint8_t *l0 = left + (size_t)y * stepL, *ln = left + (size_t)yn * stepL, *ls = left + (size_t)ys * stepL;
            const uint8_t *r0 = right + (size_t)y * stepR, *rn = right + (size_t)yn * stepR, *rs = right + (size_t)ys * stepR;
            for (int x = 0; x < W; x++) { gL[x] = gR[x] = rL[x] = rR[x] = ftzero; }   /* borders: tab(0), also for the raw channel */
            for (int x = 1; x < W - 1; x++) {
                int a = 2 * (l0[x + 1] - l0[x - 1]) + (ln[x + 1] - ln[x - 1]) + (ls[x + 1] - ls[x - 1]);
                int b = 2 * (r0[x + 1] - r0[x - 1]) + (rn[x + 1] - rn[x - 1]) + (rs[x + 1] - rs[x - 1]);
                gL[x] = imin(imax(a, -ftzero), ftzero) + ftzero;
                gR[x] = imin(imax(b, -ftzero), ftzero) + ftzero;
                rL[x] = l0[x]; rR[x] = r0[x];
            }
            int16_t* prow = pix + (size_t)y * rowCells;
            for (int xp = 0; xp < W1; xp++) {
                int x = xp + minX1;
                for (int d = 0; d < D; d++) {
                    int xr = x - (minD + d);
                    int cost = 0;
                    for (int ch = 0; ch < 2; ch++) {
                        const int* a = ch ? rL : gL; const int* b = ch ? rR : gR;
                        int u = a[x];
                        int ul = x > 0 ? (u + a[x - 1]) / 2 : u, ur = x < W - 1 ? (u + a[x + 1]) / 2 : u;
                        int u0 = imin(imin(ul, ur), u), u1 = imax(imax(ul, ur), u);
                        int v = b[xr];
                        int vl = xr > 0 ? (v + b[xr - 1]) / 2 : v, vr = xr < W - 1 ? (v + b[xr + 1]) / 2 : v;
                        int v0 = imin(imin(vl, vr), v), v1 = imax(imax(vl, vr), v);
                        int c0 = imax(0, imax(u - v1, v0 - u)), c1 = imax(0, imax(v - u1, u0 - v));
                        cost += imin(c0, c1) >> (ch ? 2 : 0);
                    }
                    prow[(size_t)xp * D + d] = (int16_t)cost;
                }
            }
        }
        free(gL);

        /* A.3 block aggregation, saturating int16 recurrences in exactly this order */
        for (int y = 0; y < H; y++) {
            const int16_t* prow = pix + (size_t)y * rowCells; int16_t* hrow = hsum + (size_t)y * rowCells;
            for (int d = 0; d < D; d++) {
                int h = sat16(prow[d] * (SW2 + 1));
                for (int k = 1; k <= SW2; k++) h = sat16(h + prow[(size_t)imin(k, W1 - 1) * D + d]);
                hrow[d] = (int16_t)h;
                for (int xp = 1; xp < W1; xp++) {
                    h = sat16(sat16(h - prow[(size_t)imax(xp - SW2 - 1, 0) * D + d]) + prow[(size_t)imin(xp + SW2, W1 - 1) * D + d]);
                    hrow[(size_t)xp * D + d] = (int16_t)h;
                }
            }
        }
        for (size_t i = 0; i < rowCells; i++) {
            int c = (int16_t)P2;
            for (int k = 0; k <= SH2; k++) c = sat16(c + sat16(hsum[(size_t)imin(k, H - 1) * rowCells + i] * (k == 0 ? SH2 + 1 : 1)));
            C[i] = (int16_t)c;
            for (int y = 1; y < H; y++) {
                int ho = hsum[(size_t)imax(y - SH2 - 1, 0) * rowCells + i], hn = hsum[(size_t)imin(y + SH2, H - 1) * rowCells + i];
                /* OpenCV evaluates the first computed column (x' = 0) as (Cprev + new) - old and every other column as
                   (Cprev - old) + new; with saturating lanes the two differ once C sticks at 32767 (pinned by the
                   satdrift golden fixture). */
                c = (i < (size_t)D) ? sat16(sat16(c + hn) - ho) : sat16(sat16(c - ho) + hn);
                C[(size_t)y * rowCells + i] = (int16_t)c;
            }
        }
        for (size_t i = 0; i < cells; i++) if (C[i] < minCseen) minCseen = C[i];
        free(pix); free(hsum);

        /* A.4 path aggregation.  Lr[b][x'+1][r][d+1], d in [-1, D]; border columns x'=-1, x'=W1 stay 0. */
        const int D2 = D + 2; const int16_t P1s = (int16_t)P1;
        const size_t colStride = (size_t)4 * D2, rowStride = (size_t)(W1 + 2) * colStride;
        int16_t* Lr = (int16_t*)malloc(2 * rowStride * sizeof(int16_t));
        int16_t* minLr = (int16_t*)malloc(2 * (size_t)(W1 + 2) * 4 * sizeof(int16_t));
        int16_t* disp2 = (int16_t*)malloc(sizeof(int16_t) * W); int16_t* disp2cost = (int16_t*)malloc(sizeof(int16_t) * W);
#define LRP(b, xp, r) (Lr + (size_t)(b) * rowStride + (size_t)((xp) + 1) * colStride + (size_t)(r) * D2 + 1)
#define MINLR(b, xp, r) (minLr[(size_t)(b) * (W1 + 2) * 4 + (size_t)((xp) + 1) * 4 + (r)])
        for (int pass = 1; pass <= npasses; pass++) {
            int y1, y2, dy, x1, x2, dx;
            if (pass == 1) { y1 = 0; y2 = H; dy = 1; x1 = 0; x2 = W1; dx = 1; }
            else { y1 = H - 1; y2 = -1; dy = -1; x1 = W1 - 1; x2 = -1; dx = -1; }
            memset(Lr, 0, 2 * rowStride * sizeof(int16_t));
            memset(minLr, 0, 2 * (size_t)(W1 + 2) * 4 * sizeof(int16_t));
            int cur = 0;
            for (int y = y1; y != y2; y += dy, cur ^= 1) {
                const int prev = cur ^ 1;
                const int16_t* Crow = C + (size_t)y * rowCells; int16_t* Srow = S + (size_t)y * rowCells;
                for (int xp = x1; xp != x2; xp += dx) {
                    const int16_t* Cp = Crow + (size_t)xp * D; int16_t* Sp = Srow + (size_t)xp * D;
                    for (int r = 0; r < 4; r++) {
                        int sb = r == 0 ? cur : prev;
                        int sx = r == 0 ? xp - dx : (r == 1 ? xp - 1 : (r == 2 ? xp : xp + 1));
                        int16_t* Lp = LRP(sb, sx, r);
                        const int16_t delta = (int16_t)(P2 + MINLR(sb, sx, r));
                        Lp[-1] = Lp[D] = ORC_MAX_COST;
                        int16_t* Lc = LRP(cur, xp, r);
                        int mn = ORC_MAX_COST;
                        for (int d = 0; d < D; d++) {
                            int m = imin(imin(Lp[d], sat16(Lp[d - 1] + P1s)), imin(sat16(Lp[d + 1] + P1s), delta));
                            int L = sat16(sat16(m - delta) + Cp[d]);
                            Lc[d] = (int16_t)L; mn = imin(mn, L);
                            Sp[d] = sat16(Sp[d] + L);
                        }
                        MINLR(cur, xp, r) = (int16_t)mn;
                        if (mn < minLseen) minLseen = mn;
                    }
                }
                if (pass != npasses) continue;
                /* A.5 WTA (+ the 5th, right-to-left path for MODE_SGBM) */
                int16_t* drow = raw + (size_t)y * W;
                for (int x = 0; x < W; x++) { disp2[x] = (int16_t)INVALID; disp2cost[x] = ORC_MAX_COST; }
                for (int xp = W1 - 1; xp >= 0; xp--) {
                    int16_t* Sp = Srow + (size_t)xp * D;
                    if (npasses == 1) {
                        const int16_t* Cp = Crow + (size_t)xp * D;
                        int16_t* Lp = LRP(cur, xp + 1, 0);
                        const int16_t delta = (int16_t)(P2 + MINLR(cur, xp + 1, 0));
                        Lp[-1] = Lp[D] = ORC_MAX_COST;
                        int16_t* Lc = LRP(cur, xp, 0);
                        int mn = ORC_MAX_COST;
                        for (int d = 0; d < D; d++) {
                            int m = imin(imin(Lp[d], sat16(Lp[d - 1] + P1s)), imin(sat16(Lp[d + 1] + P1s), delta));
                            int L = sat16(sat16(m - delta) + Cp[d]);
                            Lc[d] = (int16_t)L; mn = imin(mn, L);
                            Sp[d] = sat16(Sp[d] + L);
                        }
                        MINLR(cur, xp, 0) = (int16_t)mn;
                        if (mn < minLseen) minLseen = mn;
                    }
                    int minS = ORC_MAX_COST, best = -1;
                    for (int d = 0; d < D; d++) if (Sp[d] < minS) { minS = Sp[d]; best = d; }
                    int d;
                    for (d = 0; d < D; d++) if (Sp[d] * (100 - uniq) < minS * 100 && abs(best - d) > 1) break;
                    if (d < D) continue;
                    if (best < 0) continue;   /* equivalent to letting -1 through */
                    int x2i = xp + minX1 - best - minD;
                    if (disp2cost[x2i] > minS) { disp2cost[x2i] = (int16_t)minS; disp2[x2i] = (int16_t)(best + minD); }
                    int dd;
                    if (0 < best && best < D - 1) {
                        int den = imax(Sp[best - 1] + Sp[best + 1] - 2 * Sp[best], 1);
                        dd = best * ORC_DISP_SCALE + ((Sp[best - 1] - Sp[best + 1]) * ORC_DISP_SCALE + den) / (den * 2);
                    } else dd = best * ORC_DISP_SCALE;
                    drow[xp + minX1] = (int16_t)(dd + minD * ORC_DISP_SCALE);
                }
                for (int x = minX1; x < maxX1; x++) {
                    int d1 = drow[x];
                    if (d1 == INVALID) continue;
                    int _d = d1 >> ORC_DISP_SHIFT, d_ = (d1 + ORC_DISP_SCALE - 1) >> ORC_DISP_SHIFT;
                    int _x = x - _d, x_ = x - d_;
                    if (0 <= _x && _x < W && disp2[_x] >= minD && abs(disp2[_x] - _d) > d12 &&
                        0 <= x_ && x_ < W && disp2[x_] >= minD && abs(disp2[x_] - d_) > d12)
                        drow[x] = (int16_t)INVALID;
                }
            }
        }
#undef LRP
#undef MINLR
        if (C_out) memcpy(C_out, C, cells * sizeof(int16_t));
        if (S_out) memcpy(S_out, S, cells * sizeof(int16_t));
        free(Lr); free(minLr); free(disp2); free(disp2cost); free(C); free(S);
    }
    if (raw_out) memcpy(raw_out, raw, (size_t)W * H * sizeof(int16_t));
    /* A.6 post filters */
    orc_median3_s16(raw, disp, W, H);
    if (prm->speckleWindowSize > 0)
        orc_filter_speckles(disp, W, H, INVALID, prm->speckleWindowSize, ORC_DISP_SCALE * prm->speckleRange);
    if (stats) { stats[0] = minLseen; stats[1] = minCseen; }
    free(raw);
    return 0;
}

/* ------------------------------------------------------------------------- */
/* TrackingFrame::SetupFrame / PruneDisparityImage (TrackingFrame.cpp:25-67),  */
/* SURVEY A.8.  disp is modified in place (pruned pixels -> 0).               */
/* out_stats (optional, 4 doubles): mn, mx, sigma, ithresh.                   */
void orc_prune(int16_t* disp, int W, int H, float* dispf, uint8_t* mask, uint8_t* norm_u8_out, double* out_stats)
{
    size_t N = (size_t)W * H;
    int mn = 32767, mx = -32768;
    for (size_t i = 0; i < N; i++) { if (disp[i] < mn) mn = disp[i]; if (disp[i] > mx) mx = disp[i]; }
    /* cv::normalize(NORM_MINMAX, 0..255) -> convertTo(CV_8U, scale, shift) */
    double dmn = mn, dmx = mx;
    double scale = (255.0 - 0.0) * (dmx - dmn > DBL_EPSILON ? 1.0 / (dmx - dmn) : 0.0);
    double shift = 0.0 - dmn * scale;
    float a = (float)scale, b = (float)shift;
    uint8_t* u8 = (uint8_t*)malloc(N);
    uint64_t sum = 0, sq = 0;
    for (size_t i = 0; i < N; i++) {
        float f = fmaf((float)disp[i], a, b);
        long r = lrintf(f);                       /* round-half-even under the default rounding mode */
        uint8_t v = (uint8_t)(r < 0 ? 0 : (r > 255 ? 255 : r));
        u8[i] = v; sum += v; sq += (uint64_t)v * v;
    }
    double inv = 1.0 / (double)N;                 /* cv::meanStdDev: scale = 1./total; mean = s*scale; var = sq*scale - mean^2 */
    double mean = (double)sum * inv;
    double var = (double)sq * inv - mean * mean;
    double sigma = sqrt(var > 0 ? var : 0);
    double thr = sigma * 1.3;
    int ithr = (int)floor(thr);
    for (size_t i = 0; i < N; i++) {
        uint8_t m = u8[i] > ithr ? 255 : 0;
        if (!m) disp[i] = 0;
        float f = (float)disp[i] * (1.0f / 16.0f);
        if (f <= 0.0f) m = 0;
        if (dispf) dispf[i] = f;
        if (mask) mask[i] = m;
    }
    if (norm_u8_out) memcpy(norm_u8_out, u8, N);
    if (out_stats) { out_stats[0] = mn; out_stats[1] = mx; out_stats[2] = sigma; out_stats[3] = ithr; }
    free(u8);
}

/* ------------------------------------------------------------------------- */
/* cv::reprojectImageTo3D(disp16S, Q, handleMissingValues=true), SURVEY A.9.   */
/* xyz: float [H][W][3].                                                      */
void orc_reproject(const int16_t* disp, int W, int H, const float* Q /*4x4 row-major f32*/, int handleMissing, float* xyz)
{
    double q[16]; for (int i = 0; i < 16; i++) q[i] = (double)Q[i];
    double minDisp = FLT_MAX;
    if (handleMissing) { int mn = 32767; for (size_t i = 0; i < (size_t)W * H; i++) if (disp[i] < mn) mn = disp[i]; minDisp = (double)mn; }
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            double d = (double)(float)disp[(size_t)y * W + x];
            double h[4];
            for (int i = 0; i < 4; i++) h[i] = q[i * 4 + 0] * x + q[i * 4 + 1] * y + q[i * 4 + 2] * d + q[i * 4 + 3];
            float p0 = (float)h[0], p1 = (float)h[1], p2 = (float)h[2];
            double iw = 1.0 / h[3];
            p0 = (float)((double)p0 * iw); p1 = (float)((double)p1 * iw); p2 = (float)((double)p2 * iw);
            if (fabs(d - minDisp) <= FLT_EPSILON) p2 = 10000.0f;
            float* o = xyz + ((size_t)y * W + x) * 3; o[0] = p0; o[1] = p1; o[2] = p2;
        }
}

/* 16-byte packed output point: x,y,z f32 + b,g,r,a bytes (a=255).             */
typedef struct orc_point { float x, y, z; uint8_t b, g, r, a; } orc_point;

/* Reconstruct3D::GeneratePointCloud (Reconstruct3D.cpp:106-153): reproject,   */
/* skip Z == 10000, recentre, flip Y, then the diag(-1,1,1,1) transform.       */
int orc_generate_point_cloud(const int16_t* disp, const uint8_t* bgr, size_t bstep, int W, int H, const float* Q, orc_point* out)
{
    float* xyz = (float*)malloc((size_t)W * H * 3 * sizeof(float));
    orc_reproject(disp, W, H, Q, 1, xyz);
    int n = 0;
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            const float* c = xyz + ((size_t)y * W + x) * 3;
            if (c[2] == 10000.0f) continue;
            orc_point p;
            float px = c[0] - (W / 2.0f);
            p.y = -(c[1] - (H / 2.0f));
            p.z = c[2];
            /* pcl::transformPointCloud with P = diag(-1,1,1,1): x' = -1*x + 0*y + 0*z + 0 (f32) */
            p.x = -px;
            const uint8_t* s = bgr + (size_t)y * bstep + (size_t)x * 3;
            p.b = s[0]; p.g = s[1]; p.r = s[2]; p.a = 255;
            out[n++] = p;
        }
    free(xyz);
    return n;
}

/* Reconstruct3D::Triangulate3D (Reconstruct3D.cpp:156-216).                   */
int orc_triangulate(const float* dispf, const uint8_t* bgr, size_t bstep, const uint8_t* mask, int W, int H,
                    float f, float b, float cx, float cy, orc_point* out)
{
    int n = 0;
    for (int i = 0; i < H; i++)
        for (int j = 0; j < W; j++) {
            float d = dispf[(size_t)i * W + j];
            if (d <= 0.0) continue;
            if (mask && mask[(size_t)i * W + j] == 0) continue;
            orc_point p;
            p.z = f * b / d;
            p.x = ((float)j - cx) * (p.z / f);
            p.y = -((float)i - cy) * (p.z / f);
            const uint8_t* s = bgr + (size_t)i * bstep + (size_t)j * 3;
            p.b = s[0]; p.g = s[1]; p.r = s[2]; p.a = 255;
            out[n++] = p;
        }
    return n;
}

/* ------------------------------------------------------------------------- */
/* pcl::VoxelGrid<PointXYZRGB>::applyFilter.             */
/* Returns number of output points (ascending voxel index), or -1 when the     */
/* index would overflow int32 (PCL warns and copies the input through; the     */
/* caller gets the input copied to out and *passthrough = 1).                  */
typedef struct { int32_t idx; int32_t pt; } orc_vkey;
static int vkey_cmp(const void* a, const void* b)
{
    const orc_vkey* p = (const orc_vkey*)a; const orc_vkey* q = (const orc_vkey*)b;
    if (p->idx != q->idx) return p->idx < q->idx ? -1 : 1;
    return p->pt < q->pt ? -1 : (p->pt > q->pt);      /* make the (unstable in PCL) order deterministic: by point index */
}
int orc_voxel_grid(const orc_point* in, int n, float lx, float ly, float lz, orc_point* out, int* counts_out, int* passthrough)
{
    if (passthrough) *passthrough = 0;
    if (n <= 0) return 0;
    float mnx = FLT_MAX, mny = FLT_MAX, mnz = FLT_MAX, mxx = -FLT_MAX, mxy = -FLT_MAX, mxz = -FLT_MAX;
    for (int i = 0; i < n; i++) {
        if (in[i].x < mnx) mnx = in[i].x; if (in[i].x > mxx) mxx = in[i].x;
        if (in[i].y < mny) mny = in[i].y; if (in[i].y > mxy) mxy = in[i].y;
        if (in[i].z < mnz) mnz = in[i].z; if (in[i].z > mxz) mxz = in[i].z;
    }
    float ix = 1.0f / lx, iy = 1.0f / ly, iz = 1.0f / lz;
    int64_t dx = (int64_t)((mxx - mnx) * ix) + 1, dy = (int64_t)((mxy - mny) * iy) + 1, dz = (int64_t)((mxz - mnz) * iz) + 1;
    if (dx * dy * dz > (int64_t)INT32_MAX) {
        memcpy(out, in, (size_t)n * sizeof(orc_point));
        if (passthrough) *passthrough = 1;
        return n;
    }
    int mbx = (int)floorf(mnx * ix), mby = (int)floorf(mny * iy), mbz = (int)floorf(mnz * iz);
    int Mbx = (int)floorf(mxx * ix), Mby = (int)floorf(mxy * iy);
    int divx = Mbx - mbx + 1, divy = Mby - mby + 1;
    int mul1 = divx, mul2 = divx * divy;
    orc_vkey* keys = (orc_vkey*)malloc((size_t)n * sizeof(orc_vkey));
    for (int i = 0; i < n; i++) {
        int ijk0 = (int)floorf(in[i].x * ix) - mbx, ijk1 = (int)floorf(in[i].y * iy) - mby, ijk2 = (int)floorf(in[i].z * iz) - mbz;
        keys[i].idx = ijk0 + ijk1 * mul1 + ijk2 * mul2; keys[i].pt = i;
    }
    qsort(keys, (size_t)n, sizeof(orc_vkey), vkey_cmp);
    int m = 0;
    for (int s = 0; s < n;) {
        int e = s; float sx = 0, sy = 0, sz = 0, sr = 0, sg = 0, sb = 0, sa = 0;
        while (e < n && keys[e].idx == keys[s].idx) {
            const orc_point* p = in + keys[e].pt;
            sx += p->x; sy += p->y; sz += p->z; sr += (float)p->r; sg += (float)p->g; sb += (float)p->b; sa += (float)p->a;
            e++;
        }
        float cnt = (float)(e - s);
        orc_point o; o.x = sx / cnt; o.y = sy / cnt; o.z = sz / cnt;
        o.r = (uint8_t)(sr / cnt); o.g = (uint8_t)(sg / cnt); o.b = (uint8_t)(sb / cnt); o.a = (uint8_t)(sa / cnt);
        if (counts_out) counts_out[m] = e - s;
        out[m++] = o; s = e;
    }
    free(keys);
    return m;
}

/* ------------------------------------------------------------------------- */
/* pcl::StatisticalOutlierRemoval, exact brute-force kNN. */
/* dist_out (optional, n floats): mean distance to the k nearest neighbours.   */
/* Returns kept count; out keeps input order.  thr_out: the mu + alpha*sigma.   */
int orc_sor(const orc_point* in, int n, int k, double alpha, orc_point* out, float* dist_out, double* thr_out, uint8_t* keep_out)
{
    if (n <= 0) return 0;
    float* dist = (float*)calloc((size_t)n, sizeof(float));
    int kk = k + 1;                                   /* the query itself is among the results */
    float* best = (float*)malloc(sizeof(float) * (size_t)kk);
    int valid = 0;
    for (int i = 0; i < n; i++) {
        if (n < kk) { dist[i] = 0.0f; continue; }     /* nearestKSearch cannot return k+1 results */
        int cnt = 0;
        for (int j = 0; j < n; j++) {
            float dx = in[i].x - in[j].x, dy = in[i].y - in[j].y, dz = in[i].z - in[j].z;
            float d2 = dx * dx; d2 += dy * dy; d2 += dz * dz;
            if (cnt < kk) { int p = cnt++; while (p > 0 && best[p - 1] > d2) { best[p] = best[p - 1]; p--; } best[p] = d2; }
            else if (d2 < best[kk - 1]) { int p = kk - 1; while (p > 0 && best[p - 1] > d2) { best[p] = best[p - 1]; p--; } best[p] = d2; }
        }
        double s = 0; for (int j = 1; j < kk; j++) s += (double)sqrtf(best[j]);   /* skip j = 0: the nearest (itself / a duplicate) */
        dist[i] = (float)(s / (double)k);
        valid++;
    }
    double sum = 0, sq = 0;
    for (int i = 0; i < n; i++) { sum += dist[i]; float pr = dist[i] * dist[i]; sq += pr; }   /* PCL: float product, double accumulation */
    double mean = sum / (double)valid;
    double var = (sq - sum * sum / (double)valid) / ((double)valid - 1.0);
    double sd = sqrt(var);
    double thr = mean + alpha * sd;
    int m = 0;
    for (int i = 0; i < n; i++) {
        int keep = !((double)dist[i] > thr);
        if (keep_out) keep_out[i] = (uint8_t)keep;
        if (keep) out[m++] = in[i];
    }
    if (dist_out) memcpy(dist_out, dist, (size_t)n * sizeof(float));
    if (thr_out) *thr_out = thr;
    free(dist); free(best);
    return m;
}

/* ------------------------------------------------------------------------- */
/* cv::StereoBM::compute (PREFILTER_XSOBEL), the other branch of the reference's  */
/* block-matcher switch: Reconstruct3D.cpp:43-48 (StereoBM::create(16, 21) +      */
/* setBlockSize(window_size) + setNumDisparities), SURVEY 8(f) row n2.            */
/* Restated from OpenCV's published algorithm (modules/calib3d/src/stereobm.cpp:  */
/* prefilterXSobel, findStereoCorrespondenceBM, getValidDisparityROI); pinned     */
/* against cv2.StereoBM by tests/golden (bm_*.npz).                              */
typedef struct orc_bm_params {
    int minDisparity, numDisparities, blockSize, preFilterCap, textureThreshold, uniquenessRatio, speckleWindowSize, speckleRange;
} orc_bm_params;

static void orc_prefilter_xsobel(const uint8_t* src, size_t step, int W, int H, int cap, uint8_t* dst)
{
    /* rows are processed in pairs; a trailing odd row is filled with the neutral value `cap` */
    int y;
    for (y = 0; y < H - 1; y += 2) {
        const uint8_t* s1 = src + (size_t)y * step;
        const uint8_t* s0 = y > 0 ? s1 - step : (H > 1 ? s1 + step : s1);
        const uint8_t* s2 = y < H - 1 ? s1 + step : (H > 1 ? s1 - step : s1);
        const uint8_t* s3 = y < H - 2 ? s1 + 2 * step : s1;
        uint8_t* d0 = dst + (size_t)y * W; uint8_t* d1 = d0 + W;
        d0[0] = d0[W - 1] = d1[0] = d1[W - 1] = (uint8_t)cap;
        for (int x = 1; x < W - 1; x++) {
            int e0 = s0[x + 1] - s0[x - 1], e1 = s1[x + 1] - s1[x - 1], e2 = s2[x + 1] - s2[x - 1], e3 = s3[x + 1] - s3[x - 1];
            int v0 = e0 + e1 * 2 + e2, v1 = e1 + e2 * 2 + e3;
            d0[x] = (uint8_t)(imin(imax(v0, -cap), cap) + cap);
            d1[x] = (uint8_t)(imin(imax(v1, -cap), cap) + cap);
        }
    }
    for (; y < H; y++) memset(dst + (size_t)y * W, cap, (size_t)W);
}

int orc_stereobm_compute(const uint8_t* left, size_t stepL, const uint8_t* right, size_t stepR, int W, int H,
                         const orc_bm_params* p, int16_t* disp)
{
    const int ndisp = p->numDisparities, mindisp = p->minDisparity, wsz = p->blockSize, wsz2 = wsz / 2;
    const int cap = p->preFilterCap, FILTERED = (mindisp - 1) * 16;
    const int lofs = imax(ndisp - 1 + mindisp, 0), rofs = -imin(ndisp - 1 + mindisp, 0), width1 = W - rofs - ndisp + 1;
    for (size_t i = 0; i < (size_t)W * H; i++) disp[i] = (int16_t)FILTERED;
    if (lofs >= W || rofs >= W || width1 < 1) return 0;
    uint8_t* L = (uint8_t*)malloc((size_t)W * H); uint8_t* R = (uint8_t*)malloc((size_t)W * H);
    orc_prefilter_xsobel(left, stepL, W, H, cap, L);
    orc_prefilter_xsobel(right, stepR, W, H, cap, R);
    /* column SAD and texture sums with replicated borders, then the window sums -- direct (non-incremental) evaluation of
       what findStereoCorrespondenceBM maintains incrementally */
    int* sad = (int*)malloc(sizeof(int) * (size_t)(ndisp + 2));
    int32_t* hsad = (int32_t*)malloc(sizeof(int32_t) * (size_t)H * ndisp);
    int32_t* htext = (int32_t*)malloc(sizeof(int32_t) * (size_t)H);
    /* for minDisparity > 0 OpenCV's loop runs to lofs + x >= W (its stores spill into the masked left border of the next
       row); those columns do not exist, so the restatement stops at the image edge */
    for (int x = 0; x < width1 && lofs + x < W; x++) {
        /* hsad[y][d] = sum over window columns of |L(y, xl) - R(y, xr + d)|, xl = lofs + clamp(xx, -lofs, W-1-lofs),
           xr = rofs + clamp(xx, -rofs, W-ndisp-rofs); array index d means disparity ndisp-1-d+mindisp */
        for (int y = 0; y < H; y++) {
            int32_t* hs = hsad + (size_t)y * ndisp;
            memset(hs, 0, sizeof(int32_t) * (size_t)ndisp);
            int t = 0;
            for (int xx = x - wsz2; xx <= x + wsz2; xx++) {
                int xl = lofs + imin(imax(xx, -lofs), W - 1 - lofs);
                int xr = rofs + imin(imax(xx, -rofs), W - ndisp - rofs);
                int lval = L[(size_t)y * W + xl];
                const uint8_t* rp = R + (size_t)y * W + xr;
                for (int d = 0; d < ndisp; d++) hs[d] += abs(lval - rp[d]);
                t += abs(lval - cap);
            }
            htext[y] = t;
        }
        for (int y = 0; y < H; y++) {
            int tsum = 0;
            for (int d = 0; d < ndisp; d++) sad[d + 1] = 0;
            for (int yy = y - wsz2; yy <= y + wsz2; yy++) {
                int yc = imin(imax(yy, 0), H - 1);
                const int32_t* hs = hsad + (size_t)yc * ndisp;
                for (int d = 0; d < ndisp; d++) sad[d + 1] += hs[d];
                tsum += htext[yc];
            }
            int16_t* dp = disp + (size_t)y * W + lofs + x;
            int minsad = 0x7fffffff, mind = -1;
            for (int d = 0; d < ndisp; d++) if (sad[d + 1] < minsad) { minsad = sad[d + 1]; mind = d; }
            if (tsum < p->textureThreshold) { *dp = (int16_t)FILTERED; continue; }
            if (p->uniquenessRatio > 0) {
                int thresh = minsad + (minsad * p->uniquenessRatio / 100), d;
                for (d = 0; d < ndisp; d++) if ((d < mind - 1 || d > mind + 1) && sad[d + 1] <= thresh) break;
                if (d < ndisp) { *dp = (int16_t)FILTERED; continue; }
            }
            sad[0] = sad[2]; sad[ndisp + 1] = sad[ndisp - 1];
            int pp = sad[mind + 2], nn = sad[mind], dd = pp + nn - 2 * sad[mind + 1] + abs(pp - nn);
            *dp = (int16_t)((((ndisp - mind - 1 + mindisp) * 256 + (dd != 0 ? (pp - nn) * 256 / dd : 0)) + 15) >> 4);
        }
    }
    free(sad); free(hsad); free(htext); free(L); free(R);
    /* validDisparityRect (getValidDisparityROI with both ROIs = whole image): everything outside is FILTERED */
    {
        const int maxD = mindisp + ndisp - 1;
        const int xmin = imax(0, maxD) + wsz2, xmax = W - wsz2, ymin = wsz2, ymax = H - wsz2;
        const int ok = (xmax - xmin > 0) && (ymax - ymin > 0);
        for (int y = 0; y < H; y++)
            for (int x = 0; x < W; x++)
                if (!ok || x < xmin || x >= xmax || y < ymin || y >= ymax) disp[(size_t)y * W + x] = (int16_t)FILTERED;
    }
    if (p->speckleRange >= 0 && p->speckleWindowSize > 0)
        orc_filter_speckles(disp, W, H, FILTERED, p->speckleWindowSize, p->speckleRange);
    return 0;
}

/* ------------------------------------------------------------------------- */
/* the steps right after the path in MappingSystem.           */
/* pcl::RadiusOutlierRemoval (MappingSystem.cpp:303-307: radius 1.0, 300          */
/* neighbours; .filter commented in the reference): for dense clouds PCL asks for  */
/* the min_pts+1 nearest neighbours (the query included) and keeps the point iff   */
/* the farthest of them is within the radius (squared distance <= r^2).            */
int orc_radius_outlier_removal(const orc_point* in, int n, float radius, int min_pts, orc_point* out, uint8_t* keep_out)
{
    const float r2 = radius * radius;
    int m = 0;
    for (int i = 0; i < n; i++) {
        int cnt = 0;
        for (int j = 0; j < n && cnt <= min_pts; j++) {
            float dx = in[i].x - in[j].x, dy = in[i].y - in[j].y, dz = in[i].z - in[j].z;
            float d2 = dx * dx; d2 += dy * dy; d2 += dz * dz;
            if (d2 <= r2) cnt++;
        }
        int keep = cnt > min_pts;                      /* min_pts + 1 points (itself included) lie within the radius */
        if (keep_out) keep_out[i] = (uint8_t)keep;
        if (keep) out[m++] = in[i];
    }
    return m;
}

/* pcl::transformPointCloud(cloud, cloud, T) (MappingSystem.cpp:318; Reconstruct3D.cpp:149-150): p' = T * (x, y, z, 1),   */
/* evaluated as Eigen does for a 4x4 float matrix: ((col0*x + col1*y) + col2*z) + col3, no FMA (no -mfma in the build). */
void orc_transform_cloud(const orc_point* in, int n, const float* T /*4x4 row-major*/, orc_point* out)
{
    for (int i = 0; i < n; i++) {
        orc_point p = in[i];
        float v[3];
        for (int r = 0; r < 3; r++) {
            float a = T[r * 4 + 0] * in[i].x; float b = T[r * 4 + 1] * in[i].y; float s = a + b;
            float c = T[r * 4 + 2] * in[i].z; s = s + c; s = s + T[r * 4 + 3];
            v[r] = s;
        }
        p.x = v[0]; p.y = v[1]; p.z = v[2];
        out[i] = p;
    }
}

/* ------------------------------------------------------------------------------------------------
 * Stereo rectification of the input images: Reconstruct3D::RectifyImages
 * (cv_reconstruct/src/reconstruct/Reconstruct3D.cpp:330-361) =
 *   cv::initUndistortRectifyMap(K, D, R, P, size, CV_16SC2, map1, map2)  for each camera, then
 *   cv::remap(image, rectified, map1, map2, cv::INTER_LINEAR)            (BORDER_CONSTANT, value 0).
 * OpenCV is a third-party dependency of the reference (4.x; pinned here against cv2 4.13.0 by
 * tests/golden/rect_*.npz).  Restated from its published algorithm:
 *   - iR = inverse(P[:, :3] * R) by the 3x3 closed form (adjugate times 1/det), all in double;
 *   - per row i the homogeneous coordinate starts at (i*ir1 + ir2, i*ir4 + ir5, i*ir7 + ir8) and is advanced by
 *     REPEATED ADDITION of (ir0, ir3, ir6) per column (the running sums are part of the result's rounding);
 *   - radial/tangential/thin-prism distortion with up to 12 coefficients (k1 k2 p1 p2 k3 k4 k5 k6 s1 s2 s3 s4);
 *     the tilt model (tauX, tauY) of 14-coefficient vectors is not restated (identity), callers must pass <= 12;
 *   - fixed point: iu = round_half_even(u * 32), map1 = (iu >> 5, iv >> 5) as int16, map2 = (iv & 31) * 32 + (iu & 31).
 * remap, 8-bit, bilinear: weights w = {(32-fx)(32-fy), fx(32-fy), (32-fx)fy, fx*fy} * 32 (sum 2^15; the float table
 * OpenCV builds is exact for these values), result = (sum w_i * S_i + 2^14) >> 15; a tap outside the image reads 0.
 */
static void orc_inv3(const double* S, double* t)
{
    double d = S[0] * (S[4] * S[8] - S[5] * S[7]) - S[1] * (S[3] * S[8] - S[5] * S[6]) + S[2] * (S[3] * S[7] - S[4] * S[6]);
    if (d != 0.0) d = 1. / d;
    t[0] = (S[4] * S[8] - S[5] * S[7]) * d;
    t[1] = (S[2] * S[7] - S[1] * S[8]) * d;
    t[2] = (S[1] * S[5] - S[2] * S[4]) * d;
    t[3] = (S[5] * S[6] - S[3] * S[8]) * d;
    t[4] = (S[0] * S[8] - S[2] * S[6]) * d;
    t[5] = (S[2] * S[3] - S[0] * S[5]) * d;
    t[6] = (S[3] * S[7] - S[4] * S[6]) * d;
    t[7] = (S[1] * S[6] - S[0] * S[7]) * d;
    t[8] = (S[0] * S[4] - S[1] * S[3]) * d;
}

/* the nine doubles the per-pixel loop needs: inverse(P3x3 * R).  P has row pitch pstride (3 or 4). */
void orc_rectify_inverse(const double* R, const double* P, int pstride, double* ir)
{
    double M[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) M[i * 3 + j] = P[i * pstride + 0] * R[0 * 3 + j] + P[i * pstride + 1] * R[1 * 3 + j] + P[i * pstride + 2] * R[2 * 3 + j];
    orc_inv3(M, ir);
}

static inline int orc_round_half_even(double v) { return (int)lrint(v); }

void orc_init_undistort_rectify_map(const double* K, const double* D, int nD, const double* R, const double* P, int pstride, int W, int H,
                                    int16_t* map1 /* H*W*2 */, uint16_t* map2 /* H*W */)
{
    double ir[9], k[12] = { 0 };
    for (int i = 0; i < nD && i < 12; i++) k[i] = D[i];
    orc_rectify_inverse(R, P, pstride, ir);
    const double u0 = K[2], v0 = K[5], fx = K[0], fy = K[4];
    const double k1 = k[0], k2 = k[1], p1 = k[2], p2 = k[3], k3 = k[4], k4 = k[5], k5 = k[6], k6 = k[7], s1 = k[8], s2 = k[9], s3 = k[10], s4 = k[11];
    for (int i = 0; i < H; i++) {
        double _x = i * ir[1] + ir[2], _y = i * ir[4] + ir[5], _w = i * ir[7] + ir[8];
        for (int j = 0; j < W; j++, _x += ir[0], _y += ir[3], _w += ir[6]) {
            double w = 1. / _w, x = _x * w, y = _y * w;
            double x2 = x * x, y2 = y * y;
            double r2 = x2 + y2, _2xy = 2 * x * y;
            double kr = (1 + ((k3 * r2 + k2) * r2 + k1) * r2) / (1 + ((k6 * r2 + k5) * r2 + k4) * r2);
            double xd = (x * kr + p1 * _2xy + p2 * (r2 + 2 * x2) + s1 * r2 + s2 * r2 * r2);
            double yd = (y * kr + p1 * (r2 + 2 * y2) + p2 * _2xy + s3 * r2 + s4 * r2 * r2);
            double u = fx * xd + u0;                    /* identity tilt: vecTilt = (xd, yd, 1), invProj = 1 */
            double v = fy * yd + v0;
            int iu = orc_round_half_even(u * 32), iv = orc_round_half_even(v * 32);
            map1[((size_t)i * W + j) * 2 + 0] = (int16_t)(iu >> 5);
            map1[((size_t)i * W + j) * 2 + 1] = (int16_t)(iv >> 5);
            map2[(size_t)i * W + j] = (uint16_t)((iv & 31) * 32 + (iu & 31));
        }
    }
}

void orc_remap_bilinear_u8(const uint8_t* src, size_t sstep, int sw, int sh, int cn, const int16_t* map1, const uint16_t* map2, int W, int H,
                           uint8_t* dst, size_t dstep)
{
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            const int sx = map1[((size_t)y * W + x) * 2], sy = map1[((size_t)y * W + x) * 2 + 1];
            const int f = map2[(size_t)y * W + x] & 1023, fx = f & 31, fy = f >> 5;
            const int w[4] = { (32 - fx) * (32 - fy) * 32, fx * (32 - fy) * 32, (32 - fx) * fy * 32, fx * fy * 32 };
            for (int c = 0; c < cn; c++) {
                int acc = 0;
                for (int t = 0; t < 4; t++) {
                    const int xx = sx + (t & 1), yy = sy + (t >> 1);
                    const int s = (xx >= 0 && xx < sw && yy >= 0 && yy < sh) ? src[(size_t)yy * sstep + (size_t)xx * cn + c] : 0;
                    acc += s * w[t];
                }
                dst[(size_t)y * dstep + (size_t)x * cn + c] = (uint8_t)((acc + (1 << 14)) >> 15);
            }
        }
}

/* ---------------------------------------------------------------------------------------------------------------------
 * pcl::IterativeClosestPoint<PointXYZRGB, PointXYZRGB>::align as MappingSystem.cpp:321-341 (and the earlier
 * copy :104-127) configures it: setMaxCorrespondenceDistance(5), setMaximumIterations(50), setTransformationEpsilon(1e-8),
 * setEuclideanFitnessEpsilon(1); every other setting at PCL's default (point-to-point, TransformationEstimationSVD, no
 * reciprocal correspondences, no rejectors, identity guess).
 * Restated from PCL's registration/impl/icp.hpp, correspondence_estimation.hpp, transformation_estimation_svd.hpp and
 * default_convergence_criteria.hpp FROM MEMORY -- PCL is absent here and the reference calls are commented out / hold no
 * vectors => PARITY UNPINNED (tests cross-check this file against numpy + scipy.cKDTree instead).
 *   per iteration: nearest target point of every (already transformed) source point, kept iff d^2 <= max_dist^2 (f32 L2);
 *   fewer than 3 correspondences => stop, not converged; Umeyama / SVD without scaling on the correspondences; the source is
 *   transformed by the increment (f32, like pcl::transformPointCloud); final = increment * final; convergence:
 *     iterations >= max                                            -> converged (ITERATIONS)
 *     cos = (trace(R) - 1) / 2 >= 1 - transformation_epsilon  and  |t|^2 <= transformation_epsilon   -> converged (TRANSFORM)
 *     mse = mean d^2 of the correspondences:  |mse - prev| < 1e-12                   -> converged (ABS_MSE)
 *                                             |mse - prev| / prev < fitness_epsilon  -> converged (REL_MSE)
 *   (prev starts at DBL_MAX, so the relative test first fires in the second iteration: with the reference's epsilon of 1 the
 *   alignment ends after two iterations unless the error more than doubled.)
 * Arithmetic here: distances and point transforms in f32 without FMA, sums / covariance / SVD in f64, the increment rounded to
 * f32 before it is applied. */
static void orc_svd3(double A[3][3], double U[3][3], double S[3], double V[3][3])
{
    /* one-sided Jacobi (Hestenes): rotate column pairs of A until they are orthogonal; A = U diag(S) V^T */
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) V[i][j] = i == j;
    for (int sweep = 0; sweep < 60; sweep++) {
        double off = 0;
        for (int p = 0; p < 2; p++) for (int q = p + 1; q < 3; q++) {
            double al = 0, be = 0, ga = 0;
            for (int i = 0; i < 3; i++) { al += A[i][p] * A[i][p]; be += A[i][q] * A[i][q]; ga += A[i][p] * A[i][q]; }
            if (fabs(ga) <= 1e-300 || fabs(ga) <= 1e-17 * sqrt(al * be)) continue;
            off += fabs(ga);
            const double zeta = (be - al) / (2.0 * ga);
            const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
            const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
            for (int i = 0; i < 3; i++) {
                const double ap = A[i][p], aq = A[i][q];
                A[i][p] = c * ap - s * aq; A[i][q] = s * ap + c * aq;
                const double vp = V[i][p], vq = V[i][q];
                V[i][p] = c * vp - s * vq; V[i][q] = s * vp + c * vq;
            }
        }
        if (off == 0) break;
    }
    for (int j = 0; j < 3; j++) { S[j] = sqrt(A[0][j] * A[0][j] + A[1][j] * A[1][j] + A[2][j] * A[2][j]); }
    /* columns of U; a vanishing singular value gets the cross product of the other two (or any completion) */
    int big[3] = { 0, 1, 2 };
    for (int a = 0; a < 3; a++) for (int b2 = a + 1; b2 < 3; b2++) if (S[big[b2]] > S[big[a]]) { int t = big[a]; big[a] = big[b2]; big[b2] = t; }
    const double tiny = 1e-12 * (S[big[0]] > 0 ? S[big[0]] : 1.0);
    for (int r = 0; r < 3; r++) {
        const int j = big[r];
        if (S[j] > tiny) { for (int i = 0; i < 3; i++) U[i][j] = A[i][j] / S[j]; continue; }
        double u[3];
        if (r == 2) {       /* two good columns: complete with their cross product */
            const int a = big[0], b2 = big[1];
            u[0] = U[1][a] * U[2][b2] - U[2][a] * U[1][b2]; u[1] = U[2][a] * U[0][b2] - U[0][a] * U[2][b2]; u[2] = U[0][a] * U[1][b2] - U[1][a] * U[0][b2];
        } else if (r == 1) { /* one good column: any unit vector orthogonal to it */
            const int a = big[0];
            const int m = fabs(U[0][a]) < fabs(U[1][a]) ? (fabs(U[0][a]) < fabs(U[2][a]) ? 0 : 2) : (fabs(U[1][a]) < fabs(U[2][a]) ? 1 : 2);
            double e[3] = { 0, 0, 0 }; e[m] = 1;
            const double d = U[m][a];
            for (int i = 0; i < 3; i++) u[i] = e[i] - d * U[i][a];
        } else { u[0] = 1; u[1] = 0; u[2] = 0; }
        const double nn = sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
        for (int i = 0; i < 3; i++) U[i][j] = u[i] / nn;
    }
}
static double orc_det3(double M[3][3])
{
    return M[0][0] * (M[1][1] * M[2][2] - M[1][2] * M[2][1]) - M[0][1] * (M[1][0] * M[2][2] - M[1][2] * M[2][0]) + M[0][2] * (M[1][0] * M[2][1] - M[1][1] * M[2][0]);
}
/* Umeyama without scaling from the 17 sums {n, sum src[3], sum dst[3], sum dst*src^T [9], sum d^2}: T (row-major 4x4, f32) */
void orc_umeyama_from_sums(const double* sums, float* T)
{
    const double n = sums[0];
    double ms[3], md[3], H[3][3];
    for (int i = 0; i < 3; i++) { ms[i] = sums[1 + i] / n; md[i] = sums[4 + i] / n; }
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) H[i][j] = sums[7 + 3 * i + j] / n - md[i] * ms[j];   /* sigma = mean (d - md)(s - ms)^T */
    double A[3][3], U[3][3], S[3], V[3][3];
    memcpy(A, H, sizeof(A));
    orc_svd3(A, U, S, V);
    double sg[3] = { 1, 1, 1 };
    if (orc_det3(U) * orc_det3(V) < 0) { int m = 0; for (int j = 1; j < 3; j++) if (S[j] < S[m]) m = j; sg[m] = -1; }
    double R[3][3];
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { double v = 0; for (int k = 0; k < 3; k++) v += U[i][k] * sg[k] * V[j][k]; R[i][j] = v; }
    for (int i = 0; i < 3; i++) {
        double t = md[i];
        for (int j = 0; j < 3; j++) { T[4 * i + j] = (float)R[i][j]; t -= R[i][j] * ms[j]; }
        T[4 * i + 3] = (float)t;
    }
    T[12] = T[13] = T[14] = 0.f; T[15] = 1.f;
}
static void orc_mat4_mul_f32(const float* A, const float* B, float* C)   /* C = A * B, f32 like Eigen::Matrix4f */
{
    float R[16];
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) { float v = 0.f; for (int k = 0; k < 4; k++) v += A[4 * i + k] * B[4 * k + j]; R[4 * i + j] = v; }
    memcpy(C, R, sizeof(R));
}
/* state: 0 not converged (too few correspondences), 1 ITERATIONS, 2 TRANSFORM, 3 ABS_MSE, 4 REL_MSE */
int orc_icp(const orc_point* src, int ns, const orc_point* tgt, int nt, double max_dist, int max_iter, double trans_eps, double fit_eps,
            float* T_out, orc_point* aligned, int* iters_out, double* fitness_out, int* corr_out)
{
    orc_point* cur = (orc_point*)malloc(sizeof(orc_point) * (size_t)(ns > 0 ? ns : 1));
    memcpy(cur, src, sizeof(orc_point) * (size_t)ns);
    float Tf[16] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1 };
    const float md2 = (float)(max_dist * max_dist);
    double prev_mse = DBL_MAX;
    int state = 0, it = 0, ncorr = 0;
    for (;;) {
        double sums[17];
        for (int i = 0; i < 17; i++) sums[i] = 0;
        for (int i = 0; i < ns; i++) {
            float best = FLT_MAX; int bj = -1;
            for (int j = 0; j < nt; j++) {
                const float dx = cur[i].x - tgt[j].x, dy = cur[i].y - tgt[j].y, dz = cur[i].z - tgt[j].z;
                float d = dx * dx; d = d + dy * dy; d = d + dz * dz;
                if (d < best) { best = d; bj = j; }
            }
            if (bj < 0 || best > md2) continue;
            const double s[3] = { cur[i].x, cur[i].y, cur[i].z }, d3[3] = { tgt[bj].x, tgt[bj].y, tgt[bj].z };
            sums[0] += 1;
            for (int a = 0; a < 3; a++) { sums[1 + a] += s[a]; sums[4 + a] += d3[a]; for (int b2 = 0; b2 < 3; b2++) sums[7 + 3 * a + b2] += d3[a] * s[b2]; }
            sums[16] += (double)best;
        }
        ncorr = (int)sums[0];
        if (ncorr < 3) { state = 0; break; }
        float Ti[16];
        orc_umeyama_from_sums(sums, Ti);
        orc_transform_cloud(cur, ns, Ti, cur);
        orc_mat4_mul_f32(Ti, Tf, Tf);
        it++;
        if (it >= max_iter) { state = 1; break; }
        const double cosang = 0.5 * ((double)Ti[0] + (double)Ti[5] + (double)Ti[10] - 1.0);
        const double tsq = (double)Ti[3] * Ti[3] + (double)Ti[7] * Ti[7] + (double)Ti[11] * Ti[11];
        if (cosang >= 1.0 - trans_eps && tsq <= trans_eps) { state = 2; break; }
        const double mse = sums[16] / sums[0];
        if (fabs(mse - prev_mse) < 1e-12) { state = 3; break; }
        if (fabs(mse - prev_mse) / prev_mse < fit_eps) { state = 4; break; }
        prev_mse = mse;
    }
    /* getFitnessScore(): mean squared distance of every aligned source point to its nearest target point */
    double fs = 0; int fn = 0;
    for (int i = 0; i < ns && nt > 0; i++) {
        float best = FLT_MAX;
        for (int j = 0; j < nt; j++) {
            const float dx = cur[i].x - tgt[j].x, dy = cur[i].y - tgt[j].y, dz = cur[i].z - tgt[j].z;
            float d = dx * dx; d = d + dy * dy; d = d + dz * dz;
            if (d < best) best = d;
        }
        fs += (double)best; fn++;
    }
    if (T_out) memcpy(T_out, Tf, sizeof(Tf));
    if (aligned) memcpy(aligned, cur, sizeof(orc_point) * (size_t)ns);
    if (iters_out) *iters_out = it;
    if (fitness_out) *fitness_out = fn ? fs / fn : 0.0;
    if (corr_out) *corr_out = ncorr;
    free(cur);
    return state;
}
