// Stereo rectification of the input images on the device: Reconstruct3D::RectifyImages
// (cv_reconstruct/src/reconstruct/Reconstruct3D.cpp:330-361) = cv::initUndistortRectifyMap(.., CV_16SC2, ..) per camera
// + cv::remap(.., INTER_LINEAR) per image.  The reference rebuilds the maps for every frame; here they are built once
// per calibration (cvr_set_rectification) and stay resident, the per-frame work is the remap alone.
#include "cvr_internal.cuh"

// ---- maps ---------------------------------------------------------------------------------------
// One thread per image row.  OpenCV advances the homogeneous coordinate of a row by repeated addition
// (_x += ir[0] per column); the rounding of those running sums is part of the result, so the row is walked
// sequentially, in double, with every operation rounded on its own (no fused multiply-add).
struct RectCam { double ir[9]; double k[12]; double u0, v0, fx, fy; };

__global__ void k_rectify_maps(RectCam cam, int W, int H, short2* __restrict__ map1, uint16_t* __restrict__ map2)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= H) return;
    const double* ir = cam.ir;
    const double k1 = cam.k[0], k2 = cam.k[1], p1 = cam.k[2], p2 = cam.k[3], k3 = cam.k[4], k4 = cam.k[5], k5 = cam.k[6], k6 = cam.k[7];
    const double s1 = cam.k[8], s2 = cam.k[9], s3 = cam.k[10], s4 = cam.k[11];
    const double di = (double)i;
    double _x = __dadd_rn(__dmul_rn(di, ir[1]), ir[2]), _y = __dadd_rn(__dmul_rn(di, ir[4]), ir[5]), _w = __dadd_rn(__dmul_rn(di, ir[7]), ir[8]);
    short2* m1 = map1 + (size_t)i * W;
    uint16_t* m2 = map2 + (size_t)i * W;
    for (int j = 0; j < W; j++) {
        const double w = __ddiv_rn(1.0, _w), x = __dmul_rn(_x, w), y = __dmul_rn(_y, w);
        const double x2 = __dmul_rn(x, x), y2 = __dmul_rn(y, y);
        const double r2 = __dadd_rn(x2, y2), _2xy = __dmul_rn(__dmul_rn(2.0, x), y);
        const double num = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(k3, r2), k2), r2), k1), r2));
        const double den = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(k6, r2), k5), r2), k4), r2));
        const double kr = __ddiv_rn(num, den);
        double xd = __dadd_rn(__dmul_rn(x, kr), __dmul_rn(p1, _2xy));
        xd = __dadd_rn(xd, __dmul_rn(p2, __dadd_rn(r2, __dmul_rn(2.0, x2))));
        xd = __dadd_rn(xd, __dmul_rn(s1, r2));
        xd = __dadd_rn(xd, __dmul_rn(__dmul_rn(s2, r2), r2));
        double yd = __dadd_rn(__dmul_rn(y, kr), __dmul_rn(p1, __dadd_rn(r2, __dmul_rn(2.0, y2))));
        yd = __dadd_rn(yd, __dmul_rn(p2, _2xy));
        yd = __dadd_rn(yd, __dmul_rn(s3, r2));
        yd = __dadd_rn(yd, __dmul_rn(__dmul_rn(s4, r2), r2));
        const double u = __dadd_rn(__dmul_rn(cam.fx, xd), cam.u0), v = __dadd_rn(__dmul_rn(cam.fy, yd), cam.v0);
        const int iu = __double2int_rn(__dmul_rn(u, 32.0)), iv = __double2int_rn(__dmul_rn(v, 32.0));   // cvRound: half to even
        m1[j] = make_short2((short)(iu >> 5), (short)(iv >> 5));
        m2[j] = (uint16_t)((iv & 31) * 32 + (iu & 31));
        _x = __dadd_rn(_x, ir[0]); _y = __dadd_rn(_y, ir[3]); _w = __dadd_rn(_w, ir[6]);
    }
}

// ---- remap --------------------------------------------------------------------------------------
// cv::remap, 8-bit, INTER_LINEAR, CV_16SC2 + CV_16UC1 maps, BORDER_CONSTANT 0.  One thread per destination pixel of
// one image; blockIdx.z = 2 * frame + camera.  Weights (32-fx)(32-fy)*32 ... sum to 2^15 (OpenCV's BilinearTab_i is
// exactly this), result = (sum + 2^14) >> 15.  A tap outside the source reads the border value 0.
template <int CN>
__global__ void __launch_bounds__(256) k_remap_bilinear(const uint8_t* __restrict__ srcL, size_t lstep, size_t lfs, const uint8_t* __restrict__ srcR,
                                                        size_t rstep, size_t rfs, const short2* __restrict__ map1L, const uint16_t* __restrict__ map2L,
                                                        const short2* __restrict__ map1R, const uint16_t* __restrict__ map2R, int W, int H,
                                                        uint8_t* __restrict__ dstL, uint8_t* __restrict__ dstR, size_t dstep, size_t dfs)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= W) return;
    const int b = blockIdx.z >> 1, cam = blockIdx.z & 1;
    const uint8_t* src = cam ? srcR + (size_t)b * rfs : srcL + (size_t)b * lfs;
    const size_t sstep = cam ? rstep : lstep;
    const size_t mi = (size_t)y * W + x;
    const short2 xy = __ldg((cam ? map1R : map1L) + mi);
    const int f = __ldg((cam ? map2R : map2L) + mi) & 1023, fx = f & 31, fy = f >> 5;
    const int sx = xy.x, sy = xy.y;
    const int w00 = (32 - fx) * (32 - fy) * 32, w01 = fx * (32 - fy) * 32, w10 = (32 - fx) * fy * 32, w11 = fx * fy * 32;
    uint8_t* d = (cam ? dstR : dstL) + (size_t)b * dfs + (size_t)y * dstep + (size_t)x * CN;
    int acc[CN];
#pragma unroll
    for (int c = 0; c < CN; c++) acc[c] = 1 << 14;
    if ((unsigned)sx < (unsigned)(W - 1) && (unsigned)sy < (unsigned)(H - 1)) {          // all four taps inside
        const uint8_t* s0 = src + (size_t)sy * sstep + (size_t)sx * CN;
        const uint8_t* s1 = s0 + sstep;
#pragma unroll
        for (int c = 0; c < CN; c++) acc[c] += (int)s0[c] * w00 + (int)s0[CN + c] * w01 + (int)s1[c] * w10 + (int)s1[CN + c] * w11;
    } else {
        const bool x0 = sx >= 0 && sx < W, x1 = sx + 1 >= 0 && sx + 1 < W, y0 = sy >= 0 && sy < H, y1 = sy + 1 >= 0 && sy + 1 < H;
        const uint8_t* s0 = src + (ptrdiff_t)sy * (ptrdiff_t)sstep + (ptrdiff_t)sx * CN;
        const uint8_t* s1 = s0 + sstep;
#pragma unroll
        for (int c = 0; c < CN; c++) {
            if (x0 && y0) acc[c] += (int)s0[c] * w00;
            if (x1 && y0) acc[c] += (int)s0[CN + c] * w01;
            if (x0 && y1) acc[c] += (int)s1[c] * w10;
            if (x1 && y1) acc[c] += (int)s1[CN + c] * w11;
        }
    }
#pragma unroll
    for (int c = 0; c < CN; c++) d[c] = (uint8_t)(acc[c] >> 15);
}

// 3x3 inverse as cv::invert(DECOMP_LU) does it for 3x3 doubles (adjugate times 1/det), on the host, in double.
static void inv3(const double* S, double* t)
{
    double d = S[0] * (S[4] * S[8] - S[5] * S[7]) - S[1] * (S[3] * S[8] - S[5] * S[6]) + S[2] * (S[3] * S[7] - S[4] * S[6]);
    if (d != 0.0) d = 1. / d;
    t[0] = (S[4] * S[8] - S[5] * S[7]) * d; t[1] = (S[2] * S[7] - S[1] * S[8]) * d; t[2] = (S[1] * S[5] - S[2] * S[4]) * d;
    t[3] = (S[5] * S[6] - S[3] * S[8]) * d; t[4] = (S[0] * S[8] - S[2] * S[6]) * d; t[5] = (S[2] * S[3] - S[0] * S[5]) * d;
    t[6] = (S[3] * S[7] - S[4] * S[6]) * d; t[7] = (S[1] * S[6] - S[0] * S[7]) * d; t[8] = (S[0] * S[4] - S[1] * S[3]) * d;
}

int cvr_launch_rectify_maps(cvr_ctx* c, int cam, const double* K, const double* D, int nD, const double* R, const double* P, int pcols)
{
    RectCam rc;
    double M[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            // same association as the 3x3 matrix product: (a0*b0 + a1*b1) + a2*b2, each product rounded (host code, -ffp-contract=off)
            volatile double t0 = P[i * pcols + 0] * R[0 * 3 + j], t1 = P[i * pcols + 1] * R[1 * 3 + j], t2 = P[i * pcols + 2] * R[2 * 3 + j];
            volatile double s = t0 + t1;
            M[i * 3 + j] = s + t2;
        }
    inv3(M, rc.ir);
    for (int i = 0; i < 12; i++) rc.k[i] = i < nD ? D[i] : 0.0;
    rc.u0 = K[2]; rc.v0 = K[5]; rc.fx = K[0]; rc.fy = K[4];
    const int W = c->prm.width, H = c->prm.height;
    k_rectify_maps<<<cvr_div_up(H, 32), 32, 0, c->stream>>>(rc, W, H, c->d_rmap1[cam], c->d_rmap2[cam]);
    c->launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { c->err = std::string("k_rectify_maps launch: ") + cudaGetErrorString(e); return CVR_ERR_CUDA; }
    return CVR_OK;
}

int cvr_launch_remap(cvr_ctx* c, int B, int ch, const uint8_t* dl, size_t lstep, size_t lfs, const uint8_t* dr, size_t rstep, size_t rfs,
                     uint8_t* outL, uint8_t* outR, size_t ostep, size_t ofs)
{
    const int W = c->prm.width, H = c->prm.height;
    const dim3 g(cvr_div_up(W, 256), H, 2 * B);
#define CVR_REMAP(CN) k_remap_bilinear<CN><<<g, 256, 0, c->stream>>>(dl, lstep, lfs, dr, rstep, rfs, c->d_rmap1[0], c->d_rmap2[0], c->d_rmap1[1], c->d_rmap2[1], W, H, outL, outR, ostep, ofs)
    if (ch == 1) CVR_REMAP(1); else if (ch == 3) CVR_REMAP(3); else CVR_REMAP(4);
#undef CVR_REMAP
    c->launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { c->err = std::string("k_remap_bilinear launch: ") + cudaGetErrorString(e); return CVR_ERR_CUDA; }
    return CVR_OK;
}
