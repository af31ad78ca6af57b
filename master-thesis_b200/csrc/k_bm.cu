// k_bm.cu -- cv::StereoBM::compute (PREFILTER_XSOBEL) semantics: the other branch of the reference's block-matcher
// switch (cv_reconstruct/src/reconstruct/Reconstruct3D.cpp:43-48, :378-380; "block_matcher": "stereo_bm" is the default in
// cv_reconstruct/resources/config/config_default.json:8-12).  SURVEY 8(f) row n2.  Bit-exact against cv2.StereoBM
// (tests/golden/bm_*.npz) for minDisparity <= 0 (the reference never sets a positive one for StereoBM).
//
//   k_bm_prefilter   x-Sobel clipped to [-cap, cap] + cap, rows reflected at the image border, trailing odd row = cap
//   k_bm_hsad        per row: sliding horizontal sum over the window of |L - R| for every disparity -> int16 volume,
//                    and of the texture measure |L - cap|
//   k_bm_wta         one warp per column marching down the rows: vertical window sums in registers, first-minimum
//                    winner, texture / uniqueness rejection, OpenCV's sub-pixel formula, valid-rectangle mask
//   (speckle filter: the k_ccl_* kernels of k_sgbm.cu)
// SAD array index d stands for disparity ndisp - 1 - d + minDisparity, exactly as in OpenCV (ties go to the larger disparity).
#include "cvr_internal.cuh"


__global__ void k_bm_prefilter(const uint8_t* __restrict__ gray, size_t grayFrameStride, uint8_t* __restrict__ pre, BmDev p)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, img = blockIdx.z & 1, b = blockIdx.z >> 1;
    if (x >= p.W) return;
    const uint8_t* g = gray + (size_t)b * grayFrameStride + (size_t)img * p.H * p.Wp;
    int v = p.cap;
    const bool tail = (p.H & 1) && y == p.H - 1;                 // rows are filtered in pairs; an odd last row stays neutral
    if (!tail && x > 0 && x < p.W - 1) {
        const int ym = y > 0 ? y - 1 : (p.H > 1 ? 1 : 0), yp = y < p.H - 1 ? y + 1 : (p.H > 1 ? y - 1 : y);
        const uint8_t *r0 = g + (size_t)ym * p.Wp + x, *r1 = g + (size_t)y * p.Wp + x, *r2 = g + (size_t)yp * p.Wp + x;
        const int s = ((int)r0[1] - (int)r0[-1]) + 2 * ((int)r1[1] - (int)r1[-1]) + ((int)r2[1] - (int)r2[-1]);
        v = min(max(s, -p.cap), p.cap) + p.cap;
    }
    pre[(size_t)b * grayFrameStride + (size_t)img * p.H * p.Wp + (size_t)y * p.Wp + x] = (uint8_t)v;
}

// One CTA per (row, frame), one thread per SAD index d.  hsad[y][x][d], x in [0, width1).
__global__ void k_bm_hsad(const uint8_t* __restrict__ pre, size_t frameStride, int16_t* __restrict__ hsad, int* __restrict__ htext, BmDev p)
{
    extern __shared__ uint8_t smb[];
    uint8_t* Lr = smb;                       // [W]
    uint8_t* Rr = smb + p.W;                 // [W]
    uint8_t* ringbuf = Rr + p.W;             // [ring][Dp]  (+ one extra column of [ring] for the texture term)
    const int y = blockIdx.x, b = blockIdx.y, W = p.W, d = threadIdx.x;
    const uint8_t* gl = pre + (size_t)b * frameStride + (size_t)y * p.Wp;
    const uint8_t* gr = gl + (size_t)p.H * p.Wp;
    for (int x = threadIdx.x; x < W; x += blockDim.x) { Lr[x] = gl[x]; Rr[x] = gr[x]; }
    __syncthreads();
    const int Dp = p.Dp, wsz2 = p.wsz2, wsz = p.wsz, width1 = p.width1, ringBytes = p.ring * Dp;
    int16_t* out = hsad + (((size_t)b * p.H + y) * width1) * Dp + d;
    if (d >= p.ndisp) {
        for (int x = 0; x < width1; x++) out[(size_t)x * Dp] = 0;
        return;
    }
    uint8_t* ring = ringbuf + d;
    uint8_t* tring = ringbuf + (size_t)ringBytes;            // texture ring (thread 0 only)
    int* tout = htext + ((size_t)b * p.H + y) * W;
    int run = 0, trun = 0, wi = 0, ri = 0, twi = 0, tri = 0;
    for (int j = -wsz2; j <= width1 - 1 + wsz2; j++) {
        const int xl = p.lofs + min(max(j, -p.lofs), W - 1 - p.lofs);
        const int xr = p.rofs + min(max(j, -p.rofs), W - p.ndisp - p.rofs);
        const int lval = Lr[xl];
        const int diff = abs(lval - (int)Rr[xr + d]);
        run += diff;
        const int n = j + wsz2;                               // elements seen before this one
        if (n >= wsz) { run -= (int)ring[ri]; ri += Dp; if (ri == ringBytes) ri = 0; }
        ring[wi] = (uint8_t)diff; wi += Dp; if (wi == ringBytes) wi = 0;
        if (d == 0) {
            const int t = abs(lval - p.cap);
            trun += t;
            if (n >= wsz) { trun -= (int)tring[tri]; tri = (tri + 1 == p.ring) ? 0 : tri + 1; }
            tring[twi] = (uint8_t)t; twi = (twi + 1 == p.ring) ? 0 : twi + 1;
        }
        if (j >= wsz2) {
            out[(size_t)(j - wsz2) * Dp] = (int16_t)run;
            if (d == 0) tout[j - wsz2] = trun;
        }
    }
}

// One warp per (column x, frame).
template <int DPL>
__global__ void __launch_bounds__(128) k_bm_wta(const int16_t* __restrict__ hsad, const int* __restrict__ htext, int16_t* __restrict__ disp, BmDev p, int B)
{
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= B * p.width1) return;
    const int b = warp / p.width1, x = warp % p.width1;
    const int H = p.H, W = p.W, Dp = p.Dp, nd = p.ndisp, wsz2 = p.wsz2;
    const size_t rowStride = (size_t)p.width1 * Dp;
    const int16_t* hs = hsad + ((size_t)b * H * p.width1 + x) * Dp + lane * DPL;
    const int* ht = htext + (size_t)b * H * W + x;
    int16_t* out = disp + (size_t)b * H * W + p.lofs + x;
    const int d0 = lane * DPL;
    auto ldrow = [&](int yy, int (&v)[DPL]) {
        const int16_t* q = hs + (size_t)yy * rowStride;
#pragma unroll
        for (int i = 0; i < DPL; i += 2) { const unsigned w = *(const unsigned*)(q + i); v[i] = (int)(w & 0xffffu); v[i + 1] = (int)(w >> 16); }
    };
    int sad[DPL], tmp[DPL];
    ldrow(0, tmp);
#pragma unroll
    for (int i = 0; i < DPL; i++) sad[i] = tmp[i] * (wsz2 + 1);
    int tsum = ht[0] * (wsz2 + 1);
    for (int yy = 1; yy <= wsz2; yy++) {
        const int yc = min(yy, H - 1);
        ldrow(yc, tmp);
#pragma unroll
        for (int i = 0; i < DPL; i++) sad[i] += tmp[i];
        tsum += ht[(size_t)yc * W];
    }
    // valid rectangle (getValidDisparityROI with full-image ROIs)
    const int maxD = p.minD + nd - 1;
    const int xmin = max(0, maxD) + wsz2, xmax = W - wsz2, ymin = wsz2, ymax = H - wsz2;
    const int xi = p.lofs + x;
    const bool colOK = xi >= xmin && xi < xmax && (xmax - xmin > 0) && (ymax - ymin > 0) && xi < W;
    for (int y = 0; y < H; y++) {
        int res = p.FILTERED;
        if (colOK && y >= ymin && y < ymax && tsum >= p.textureThreshold) {          // warp-uniform
            int key = 0x7fffffff;
#pragma unroll
            for (int i = 0; i < DPL; i++)
                if (d0 + i < nd) key = min(key, (sad[i] << 8) | (d0 + i));
            key = __reduce_min_sync(0xffffffffu, key);
            const int minsad = key >> 8, mind = key & 255;
            bool bad = false;
            if (p.uniq > 0) {
                const int thresh = minsad + (minsad * p.uniq / 100);
#pragma unroll
                for (int i = 0; i < DPL; i++) {
                    const int d = d0 + i;
                    if (d < nd && (d < mind - 1 || d > mind + 1) && sad[i] <= thresh) bad = true;
                }
            }
            if (!__any_sync(0xffffffffu, bad)) {
                // p = sad[mind+1], n = sad[mind-1] with sad[-1] = sad[1], sad[ndisp] = sad[ndisp-2]
                const int ip = (mind + 1 < nd) ? mind + 1 : nd - 2, in = (mind - 1 >= 0) ? mind - 1 : 1;
                int pv = 0, nv = 0;
#pragma unroll
                for (int i = 0; i < DPL; i++) { pv += (d0 + i == ip) ? sad[i] : 0; nv += (d0 + i == in) ? sad[i] : 0; }
                pv = __reduce_add_sync(0xffffffffu, pv);
                nv = __reduce_add_sync(0xffffffffu, nv);
                const int dd = pv + nv - 2 * minsad + abs(pv - nv);
                res = (((nd - mind - 1 + p.minD) * 256 + (dd != 0 ? (pv - nv) * 256 / dd : 0)) + 15) >> 4;
            }
        }
        if (lane == 0 && xi < W) out[(size_t)y * W] = (int16_t)res;
        if (y + 1 < H) {                                                             // slide the vertical window to row y + 1
            const int ya = min(y + 1 + wsz2, H - 1), ys = max(y - wsz2, 0);
            int add[DPL], sub[DPL];
            ldrow(ya, add); ldrow(ys, sub);
#pragma unroll
            for (int i = 0; i < DPL; i++) sad[i] += add[i] - sub[i];
            tsum += ht[(size_t)ya * W] - ht[(size_t)ys * W];
        }
    }
}

__global__ void k_bm_fill(int16_t* p, size_t n, int v)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = (int16_t)v;
}

int cvr_launch_speckle(cvr_ctx* c, int B, int newVal, int maxSize, int maxDiff);   // k_sgbm.cu

int cvr_launch_bm(cvr_ctx* c, int B)
{
    const BmDev& p = c->bm;
    const size_t npix = (size_t)B * p.W * p.H;
    k_bm_fill<<<(unsigned)cvr_div_up_sz(npix, 256), 256, 0, c->stream>>>(c->d_med, npix, p.FILTERED);
    LAUNCH_CHECK(c);
    if (p.lofs >= p.W || p.rofs >= p.W || p.width1 < 1) return CVR_OK;
    const size_t gfs = (size_t)2 * p.H * p.Wp;
    k_bm_prefilter<<<dim3(cvr_div_up(p.W, 256), p.H, 2 * B), 256, 0, c->stream>>>(c->d_gray, gfs, c->d_bmpre, p);
    LAUNCH_CHECK(c);
    const size_t smem = (size_t)2 * p.W + (size_t)p.ring * p.Dp + p.ring;
    if (smem > 48 * 1024) CVR_CUDA(c, cudaFuncSetAttribute(k_bm_hsad, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_bm_hsad<<<dim3(p.H, B), p.Dp, smem, c->stream>>>(c->d_bmpre, gfs, c->d_volA, c->d_label, p);
    LAUNCH_CHECK(c);
    cvr_stage_mark(c, "bm_hsad");
    const int blocks = cvr_div_up(B * p.width1, 4);
    switch (p.Dp / 32) {
        case 2: k_bm_wta<2><<<blocks, 128, 0, c->stream>>>(c->d_volA, c->d_label, c->d_med, p, B); break;
        case 4: k_bm_wta<4><<<blocks, 128, 0, c->stream>>>(c->d_volA, c->d_label, c->d_med, p, B); break;
        case 6: k_bm_wta<6><<<blocks, 128, 0, c->stream>>>(c->d_volA, c->d_label, c->d_med, p, B); break;
        case 8: k_bm_wta<8><<<blocks, 128, 0, c->stream>>>(c->d_volA, c->d_label, c->d_med, p, B); break;
        default: c->err = "unsupported padded disparity count"; return CVR_ERR_UNSUPPORTED;
    }
    LAUNCH_CHECK(c);
    cvr_stage_mark(c, "bm_wta");
    if (c->prm.bm_speckle_range >= 0 && c->prm.bm_speckle_window_size > 0) {
        int rc = cvr_launch_speckle(c, B, p.FILTERED, c->prm.bm_speckle_window_size, c->prm.bm_speckle_range);
        if (rc) return rc;
        cvr_stage_mark(c, "speckle");
    }
    return CVR_OK;
}
