// k_cloud.cu -- the glue stages between disparity and point cloud:
//   prune       TrackingFrame::SetupFrame / PruneDisparityImage  (cv_reconstruct/src/system/TrackingFrame.cpp:25-67)
//   reproject   Reconstruct3D::GeneratePointCloud                (cv_reconstruct/src/reconstruct/Reconstruct3D.cpp:106-153)
//   triangulate Reconstruct3D::Triangulate3D                     (cv_reconstruct/src/reconstruct/Reconstruct3D.cpp:156-216)
// Both cloud generators emit surviving pixels in row-major order (the reference's push_back order) through an
// ordered stream compaction: per-1024-pixel block counts -> per-frame exclusive scan -> write.
// Floating point follows the reference's evaluation order with explicit round-to-nearest intrinsics so that
// nvcc cannot contract a*b+c into an FMA where the CPU code has none.
#include "cvr_internal.cuh"
#include "cvr_scan.cuh"
#include <float.h>
#include <mutex>


// ------------------------------------------------------------------------------------------------
// prune
__global__ void k_prune_init(PruneStats* st, int B)
{
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    st[b].mn = 32767; st[b].mx = -32768; st[b].sum = 0; st[b].sq = 0; st[b].a = 0; st[b].b = 0; st[b].ithr = 0;
}
__global__ void k_prune_minmax(const int16_t* __restrict__ disp, int N, PruneStats* st)
{
    const int b = blockIdx.y;
    const int16_t* d = disp + (size_t)b * N;
    int mn = 32767, mx = -32768;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) { int v = d[i]; mn = min(mn, v); mx = max(mx, v); }
    mn = __reduce_min_sync(0xffffffffu, mn); mx = __reduce_max_sync(0xffffffffu, mx);
    if ((threadIdx.x & 31) == 0) { atomicMin(&st[b].mn, mn); atomicMax(&st[b].mx, mx); }
}
// cv::normalize(NORM_MINMAX, 0, 255, CV_8U): scale/shift in f64, applied as one f32 FMA, round-half-even, saturate.
__global__ void k_prune_norm(const int16_t* __restrict__ disp, int N, PruneStats* st, uint8_t* __restrict__ u8)
{
    const int b = blockIdx.y;
    const int16_t* d = disp + (size_t)b * N;
    uint8_t* o = u8 + (size_t)b * N;
    const double dmn = (double)st[b].mn, dmx = (double)st[b].mx;
    const double rng = __dsub_rn(dmx, dmn);
    const double scale = __dmul_rn(255.0, (rng > DBL_EPSILON ? __ddiv_rn(1.0, rng) : 0.0));
    const double shift = __dsub_rn(0.0, __dmul_rn(dmn, scale));
    const float fa = (float)scale, fb = (float)shift;
    unsigned long long sum = 0, sq = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
        const float f = __fmaf_rn((float)d[i], fa, fb);
        int r = __float2int_rn(f);
        r = min(max(r, 0), 255);
        o[i] = (uint8_t)r;
        sum += (unsigned)r; sq += (unsigned)(r * r);
    }
    __shared__ unsigned long long ssum[32], ssq[32];
#pragma unroll
    for (int k = 16; k > 0; k >>= 1) { sum += __shfl_down_sync(0xffffffffu, sum, k); sq += __shfl_down_sync(0xffffffffu, sq, k); }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { ssum[w] = sum; ssq[w] = sq; }
    __syncthreads();
    if (w == 0) {
        const int nw = blockDim.x >> 5;
        sum = lane < nw ? ssum[lane] : 0; sq = lane < nw ? ssq[lane] : 0;
#pragma unroll
        for (int k = 16; k > 0; k >>= 1) { sum += __shfl_down_sync(0xffffffffu, sum, k); sq += __shfl_down_sync(0xffffffffu, sq, k); }
        if (lane == 0) { atomicAdd(&st[b].sum, sum); atomicAdd(&st[b].sq, sq); }
    }
}
// cv::meanStdDev (population sigma, f64) -> cv::threshold(u8 > floor(sigma*factor)) -> zero pruned disparities ->
// convertTo(CV_32F, 1/16) -> mask &= disparity > 0.
// disp / pruned and u8 / mask may alias (in-place use): every element is read before the same thread writes it.
__global__ void k_prune_apply(const int16_t* disp, const uint8_t* u8, int N, const PruneStats* __restrict__ st,
                              double factor, int16_t* pruned, float* __restrict__ dispf, uint8_t* mask)
{
    const int b = blockIdx.y;
    const double inv = __ddiv_rn(1.0, (double)N);
    const double mean = __dmul_rn((double)st[b].sum, inv);
    const double var = __dsub_rn(__dmul_rn((double)st[b].sq, inv), __dmul_rn(mean, mean));
    const double sigma = sqrt(var > 0.0 ? var : 0.0);
    const int ithr = (int)floor(__dmul_rn(sigma, factor));
    const size_t o = (size_t)b * N;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
        const int keep = (int)u8[o + i] > ithr;
        const int v = keep ? (int)disp[o + i] : 0;
        const float f = __fmul_rn((float)v, 0.0625f);
        if (pruned) pruned[o + i] = (int16_t)v;
        if (dispf) dispf[o + i] = f;
        if (mask) mask[o + i] = (keep && f > 0.0f) ? 255 : 0;
    }
}

int cvr_launch_prune(cvr_ctx* c, int B, const int16_t* disp_in)
{
    const int N = c->sg.W * c->sg.H;
    const int gx = min(cvr_div_up(N, 256), 296);
    k_prune_init<<<cvr_div_up(B, 128), 128, 0, c->stream>>>(c->d_pstats, B);
    LAUNCH_CHECK(c);
    k_prune_minmax<<<dim3(gx, B), 256, 0, c->stream>>>(disp_in, N, c->d_pstats);
    LAUNCH_CHECK(c);
    // the normalised u8 image lives in d_mask until k_prune_apply overwrites it element-wise (same index read then written)
    k_prune_norm<<<dim3(gx, B), 256, 0, c->stream>>>(disp_in, N, c->d_pstats, c->d_mask);
    LAUNCH_CHECK(c);
    k_prune_apply<<<dim3(gx, B), 256, 0, c->stream>>>(disp_in, c->d_mask, N, c->d_pstats, c->cl.sigma_factor, c->d_pruned, c->d_dispf, c->d_mask);
    LAUNCH_CHECK(c);
    cvr_stage_mark(c, "prune");
    return CVR_OK;
}

// ------------------------------------------------------------------------------------------------
// The batch path does not materialise the pruned disparity, the float disparity or the mask: two reduction passes over
// the final disparity give the frame's PruneStats (min/max, then sum and sum of squares of the normalised 8-bit image),
// and the cloud kernel evaluates normalize / threshold / convertTo per pixel on the fly (k_cloud_select<.., FUSED>).
// 16-byte loads over the aligned body of the frame, scalar head and tail.
template <typename F1, typename F8>
__device__ __forceinline__ void s16_stream(const int16_t* p, int n, int tid, int nthreads, F1 f1, F8 f8)
{
    int head = (int)(((16 - ((size_t)p & 15)) & 15) >> 1);
    if (head > n) head = n;
    const int nv = (n - head) >> 3;
    for (int i = tid; i < head; i += nthreads) f1((int)p[i]);
    const uint4* v = (const uint4*)(p + head);
    for (int i = tid; i < nv; i += nthreads) f8(v[i]);
    for (int i = head + nv * 8 + tid; i < n; i += nthreads) f1((int)p[i]);
}
__global__ void __launch_bounds__(256) k_prune_minmax8(const int16_t* __restrict__ disp, int N, PruneStats* st)
{
    const int b = blockIdx.y;
    unsigned mn2 = 0x7fff7fffu, mx2 = 0x80008000u;
    int mn = 32767, mx = -32768;
    s16_stream(disp + (size_t)b * N, N, blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x,
               [&](int v) { mn = min(mn, v); mx = max(mx, v); },
               [&](const uint4& q) {
                   mn2 = __vmins2(__vmins2(mn2, q.x), __vmins2(q.y, __vmins2(q.z, q.w)));
                   mx2 = __vmaxs2(__vmaxs2(mx2, q.x), __vmaxs2(q.y, __vmaxs2(q.z, q.w)));
               });
    mn = min(mn, min((int)(short)(mn2 & 0xffffu), (int)mn2 >> 16));
    mx = max(mx, max((int)(short)(mx2 & 0xffffu), (int)mx2 >> 16));
    mn = __reduce_min_sync(0xffffffffu, mn); mx = __reduce_max_sync(0xffffffffu, mx);
    if ((threadIdx.x & 31) == 0) { atomicMin(&st[b].mn, mn); atomicMax(&st[b].mx, mx); }
}
// scale / shift of cv::normalize(NORM_MINMAX, 0, 255) as the two f32 constants of its single-rounding FMA
__device__ __forceinline__ void prune_scale(const PruneStats& s, float& fa, float& fb)
{
    const double dmn = (double)s.mn, dmx = (double)s.mx;
    const double rng = __dsub_rn(dmx, dmn);
    const double scale = __dmul_rn(255.0, (rng > DBL_EPSILON ? __ddiv_rn(1.0, rng) : 0.0));
    fa = (float)scale; fb = (float)__dsub_rn(0.0, __dmul_rn(dmn, scale));
}
__device__ __forceinline__ int prune_u8(int d, float fa, float fb) { return min(max(__float2int_rn(__fmaf_rn((float)d, fa, fb)), 0), 255); }
__device__ __forceinline__ int prune_threshold(const PruneStats& s, int N, double factor)
{
    const double inv = __ddiv_rn(1.0, (double)N);
    const double mean = __dmul_rn((double)s.sum, inv);
    const double var = __dsub_rn(__dmul_rn((double)s.sq, inv), __dmul_rn(mean, mean));
    return (int)floor(__dmul_rn(sqrt(var > 0.0 ? var : 0.0), factor));
}
__global__ void __launch_bounds__(256) k_prune_stats8(const int16_t* __restrict__ disp, int N, PruneStats* st)
{
    const int b = blockIdx.y;
    float fa, fb;
    prune_scale(st[b], fa, fb);
    unsigned sum = 0; unsigned long long sq = 0;                       // a thread sees < 2^24 pixels: the 8-bit sum fits 32 bits
    auto one = [&](int v) { const unsigned r = (unsigned)prune_u8(v, fa, fb); sum += r; sq += r * r; };
    s16_stream(disp + (size_t)b * N, N, blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x, one,
               [&](const uint4& q) {
                   const unsigned w[4] = { q.x, q.y, q.z, q.w };
#pragma unroll
                   for (int j = 0; j < 4; j++) { one((int)(short)(w[j] & 0xffffu)); one((int)w[j] >> 16); }
               });
    unsigned long long s64 = sum;
    __shared__ unsigned long long ssum[8], ssq[8];
#pragma unroll
    for (int k = 16; k > 0; k >>= 1) { s64 += __shfl_down_sync(0xffffffffu, s64, k); sq += __shfl_down_sync(0xffffffffu, sq, k); }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { ssum[w] = s64; ssq[w] = sq; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 8; k++) { s64 += ssum[k]; sq += ssq[k]; }
        atomicAdd(&st[b].sum, s64); atomicAdd(&st[b].sq, sq);
    }
}
// per-frame records a batch of the filter chain starts from: prune statistics and the two bounding boxes (cloud -> voxel
// grid, voxel centroids -> outlier search grid) that the producing kernels accumulate
__global__ void k_chain_init(PruneStats* st, VoxBox* boxA, VoxBox* boxB, int B)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    if (st) { st[b].mn = 32767; st[b].mx = -32768; st[b].sum = 0; st[b].sq = 0; st[b].a = 0; st[b].b = 0; st[b].ithr = 0; }
    for (int k = 0; k < 3; k++) {
        if (boxA) { boxA[b].mn[k] = 0xffffffffu; boxA[b].mx[k] = 0u; }
        if (boxB) { boxB[b].mn[k] = 0xffffffffu; boxB[b].mx[k] = 0u; }
    }
}
int cvr_launch_chain_init(cvr_ctx* c, int B, bool stats, bool boxA, bool boxB)
{
    k_chain_init<<<cvr_div_up(B, 128), 128, 0, c->stream>>>(stats ? c->d_pstats : nullptr, boxA ? c->d_boxv : nullptr, boxB ? c->d_boxs : nullptr, B);
    LAUNCH_CHECK(c);
    return CVR_OK;
}
// prune statistics only (batch path); d_pstats initialised by cvr_launch_chain_init
int cvr_launch_prune_stats(cvr_ctx* c, int B, const int16_t* disp_in)
{
    const int N = c->sg.W * c->sg.H;
    const int gx = cvr_frame_blocks(c, B, (size_t)N / 8 + 1, 256, 8);
    k_prune_minmax8<<<dim3(gx, B), 256, 0, c->stream>>>(disp_in, N, c->d_pstats);
    LAUNCH_CHECK(c);
    k_prune_stats8<<<dim3(gx, B), 256, 0, c->stream>>>(disp_in, N, c->d_pstats);
    LAUNCH_CHECK(c);
    cvr_stage_mark(c, "prune");
    return CVR_OK;
}

// ------------------------------------------------------------------------------------------------
// cloud generation
struct CloudArgs {
    const int16_t* disp16;   // reproject input / fused input (final disparity) [B][H][W]
    const float* dispf;      // triangulate input [B][H][W]
    const uint8_t* mask;     // optional [B][H][W]
    const uint8_t* bgr; size_t bstep, bfs; int ch;
    int W, H;
    CloudDev cl;
    const int* minDisp;      // reproject: per-frame min disparity (handleMissingValues)
    const PruneStats* pstats;   // FUSED: the frame's prune statistics
};

// FUSED (triangulate only): disparity, mask and float conversion of TrackingFrame::SetupFrame evaluated here from the
// final disparity: keep = u8 > ithr; v = keep ? d : 0; f = v / 16; mask = keep && f > 0; a point iff f > 0 && mask.
template <int MODE, bool FUSED>
__device__ __forceinline__ bool make_point(const CloudArgs& a, int b, int i, cvr_point& pt, float fa, float fb, int ithr)
{
    const int W = a.W, y = i / W, x = i - y * W;
    const size_t o = (size_t)b * W * a.H + i;
    if (MODE == CVR_CLOUD_TRIANGULATE) {
        float d;
        if (FUSED) {
            const int d16 = a.disp16[o];
            if (!(prune_u8(d16, fa, fb) > ithr)) return false;
            d = __fmul_rn((float)d16, 0.0625f);
            if (!(d > 0.0f)) return false;
        } else {
            d = a.dispf[o];
            if (!(d > 0.0f)) return false;                    // `if (d <= 0.0) continue;` (NaN falls through in the reference; disparities are finite)
            if (a.mask && a.mask[o] == 0) return false;
        }
        const float z = __fdiv_rn(__fmul_rn(a.cl.f, a.cl.b), d);
        const float zf = __fdiv_rn(z, a.cl.f);
        pt.z = z;
        pt.x = __fmul_rn(__fsub_rn((float)x, a.cl.cx), zf);
        pt.y = __fmul_rn(-__fsub_rn((float)y, a.cl.cy), zf);
    } else {
        const double d = (double)(float)a.disp16[o];
        const float* Q = a.cl.Q;
        double h[4];
#pragma unroll
        for (int r = 0; r < 4; r++)
            h[r] = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn((double)Q[r * 4 + 0], (double)x), __dmul_rn((double)Q[r * 4 + 1], (double)y)),
                                       __dmul_rn((double)Q[r * 4 + 2], d)), (double)Q[r * 4 + 3]);
        const double iw = __ddiv_rn(1.0, h[3]);
        float p0 = (float)__dmul_rn((double)(float)h[0], iw);
        float p1 = (float)__dmul_rn((double)(float)h[1], iw);
        float p2 = (float)__dmul_rn((double)(float)h[2], iw);
        if (fabs(d - (double)a.minDisp[b]) <= (double)FLT_EPSILON) p2 = 10000.0f;
        if (p2 == 10000.0f) return false;                     // MISSING_DISPARITY_Z, Reconstruct3D.cpp:124
        const float px = __fsub_rn(p0, __fdiv_rn((float)W, 2.0f));
        pt.x = -px;                                           // transformPointCloud(diag(-1,1,1,1))
        pt.y = -__fsub_rn(p1, __fdiv_rn((float)a.H, 2.0f));
        pt.z = p2;
    }
    const uint8_t* s = a.bgr + (size_t)b * a.bfs + (size_t)y * a.bstep + (size_t)x * a.ch;
    if (a.ch == 1) { pt.b = pt.g = pt.r = s[0]; } else { pt.b = s[0]; pt.g = s[1]; pt.r = s[2]; }
    pt.a = 255;
    return true;
}

__global__ void k_min_s16(const int16_t* __restrict__ disp, int N, int* out)
{
    const int b = blockIdx.y;
    int mn = 32767;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) mn = min(mn, (int)disp[(size_t)b * N + i]);
    mn = __reduce_min_sync(0xffffffffu, mn);
    if ((threadIdx.x & 31) == 0) atomicMin(&out[b], mn);
}
__global__ void k_fill_int(int* p, int n, int v) { int i = blockIdx.x * blockDim.x + threadIdx.x; if (i < n) p[i] = v; }

// Surviving pixels in row-major order (the reference's push_back order) by the single-pass ordered selection of
// cvr_scan.cuh; the kernel also accumulates the cloud's bounding box for the voxel grid that follows.
#define CLOUD_ITEMS 8                                              // pixels per thread: a chunk of the cloud selection is 2048 pixels
#define CLOUD_CHUNK (CVR_SEL_THREADS * CLOUD_ITEMS)
template <int MODE, bool FUSED>
__global__ void __launch_bounds__(CVR_SEL_THREADS) k_cloud_select(CloudArgs a, int* sel, size_t selStride, cvr_point* __restrict__ out, size_t outStride,
                                                                  int* __restrict__ nout, VoxBox* box, double factor)
{
    __shared__ int ws[33];
    __shared__ int sh;
    __shared__ float sfa, sfb;
    __shared__ int sthr;
    const int b = blockIdx.y, N = a.W * a.H, nch = (N + CLOUD_CHUNK - 1) / CLOUD_CHUNK;
    if (FUSED && threadIdx.x == 0) {
        const PruneStats s = a.pstats[b];
        float fa, fb;
        prune_scale(s, fa, fb);
        sfa = fa; sfb = fb; sthr = prune_threshold(s, N, factor);
    }
    __syncthreads();
    const float fa = FUSED ? sfa : 0.f, fb = FUSED ? sfb : 0.f;
    const int ithr = FUSED ? sthr : 0;
    unsigned* status = (unsigned*)sel + (size_t)b * selStride;
    int* ticket = (int*)status + (selStride - 1);
    unsigned mn[3] = { 0xffffffffu, 0xffffffffu, 0xffffffffu }, mx[3] = { 0, 0, 0 };
    bool any = false;
    for (;;) {
        const int chunk = sel_next_chunk(ticket, &sh);
        if (chunk >= nch) break;
        const int i0 = chunk * CLOUD_CHUNK + threadIdx.x * CLOUD_ITEMS;
        cvr_point pt[CLOUD_ITEMS];
        int f[CLOUD_ITEMS], cnt = 0;
#pragma unroll
        for (int k = 0; k < CLOUD_ITEMS; k++) {
            f[k] = (i0 + k < N) ? (int)make_point<MODE, FUSED>(a, b, i0 + k, pt[k], fa, fb, ithr) : 0;
            cnt += f[k];
        }
        int tot;
        const int e = block_excl_scan(cnt, ws, &tot);
        const int base = sel_chain_prefix(status, chunk, tot, &sh);
        cvr_point* o = out + (size_t)b * outStride + base + e;
#pragma unroll
        for (int k = 0; k < CLOUD_ITEMS; k++)
            if (f[k]) {
                *(uint4*)o = *(const uint4*)&pt[k]; o++;
                const unsigned ox = f2ord(pt[k].x), oy = f2ord(pt[k].y), oz = f2ord(pt[k].z);
                mn[0] = min(mn[0], ox); mx[0] = max(mx[0], ox); mn[1] = min(mn[1], oy); mx[1] = max(mx[1], oy); mn[2] = min(mn[2], oz); mx[2] = max(mx[2], oz);
                any = true;
            }
        if (chunk == nch - 1 && threadIdx.x == 0) nout[b] = base + tot;
    }
    if (box && __any_sync(0xffffffffu, any)) {
#pragma unroll
        for (int k = 0; k < 3; k++) { mn[k] = __reduce_min_sync(0xffffffffu, mn[k]); mx[k] = __reduce_max_sync(0xffffffffu, mx[k]); }
        if ((threadIdx.x & 31) == 0)
            for (int k = 0; k < 3; k++) { atomicMin(&box[b].mn[k], mn[k]); atomicMax(&box[b].mx[k], mx[k]); }
    }
}

// fused = true: triangulate straight from the final disparity `disp16` and the frame's PruneStats (batch path);
// d_boxv must have been initialised (cvr_launch_chain_init): the cloud's bounding box lands there.
int cvr_launch_cloud(cvr_ctx* c, int B, int mode, const int16_t* disp16, const float* dispf, const uint8_t* mask,
                     const uint8_t* bgr, size_t bstep, size_t bfs, int channels, bool fused)
{
    const int W = c->sg.W, H = c->sg.H, N = W * H;
    const size_t selStride = cvr_sel_stride(c->maxN);
    CloudArgs a;
    a.disp16 = disp16; a.dispf = dispf; a.mask = mask; a.bgr = bgr; a.bstep = bstep; a.bfs = bfs; a.ch = channels;
    a.W = W; a.H = H; a.cl = c->cl; a.minDisp = c->d_flags; a.pstats = c->d_pstats;
    int* sel = c->d_sel;                                               // selection 0 of the chain
    if (!fused) { const int rc0 = cvr_launch_chain_init(c, B, false, true, false); if (rc0) return rc0; }   // stand-alone call: nobody prepared the box
    CVR_CUDA(c, cudaMemsetAsync(sel, 0, (size_t)B * selStride * sizeof(int), c->stream));
    const int nch = cvr_div_up(N, CLOUD_CHUNK);
    const dim3 g(cvr_frame_blocks(c, B, (size_t)nch, 1, 8), B);
    const double factor = c->cl.sigma_factor;
    if (mode == CVR_CLOUD_REPROJECT) {
        k_fill_int<<<cvr_div_up(B, 128), 128, 0, c->stream>>>(c->d_flags, B, 32767);
        LAUNCH_CHECK(c);
        k_min_s16<<<dim3(min(cvr_div_up(N, 256), 296), B), 256, 0, c->stream>>>(disp16, N, c->d_flags);
        LAUNCH_CHECK(c);
        k_cloud_select<CVR_CLOUD_REPROJECT, false><<<g, CVR_SEL_THREADS, 0, c->stream>>>(a, sel, selStride, c->d_cloud, (size_t)c->maxN, c->d_ncloud, c->d_boxv, factor);
    } else if (fused) {
        k_cloud_select<CVR_CLOUD_TRIANGULATE, true><<<g, CVR_SEL_THREADS, 0, c->stream>>>(a, sel, selStride, c->d_cloud, (size_t)c->maxN, c->d_ncloud, c->d_boxv, factor);
    } else {
        k_cloud_select<CVR_CLOUD_TRIANGULATE, false><<<g, CVR_SEL_THREADS, 0, c->stream>>>(a, sel, selStride, c->d_cloud, (size_t)c->maxN, c->d_ncloud, c->d_boxv, factor);
    }
    LAUNCH_CHECK(c);
    cvr_stage_mark(c, "cloud");
    return CVR_OK;
}

// ------------------------------------------------------------------------------------------------
// Allocation-free prune (cvr_prune_noalloc, include/cvrecon_b200.h).  Same four kernels as cvr_launch_prune, but
// the scratch they need (the min/max/sum record and the normalised 8-bit image) lives in a fixed array that is
// part of the library image, so the call never asks the driver for memory.
#define CVR_NOALLOC_STATS_BYTES 256
static __device__ __align__(256) unsigned char g_prune_scratch[CVR_NOALLOC_STATS_BYTES + CVR_NOALLOC_MAX_PIXELS];

// g_prune_scratch as written in host code is only a host-side handle for the symbol; the device address has to be
// asked for, once per device, with that device current.
#define CVR_NOALLOC_MAX_DEVICES 64
static unsigned char* g_prune_scratch_dev[CVR_NOALLOC_MAX_DEVICES];
static std::mutex g_prune_scratch_mtx;

extern "C" int cvr_prune_noalloc(const int16_t* d_disp, int width, int height, double sigma_factor,
                                 int16_t* d_pruned, float* d_disp_f32, uint8_t* d_mask, void* stream)
{
    if (!d_disp || width <= 0 || height <= 0) { cvr_set_thread_error("cvr_prune_noalloc: null disparity or non-positive size"); return CVR_ERR_INVALID_ARG; }
    const long long npix = (long long)width * height;
    if (npix > (long long)CVR_NOALLOC_MAX_PIXELS) {
        cvr_set_thread_error("cvr_prune_noalloc: width*height exceeds CVR_NOALLOC_MAX_PIXELS (the built-in scratch holds 4 Mi pixels)");
        return CVR_ERR_CAPACITY;
    }
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= CVR_NOALLOC_MAX_DEVICES) { cvr_set_thread_error("cvr_prune_noalloc: no current CUDA device (or device index >= 64)"); return CVR_ERR_NO_DEVICE; }
    unsigned char* scratch;
    {
        std::lock_guard<std::mutex> lk(g_prune_scratch_mtx);
        if (!g_prune_scratch_dev[dev]) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_prune_scratch);
            if (e != cudaSuccess || !p) {
                cvr_set_thread_error((std::string("cvr_prune_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e)).c_str());
                return CVR_ERR_CUDA;
            }
            g_prune_scratch_dev[dev] = (unsigned char*)p;
        }
        scratch = g_prune_scratch_dev[dev];
    }
    PruneStats* st = (PruneStats*)scratch;
    uint8_t* u8 = scratch + CVR_NOALLOC_STATS_BYTES;
    cudaStream_t s = (cudaStream_t)stream;
    const int N = (int)npix;
    const int gx = min(cvr_div_up(N, 256), 296);
    // the record keeps the previous call's min/max/sums: reset it first; the 8-bit image is fully rewritten before use
    (void)cudaGetLastError();                                      // an earlier, unrelated error of the caller is not this call's failure
    k_prune_init<<<1, 32, 0, s>>>(st, 1);
    k_prune_minmax<<<dim3(gx, 1), 256, 0, s>>>(d_disp, N, st);
    k_prune_norm<<<dim3(gx, 1), 256, 0, s>>>(d_disp, N, st, u8);
    k_prune_apply<<<dim3(gx, 1), 256, 0, s>>>(d_disp, u8, N, st, sigma_factor, d_pruned, d_disp_f32, d_mask);
    e = cudaGetLastError();
    if (e != cudaSuccess) { cvr_set_thread_error((std::string("cvr_prune_noalloc: kernel launch failed: ") + cudaGetErrorString(e)).c_str()); return CVR_ERR_CUDA; }
    return CVR_OK;
}
