// k_sgbm.cu -- semi-global block matching with cv::StereoSGBM (MODE_SGBM / MODE_HH) semantics,
// bit-exact including the int16 saturation regime.  Replaces the OpenCV calls made by
// Reconstruct3D::GenerateDisparityMap (cv_reconstruct/src/reconstruct/Reconstruct3D.cpp:72-84).
//
// Kernel set (one launch covers all B frames of a batch):
//   k_gray        BGR(A) -> gray                                   (A.1)
//   k_pix_hsum    clipped x-Sobel + Birchfield-Tomasi pixel cost,   (A.2, A.3 horizontal)
//                 horizontal block sum -> hsum volume
//   k_vsum        saturating vertical block sum + P2 bias -> C      (A.3 vertical)
//   k_path_h      horizontal path (-> or <-), one warp per row      (A.4 r=0 / 5th path)
//   k_path_v      the three vertical/diagonal paths fused, a        (A.4 r=1..3)
//                 thread-block cluster per frame marching over rows
//   (WTA fused into the last path kernel)                           (A.5 first half)
//   k_lrcheck     right-view map by packed atomicMin + LR check     (A.5 second half)
//   k_median3     3x3 median                                        (A.6)
//   k_ccl_*       speckle filter as union-find labelling            (A.7)
//
// Volumes are int16 [B][H][W1][Dp], Dp = D padded to a multiple of 64, so a warp owns one (x') column with
// DPL = Dp/32 consecutive disparities per lane; the min over disparities is one CREDUX warp reduction.
#include "cvr_internal.cuh"
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

#define FULL 0xffffffffu

// ------------------------------------------------------------------------------------------------
// A.1 gray
__global__ void k_gray(const uint8_t* __restrict__ src, size_t step, size_t fstride, int ch, uint8_t* __restrict__ dst,
                       int Wp, size_t dstFrameStride, int W, int H)
{
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
    if (x >= W) return;
    const uint8_t* s = src + (size_t)b * fstride + (size_t)y * step + (size_t)x * ch;
    uint8_t g;
    if (ch == 1) g = s[0];
    else g = (uint8_t)((3735 * s[0] + 19235 * s[1] + 9798 * s[2] + 16384) >> 15);
    dst[(size_t)b * dstFrameStride + (size_t)y * Wp + x] = g;
}

// BGR -> gray, four pixels per thread, both cameras in one launch (blockIdx.z = 2 * frame + camera): the 12 source bytes
// come from four aligned 32-bit words (funnel-shifted by the row's misalignment), the result is one 32-bit store (the gray
// pitch Wp is a multiple of 16).  A thread whose words would run past the end of the frame's last row takes the byte form.
__global__ void __launch_bounds__(128) k_gray3_x4(const uint8_t* __restrict__ srcL, size_t lstep, size_t lfs, const uint8_t* __restrict__ srcR, size_t rstep,
                                                  size_t rfs, uint8_t* __restrict__ dst, int Wp, size_t dstFrameStride, int W, int H)
{
    const int x = 4 * (blockIdx.x * blockDim.x + threadIdx.x), y = blockIdx.y, b = blockIdx.z >> 1, cam = blockIdx.z & 1;
    if (x >= W) return;
    const uint8_t* row = (cam ? srcR + (size_t)b * rfs + (size_t)y * rstep : srcL + (size_t)b * lfs + (size_t)y * lstep);
    const uint8_t* s = row + (size_t)x * 3;
    uint8_t* d = dst + (size_t)b * dstFrameStride + (size_t)cam * H * Wp + (size_t)y * Wp + x;
    unsigned g[4];
    if (x + 4 <= W && !(y == H - 1 && x + 8 > W)) {            // the 16 bytes read below stay inside this frame
        const unsigned sh = ((unsigned)(size_t)s & 3u) * 8u;
        const unsigned* wp = (const unsigned*)((size_t)s & ~(size_t)3);
        const unsigned a0 = wp[0], a1 = wp[1], a2 = wp[2], a3 = wp[3];
        const unsigned w0 = __funnelshift_r(a0, a1, sh), w1 = __funnelshift_r(a1, a2, sh), w2 = __funnelshift_r(a2, a3, sh);   // bytes 0..11
        const unsigned px[4][3] = { { w0 & 255u, (w0 >> 8) & 255u, (w0 >> 16) & 255u }, { w0 >> 24, w1 & 255u, (w1 >> 8) & 255u },
                                    { (w1 >> 16) & 255u, w1 >> 24, w2 & 255u }, { (w2 >> 8) & 255u, (w2 >> 16) & 255u, w2 >> 24 } };
#pragma unroll
        for (int k = 0; k < 4; k++) g[k] = (3735u * px[k][0] + 19235u * px[k][1] + 9798u * px[k][2] + 16384u) >> 15;
        *(unsigned*)d = g[0] | (g[1] << 8) | (g[2] << 16) | (g[3] << 24);
        return;
    }
    for (int k = 0; k < 4 && x + k < W; k++) d[k] = (uint8_t)((3735 * s[3 * k] + 19235 * s[3 * k + 1] + 9798 * s[3 * k + 2] + 16384) >> 15);
}

// ------------------------------------------------------------------------------------------------
// A.2 + A.3 (horizontal).  One CTA per (row, frame), one thread per disparity.
// Per-pixel packed descriptors, G channel (clipped Sobel) in the low half, raw intensity in the high half:
//   left : (U, U0, -U, -U1)     right: (V, V0, -V, -V1)
// so that the four BT differences are four VIADD.16x2 and the two max(.,.,0) are VIMNMX3.S16x2.RELU.
__device__ __forceinline__ unsigned pk_neg(unsigned a) { return __vneg2(a); }

// grid.x = H * nseg: a CTA owns one segment [s, e) of the computed columns of one row, which keeps the per-pixel
// descriptors of just that segment (plus the block-sum halo and the disparity range) in shared memory, so that many
// CTAs fit on an SM.  The horizontal block sum is a sliding window over j = s - SW2 .. e - 1 + SW2 with the column index
// clamped to [0, W1-1] exactly like OpenCV's clamped recurrence (closed form: provably unsaturated, see cvr_create).
// Right-image descriptors are read at a stride of two entries (a thread owns two adjacent disparities), which would put
// the 8 lanes of a 128-bit load phase on 4 bank groups.  They are therefore kept as two arrays, entries of even and of odd
// index: for a given column all threads read the same parity, i.e. consecutive entries of one array (conflict free), and
// of the four descriptors two neighbouring columns need, two are the same entry.
__global__ void k_pix_hsum(const uint8_t* __restrict__ gray, size_t grayFrameStride, int16_t* __restrict__ hsum, SgbmDev p, int nseg)
{
    extern __shared__ uint4 sm4[];
    const int y = blockIdx.x / nseg, seg = blockIdx.x % nseg, b = blockIdx.y, W = p.W, H = p.H;
    const int W1 = p.W1, Dp = p.Dp, SW2 = p.SW2;
    const int s = (int)(((long long)W1 * seg) / nseg), e = (int)(((long long)W1 * (seg + 1)) / nseg);
    const int ps = max(s - SW2, 0), pe = min(e - 1 + SW2, W1 - 1);                 // computed columns whose pixel cost is needed
    const int xl0 = ps + p.minX1, xl1 = pe + p.minX1;                              // left pixels needed: [xl0, xl1]
    const int xr0 = max(xl0 - p.minD - (p.D - 1), 0), xr1 = min(xl1 - p.minD, W - 1);   // right pixels needed: [xr0, xr1]
    const int nL = xl1 - xl0 + 1, nR = xr1 - xr0 + 1;
    uint4* Lpk = sm4;                                   // [nL]
    uint4* Rpk = sm4 + nL;                              // [nR]
    const int nRh = (nR + 3) >> 1;                      // entries per parity array (index -1 is never read: see offR below)
    uint4* const Rev = Rpk;                             // R[2j]
    uint4* const Rod = Rpk + nRh;                       // R[2j + 1]
    unsigned* aL = (unsigned*)(Rpk + 2 * nRh);          // [nL + 2]  packed (G | R << 16) of left pixels xl0-1 .. xl1+1
    unsigned* aR = aL + nL + 2;                         // [nR + 2]
    uint8_t* ringbuf = (uint8_t*)(aR + nR + 2);
    const int tid = threadIdx.x, nt = blockDim.x;
    const uint8_t* gl = gray + (size_t)b * grayFrameStride;
    const uint8_t* gr = gl + (size_t)H * p.Wp;
    const int yn = max(y - 1, 0), ys = min(y + 1, H - 1);
    const int ft = p.ftzero;
    const unsigned border = (unsigned)ft | ((unsigned)ft << 16);
    for (int i = tid; i < nL + 2 + nR + 2; i += nt) {
        const bool right = i >= nL + 2;
        const int x = right ? xr0 - 1 + (i - nL - 2) : xl0 - 1 + i;
        const uint8_t* g = right ? gr : gl;
        unsigned v = border;
        if (x > 0 && x < W - 1) {
            const uint8_t *r0 = g + (size_t)y * p.Wp + x, *rn = g + (size_t)yn * p.Wp + x, *rs = g + (size_t)ys * p.Wp + x;
            int a = 2 * ((int)r0[1] - (int)r0[-1]) + ((int)rn[1] - (int)rn[-1]) + ((int)rs[1] - (int)rs[-1]);
            a = min(max(a, -ft), ft) + ft;
            v = (unsigned)a | ((unsigned)r0[0] << 16);
        }
        (right ? aR : aL)[right ? i - nL - 2 : i] = v;      // columns outside the image never matter (guarded below)
    }
    __syncthreads();
    for (int i = tid; i < nL + nR; i += nt) {
        const bool right = i >= nL;
        const int j = right ? i - nL : i;
        const int x = (right ? xr0 : xl0) + j;
        const unsigned* a = (right ? aR : aL) + j + 1;
        const unsigned u = a[0];
        const unsigned ul = x > 0 ? (((u + a[-1]) >> 1) & 0x7fff7fffu) : u;
        const unsigned ur = x < W - 1 ? (((u + a[1]) >> 1) & 0x7fff7fffu) : u;
        const unsigned u0 = __vimin3_s16x2(ul, ur, u), u1 = __vimax3_s16x2(ul, ur, u);
        (right ? ((j & 1) ? Rod : Rev) : Lpk)[right ? (j >> 1) : j] = make_uint4(u, u0, pk_neg(u), pk_neg(u1));
    }
    __syncthreads();

    // ---- sliding block sums.  The CTA's Dp threads form two groups of Dp/2; each group owns one half of the segment and
    // each thread two adjacent disparities (d, d+1), whose costs, running sums and outputs travel as packed 16-bit pairs:
    // per pair of elements one broadcast load of the left descriptor, two of the right ones, one 16-bit ring read and
    // write, one 32-bit store -- the kernel is bound by shared-memory wavefronts (ncu: 98 % of the LSU data pipe).
    const int half = Dp >> 1, grp = tid / half, q = tid - grp * half, d = 2 * q;
    const int mid = s + ((e - s + 1) >> 1);
    const int hs_ = grp ? mid : s, he_ = grp ? e : mid;                        // this group's output columns [hs_, he_)
    if (hs_ >= he_) return;
    unsigned* out = (unsigned*)(hsum + (((size_t)b * H + y) * W1) * Dp) + q;    // pair q of column 0; a column is Dp/2 words
    if (d >= p.D) {                                                           // padding disparities (D is even: a pair is all padding or none)
        for (int xp = hs_; xp < he_; xp++) out[(size_t)xp * half] = 0u;
        return;
    }
    const int offR = p.minX1 - p.minD - d - xr0;        // right descriptor index of computed column xp for disparity d: xp + offR (d + 1: one less)
    const int offL = p.minX1 - xl0;
    auto cost1 = [](const uint4& l, const uint4& r) -> unsigned {
        const unsigned c0 = __vimax3_s16x2_relu(__vadd2(l.x, r.w), __vadd2(r.y, l.z), 0u);   // max(0, u - v1, v0 - u)
        const unsigned c1 = __vimax3_s16x2_relu(__vadd2(r.x, l.w), __vadd2(l.y, r.z), 0u);   // max(0, v - u1, u0 - v)
        const unsigned m = __vmins2(c0, c1);
        return (m & 0xffffu) + (m >> 18);                                     // Sobel-channel cost + (intensity cost >> 2), <= 255
    };
    auto pixcost2 = [&](int xp) -> unsigned {                                 // the costs of (d, d + 1) as a packed 16-bit pair
        const uint4 l = Lpk[xp + offL];
        const int i = xp + offR;                                              // right descriptor of disparity d; d + 1 uses i - 1
        const uint4 ra = (i & 1) ? Rod[i >> 1] : Rev[i >> 1], rb = (i & 1) ? Rev[i >> 1] : Rod[(i >> 1) - 1];
        return cost1(l, ra) | (cost1(l, rb) << 16);
    };
    // ring of the last bs + 1 cost pairs per thread: p.ring slots (a power of two), slot s of this thread at ring[s * half]
    unsigned* ring = (unsigned*)ringbuf + (size_t)grp * p.ring * half + q;
    const int mask = p.ring - 1, bs = p.bs;
    unsigned run = 0;                                                         // two running sums (each < 2^15: provably unsaturated)
    const int j0 = hs_ - SW2, j1 = he_ - 1 + SW2;
    // warm-up: the first bs - 1 window elements (clamped at the left image border), nothing to emit or to drop yet
    int j = j0, ws = 0;                                                       // ws: the ring slot written next
    for (; j < hs_ + SW2; j++) {
        const unsigned cp = pixcost2(min(max(j, 0), W1 - 1));
        run += cp;
        ring[ws * half] = cp;
        ws = (ws + 1) & mask;
    }
    // steady state: one new element, one output; from the second output on the element leaving the window is dropped.
    unsigned* o = out + (size_t)hs_ * half;
    {
        const unsigned cp = pixcost2(min(j, W1 - 1));
        run += cp;
        ring[ws * half] = cp; ws = (ws + 1) & mask;
        *o = run; o += half; j++;
    }
    const int jmid = min(j1, W1 - 1);
    auto emit = [&](unsigned cp) {                                            // one steady-state step with the cost pair of element j
        run = run + cp - ring[((ws - bs) & mask) * half];                     // the element written bs steps ago leaves the window
        ring[ws * half] = cp;
        *o = run; o += half;
        ws = (ws + 1) & mask;
    };
    if (j <= jmid && ((j + offR) & 1)) { emit(pixcost2(j)); j++; }            // make the right index even
    {
        const uint4* lp = Lpk + (j + offL);
        const uint4* ev = Rev + ((j + offR) >> 1);
        const uint4* od = Rod + ((j + offR) >> 1);
        for (; j + 1 <= jmid; j += 2, lp += 2, ev++, od++) {                  // two columns per turn: right indices 2m and 2m + 1
            const uint4 l0 = lp[0], l1 = lp[1], e0 = ev[0], o0 = od[-1], o1 = od[0];
            emit(cost1(l0, e0) | (cost1(l0, o0) << 16));                      // column j    : d -> R[2m],     d + 1 -> R[2m - 1]
            emit(cost1(l1, o1) | (cost1(l1, e0) << 16));                      // column j + 1: d -> R[2m + 1], d + 1 -> R[2m]
        }
    }
    if (j <= jmid) { emit(pixcost2(j)); j++; }
    if (j <= j1) {                                                            // right image border: the last column repeats
        const unsigned cl = pixcost2(W1 - 1);
        for (; j <= j1; j++) emit(cl);
    }
}

// ------------------------------------------------------------------------------------------------
// k_pix_hsum4: as k_pix_hsum, FOUR adjacent disparities per thread (two packed pairs).  The per-column work that does not
// depend on the disparity (the broadcast load of the left descriptor, loop and address arithmetic, the ring slot) is shared
// by twice as many disparities, and the right-image descriptors are reused from column to column: the thread of disparities
// d .. d+3 needs R[i], R[i-1], R[i-2], R[i-3] (i = right pixel of disparity d) and one column later R[i+1], R[i], R[i-1],
// R[i-2] -- ONE new descriptor per column, kept in four registers whose roles rotate (the column loop is unrolled by four so the
// rotation is register renaming).  The descriptors are stored as four arrays by index mod 4: at a given column every thread
// reads the same array at consecutive entries (conflict-free 16-byte loads).  Shared-memory wavefronts per column and 64
// disparities: 9 -> 4.5; the CTA has Dp / 2 threads (two groups of Dp / 4, each owning half of the segment).
__global__ void k_pix_hsum4(const uint8_t* __restrict__ gray, size_t grayFrameStride, int16_t* __restrict__ hsum, SgbmDev p, int nseg)
{
    extern __shared__ uint4 sm4[];
    const int y = blockIdx.x / nseg, seg = blockIdx.x % nseg, b = blockIdx.y, W = p.W, H = p.H;
    const int W1 = p.W1, Dp = p.Dp, SW2 = p.SW2;
    const int s = (int)(((long long)W1 * seg) / nseg), e = (int)(((long long)W1 * (seg + 1)) / nseg);
    const int ps = max(s - SW2, 0), pe = min(e - 1 + SW2, W1 - 1);                 // computed columns whose pixel cost is needed
    const int xl0 = ps + p.minX1, xl1 = pe + p.minX1;                              // left pixels needed: [xl0, xl1]
    const int xr0 = max(xl0 - p.minD - (p.D - 1), 0), xr1 = min(xl1 - p.minD, W - 1);   // right pixels needed: [xr0, xr1]
    const int nL = xl1 - xl0 + 1, nR = xr1 - xr0 + 1;
    const int nR4 = (nR + 3) / 4 + 1;                   // entries per residue array
    uint4* Lpk = sm4;                                   // [nL]
    uint4* R4 = sm4 + nL;                               // [4][nR4]: R4[k][j] = descriptor of right pixel xr0 + 4 j + k
    unsigned* aL = (unsigned*)(R4 + 4 * nR4);           // [nL + 2]  packed (G | R << 16) of left pixels xl0-1 .. xl1+1
    unsigned* aR = aL + nL + 2;                         // [nR + 2]
    // the ring of the last bs + 1 cost quadruples (one byte per cost: <= 2 * ftzero + 63 <= 189) reuses the unpacked-pixel staging,
    // which is dead once the descriptors are built
    unsigned* ringbuf = aL;
    const int tid = threadIdx.x, nt = blockDim.x;
    const uint8_t* gl = gray + (size_t)b * grayFrameStride;
    const uint8_t* gr = gl + (size_t)H * p.Wp;
    const int yn = max(y - 1, 0), ys = min(y + 1, H - 1);
    const int ft = p.ftzero;
    const unsigned border = (unsigned)ft | ((unsigned)ft << 16);
    for (int i = tid; i < nL + 2 + nR + 2; i += nt) {
        const bool right = i >= nL + 2;
        const int x = right ? xr0 - 1 + (i - nL - 2) : xl0 - 1 + i;
        const uint8_t* g = right ? gr : gl;
        unsigned v = border;
        if (x > 0 && x < W - 1) {
            const uint8_t *r0 = g + (size_t)y * p.Wp + x, *rn = g + (size_t)yn * p.Wp + x, *rs = g + (size_t)ys * p.Wp + x;
            int a = 2 * ((int)r0[1] - (int)r0[-1]) + ((int)rn[1] - (int)rn[-1]) + ((int)rs[1] - (int)rs[-1]);
            a = min(max(a, -ft), ft) + ft;
            v = (unsigned)a | ((unsigned)r0[0] << 16);
        }
        (right ? aR : aL)[right ? i - nL - 2 : i] = v;      // columns outside the image never matter (guarded below)
    }
    __syncthreads();
    for (int i = tid; i < nL + nR; i += nt) {
        const bool right = i >= nL;
        const int j = right ? i - nL : i;
        const int x = (right ? xr0 : xl0) + j;
        const unsigned* a = (right ? aR : aL) + j + 1;
        const unsigned u = a[0];
        const unsigned ul = x > 0 ? (((u + a[-1]) >> 1) & 0x7fff7fffu) : u;
        const unsigned ur = x < W - 1 ? (((u + a[1]) >> 1) & 0x7fff7fffu) : u;
        const unsigned u0 = __vimin3_s16x2(ul, ur, u), u1 = __vimax3_s16x2(ul, ur, u);
        const uint4 dsc = make_uint4(u, u0, pk_neg(u), pk_neg(u1));
        if (right) R4[(j & 3) * nR4 + (j >> 2)] = dsc; else Lpk[j] = dsc;
    }
    __syncthreads();

    const int quarter = Dp >> 2, grp = tid / quarter, q = tid - grp * quarter, d = 4 * q;
    const int mid = s + ((e - s + 1) >> 1);
    const int hs_ = grp ? mid : s, he_ = grp ? e : mid;                        // this group's output columns [hs_, he_)
    if (hs_ >= he_) return;
    const int half = Dp >> 1;                                                  // words per output column
    uint2* out = (uint2*)((unsigned*)(hsum + (((size_t)b * H + y) * W1) * Dp) + 2 * q);   // this thread's two words of column 0
    if (d >= p.D) {                                                            // padding disparities (D is a multiple of 16: all four or none)
        for (int xp = hs_; xp < he_; xp++) out[(size_t)xp * (half / 2)] = make_uint2(0u, 0u);
        return;
    }
    const int offR = p.minX1 - p.minD - d - xr0;        // right descriptor index of computed column xp for disparity d: xp + offR (d + t: t less)
    const int offL = p.minX1 - xl0;
    auto cost1 = [](const uint4& l, const uint4& r) -> unsigned {
        const unsigned c0 = __vimax3_s16x2_relu(__vadd2(l.x, r.w), __vadd2(r.y, l.z), 0u);   // max(0, u - v1, v0 - u)
        const unsigned c1 = __vimax3_s16x2_relu(__vadd2(r.x, l.w), __vadd2(l.y, r.z), 0u);   // max(0, v - u1, u0 - v)
        const unsigned m = __vmins2(c0, c1);
        return (m & 0xffffu) + (m >> 18);                                     // Sobel-channel cost + (intensity cost >> 2), <= 255
    };
    auto rdesc = [&](int i) -> uint4 { return R4[(i & 3) * nR4 + (i >> 2)]; };
    auto pixcost4 = [&](int xp) -> uint2 {                                     // generic form: the four descriptors fetched afresh
        const uint4 l = Lpk[xp + offL];
        const int i = xp + offR;
        return make_uint2(cost1(l, rdesc(i)) | (cost1(l, rdesc(i - 1)) << 16), cost1(l, rdesc(i - 2)) | (cost1(l, rdesc(i - 3)) << 16));
    };
    // ring of the last bs + 1 cost quadruples per thread: p.ring slots (a power of two), slot s of this thread at ring[s * quarter]
    unsigned* ring = ringbuf + (size_t)grp * p.ring * quarter + q;
    auto pack8 = [](const uint2 cp) -> unsigned { return __byte_perm(cp.x, cp.y, 0x6420); };          // four costs, a byte each
    const int mask = p.ring - 1, bs = p.bs;
    unsigned run0 = 0, run1 = 0;                                               // four running sums (each < 2^15: provably unsaturated)
    const int j0 = hs_ - SW2, j1 = he_ - 1 + SW2;
    int j = j0, ws = 0;                                                        // ws: the ring slot written next
    for (; j < hs_ + SW2; j++) {                                               // warm-up: the first bs - 1 window elements
        const uint2 cp = pixcost4(min(max(j, 0), W1 - 1));
        run0 += cp.x; run1 += cp.y;
        ring[ws * quarter] = pack8(cp);
        ws = (ws + 1) & mask;
    }
    uint2* o = out + (size_t)hs_ * (half / 2);
    {
        const uint2 cp = pixcost4(min(j, W1 - 1));
        run0 += cp.x; run1 += cp.y;
        ring[ws * quarter] = pack8(cp); ws = (ws + 1) & mask;
        *o = make_uint2(run0, run1); o += half / 2; j++;
    }
    const int jmid = min(j1, W1 - 1);
    auto emit = [&](const uint2 cp) {                                          // one steady-state step with the cost quadruple of element j
        const unsigned old = ring[((ws - bs) & mask) * quarter];               // the element written bs steps ago leaves the window
        run0 = run0 + cp.x - __byte_perm(old, 0u, 0x4140); run1 = run1 + cp.y - __byte_perm(old, 0u, 0x4342);
        ring[ws * quarter] = pack8(cp);
        *o = make_uint2(run0, run1); o += half / 2;
        ws = (ws + 1) & mask;
    };
    while (j <= jmid && ((j + offR) & 3) != 0) { emit(pixcost4(j)); j++; }    // make the right index of disparity d a multiple of four
    if (j + 3 <= jmid) {
        const int i0 = j + offR;                                               // multiple of 4
        uint4 ra = rdesc(i0), rb = rdesc(i0 - 1), rc = rdesc(i0 - 2), rd = rdesc(i0 - 3);   // R[i0], R[i0-1], R[i0-2], R[i0-3]
        const uint4* lp = Lpk + (j + offL);
        const uint4* r1 = R4 + 1 * nR4 + (i0 >> 2);                            // R[i0 + 1], then + 4 per turn
        const uint4* r2 = R4 + 2 * nR4 + (i0 >> 2);
        const uint4* r3 = R4 + 3 * nR4 + (i0 >> 2);
        const uint4* r0n = R4 + 0 * nR4 + (i0 >> 2) + 1;                       // R[i0 + 4]
        for (;;) {
            const uint4 l0 = lp[0], l1 = lp[1], l2 = lp[2], l3 = lp[3];
            emit(make_uint2(cost1(l0, ra) | (cost1(l0, rb) << 16), cost1(l0, rc) | (cost1(l0, rd) << 16)));     // column j    : d..d+3 <- R[i0], R[i0-1], R[i0-2], R[i0-3]
            rd = *r1;                                                                                            // R[i0+1]
            emit(make_uint2(cost1(l1, rd) | (cost1(l1, ra) << 16), cost1(l1, rb) | (cost1(l1, rc) << 16)));     // column j + 1
            rc = *r2;                                                                                            // R[i0+2]
            emit(make_uint2(cost1(l2, rc) | (cost1(l2, rd) << 16), cost1(l2, ra) | (cost1(l2, rb) << 16)));     // column j + 2
            rb = *r3;                                                                                            // R[i0+3]
            emit(make_uint2(cost1(l3, rb) | (cost1(l3, rc) << 16), cost1(l3, rd) | (cost1(l3, ra) << 16)));     // column j + 3
            j += 4;
            if (j + 3 > jmid) break;
            ra = *r0n;                                                                                           // R[i0+4]: the next turn's R[i0]
            // roles for the next turn: R[i0'] = ra, R[i0'-1] = rb (R[i0+3]), R[i0'-2] = rc (R[i0+2]), R[i0'-3] = rd (R[i0+1])
            lp += 4; r1++; r2++; r3++; r0n++;
        }
    }
    for (; j <= jmid; j++) emit(pixcost4(j));
    if (j <= j1) {                                                            // right image border: the last column repeats
        const uint2 cl = pixcost4(W1 - 1);
        for (; j <= j1; j++) emit(cl);
    }
}

// ------------------------------------------------------------------------------------------------
// A.3 (vertical): C[0] = s16(P2) (+) hsum[0]*(SH2+1) (+) hsum[1..SH2];  C[y] = sat(sat(C[y-1] - old) + new)
// (x' = 0: sat(sat(C[y-1] + new) - old)).
// One thread per (frame, x', 4 disparities) marching down the rows.
__device__ __forceinline__ int sat16i(int v) { return min(max(v, -32768), 32767); }
__device__ __forceinline__ int lo16(unsigned w) { return (int)(short)(w & 0xffffu); }
__device__ __forceinline__ int hi16(unsigned w) { return ((int)w) >> 16; }
__device__ __forceinline__ unsigned pack16(int lo, int hi) { return ((unsigned)lo & 0xffffu) | ((unsigned)hi << 16); }

// Saturating int16 add / subtract of a NON-NEGATIVE operand, two lanes per word: only one side can saturate, which leaves
// three native packed instructions each (the generic __vaddss2 / __vsubss2 expand to six).
//   sat(a + n) = min(a, 32767 - n) + n,  32767 - n = n ^ 0x7fff     sat(a - b) = max(a, b - 32768) - b,  b - 32768 = b ^ 0x8000
__device__ __forceinline__ unsigned sat_add_pos2(unsigned a, unsigned n) { return __vadd2(__vmins2(a, n ^ 0x7fff7fffu), n); }
__device__ __forceinline__ unsigned sat_sub_pos2(unsigned a, unsigned b) { return __vsub2(__vmaxs2(a, b ^ 0x80008000u), b); }

// One thread per (frame, x', 8 disparities) marching down the rows.  The "old" row (y - SH2 - 1) was read as the "new"
// row bs rows earlier and is simply read again -- from the L2, IF the rows between the two reads fit there.  With one row in
// flight per thread and full occupancy they do not (300 k resident threads x 28 rows x 16 B = 136 MB: ncu showed 288 MB of DRAM
// traffic per KITTI frame for 214 MB algorithmic).  Here a thread keeps the loads of P = 16 rows in flight (2 P uint4 registers:
// 152 registers per thread, so only 12 warps per SM are resident): the same bytes are in flight with a fifth of the columns,
// and the windows of all resident threads, (28 + P) rows each, shrink to ~40 MB and stay in the L2.  Measured on B200, KITTI
// batch 110 (us per frame): P = 1 48.3, P = 8 40.2, P = 16 35.9, P = 24 43.4, P = 32 53.5; capping the occupancy further with
// a shared-memory pad (CVR_VSUM_OCC) changes nothing at P = 16 and hurts at P = 8.
template <int P>
__global__ void __launch_bounds__(128) k_vsum(const int16_t* __restrict__ hsum, int16_t* __restrict__ C, SgbmDev p, int B, int* flags)
{
    const size_t rowWords = (size_t)p.W1 * p.Dp / 8;          // uint4 words per row (Dp is a multiple of 32)
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= rowWords * B) return;
    const int b = (int)(t / rowWords);
    const size_t i = t % rowWords;
    const int H = p.H, SH2 = p.SH2;
    const uint4* hs = (const uint4*)hsum + (size_t)b * H * rowWords + i;
    uint4* cs = (uint4*)C + (size_t)b * H * rowWords + i;
    // OpenCV evaluates the first computed column (x' = 0) as (Cprev + new) - old and all others as (Cprev - old) + new;
    // in saturating int16 lanes the two differ once C sticks at 32767 (pinned by tests/golden/satdrift_d32_bs27.npz).
    const bool first = i < (size_t)p.Dp / 8;
    int c[8];
#pragma unroll
    for (int j = 0; j < 8; j++) c[j] = p.P2s;
    for (int k = 0; k <= SH2; k++) {
        const uint4 h = hs[(size_t)min(k, H - 1) * rowWords];
        const int sc = (k == 0) ? SH2 + 1 : 1;
        const int hv[8] = { lo16(h.x), hi16(h.x), lo16(h.y), hi16(h.y), lo16(h.z), hi16(h.z), lo16(h.w), hi16(h.w) };
#pragma unroll
        for (int j = 0; j < 8; j++) c[j] = sat16i(c[j] + sat16i(hv[j] * sc));
    }
    // hsum is a sum of non-negative pixel costs, so the row recurrence only ever adds / subtracts non-negative values
    unsigned w[4] = { pack16(c[0], c[1]), pack16(c[2], c[3]), pack16(c[4], c[5]), pack16(c[6], c[7]) };
    unsigned negw = w[0] | w[1] | w[2] | w[3];
    cs[0] = make_uint4(w[0], w[1], w[2], w[3]);
    for (int y0 = 1; y0 < H; y0 += P) {
        uint4 ho[P], hn[P];
#pragma unroll
        for (int k = 0; k < P; k++) {                                   // all loads of the next P rows first
            const int y = min(y0 + k, H - 1);
            ho[k] = __ldcs(hs + (size_t)max(y - SH2 - 1, 0) * rowWords);  // last use of that row: evict-first
            hn[k] = hs[(size_t)min(y + SH2, H - 1) * rowWords];           // read again bs rows later
        }
#pragma unroll
        for (int k = 0; k < P; k++) {
            if (y0 + k >= H) break;
            const unsigned o[4] = { ho[k].x, ho[k].y, ho[k].z, ho[k].w }, n[4] = { hn[k].x, hn[k].y, hn[k].z, hn[k].w };
#pragma unroll
            for (int j = 0; j < 4; j++) {
                w[j] = first ? sat_sub_pos2(sat_add_pos2(w[j], n[j]), o[j]) : sat_add_pos2(sat_sub_pos2(w[j], o[j]), n[j]);
                negw |= w[j];
            }
            __stcs(cs + (size_t)(y0 + k) * rowWords, make_uint4(w[0], w[1], w[2], w[3]));   // streamed out, not re-read here
        }
    }
    if (negw & 0x80008000u) atomicOr(&flags[b], 1);   // a negative C exists in this frame: the packed path kernels take their exact t + C form
}

// ------------------------------------------------------------------------------------------------
// Path-cost update of A.4 for one disparity, int32 lanes with the int16 saturations made explicit:
//   m = min(Lp[d], sat(Lp[d-1]+P1s), sat(Lp[d+1]+P1s), delta);  L = sat(sat(m - delta) + C)
// min(sat(a+k), sat(b+k)) == sat(min(a,b)+k) (monotone); the upper clamp of that term is absorbed by
// min(., Lp[d]) (Lp[d] <= 32767); m <= delta so m - delta <= 0 and only lower clamps remain.
__device__ __forceinline__ int path_cost(int lm1, int l0, int lp1, int delta, int P1s, int c)
{
    int n = max(min(lm1, lp1) + P1s, -32768);
    int m = min(min(n, l0), delta);
    int t = max(m - delta, -32768);
    return max(t + c, -32768);
}

template <int DPL>
__device__ __forceinline__ void load_vec(const int16_t* p, int (&v)[DPL])
{
    // DPL consecutive int16 at p (p is DPL*2-byte aligned; DPL in {2,4,6,8})
    const unsigned* w = (const unsigned*)p;
    if (DPL == 2) { unsigned a = w[0]; v[0] = lo16(a); v[1] = hi16(a); }
    else if (DPL == 4) { uint2 a = *(const uint2*)p; v[0] = lo16(a.x); v[1] = hi16(a.x); v[2] = lo16(a.y); v[3] = hi16(a.y); }
    else if (DPL == 8) { uint4 a = *(const uint4*)p; v[0] = lo16(a.x); v[1] = hi16(a.x); v[2] = lo16(a.y); v[3] = hi16(a.y);
                         v[4] = lo16(a.z); v[5] = hi16(a.z); v[6] = lo16(a.w); v[7] = hi16(a.w); }
    else {
#pragma unroll
        for (int i = 0; i < DPL / 2; i++) { unsigned a = w[i]; v[2 * i] = lo16(a); v[2 * i + 1] = hi16(a); }
    }
}
template <int DPL>
__device__ __forceinline__ void load_vec_cg(const int16_t* p, int (&v)[DPL])
{
    const unsigned* w = (const unsigned*)p;
    if (DPL == 4) { uint2 a = __ldcg((const uint2*)p); v[0] = lo16(a.x); v[1] = hi16(a.x); v[2] = lo16(a.y); v[3] = hi16(a.y); }
    else if (DPL == 8) { uint4 a = __ldcg((const uint4*)p); v[0] = lo16(a.x); v[1] = hi16(a.x); v[2] = lo16(a.y); v[3] = hi16(a.y);
                         v[4] = lo16(a.z); v[5] = hi16(a.z); v[6] = lo16(a.w); v[7] = hi16(a.w); }
    else {
#pragma unroll
        for (int i = 0; i < DPL / 2; i++) { unsigned a = __ldcg(w + i); v[2 * i] = lo16(a); v[2 * i + 1] = hi16(a); }
    }
}
template <int DPL>
__device__ __forceinline__ void store_vec(int16_t* p, const int (&v)[DPL])
{
    unsigned* w = (unsigned*)p;
    if (DPL == 4) *(uint2*)p = make_uint2(pack16(v[0], v[1]), pack16(v[2], v[3]));
    else if (DPL == 8) *(uint4*)p = make_uint4(pack16(v[0], v[1]), pack16(v[2], v[3]), pack16(v[4], v[5]), pack16(v[6], v[7]));
    else {
#pragma unroll
        for (int i = 0; i < DPL / 2; i++) w[i] = pack16(v[2 * i], v[2 * i + 1]);
    }
}

// A.5 (first half) for one column held by a warp: S[i] is the aggregated cost of disparity lane*DPL+i.
// Writes aux = { (minS << 16) | best, (S[best-1] << 16) | (S[best+1] & 0xffff) } for accepted pixels ({0x7fffffff, 0}
// otherwise); the sub-pixel division itself is done one thread per pixel in k_lrcheck.
template <int DPL>
__device__ __forceinline__ void wta_column(const int (&S)[DPL], int lane, const SgbmDev& p, int16_t* rawPix, int2* auxPix)
{
    (void)rawPix;
    const int d0 = lane * DPL;
    int key = 0x7fffffff;
#pragma unroll
    for (int i = 0; i < DPL; i++)
        if (d0 + i < p.D) key = min(key, (int)(((unsigned)S[i] << 16) | (unsigned)(d0 + i)));
    key = __reduce_min_sync(FULL, key);
    const int minS = key >> 16, best = key & 0xffff;
    bool bad = false;
    const int lim = minS * 100, mul = 100 - p.uniq;
#pragma unroll
    for (int i = 0; i < DPL; i++) {
        const int d = d0 + i;
        if (d < p.D && S[i] * mul < lim && abs(best - d) > 1) bad = true;
    }
    const bool reject = __any_sync(FULL, bad) || (minS == CVR_MAX_COST);
    int2 aux = make_int2(0x7fffffff, 0);
    if (!reject) {               // warp-uniform
        int a = 0, c = 0;
        if (0 < best && best < p.D - 1) {
#pragma unroll
            for (int i = 0; i < DPL; i++) {
                a += (d0 + i == best - 1) ? S[i] : 0;
                c += (d0 + i == best + 1) ? S[i] : 0;
            }
            a = __reduce_add_sync(FULL, a);
            c = __reduce_add_sync(FULL, c);
        }
        aux = make_int2((int)(((unsigned)minS << 16) | (unsigned)best), (int)(((unsigned)a << 16) | ((unsigned)c & 0xffffu)));
    }
    if (lane == 0) *auxPix = aux;
}

// ------------------------------------------------------------------------------------------------
// Horizontal path: one warp per (frame, row); sequential over x', state in registers.
//   REVERSE = false:  ->  (A.4 r = 0)        REVERSE = true:  <-  (MODE_SGBM 5th path / MODE_HH pass-2 r = 0)
//   accumulate: S = sat(S + L) (else S = L, the first path written);  WTA: winner-take-all on the final S.
template <int DPL, bool REVERSE, bool WTA>
__global__ void __launch_bounds__(128) k_path_h(const int16_t* __restrict__ C, int16_t* __restrict__ S, int16_t* __restrict__ raw,
                                                int2* __restrict__ aux, SgbmDev p, int B, int accumulate, int storeS, int* flags)
{
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= B * p.H) return;
    const int b = warp / p.H, y = warp % p.H;
    const int W1 = p.W1, Dp = p.Dp, D = p.D;
    const size_t rowBase = (((size_t)b * p.H + y) * W1) * Dp + (size_t)lane * DPL;
    const int16_t* Crow = C + rowBase;
    int16_t* Srow = S + rowBase;
    const int d0 = lane * DPL;
    int Lp[DPL];
#pragma unroll
    for (int i = 0; i < DPL; i++) Lp[i] = (d0 + i < D) ? 0 : CVR_MAX_COST;
    int minPrev = 0;
    const int P1s = p.P1s, P2 = p.P2;
    int xp = REVERSE ? W1 - 1 : 0;
    const int dx = REVERSE ? -1 : 1;
    int cN[DPL], sN[DPL];
    load_vec<DPL>(Crow + (size_t)xp * Dp, cN);
    if (accumulate) load_vec<DPL>(Srow + (size_t)xp * Dp, sN);
    for (int step = 0; step < W1; step++, xp += dx) {
        int c[DPL], s[DPL];
#pragma unroll
        for (int i = 0; i < DPL; i++) { c[i] = cN[i]; s[i] = accumulate ? sN[i] : 0; }
        if (step + 1 < W1) {
            load_vec<DPL>(Crow + (size_t)(xp + dx) * Dp, cN);
            if (accumulate) load_vec<DPL>(Srow + (size_t)(xp + dx) * Dp, sN);
        }
        const int delta = (int)(short)(P2 + minPrev);
        int left = __shfl_up_sync(FULL, Lp[DPL - 1], 1);
        int right = __shfl_down_sync(FULL, Lp[0], 1);
        if (lane == 0) left = CVR_MAX_COST;
        if (lane == 31) right = CVR_MAX_COST;
        int L[DPL];
        int mn = CVR_MAX_COST;
#pragma unroll
        for (int i = 0; i < DPL; i++) {
            const int lm1 = (i == 0) ? left : Lp[i - 1];
            const int lp1 = (i == DPL - 1) ? right : Lp[i + 1];
            int v = path_cost(lm1, Lp[i], lp1, delta, P1s, c[i]);
            if (d0 + i >= D) v = CVR_MAX_COST;
            L[i] = v;
            mn = min(mn, v);
            s[i] = sat16i(s[i] + v);
        }
        minPrev = __reduce_min_sync(FULL, mn);
        if (minPrev < 0 && lane == 0) atomicOr(&flags[b], 2);       // tells the packed kernels that S may hold negative values
#pragma unroll
        for (int i = 0; i < DPL; i++) Lp[i] = L[i];
        if (!WTA || storeS) store_vec<DPL>(Srow + (size_t)xp * Dp, s);
        if (WTA) wta_column<DPL>(s, lane, p, raw + ((size_t)b * p.H + y) * p.W + xp + p.minX1, aux + ((size_t)b * p.H + y) * W1 + xp);
    }
}

// ------------------------------------------------------------------------------------------------
// Vertical path state: lr [B][2][W1+2][3][Dp] int16, minlr [B][2][W1+2][4] int.  Border columns and the
// buffer read by the first row hold the "zero state": L = 0 for d < D (MAX_COST in the padding), minL = 0.
__global__ void k_lr_init(int16_t* lr, int* minlr, SgbmDev p, int B)
{
    const size_t n = (size_t)B * 2 * (p.W1 + 2) * 3 * p.Dp;
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) lr[i] = ((int)(i % p.Dp) < p.D) ? 0 : CVR_MAX_COST;
    const size_t m = (size_t)B * 2 * (p.W1 + 2) * 4;
    if (i < m) minlr[i] = 0;
}

// The three paths that cross rows (A.4 r = 1: from (x'-1), r = 2: from (x'), r = 3: from (x'+1) of the previous
// row), fused.  A cluster of CL CTAs owns one frame; CTA `rank` owns the column strip [c0, c1); the per-row
// hand-over of the strip edges goes through the L2-resident state arrays, ordered by one cluster barrier per row.
// UP = false: rows top-down (pass 1);  UP = true: bottom-up (MODE_HH pass 2).  S = sat(S + L1), + L2, + L3.
template <int DPL, bool UP, bool WTA, int NT>
__global__ void __launch_bounds__(NT) k_path_v(const int16_t* __restrict__ C, int16_t* __restrict__ S, int16_t* lr, int* minlr,
                                               int16_t* __restrict__ raw, int2* __restrict__ aux, SgbmDev p, int CL, int storeS, int* flags)
{
    const int b = blockIdx.y, rank = blockIdx.x;
    const int W1 = p.W1, Dp = p.Dp, D = p.D, H = p.H;
    const int c0 = (int)(((long long)W1 * rank) / CL), c1 = (int)(((long long)W1 * (rank + 1)) / CL);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, NW = NT / 32;
    const int d0 = lane * DPL;
    const size_t colStride = (size_t)3 * Dp, bufStride = (size_t)(W1 + 2) * colStride;
    int16_t* lrb = lr + (size_t)b * 2 * bufStride + (size_t)lane * DPL;
    int* mlb = minlr + (size_t)b * 2 * (W1 + 2) * 4;
    const int P1s = p.P1s, P2 = p.P2;
    for (int t = 0; t < H; t++) {
        const int y = UP ? H - 1 - t : t;
        const int cur = t & 1, prev = cur ^ 1;
        const size_t rowBase = (((size_t)b * H + y) * W1) * Dp + (size_t)lane * DPL;
        const int16_t* lrPrev = lrb + (size_t)prev * bufStride;
        int16_t* lrCur = lrb + (size_t)cur * bufStride;
        const int* mlPrev = mlb + (size_t)prev * (W1 + 2) * 4;
        int* mlCur = mlb + (size_t)cur * (W1 + 2) * 4;
        for (int xp = c0 + warp; xp < c1; xp += NW) {
            int c[DPL], s[DPL];
            load_vec<DPL>(C + rowBase + (size_t)xp * Dp, c);
            load_vec<DPL>(S + rowBase + (size_t)xp * Dp, s);
            int mins[3];
#pragma unroll
            for (int r = 0; r < 3; r++) {
                const int sx = xp + r;                      // state column index (x'+1) of the source: x'-1, x', x'+1
                int Lp[DPL];
                load_vec_cg<DPL>(lrPrev + (size_t)sx * colStride + (size_t)r * Dp, Lp);
                const int delta = (int)(short)(P2 + __ldcg(mlPrev + sx * 4 + r));
                int left = __shfl_up_sync(FULL, Lp[DPL - 1], 1);
                int right = __shfl_down_sync(FULL, Lp[0], 1);
                if (lane == 0) left = CVR_MAX_COST;
                if (lane == 31) right = CVR_MAX_COST;
                int L[DPL];
                int mn = CVR_MAX_COST;
#pragma unroll
                for (int i = 0; i < DPL; i++) {
                    const int lm1 = (i == 0) ? left : Lp[i - 1];
                    const int lp1 = (i == DPL - 1) ? right : Lp[i + 1];
                    int v = path_cost(lm1, Lp[i], lp1, delta, P1s, c[i]);
                    if (d0 + i >= D) v = CVR_MAX_COST;
                    L[i] = v;
                    mn = min(mn, v);
                    s[i] = sat16i(s[i] + v);
                }
                mins[r] = __reduce_min_sync(FULL, mn);
                if (mins[r] < 0 && lane == 0) atomicOr(&flags[b], 2);
                store_vec<DPL>(lrCur + (size_t)(xp + 1) * colStride + (size_t)r * Dp, L);
            }
            if (lane < 3) mlCur[(xp + 1) * 4 + lane] = (lane == 0) ? mins[0] : (lane == 1 ? mins[1] : mins[2]);
            if (!WTA || storeS) store_vec<DPL>(S + rowBase + (size_t)xp * Dp, s);
            if (WTA) wta_column<DPL>(s, lane, p, raw + ((size_t)b * H + y) * p.W + xp + p.minX1, aux + ((size_t)b * H + y) * W1 + xp);
        }
        if (CL > 1) cg::this_cluster().sync(); else __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------
// A.5 second half.  One CTA per (row, frame).  disp2[x2] is the `best` of the pixel with the smallest minS
// among those mapping to x2, the largest x' winning ties (the sequential right-to-left scan with strict '>').
__global__ void k_lrcheck(int16_t* __restrict__ raw, const int2* __restrict__ aux, int16_t* __restrict__ disp, SgbmDev p, int keepRaw)
{
    extern __shared__ unsigned key[];                   // [W] right-view keys, then [W] raw disparities of this row
    int* rawsm = (int*)(key + p.W);
    const int y = blockIdx.x, b = blockIdx.y, W = p.W, W1 = p.W1;
    const size_t row = (size_t)b * p.H + y;
    const int2* auxr = aux + row * W1;
    int16_t* rawr = raw + row * W;
    int16_t* out = disp + row * W;
    for (int x = threadIdx.x; x < W; x += blockDim.x) { key[x] = 0xffffffffu; rawsm[x] = p.INVALID; }
    __syncthreads();
    for (int xp = threadIdx.x; xp < W1; xp += blockDim.x) {
        const int2 a = auxr[xp];
        if (a.x == 0x7fffffff) continue;
        const int minS = a.x >> 16, best = a.x & 0xffff;
        // sub-pixel interpolation (A.5): C++ '/' truncates toward zero
        int dd = best * CVR_DISP_SCALE;
        if (0 < best && best < p.D - 1) {
            const int sa = a.y >> 16, sc = (int)(short)(a.y & 0xffff);
            const int den = max(sa + sc - 2 * minS, 1);
            dd += ((sa - sc) * CVR_DISP_SCALE + den) / (den * 2);
        }
        rawsm[xp + p.minX1] = dd + p.minD * CVR_DISP_SCALE;
        const int x2 = xp + p.minX1 - best - p.minD;
        // minS == 32767 is always rejected upstream, so (minS + 32768) <= 65534 and a real key never equals the sentinel
        atomicMin(&key[x2], ((unsigned)(minS + 32768) << 16) | (unsigned)(0xffff - xp));
    }
    __syncthreads();
    for (int x = threadIdx.x; x < W; x += blockDim.x) {
        int v = p.INVALID;
        const int d1 = rawsm[x];
        if (keepRaw) rawr[x] = (int16_t)d1;             // only kept for cvr_debug_fetch(2) (CVR_KEEP_S=1)
        if (x >= p.minX1 && x < p.maxX1) {
            v = d1;
            if (d1 != p.INVALID) {
                const int _d = d1 >> 4, d_ = (d1 + CVR_DISP_SCALE - 1) >> 4;
                const int _x = x - _d, x_ = x - d_;
                bool f1 = false, f2 = false;
                if (0 <= _x && _x < W) {
                    // an unset right-view entry holds INVALID_DISP_SCALED in OpenCV, which passes `>= minD` when minD > 1
                    const unsigned k = key[_x];
                    const int d2 = (k != 0xffffffffu) ? (auxr[0xffff - (k & 0xffff)].x & 0xffff) + p.minD : p.INVALID;
                    f1 = d2 >= p.minD && abs(d2 - _d) > p.d12;
                }
                if (0 <= x_ && x_ < W) {
                    const unsigned k = key[x_];
                    const int d2 = (k != 0xffffffffu) ? (auxr[0xffff - (k & 0xffff)].x & 0xffff) + p.minD : p.INVALID;
                    f2 = d2 >= p.minD && abs(d2 - d_) > p.d12;
                }
                if (f1 && f2) v = p.INVALID;
            }
        }
        out[x] = (int16_t)v;
    }
}

// ------------------------------------------------------------------------------------------------
// A.6: 3x3 median, BORDER_REPLICATE; INVALID participates as an ordinary value.
__device__ __forceinline__ void srt(int& a, int& b) { int t = min(a, b); b = max(a, b); a = t; }
__global__ void k_median3(const int16_t* __restrict__ src, int16_t* __restrict__ dst, int W, int H)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
    if (x >= W) return;
    const int16_t* s = src + (size_t)b * W * H;
    int v[9];
#pragma unroll
    for (int dy = -1; dy <= 1; dy++) {
        const int yy = min(max(y + dy, 0), H - 1);
#pragma unroll
        for (int dx = -1; dx <= 1; dx++) v[(dy + 1) * 3 + dx + 1] = s[(size_t)yy * W + min(max(x + dx, 0), W - 1)];
    }
    // 19-exchange median-of-9 network
    srt(v[1], v[2]); srt(v[4], v[5]); srt(v[7], v[8]); srt(v[0], v[1]); srt(v[3], v[4]); srt(v[6], v[7]);
    srt(v[1], v[2]); srt(v[4], v[5]); srt(v[7], v[8]); srt(v[0], v[3]); srt(v[5], v[8]); srt(v[4], v[7]);
    srt(v[3], v[6]); srt(v[1], v[4]); srt(v[2], v[5]); srt(v[4], v[7]); srt(v[4], v[2]); srt(v[6], v[4]);
    srt(v[4], v[2]);
    dst[(size_t)b * W * H + (size_t)y * W + x] = (int16_t)v[4];
}

// Two horizontally adjacent pixels per thread (even image width): aligned 32-bit loads, the 19-exchange network on
// packed int16 pairs (VIMNMX.S16x2), one 32-bit store.  Pairs that touch the left/right border take the scalar form.
__device__ __forceinline__ void srt2(unsigned& a, unsigned& b) { const unsigned t = __vmins2(a, b); b = __vmaxs2(a, b); a = t; }
__device__ __forceinline__ int median9(int (&v)[9])
{
    srt(v[1], v[2]); srt(v[4], v[5]); srt(v[7], v[8]); srt(v[0], v[1]); srt(v[3], v[4]); srt(v[6], v[7]);
    srt(v[1], v[2]); srt(v[4], v[5]); srt(v[7], v[8]); srt(v[0], v[3]); srt(v[5], v[8]); srt(v[4], v[7]);
    srt(v[3], v[6]); srt(v[1], v[4]); srt(v[2], v[5]); srt(v[4], v[7]); srt(v[4], v[2]); srt(v[6], v[4]);
    srt(v[4], v[2]);
    return v[4];
}
__global__ void __launch_bounds__(128) k_median3_x2(const int16_t* __restrict__ src, int16_t* __restrict__ dst, int W, int H)
{
    const int x = 2 * (blockIdx.x * blockDim.x + threadIdx.x), y = blockIdx.y, b = blockIdx.z;
    if (x >= W) return;
    const int16_t* s = src + (size_t)b * W * H;
    int16_t* d = dst + (size_t)b * W * H + (size_t)y * W + x;
    const int ym = max(y - 1, 0), yp = min(y + 1, H - 1);
    if (x >= 2 && x + 3 < W) {
        unsigned v[9];
        const int rows[3] = { ym, y, yp };
#pragma unroll
        for (int r = 0; r < 3; r++) {
            const unsigned* rp = (const unsigned*)(s + (size_t)rows[r] * W + x);       // (v[x], v[x+1]); its neighbours hold x-2..x-1 and x+2..x+3
            const unsigned w0 = rp[-1], w1 = rp[0], w2 = rp[1];
            v[3 * r + 0] = __byte_perm(w0, w1, 0x5432);                                 // (v[x-1], v[x])   : left neighbours of the pair
            v[3 * r + 1] = w1;                                                          // (v[x],   v[x+1])
            v[3 * r + 2] = __byte_perm(w1, w2, 0x5432);                                 // (v[x+1], v[x+2]) : right neighbours
        }
        srt2(v[1], v[2]); srt2(v[4], v[5]); srt2(v[7], v[8]); srt2(v[0], v[1]); srt2(v[3], v[4]); srt2(v[6], v[7]);
        srt2(v[1], v[2]); srt2(v[4], v[5]); srt2(v[7], v[8]); srt2(v[0], v[3]); srt2(v[5], v[8]); srt2(v[4], v[7]);
        srt2(v[3], v[6]); srt2(v[1], v[4]); srt2(v[2], v[5]); srt2(v[4], v[7]); srt2(v[4], v[2]); srt2(v[6], v[4]);
        srt2(v[4], v[2]);
        *(unsigned*)d = v[4];
        return;
    }
    for (int k = 0; k < 2 && x + k < W; k++) {
        int v[9];
#pragma unroll
        for (int dy = 0; dy < 3; dy++) {
            const int yy = dy == 0 ? ym : (dy == 1 ? y : yp);
#pragma unroll
            for (int dx = -1; dx <= 1; dx++) v[dy * 3 + dx + 1] = s[(size_t)yy * W + min(max(x + k + dx, 0), W - 1)];
        }
        d[k] = (int16_t)median9(v);
    }
}

// ------------------------------------------------------------------------------------------------
// A.7: speckle filter = connected components (4-neighbours, |dv| <= maxDiff, value != newVal) of size
// <= maxSize are erased.  Union-find with atomicMin roots; the result does not depend on the merge order.
__device__ __forceinline__ int uf_find(const int* L, int i)
{
    int j;
    while ((j = L[i]) != i) i = j;
    return i;
}
__device__ __forceinline__ void uf_union(int* L, int a, int b)
{
    bool done;
    do {
        a = uf_find(L, a); b = uf_find(L, b);
        if (a < b) { int old = atomicMin(&L[b], a); done = (old == b); b = old; }
        else if (b < a) { int old = atomicMin(&L[a], b); done = (old == a); a = old; }
        else done = true;
    } while (!done);
}
// Run-based labelling.  A horizontal run is a maximal stretch of valid pixels whose neighbours differ by <= maxDiff; its
// first pixel (the "start") represents it: label[start] = start, label[member] = start | CCL_MEMBER, and the run's length is
// left in size[start] (no other entry of size[] is ever read).  Found with two ballots per 32 pixels.  Unions are then only
// needed between vertically connected runs, once per run pair (at the first column of their overlap), and they only ever
// relabel starts (roots are starts).  Component sizes are summed over RUNS -- one atomic per run that is not its
// component's root -- instead of over pixels.
#define CCL_MEMBER 0x40000000
__global__ void __launch_bounds__(128) k_ccl_runs(const int16_t* __restrict__ img, int* __restrict__ label, int* __restrict__ size, int W, int H, int B,
                                                  int newVal, int maxDiff)
{
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (row >= B * H) return;
    const size_t base = (size_t)row * W;
    const int rowOff = (row % H) * W;
    int* const sizeRow = size + base;
    int carry = -1;                                           // start of the run that reaches into this chunk
    int prevV = newVal;
    for (int x0 = 0; x0 < W; x0 += 32) {
        const int x = x0 + lane;
        const int v = x < W ? (int)img[base + x] : newVal;
        int left = __shfl_up_sync(0xffffffffu, v, 1);
        if (lane == 0) left = prevV;
        const bool valid = v != newVal;
        const bool conn = valid && left != newVal && abs(v - left) <= maxDiff;       // connected to x-1 (false at x == 0: prevV starts as newVal)
        const bool isStart = valid && !conn;
        const unsigned starts = __ballot_sync(0xffffffffu, isStart);
        const unsigned vmask = __ballot_sync(0xffffffffu, valid);
        const unsigned le = starts & (0xffffffffu >> (31 - lane));
        const int s = le ? x0 + 31 - __clz(le) : carry;
        if (x < W) label[base + x] = !valid ? -1 : (isStart ? rowOff + x : ((rowOff + s) | CCL_MEMBER));
        // the first member of a run inside this chunk counts the run's members here: up to the next start or invalid pixel
        if (valid && (isStart || lane == 0)) {
            const unsigned stops = (starts | ~vmask) & ~((2u << lane) - 1u);
            const int cnt = (stops ? __ffs(stops) - 1 : 32) - lane;
            if (isStart) sizeRow[x] = cnt;
            else atomicAdd(&sizeRow[carry], cnt);                                     // continues a run begun in an earlier chunk (same warp: ordered)
        }
        __syncwarp();
        // the run continuing into the next chunk starts at the last start of this chunk, or is the carried one
        const int lastS = starts ? x0 + 31 - __clz(starts) : carry;
        carry = __shfl_sync(0xffffffffu, valid ? lastS : -1, 31);
        prevV = __shfl_sync(0xffffffffu, v, 31);
    }
}
__global__ void k_ccl_vmerge(const int16_t* __restrict__ img, int* label, int W, int H, int newVal, int maxDiff)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
    if (x >= W || y + 1 >= H) return;
    const size_t fo = (size_t)b * W * H;
    const int i = y * W + x;
    const int v = img[fo + i], q = img[fo + i + W];
    if (v == newVal || q == newVal || abs(v - q) > maxDiff) return;
    if (x > 0) {
        // the same pair of runs was already united at x-1 if both runs continue from x-1 and x-1 is vertically connected too
        const int vl = img[fo + i - 1], ql = img[fo + i + W - 1];
        if (vl != newVal && ql != newVal && abs(v - vl) <= maxDiff && abs(q - ql) <= maxDiff && abs(vl - ql) <= maxDiff) return;
    }
    int* L = label + fo;
    uf_union(L, L[i] & ~CCL_MEMBER, L[i + W] & ~CCL_MEMBER);
}
// The same for even image widths, a warp per pair of rows: every lane takes two adjacent pixels of both rows with aligned
// 32-bit loads and gets the pixel to its left from the neighbouring lane.
__global__ void __launch_bounds__(128) k_ccl_vmerge2(const int16_t* __restrict__ img, int* label, int W, int H, int B, int newVal, int maxDiff)
{
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (row >= B * H) return;
    const int y = row % H;
    if (y + 1 >= H) return;
    const size_t fo = (size_t)(row / H) * W * H;
    const unsigned* up = (const unsigned*)(img + fo + (size_t)y * W);          // W even: every row starts 4-byte aligned
    const unsigned* dn = (const unsigned*)(img + fo + (size_t)(y + 1) * W);
    int* L = label + fo;
    const int W2 = W >> 1;
    int carryV = newVal, carryQ = newVal;                                       // the pixel left of this chunk (none at x = 0)
    for (int x0 = 0; x0 < W2; x0 += 32) {
        const int xw = x0 + lane;
        const unsigned a = xw < W2 ? up[xw] : ((unsigned)(newVal & 0xffff) * 0x10001u), c = xw < W2 ? dn[xw] : ((unsigned)(newVal & 0xffff) * 0x10001u);
        const int v0 = (int)(short)(a & 0xffffu), v1 = (int)a >> 16, q0 = (int)(short)(c & 0xffffu), q1 = (int)c >> 16;
        int vl = __shfl_up_sync(0xffffffffu, v1, 1), ql = __shfl_up_sync(0xffffffffu, q1, 1);
        if (lane == 0) { vl = carryV; ql = carryQ; }
        carryV = __shfl_sync(0xffffffffu, v1, 31); carryQ = __shfl_sync(0xffffffffu, q1, 31);
        if (xw >= W2) continue;
        const int x = 2 * xw, i = y * W + x;
        // pixel x: united here unless the pair (x-1) already united the same two runs
        if (v0 != newVal && q0 != newVal && abs(v0 - q0) <= maxDiff &&
            !(vl != newVal && ql != newVal && abs(v0 - vl) <= maxDiff && abs(q0 - ql) <= maxDiff && abs(vl - ql) <= maxDiff))
            uf_union(L, L[i] & ~CCL_MEMBER, L[i + W] & ~CCL_MEMBER);
        if (v1 != newVal && q1 != newVal && abs(v1 - q1) <= maxDiff &&
            !(v0 != newVal && q0 != newVal && abs(v1 - v0) <= maxDiff && abs(q1 - q0) <= maxDiff && abs(v0 - q0) <= maxDiff))
            uf_union(L, L[i + 1] & ~CCL_MEMBER, L[i + 1 + W] & ~CCL_MEMBER);
    }
}
// run starts that are not roots: add the run's length to the root's and point straight at it (roots stay fixed points,
// so concurrent flattening is safe).  Both passes below walk the label array of the whole batch as one flat array, four
// labels per thread (16-byte loads; the frame of an element is recovered by a division only where it is needed).
__global__ void __launch_bounds__(256) k_ccl_count(int* label, int* size, int N, size_t total)
{
    const size_t g0 = 4 * ((size_t)blockIdx.x * blockDim.x + threadIdx.x);
    if (g0 >= total) return;
    int l4[4] = { -1, -1, -1, -1 };
    if (g0 + 3 < total) { const int4 q = *(const int4*)(label + g0); l4[0] = q.x; l4[1] = q.y; l4[2] = q.z; l4[3] = q.w; }
    else for (int k = 0; g0 + k < total; k++) l4[k] = label[g0 + k];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int l = l4[k];
        if (l < 0 || (l & CCL_MEMBER)) continue;              // invalid pixel or member of a run
        const size_t g = g0 + k, fo = (g / (size_t)N) * (size_t)N;
        const int i = (int)(g - fo);
        if (l == i) continue;                                 // a root
        int* L = label + fo;
        const int r = uf_find(L, l);
        L[i] = r;
        atomicAdd(&size[fo + r], size[fo + i]);
    }
}
__global__ void __launch_bounds__(256) k_ccl_apply(int16_t* img, const int* __restrict__ label, const int* __restrict__ size, int N, size_t total, int newVal, int maxSize)
{
    const size_t g0 = 4 * ((size_t)blockIdx.x * blockDim.x + threadIdx.x);
    if (g0 >= total) return;
    int l4[4] = { -1, -1, -1, -1 };
    if (g0 + 3 < total) { const int4 q = *(const int4*)(label + g0); l4[0] = q.x; l4[1] = q.y; l4[2] = q.z; l4[3] = q.w; }
    else for (int k = 0; g0 + k < total; k++) l4[k] = label[g0 + k];
    if ((l4[0] & l4[1] & l4[2] & l4[3]) < 0 && l4[0] < 0 && l4[1] < 0 && l4[2] < 0 && l4[3] < 0) return;
    size_t fo = (g0 / (size_t)N) * (size_t)N;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int l = l4[k];
        if (l < 0) continue;
        const size_t g = g0 + k;
        if (g >= fo + (size_t)N) fo += (size_t)N;             // the four elements straddle a frame boundary
        const int r = (l & CCL_MEMBER) ? label[fo + (size_t)(l & ~CCL_MEMBER)] : l;   // a start's label is its root after k_ccl_count
        if (size[fo + r] <= maxSize) img[g] = (int16_t)newVal;
    }
}

__global__ void k_fill_s16(int16_t* p, size_t n, int v)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = (int16_t)v;
}

#include "k_paths_packed.cuh"

// ------------------------------------------------------------------------------------------------
// host-side launchers

int cvr_launch_gray(cvr_ctx* c, int B, const uint8_t* left, size_t lstep, size_t lfs, const uint8_t* right, size_t rstep,
                    size_t rfs, int channels)
{
    const SgbmDev& p = c->sg;
    const size_t gfs = (size_t)2 * p.H * p.Wp;
    if (channels == 3) {
        k_gray3_x4<<<dim3(cvr_div_up(p.W, 512), p.H, 2 * B), 128, 0, c->stream>>>(left, lstep, lfs, right, rstep, rfs, c->d_gray, p.Wp, gfs, p.W, p.H);
        LAUNCH_CHECK(c);
        return CVR_OK;
    }
    dim3 g(cvr_div_up(p.W, 256), p.H, B);
    k_gray<<<g, 256, 0, c->stream>>>(left, lstep, lfs, channels, c->d_gray, p.Wp, gfs, p.W, p.H);
    LAUNCH_CHECK(c);
    k_gray<<<g, 256, 0, c->stream>>>(right, rstep, rfs, channels, c->d_gray + (size_t)p.H * p.Wp, p.Wp, gfs, p.W, p.H);
    LAUNCH_CHECK(c);
    return CVR_OK;
}

struct VCfg { bool ok; int CL, NT; VState vs; size_t smem; double cost; };   // cost: waves x columns per CTA of the whole batch

static size_t vcfg_smem(const SgbmDev& p, int CL, int NT, VState* vs)
{
    const int NW = NT / 32;
    vs->colsMax = (p.W1 + CL - 1) / CL + 1; vs->colWords = 3 * (p.Dp / 2 + 4); vs->inboxWords = p.Dp / 2 + 4; vs->sel = 0;
    return 4 * ((size_t)vs->colsMax * vs->colWords + (size_t)2 * NW * 2 * vs->inboxWords) + (size_t)2 * NW * 2 * 8;
}

// Geometry of the vertical path kernel: the largest CTA that fits the register budget, then the cluster size.
// A cluster must sit inside one GPC, so the number of clusters that run at once depends on the cluster size in a way
// only the occupancy query knows (measured on this B200: 15 clusters of 8, 22 of 6, 26 of 5); a frame takes time
// proportional to the columns per CTA, and the batch runs in ceil(B / concurrent clusters) waves.  The cluster size
// with the smallest waves x columns-per-CTA wins.  The query results are cached per context (first launch only).
template <typename K>
static VCfg pick_vcfg(cvr_ctx* c, int B, K kernel1024, K kernel512, bool v4 = false, int v4nt = 512)
{
    const SgbmDev& p = c->sg;
    VCfg r; r.ok = false; r.CL = 1; r.NT = 0; r.smem = 0; r.cost = 0; r.vs.colsMax = r.vs.colWords = r.vs.inboxWords = r.vs.sel = 0;
    const int nts[3] = { 1024, 512, 256 };
    const char* fcl = getenv("CVR_PATHV_CL");
    const int forced = fcl ? atoi(fcl) : 0;
    for (int ni = (c->pathv_max_nt <= 256 ? 2 : (c->pathv_max_nt <= 512 ? 1 : 0)); ni < 3; ni++) {
        const int NT = nts[ni], NW = NT / 32;
        if (NT > 512 && p.Dp > 128) continue;                          // wide disparity ranges need > 64 registers per thread
        if (v4 && NT != v4nt) continue;                                // k_path_v4 is built for one CTA size
        double best = 0;
        for (int CL = 16; CL >= 1; CL--) {                              // 9..16: non-portable cluster sizes (wide disparity ranges at HD need them)
            if (forced && CL != forced) continue;
            if (CL * NW > p.W1) continue;                              // every warp must own at least one column
            VState vs;
            const size_t smem = vcfg_smem(p, CL, NT, &vs);
            if (smem > (size_t)c->smem_optin) continue;
            int& act = v4 ? c->pathv4_active[CL] : c->pathv_active[ni][CL];
            if (act == 0) {                                            // not asked yet
                K kernel = NT > 512 ? kernel1024 : kernel512;
                act = -1;
                if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) == cudaSuccess &&
                    (CL <= 8 || cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess)) {
                    cudaLaunchConfig_t cfg = {};
                    cfg.gridDim = dim3(CL, 64, 1); cfg.blockDim = dim3(NT, 1, 1); cfg.dynamicSmemBytes = smem; cfg.stream = c->stream;
                    cudaLaunchAttribute at[1];
                    at[0].id = cudaLaunchAttributeClusterDimension;
                    at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
                    cfg.attrs = at; cfg.numAttrs = 1;
                    int n = 0;
                    if (cudaOccupancyMaxActiveClusters(&n, kernel, &cfg) == cudaSuccess && n > 0) act = n;
                }
                (void)cudaGetLastError();
                if (getenv("CVR_VCFG_DEBUG")) fprintf(stderr, "[vcfg] NT=%d CL=%d smem=%zu concurrent clusters=%d\n", NT, CL, smem, act);
            }
            if (act <= 0) continue;
            const double cost = (double)((B + act - 1) / act) * (double)(vs.colsMax + 8);
            if (!r.ok || cost < best) { best = cost; r.ok = true; r.CL = CL; r.NT = NT; r.vs = vs; r.smem = smem; r.cost = cost; }
        }
        if (r.ok) return r;
    }
    return r;
}

template <typename K>
static cudaError_t launch_cluster(K kernel, int CL, int B, int NT, size_t smem, cudaStream_t st, const int16_t* C, int16_t* S, int16_t* raw, int2* aux,
                                  SgbmDev p, int storeS, int* flags, VState vs)
{
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    if (CL > 8) {
        e = cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (e != cudaSuccess) return e;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(CL, B, 1); cfg.blockDim = dim3(NT, 1, 1); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, C, S, raw, aux, p, CL, storeS, flags, vs);
}

// A row of the horizontal path is one long sequential task, so a launch runs in waves of (resident CTAs) equal tasks and a
// part-filled last wave costs as much as a full one (2.2 waves of 8 CTAs per SM run as 3).  Fewer resident CTAs per SM can
// make the waves come out even at the same total throughput: pick the CTA count per SM (>= 4) that minimises
// waves x CTAs-per-SM and enforce it with a dynamic shared-memory pad.  Returns the pad in bytes.
template <typename K>
static size_t wave_balanced_pad_query(cvr_ctx* c, K kernel, int threads, int nCTA, int forced);
template <typename K>
static size_t wave_balanced_pad(cvr_ctx* c, K kernel, int threads, int nCTA)
{
    static const int forced = getenv("CVR_H3_OCC") ? atoi(getenv("CVR_H3_OCC")) : 0;   // tuning hook: CTAs per SM
    struct Memo { const void* k; int n; int dev; size_t pad; };
    static std::vector<Memo> memo;                                     // (a context is driven by one thread; contexts of one process share the answers)
    for (const Memo& m : memo) if (m.k == (const void*)kernel && m.n == nCTA && m.dev == c->device) return m.pad;
    const size_t pad = wave_balanced_pad_query(c, kernel, threads, nCTA, forced);
    memo.push_back(Memo{ (const void*)kernel, nCTA, c->device, pad });
    return pad;
}
template <typename K>
static size_t wave_balanced_pad_query(cvr_ctx* c, K kernel, int threads, int nCTA, int forced)
{
    int omax = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&omax, kernel, threads, 0) != cudaSuccess || omax < 1) { (void)cudaGetLastError(); return 0; }
    int best = omax; long long bestCost = -1;
    for (int o = omax; o >= (omax < 4 ? 1 : 4); o--) {
        const long long waves = ((long long)nCTA + (long long)o * c->sm_count - 1) / ((long long)o * c->sm_count);
        const long long cost = waves * o;
        if (bestCost < 0 || cost < bestCost) { bestCost = cost; best = o; }
    }
    if (forced > 0 && forced <= omax) best = forced;
    if (best >= omax) return 0;
    cudaFuncAttributes fa;
    if (cudaFuncGetAttributes(&fa, kernel) != cudaSuccess) { (void)cudaGetLastError(); return 0; }
    size_t pad = 0;
    for (int it = 0; it < 64; it++) {                                  // grow the pad until the occupancy query answers `best`
        pad += 2048;
        if (fa.sharedSizeBytes + pad > 48 * 1024 &&
            cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pad) != cudaSuccess) { (void)cudaGetLastError(); return 0; }
        int o = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, kernel, threads, pad) != cudaSuccess) { (void)cudaGetLastError(); return 0; }
        if (o <= best) return o == best ? pad : (pad > 2048 ? pad - 2048 : 0);
    }
    return 0;
}

template <int DPL, bool HASPAD, bool NN>
static int launch_paths(cvr_ctx* c, int B, int storeS)
{
    const SgbmDev& p = c->sg;
    const int hblocks = cvr_div_up(B * p.H, 4);
    const bool packed = !c->force_generic && p.P1s >= 0;               // P1s < 0 (P1 in [32768, 65535] mod 65536): int32 kernels
    // k_path_v4: lanes per column, words per lane, threads per CTA (measured and rejected at D = 128: 16 lanes per column in
    // 1024-thread CTAs -- 64 registers per thread spill, 90 us per frame against 61)
    constexpr int LPRV = DPL <= 4 ? 8 : 16, NWDV = DPL * 16 / LPRV, NTV4 = 512;
    const bool useV4 = packed && c->path_v_variant == 4;
    const VCfg vc = !packed ? VCfg{ false, 1, 0, { 0, 0, 0, 0 }, 0, 0 }
                    : useV4 ? pick_vcfg(c, B, k_path_v4<NWDV, LPRV, false, false, HASPAD, NN, NTV4>, k_path_v4<NWDV, LPRV, false, false, HASPAD, NN, NTV4>, true, NTV4)
                            : pick_vcfg(c, B, k_path_v3<DPL, false, false, HASPAD, NN, 1024>, k_path_v3<DPL, false, false, HASPAD, NN, 512>);
    int* fl = c->d_sgflags;
    // generic vertical kernel geometry: a cluster of CL CTAs per frame; fewer, larger clusters when the batch is small
    constexpr int NT = (DPL <= 4) ? 512 : 256;
    int CL = 1;
    while (CL < 8 && B * CL * 2 <= c->sm_count * 2 && (p.W1 / (CL * 2)) >= NT / 32) CL *= 2;
    const size_t lrN = (size_t)B * 2 * (p.W1 + 2) * 3 * p.Dp;
    auto launch_v = [&](bool up, bool wta, int sel = 0) -> int {
        cudaError_t e;
        if (vc.ok) {
            if (c->path_v_variant == 2) {          // CVR_PATH_KERNELS=v2: per-row cluster barrier instead of point-to-point hand-over
                if (!up) e = launch_cluster(k_path_v2<DPL, false, false, HASPAD>, vc.CL, B, vc.NT, vc.smem, c->stream, c->d_volC, c->d_volA, c->d_raw, c->d_aux, p, storeS, fl, vc.vs);
                else if (wta) e = launch_cluster(k_path_v2<DPL, true, true, HASPAD>, vc.CL, B, vc.NT, vc.smem, c->stream, c->d_volC, c->d_volA, c->d_raw, c->d_aux, p, storeS, fl, vc.vs);
                else e = launch_cluster(k_path_v2<DPL, true, false, HASPAD>, vc.CL, B, vc.NT, vc.smem, c->stream, c->d_volC, c->d_volA, c->d_raw, c->d_aux, p, storeS, fl, vc.vs);
            } else if (useV4) {
                VState vsel = vc.vs; vsel.sel = sel;
                if (!up) e = launch_cluster(k_path_v4<NWDV, LPRV, false, false, HASPAD, NN, NTV4>, vc.CL, B, vc.NT, vc.smem, c->stream, c->d_volC, c->d_volA, c->d_raw, c->d_aux, p, storeS, fl, vsel);
                else if (wta) e = launch_cluster(k_path_v4<NWDV, LPRV, true, true, HASPAD, NN, NTV4>, vc.CL, B, vc.NT, vc.smem, c->stream, c->d_volC, c->d_volA, c->d_raw, c->d_aux, p, storeS, fl, vsel);
                else e = launch_cluster(k_path_v4<NWDV, LPRV, true, false, HASPAD, NN, NTV4>, vc.CL, B, vc.NT, vc.smem, c->stream, c->d_volC, c->d_volA, c->d_raw, c->d_aux, p, storeS, fl, vsel);
            } else if (vc.NT > 512) {
                if (!up) e = launch_cluster(k_path_v3<DPL, false, false, HASPAD, NN, 1024>, vc.CL, B, vc.NT, vc.smem, c->stream, c->d_volC, c->d_volA, c->d_raw, c->d_aux, p, storeS, fl, vc.vs);
                else if (wta) e = launch_cluster(k_path_v3<DPL, true, true, HASPAD, NN, 1024>, vc.CL, B, vc.NT, vc.smem, c->stream, c->d_volC, c->d_volA, c->d_raw, c->d_aux, p, storeS, fl, vc.vs);
                else e = launch_cluster(k_path_v3<DPL, true, false, HASPAD, NN, 1024>, vc.CL, B, vc.NT, vc.smem, c->stream, c->d_volC, c->d_volA, c->d_raw, c->d_aux, p, storeS, fl, vc.vs);
            } else {
                if (!up) e = launch_cluster(k_path_v3<DPL, false, false, HASPAD, NN, 512>, vc.CL, B, vc.NT, vc.smem, c->stream, c->d_volC, c->d_volA, c->d_raw, c->d_aux, p, storeS, fl, vc.vs);
                else if (wta) e = launch_cluster(k_path_v3<DPL, true, true, HASPAD, NN, 512>, vc.CL, B, vc.NT, vc.smem, c->stream, c->d_volC, c->d_volA, c->d_raw, c->d_aux, p, storeS, fl, vc.vs);
                else e = launch_cluster(k_path_v3<DPL, true, false, HASPAD, NN, 512>, vc.CL, B, vc.NT, vc.smem, c->stream, c->d_volC, c->d_volA, c->d_raw, c->d_aux, p, storeS, fl, vc.vs);
            }
            c->launches++;
            if (e != cudaSuccess) { c->err = std::string("k_path_v2 launch: ") + cudaGetErrorString(e); return CVR_ERR_CUDA; }
            return CVR_OK;
        }
        k_lr_init<<<(unsigned)cvr_div_up_sz(lrN, 256), 256, 0, c->stream>>>(c->d_lr, c->d_minlr, p, B);
        LAUNCH_CHECK(c);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(CL, B, 1);
        cfg.blockDim = dim3(NT, 1, 1);
        cfg.dynamicSmemBytes = 0;
        cfg.stream = c->stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        if (!up) e = cudaLaunchKernelEx(&cfg, k_path_v<DPL, false, false, NT>, (const int16_t*)c->d_volC, c->d_volA, c->d_lr, c->d_minlr, c->d_raw, c->d_aux, p, CL, storeS, fl);
        else if (wta) e = cudaLaunchKernelEx(&cfg, k_path_v<DPL, true, true, NT>, (const int16_t*)c->d_volC, c->d_volA, c->d_lr, c->d_minlr, c->d_raw, c->d_aux, p, CL, storeS, fl);
        else e = cudaLaunchKernelEx(&cfg, k_path_v<DPL, true, false, NT>, (const int16_t*)c->d_volC, c->d_volA, c->d_lr, c->d_minlr, c->d_raw, c->d_aux, p, CL, storeS, fl);
        c->launches++;
        if (e != cudaSuccess) { c->err = std::string("k_path_v launch: ") + cudaGetErrorString(e); return CVR_ERR_CUDA; }
        return CVR_OK;
    };
    auto launch_h = [&](bool rev, bool wta, int accumulate, int store) -> int {
        if (packed && (rev ? c->path_h_rev : c->path_h_fwd) == 3) {
            // several rows per warp: 8 lanes per row up to 128 disparities, 16 lanes beyond (k_paths_packed.cuh)
            constexpr int LPR = DPL <= 4 ? 8 : 16;
            constexpr int NWD3 = DPL * 16 / LPR;
            const int wpf = cvr_div_up(p.H, 32 / LPR);
            const int blocks = cvr_div_up(B * wpf, CVR_H3_WARPS);
            constexpr int NT3 = CVR_H3_WARPS * 32;
            if (!rev) { const size_t pad = wave_balanced_pad(c, k_path_h3<NWD3, LPR, false, false, HASPAD, NN>, NT3, blocks);
                        k_path_h3<NWD3, LPR, false, false, HASPAD, NN><<<blocks, NT3, pad, c->stream>>>(c->d_volC, c->d_volA, c->d_aux, p, B, accumulate, store, fl); }
            else if (wta) { const size_t pad = wave_balanced_pad(c, k_path_h3<NWD3, LPR, true, true, HASPAD, NN>, NT3, blocks);
                            k_path_h3<NWD3, LPR, true, true, HASPAD, NN><<<blocks, NT3, pad, c->stream>>>(c->d_volC, c->d_volA, c->d_aux, p, B, accumulate, store, fl); }
            else { const size_t pad = wave_balanced_pad(c, k_path_h3<NWD3, LPR, true, false, HASPAD, NN>, NT3, blocks);
                   k_path_h3<NWD3, LPR, true, false, HASPAD, NN><<<blocks, NT3, pad, c->stream>>>(c->d_volC, c->d_volA, c->d_aux, p, B, accumulate, store, fl); }
        } else if (packed) {
            if (!rev) k_path_h2<DPL, false, false, HASPAD, NN><<<hblocks, 128, 0, c->stream>>>(c->d_volC, c->d_volA, c->d_raw, c->d_aux, p, B, accumulate, store, fl);
            else if (wta) k_path_h2<DPL, true, true, HASPAD, NN><<<hblocks, 128, 0, c->stream>>>(c->d_volC, c->d_volA, c->d_raw, c->d_aux, p, B, accumulate, store, fl);
            else k_path_h2<DPL, true, false, HASPAD, NN><<<hblocks, 128, 0, c->stream>>>(c->d_volC, c->d_volA, c->d_raw, c->d_aux, p, B, accumulate, store, fl);
        } else {
            if (!rev) k_path_h<DPL, false, false><<<hblocks, 128, 0, c->stream>>>(c->d_volC, c->d_volA, c->d_raw, c->d_aux, p, B, accumulate, store, fl);
            else if (wta) k_path_h<DPL, true, true><<<hblocks, 128, 0, c->stream>>>(c->d_volC, c->d_volA, c->d_raw, c->d_aux, p, B, accumulate, store, fl);
            else k_path_h<DPL, true, false><<<hblocks, 128, 0, c->stream>>>(c->d_volC, c->d_volA, c->d_raw, c->d_aux, p, B, accumulate, store, fl);
        }
        LAUNCH_CHECK(c);
        return CVR_OK;
    };
    // pass 1: -> then the three downward paths
    int rc = launch_h(false, false, 0, 1);
    if (rc) return rc;
    cvr_stage_mark(c, "path_h_fwd");
    rc = launch_v(false, false);
    if (rc) return rc;
    cvr_stage_mark(c, "path_v_down");
    static const bool hhExactOrder = getenv("CVR_HH_ORDER") && strcmp(getenv("CVR_HH_ORDER"), "exact") == 0;   // test hook
    if (p.mode == CVR_MODE_HH && useV4 && vc.ok && !hhExactOrder) {    // (vc.ok: k_path_v4 serves this shape; the fallback kernels know no selection)
        // MODE_HH, second pass.  OpenCV adds the paths to S in the order <- , up-left, up, up-right and takes the winner after
        // the last one; with saturating sums that order matters as soon as a path cost is negative.  While none is (the regime
        // flags of the frame are clear: every reference configuration on natural input), the sums are min(sum, 32767) in any
        // order, and the cheaper order runs: the three upward paths first, then <- with the winner-take-all in the kernel that
        // does it best (k_path_h3: 36 us per KITTI frame against 60+ in the vertical kernel, and S is not written again).
        // Frames whose flags are set at the start of a kernel are skipped by these launches (selection 1) ...
        rc = launch_v(true, false, 1);
        if (rc) return rc;
        cvr_stage_mark(c, "path_v_up");
        rc = launch_h(true, true, 1, storeS | (1 << 8));
        if (rc) return rc;
        cvr_stage_mark(c, "path_h_rev_wta");
        // ... and every frame with a flag set by now -- from the start, or raised on the way, which voids what the launches
        // above made of it -- is done again from the first path on in OpenCV's order (selection 2; frames without flags leave
        // these kernels at once).
        rc = launch_h(false, false, 0, 1 | (2 << 8));
        if (rc) return rc;
        rc = launch_v(false, false, 2);
        if (rc) return rc;
        rc = launch_h(true, false, 1, 1 | (2 << 8));
        if (rc) return rc;
        rc = launch_v(true, true, 2);
        if (rc) return rc;
        cvr_stage_mark(c, "path_exact_order");
    } else if (p.mode == CVR_MODE_HH) {
        rc = launch_h(true, false, 1, 1);
        if (rc) return rc;
        cvr_stage_mark(c, "path_h_rev");
        rc = launch_v(true, true);
        if (rc) return rc;
        cvr_stage_mark(c, "path_v_up_wta");
    } else {
        rc = launch_h(true, true, 1, storeS);
        if (rc) return rc;
        cvr_stage_mark(c, "path_h_rev_wta");
    }
    return CVR_OK;
}

// cv::filterSpeckles(d_med, newVal, maxSize, maxDiff) for every frame of the batch, in place.
int cvr_launch_speckle(cvr_ctx* c, int B, int newVal, int maxSize, int maxDiff)
{
    const int W = c->prm.width, H = c->prm.height, N = W * H;
    k_ccl_runs<<<cvr_div_up(B * H, 4), 128, 0, c->stream>>>(c->d_med, c->d_label, c->d_size, W, H, B, newVal, maxDiff);
    LAUNCH_CHECK(c);
    if ((W & 1) == 0) k_ccl_vmerge2<<<cvr_div_up(B * H, 4), 128, 0, c->stream>>>(c->d_med, c->d_label, W, H, B, newVal, maxDiff);
    else k_ccl_vmerge<<<dim3(cvr_div_up(W, 128), H, B), 128, 0, c->stream>>>(c->d_med, c->d_label, W, H, newVal, maxDiff);
    LAUNCH_CHECK(c);
    const size_t total = (size_t)B * N;
    const unsigned g4 = (unsigned)cvr_div_up_sz(total, 1024);
    k_ccl_count<<<g4, 256, 0, c->stream>>>(c->d_label, c->d_size, N, total);
    LAUNCH_CHECK(c);
    k_ccl_apply<<<g4, 256, 0, c->stream>>>(c->d_med, c->d_label, c->d_size, N, total, newVal, maxSize);
    LAUNCH_CHECK(c);
    return CVR_OK;
}

int cvr_launch_sgbm(cvr_ctx* c, int B)
{
    const SgbmDev& p = c->sg;
    const size_t npix = (size_t)B * p.W * p.H;
    const int storeS = c->keep_S ? 1 : 0;   // the final S is only written out for cvr_debug_fetch (CVR_KEEP_S=1): 1/8 of the stage's traffic
    if (p.W1 <= 0) {
        k_fill_s16<<<(unsigned)cvr_div_up_sz(npix, 256), 256, 0, c->stream>>>(c->d_med, npix, p.INVALID);
        LAUNCH_CHECK(c);
        return CVR_OK;
    }
    CVR_CUDA(c, cudaMemsetAsync(c->d_sgflags, 0, sizeof(int) * B, c->stream));
    // (d_raw needs no clearing: k_lrcheck writes every pixel of it)
    {
        static const int segLen = getenv("CVR_HSUM_SEG") ? atoi(getenv("CVR_HSUM_SEG")) : 560;   // tuning hook (B200, KITTI: 200 -> 38.9, 320 -> 33.5, 560 -> 33.0 us per frame)
        int nseg = (p.W1 + segLen / 2) / segLen;
        if (nseg < 1) nseg = 1;
        const int segMax = cvr_div_up(p.W1, nseg) + 1;
        const size_t nLmax = (size_t)segMax + 2 * p.SW2 + 2, nRmax = nLmax + p.D;
        static const int pixVariant = getenv("CVR_PIX_HSUM") ? atoi(getenv("CVR_PIX_HSUM")) : 4;   // 4 disparities per thread (default) or 2
        if (pixVariant == 4) {
            // descriptors (left + four residue arrays of the right ones) + unpacked pixels + two rings (one per thread group) of cost quadruples
            const size_t stage = (nLmax + nRmax + 6) * sizeof(unsigned), ringB = (size_t)2 * p.ring * p.Dp;   // the ring (one byte per cost) reuses the staging area
            const size_t smem = (nLmax + nRmax + 20) * sizeof(uint4) + (stage > ringB ? stage : ringB) + 16;
            if (smem > 48 * 1024) CVR_CUDA(c, cudaFuncSetAttribute(k_pix_hsum4, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            k_pix_hsum4<<<dim3(p.H * nseg, B), p.Dp / 2, smem, c->stream>>>(c->d_gray, (size_t)2 * p.H * p.Wp, c->d_volA, p, nseg);
        } else {
        const size_t smem = (nLmax + nRmax + 4) * (sizeof(uint4) + sizeof(unsigned)) + 16 + (size_t)4 * p.ring * p.Dp;   // descriptors + two rings (one per thread group) of packed cost pairs
        // the opt-in is per function and device, shared by every context on that device: set it before each launch that needs it
        if (smem > 48 * 1024) CVR_CUDA(c, cudaFuncSetAttribute(k_pix_hsum, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_pix_hsum<<<dim3(p.H * nseg, B), p.Dp, smem, c->stream>>>(c->d_gray, (size_t)2 * p.H * p.Wp, c->d_volA, p, nseg);
        }
        LAUNCH_CHECK(c);
        cvr_stage_mark(c, "pix_hsum");
    }
    {
        const size_t nthreads = (size_t)p.W1 * p.Dp / 8 * B;
        // low occupancy + deep prefetch (see k_vsum); tuning hooks: CVR_VSUM_OCC blocks per SM (0 = no cap), CVR_VSUM_P rows ahead (1, 4, 8)
        static const int occ = getenv("CVR_VSUM_OCC") ? atoi(getenv("CVR_VSUM_OCC")) : 0;
        static const int P = getenv("CVR_VSUM_P") ? atoi(getenv("CVR_VSUM_P")) : 16;
        const size_t pad = occ > 0 ? (size_t)(c->smem_optin / occ - 2048) & ~(size_t)255 : 0;
        const unsigned nb = (unsigned)cvr_div_up_sz(nthreads, 128);
        if (P >= 16) {
            if (pad > 48 * 1024) CVR_CUDA(c, cudaFuncSetAttribute(k_vsum<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pad));
            k_vsum<16><<<nb, 128, pad, c->stream>>>(c->d_volA, c->d_volC, p, B, c->d_sgflags);
        } else if (P >= 8) {
            if (pad > 48 * 1024) CVR_CUDA(c, cudaFuncSetAttribute(k_vsum<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pad));
            k_vsum<8><<<nb, 128, pad, c->stream>>>(c->d_volA, c->d_volC, p, B, c->d_sgflags);
        } else if (P >= 4) {
            if (pad > 48 * 1024) CVR_CUDA(c, cudaFuncSetAttribute(k_vsum<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pad));
            k_vsum<4><<<nb, 128, pad, c->stream>>>(c->d_volA, c->d_volC, p, B, c->d_sgflags);
        } else {
            if (pad > 48 * 1024) CVR_CUDA(c, cudaFuncSetAttribute(k_vsum<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pad));
            k_vsum<1><<<nb, 128, pad, c->stream>>>(c->d_volA, c->d_volC, p, B, c->d_sgflags);
        }
        LAUNCH_CHECK(c);
        cvr_stage_mark(c, "vsum");
    }
    int rc = cvr_path_wait_aux(c);                                     // the previous pipelined batch's LR check has read d_aux
    if (rc) return rc;
    const bool nn = !c->force_no_nn && p.P2s >= 0 && p.P1s >= p.P2s;
    switch (p.Dp / 32) {
        // NN: 0 <= (short)P2 <= (short)P1 -- the neighbour terms of the path update can never win (see col_fast); this is the
        // regime of the reference's default config (block_size 27: P2 = 69984 wraps to 4448, P1 = 17496)
#define CVR_DISPATCH(DPL_)                                                                                                   \
    case DPL_:                                                                                                               \
        if (nn) rc = (p.D == p.Dp) ? launch_paths<DPL_, false, true>(c, B, storeS) : launch_paths<DPL_, true, true>(c, B, storeS);   \
        else rc = (p.D == p.Dp) ? launch_paths<DPL_, false, false>(c, B, storeS) : launch_paths<DPL_, true, false>(c, B, storeS);    \
        break;
        CVR_DISPATCH(2) CVR_DISPATCH(4) CVR_DISPATCH(6) CVR_DISPATCH(8)
#undef CVR_DISPATCH
        default: c->err = "unsupported padded disparity count"; return CVR_ERR_UNSUPPORTED;
    }
    if (rc) return rc;
    rc = cvr_tail_split(c);                                            // pipelined batches: the rest goes to the tail stream
    if (rc) return rc;
    const size_t lrSmem = (size_t)p.W * 2 * sizeof(int);              // validate() keeps this within the opt-in limit (width <= 16384)
    if (lrSmem > 48 * 1024) CVR_CUDA(c, cudaFuncSetAttribute(k_lrcheck, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lrSmem));
    k_lrcheck<<<dim3(p.H, B), 256, lrSmem, c->stream>>>(c->d_raw, c->d_aux, c->d_disp, p, c->keep_S ? 1 : 0);
    LAUNCH_CHECK(c);
    if ((p.W & 1) == 0) k_median3_x2<<<dim3(cvr_div_up(p.W, 256), p.H, B), 128, 0, c->stream>>>(c->d_disp, c->d_med, p.W, p.H);
    else k_median3<<<dim3(cvr_div_up(p.W, 128), p.H, B), 128, 0, c->stream>>>(c->d_disp, c->d_med, p.W, p.H);
    LAUNCH_CHECK(c);
    rc = cvr_tail_aux_consumed(c);
    if (rc) return rc;
    cvr_stage_mark(c, "lr_median");
    if (c->prm.sgbm.speckle_window_size > 0) {
        rc = cvr_launch_speckle(c, B, p.INVALID, c->prm.sgbm.speckle_window_size, CVR_DISP_SCALE * c->prm.sgbm.speckle_range);
        if (rc) return rc;
        cvr_stage_mark(c, "speckle");
    }
    return CVR_OK;
}

// The batch size near `around` (within a quarter of it, at most max_batch) at which the vertical path kernel wastes the
// least: its clusters run in waves of as many as fit the GPU at once, so a batch that is a multiple of that count has no
// part-filled last wave.  Falls back to `around` where the kernel is not used (StereoBM, int32 fallback kernels).
template <int DPL, bool HASPAD, bool NN>
static int suggest_batch(cvr_ctx* c, int around)
{
    const int lo = around - around / 4 > 1 ? around - around / 4 : 1, hi = around + around / 4 < c->maxB ? around + around / 4 : c->maxB;
    int bestB = around < c->maxB ? around : c->maxB;
    double bestCost = -1;
    for (int B = lo; B <= hi; B++) {
        constexpr int LPRV = DPL <= 4 ? 8 : 16, NWDV = DPL * 16 / LPRV, NTV4 = 512;
        const VCfg vc = c->path_v_variant == 4 ? pick_vcfg(c, B, k_path_v4<NWDV, LPRV, false, false, HASPAD, NN, NTV4>, k_path_v4<NWDV, LPRV, false, false, HASPAD, NN, NTV4>, true, NTV4)
                                               : pick_vcfg(c, B, k_path_v3<DPL, false, false, HASPAD, NN, 1024>, k_path_v3<DPL, false, false, HASPAD, NN, 512>);
        if (!vc.ok) return bestB;
        const double perFrame = vc.cost / B;
        const bool closer = (B > around ? B - around : around - B) < (bestB > around ? bestB - around : around - bestB);
        if (bestCost < 0 || perFrame < bestCost * (1.0 - 1e-9) || (perFrame <= bestCost * (1.0 + 1e-9) && closer)) { bestCost = perFrame; bestB = B; }
    }
    return bestB;
}

int cvr_sgbm_suggest_batch(cvr_ctx* c, int around)
{
    const SgbmDev& p = c->sg;
    if (around < 1) around = 1;
    const int fallback = around < c->maxB ? around : c->maxB;
    if (c->prm.block_matcher != CVR_MATCHER_SGBM || p.W1 <= 0 || c->force_generic || p.P1s < 0) return fallback;
    const bool nn = !c->force_no_nn && p.P2s >= 0 && p.P1s >= p.P2s;
    const bool pad = p.D != p.Dp;
    switch (p.Dp / 32) {
#define CVR_SUGGEST(DPL_)                                                                                                \
    case DPL_:                                                                                                           \
        if (nn) return pad ? suggest_batch<DPL_, true, true>(c, around) : suggest_batch<DPL_, false, true>(c, around);    \
        return pad ? suggest_batch<DPL_, true, false>(c, around) : suggest_batch<DPL_, false, false>(c, around);
        CVR_SUGGEST(2) CVR_SUGGEST(4) CVR_SUGGEST(6) CVR_SUGGEST(8)
#undef CVR_SUGGEST
        default: return fallback;
    }
}
