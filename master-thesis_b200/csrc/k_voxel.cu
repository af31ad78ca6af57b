// k_voxel.cu -- pcl::VoxelGrid<PointXYZRGB>::applyFilter for the call site
// cv_reconstruct/src/system/MappingSystem.cpp:297-300 (setLeafSize(0.2,0.2,0.2)).
//   bounding box -> int32 voxel index per point -> stable LSD radix sort of (index, point id) ->
//   segment heads -> one thread per voxel accumulates the centroid in sorted (= point id) order in f32.
// The sort is hand-written: 8-bit digits, one warp per 512-element run, __match_any_sync ranking (stable).
#include "cvr_internal.cuh"
#include "cvr_scan.cuh"
#include <float.h>


#define RS_WARP_ELEMS 1024           // elements ranked by one warp per pass
#define RS_BLOCK_WARPS 8

__global__ void k_vox_setup(const VoxBox* box, const int* nPtr, VoxelHdr* hdr, float lx, float ly, float lz, int B)
{
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    VoxelHdr h;
    h.n = nPtr[b]; h.m = 0; h.passthrough = 0; h.maxkey = 0; h.nruns = 0; h.pad = 0;
    const float inv[3] = { __fdiv_rn(1.0f, lx), __fdiv_rn(1.0f, ly), __fdiv_rn(1.0f, lz) };
    long long cells = 1;
    for (int k = 0; k < 3; k++) {
        h.mn[k] = h.n > 0 ? ord2f(box[b].mn[k]) : 0.f; h.mx[k] = h.n > 0 ? ord2f(box[b].mx[k]) : 0.f;
        const long long dk = (long long)__fmul_rn(__fsub_rn(h.mx[k], h.mn[k]), inv[k]) + 1;
        cells *= dk;                                       // PCL multiplies three int64 (each < 2^40 here would overflow only for absurd clouds)
        h.minb[k] = (int)floorf(__fmul_rn(h.mn[k], inv[k]));
        const int maxb = (int)floorf(__fmul_rn(h.mx[k], inv[k]));
        h.div[k] = maxb - h.minb[k] + 1;
    }
    if (cells > 2147483647LL || cells <= 0) h.passthrough = 1;
    h.maxkey = h.passthrough ? 0 : (int)(cells - 1);          // keys are < cells: radix passes above its top bit are skipped
    hdr[b] = h;
}
__global__ void k_vox_keys(const cvr_point* __restrict__ pts, size_t stride, const VoxelHdr* __restrict__ hdr, float lx, float ly, float lz,
                           uint32_t* __restrict__ pkeys, size_t kvStride)
{
    const int b = blockIdx.y;
    const VoxelHdr h = hdr[b];
    const float ix = __fdiv_rn(1.0f, lx), iy = __fdiv_rn(1.0f, ly), iz = __fdiv_rn(1.0f, lz);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < h.n; i += gridDim.x * blockDim.x) {
        const float4 q = *(const float4*)(pts + (size_t)b * stride + i);
        const int i0 = (int)floorf(__fmul_rn(q.x, ix)) - h.minb[0];
        const int i1 = (int)floorf(__fmul_rn(q.y, iy)) - h.minb[1];
        const int i2 = (int)floorf(__fmul_rn(q.z, iz)) - h.minb[2];
        pkeys[(size_t)b * kvStride + i] = h.passthrough ? 0u : (uint32_t)(i0 + i1 * h.div[0] + i2 * h.div[0] * h.div[1]);
    }
}
// Points arrive in pixel order, and neighbouring pixels mostly fall into the same voxel: the cloud is a sequence of RUNS of
// consecutive points with equal voxel index, several times fewer than points.  Only the runs are sorted (stable, by index);
// a voxel is then a stretch of sorted runs, its members the points of those runs in ascending point order -- exactly the
// member order a stable sort of the points would give, which is the order the centroid sums are defined in.
// Run heads by the single-pass ordered selection: run r starts at point runStart[r] and has voxel index runKey[r].
__global__ void __launch_bounds__(CVR_SEL_THREADS) k_vox_runs(const uint32_t* __restrict__ pkeys, size_t kvStride, VoxelHdr* hdr, int* sel, size_t selStride,
                                                              uint32_t* __restrict__ runKey, uint32_t* __restrict__ runId, int* __restrict__ runStart)
{
    __shared__ int ws[33];
    __shared__ int sh;
    const int b = blockIdx.y, n = hdr[b].n, nch = (n + CVR_SEL_BLOCK - 1) / CVR_SEL_BLOCK;
    const bool pass = hdr[b].passthrough != 0;
    const uint32_t* k = pkeys + (size_t)b * kvStride;
    uint32_t* rk = runKey + (size_t)b * kvStride; uint32_t* ri = runId + (size_t)b * kvStride;
    int* rs = runStart + (size_t)b * kvStride;
    unsigned* status = (unsigned*)sel + (size_t)b * selStride;
    int* ticket = (int*)status + (selStride - 1);
    if (nch == 0) { if (blockIdx.x == 0 && threadIdx.x == 0) hdr[b].nruns = 0; return; }
    for (;;) {
        const int chunk = sel_next_chunk(ticket, &sh);
        if (chunk >= nch) break;
        const int i0 = chunk * CVR_SEL_BLOCK + threadIdx.x * CVR_SEL_ITEMS;
        int f[CVR_SEL_ITEMS], cnt = 0;
        uint32_t key[CVR_SEL_ITEMS];
        uint32_t prev = (i0 > 0 && i0 - 1 < n) ? k[i0 - 1] : 0u;
#pragma unroll
        for (int j = 0; j < CVR_SEL_ITEMS; j++) {
            const int i = i0 + j;
            key[j] = i < n ? k[i] : 0u;
            f[j] = (i < n) && (pass || i == 0 || key[j] != prev);
            prev = key[j]; cnt += f[j];
        }
        int tot;
        const int e = block_excl_scan(cnt, ws, &tot);
        const int base = sel_chain_prefix(status, chunk, tot, &sh);
        int o = base + e;
#pragma unroll
        for (int j = 0; j < CVR_SEL_ITEMS; j++) if (f[j]) { rk[o] = key[j]; ri[o] = (uint32_t)o; rs[o] = i0 + j; o++; }
        if (chunk == nch - 1 && threadIdx.x == 0) hdr[b].nruns = base + tot;
    }
}

// ---- stable LSD radix sort, 8 bits per pass ------------------------------------------------------
// hist layout per frame: [run][256] (run-major).  k_rs_scan turns it, per digit, into the global start of that
// run's elements with that digit.  Passes whose digit lies above the top bit of the largest possible key
// (VoxelHdr::maxkey) only copy the pairs through, so the ping-pong order stays fixed without a host round trip.
__device__ __forceinline__ bool rs_pass_needed(const VoxelHdr* hdr, int b, int shift) { return !hdr || ((unsigned)hdr[b].maxkey >> shift) != 0 || shift == 0; }
// The needed passes are a prefix of the four (digits above the key range are all zero), every needed pass flips the
// ping-pong buffer and a skipped one leaves the pairs in place: the sorted pairs of frame b are in buffer (#needed & 1).
__device__ __forceinline__ int rs_result_buf(const VoxelHdr* hdr, int b)
{
    const unsigned mk = (unsigned)hdr[b].maxkey;
    return (1 + (mk >> 8 ? 1 : 0) + (mk >> 16 ? 1 : 0) + (mk >> 24 ? 1 : 0)) & 1;
}

__global__ void __launch_bounds__(RS_BLOCK_WARPS * 32) k_rs_hist(const uint32_t* __restrict__ keys, size_t kvStride, const int* __restrict__ nPtr,
                                                                 int nOff, int shift, int* __restrict__ hist, size_t histStride, int nRuns,
                                                                 const VoxelHdr* __restrict__ hdr)
{
    __shared__ int sh[RS_BLOCK_WARPS][256];
    const int b = blockIdx.y, n = nPtr[b * nOff];
    if (!rs_pass_needed(hdr, b, shift)) return;
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t* kk = keys + (size_t)b * kvStride;
    for (int run = blockIdx.x * RS_BLOCK_WARPS + w; run < nRuns; run += gridDim.x * RS_BLOCK_WARPS) {   // a warp per 1024-element run
        const int beg = run * RS_WARP_ELEMS, end = min(beg + RS_WARP_ELEMS, n);
        if (beg >= end) break;
        for (int k = lane; k < 256; k += 32) sh[w][k] = 0;
        __syncwarp();
        for (int i = beg + lane; i < end; i += 32) atomicAdd(&sh[w][(kk[i] >> shift) & 255], 1);
        __syncwarp();
        int* h = hist + (size_t)b * histStride + (size_t)run * 256;
        for (int k = lane; k < 256; k += 32) h[k] = sh[w][k];
        __syncwarp();
    }
}
// one CTA of 256 threads per frame: thread d walks the runs of digit d (coalesced 1 KB rows)
__global__ void __launch_bounds__(256) k_rs_scan(int* __restrict__ hist, size_t histStride, const int* __restrict__ nPtr, int nOff, int shift,
                                                 const VoxelHdr* __restrict__ hdr)
{
    __shared__ int ws[33];
    const int b = blockIdx.x, n = nPtr[b * nOff], d = threadIdx.x;
    if (!rs_pass_needed(hdr, b, shift)) return;
    const int nr = (n + RS_WARP_ELEMS - 1) / RS_WARP_ELEMS;
    int* h = hist + (size_t)b * histStride + d;
    int sum = 0;
    for (int r = 0; r < nr; r++) { const int v = h[(size_t)r * 256]; h[(size_t)r * 256] = sum; sum += v; }
    const int base = block_excl_scan(sum, ws, nullptr);           // start of digit d in the output
    for (int r = 0; r < nr; r++) h[(size_t)r * 256] += base;
}
__global__ void __launch_bounds__(RS_BLOCK_WARPS * 32) k_rs_scatter(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                                                                    uint32_t* __restrict__ okeys, uint32_t* __restrict__ ovals, size_t kvStride,
                                                                    const int* __restrict__ nPtr, int nOff, int shift, const int* __restrict__ hist,
                                                                    size_t histStride, int nRuns, const VoxelHdr* __restrict__ hdr)
{
    __shared__ int base[RS_BLOCK_WARPS][256];
    const int b = blockIdx.y, n = nPtr[b * nOff];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t* kk = keys + (size_t)b * kvStride; const uint32_t* vv = vals + (size_t)b * kvStride;
    uint32_t* ok = okeys + (size_t)b * kvStride; uint32_t* ov = ovals + (size_t)b * kvStride;
    if (!rs_pass_needed(hdr, b, shift)) return;                        // the pairs stay where they are: see rs_result_buf
    for (int run = blockIdx.x * RS_BLOCK_WARPS + w; run < nRuns; run += gridDim.x * RS_BLOCK_WARPS) {
        const int beg = run * RS_WARP_ELEMS, end = min(beg + RS_WARP_ELEMS, n);
        if (beg >= end) break;
        const int* h = hist + (size_t)b * histStride + (size_t)run * 256;
        for (int k = lane; k < 256; k += 32) base[w][k] = h[k];
        __syncwarp();
        for (int i0 = beg; i0 < end; i0 += 32) {
            const int i = i0 + lane;
            const bool act = i < end;
            const unsigned amask = __ballot_sync(0xffffffffu, act);
            if (act) {
                const uint32_t key = kk[i], val = vv[i];
                const int dg = (key >> shift) & 255;
                const unsigned peers = __match_any_sync(amask, dg);
                const int rank = __popc(peers & ((1u << lane) - 1));
                const int dst = base[w][dg] + rank;
                ok[dst] = key; ov[dst] = val;
                __syncwarp(amask);
                if (rank == 0) base[w][dg] += __popc(peers);
            }
            __syncwarp();
        }
    }
}

// ---- segments + centroids -------------------------------------------------------------------------
// Segment heads of the sorted keys (first index of every voxel) by the single-pass ordered selection.
__global__ void __launch_bounds__(CVR_SEL_THREADS) k_seg_select(const uint32_t* __restrict__ keys0, const uint32_t* __restrict__ keys1, size_t kvStride, VoxelHdr* hdr, int* sel, size_t selStride,
                                                                int* __restrict__ segstart, size_t segStride)
{
    __shared__ int ws[33];
    __shared__ int sh;
    const int b = blockIdx.y, n = hdr[b].nruns, nch = (n + CVR_SEL_BLOCK - 1) / CVR_SEL_BLOCK;   // over the SORTED RUNS
    const bool pass = hdr[b].passthrough != 0;
    const uint32_t* k = (rs_result_buf(hdr, b) ? keys1 : keys0) + (size_t)b * kvStride;
    int* ss = segstart + (size_t)b * segStride;
    unsigned* status = (unsigned*)sel + (size_t)b * selStride;
    int* ticket = (int*)status + (selStride - 1);
    if (nch == 0) { if (blockIdx.x == 0 && threadIdx.x == 0) { ss[0] = 0; hdr[b].m = 0; } return; }
    for (;;) {
        const int chunk = sel_next_chunk(ticket, &sh);
        if (chunk >= nch) break;
        const int i0 = chunk * CVR_SEL_BLOCK + threadIdx.x * CVR_SEL_ITEMS;
        int f[CVR_SEL_ITEMS], cnt = 0;
        uint32_t prev = (i0 > 0 && i0 - 1 < n) ? k[i0 - 1] : 0u;
#pragma unroll
        for (int j = 0; j < CVR_SEL_ITEMS; j++) {
            const int i = i0 + j;
            const uint32_t cur = i < n ? k[i] : 0u;
            f[j] = (i < n) && (pass || i == 0 || cur != prev);
            prev = cur; cnt += f[j];
        }
        int tot;
        const int e = block_excl_scan(cnt, ws, &tot);
        const int base = sel_chain_prefix(status, chunk, tot, &sh);
        int* o = ss + base + e;
#pragma unroll
        for (int j = 0; j < CVR_SEL_ITEMS; j++) if (f[j]) *o++ = i0 + j;
        if (chunk == nch - 1 && threadIdx.x == 0) { ss[base + tot] = n; hdr[b].m = base + tot; }
    }
}
// One thread per voxel: PCL's CentroidPoint accumulates x,y,z and r,g,b,a as floats in member order, divides by
// the count, truncates the colours.  Member order = ascending point index (stable sort), as in the oracle.
// Also accumulates the bounding box of the centroids (the outlier filter that follows builds its search grid from it).
__global__ void __launch_bounds__(128) k_vox_centroid(const cvr_point* __restrict__ pts, size_t stride, const uint32_t* __restrict__ vals0, const uint32_t* __restrict__ vals1, size_t kvStride,
                               const VoxelHdr* __restrict__ hdr, const int* __restrict__ segstart, size_t segStride, const int* __restrict__ runStart,
                               cvr_point* __restrict__ out, size_t outStride, int* __restrict__ mOut, VoxBox* box)
{
    const int b = blockIdx.y;
    const VoxelHdr h = hdr[b];
    if (blockIdx.x == 0 && threadIdx.x == 0) mOut[b] = h.m;
    const cvr_point* p = pts + (size_t)b * stride;
    const int* ss = segstart + (size_t)b * segStride;
    const uint32_t* v = (rs_result_buf(hdr, b) ? vals1 : vals0) + (size_t)b * kvStride;   // sorted run ids
    const int* rs = runStart + (size_t)b * kvStride;
    unsigned mn[3] = { 0xffffffffu, 0xffffffffu, 0xffffffffu }, mx[3] = { 0, 0, 0 };
    bool any = false;
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < h.m; j += gridDim.x * blockDim.x) {
        cvr_point o;
        if (h.passthrough) o = p[j];
        else {
            const int s = ss[j], e = ss[j + 1];                        // the voxel's stretch of sorted runs
            float sx = 0, sy = 0, sz = 0, sr = 0, sg = 0, sb = 0, sa = 0;
            int cntPts = 0;
            auto acc = [&](const uint4& w) {                           // one member, in member order (the sums are sequential f32 adds)
                sx = __fadd_rn(sx, __uint_as_float(w.x)); sy = __fadd_rn(sy, __uint_as_float(w.y)); sz = __fadd_rn(sz, __uint_as_float(w.z));
                sb = __fadd_rn(sb, (float)(w.w & 255u)); sg = __fadd_rn(sg, (float)((w.w >> 8) & 255u)); sr = __fadd_rn(sr, (float)((w.w >> 16) & 255u));
                sa = __fadd_rn(sa, (float)(w.w >> 24));
            };
            const uint4* p4 = (const uint4*)p;
            for (int r = s; r < e; r++) {                              // runs in ascending run id = ascending point index (stable sort)
                const int rid = (int)v[r];
                const int pb = rs[rid], pe = rid + 1 < h.nruns ? rs[rid + 1] : h.n;
                int i = pb;
                for (; i + 4 <= pe; i += 4) {                          // a run's points are contiguous: four loads in flight
                    const uint4 q0 = p4[i], q1 = p4[i + 1], q2 = p4[i + 2], q3 = p4[i + 3];
                    acc(q0); acc(q1); acc(q2); acc(q3);
                }
                for (; i < pe; i++) acc(p4[i]);
                cntPts += pe - pb;
            }
            const float cnt = (float)cntPts;
            o.x = __fdiv_rn(sx, cnt); o.y = __fdiv_rn(sy, cnt); o.z = __fdiv_rn(sz, cnt);
            o.r = (uint8_t)__fdiv_rn(sr, cnt); o.g = (uint8_t)__fdiv_rn(sg, cnt); o.b = (uint8_t)__fdiv_rn(sb, cnt); o.a = (uint8_t)__fdiv_rn(sa, cnt);
        }
        out[(size_t)b * outStride + j] = o;
        const unsigned ox = f2ord(o.x), oy = f2ord(o.y), oz = f2ord(o.z);
        mn[0] = min(mn[0], ox); mx[0] = max(mx[0], ox); mn[1] = min(mn[1], oy); mx[1] = max(mx[1], oy); mn[2] = min(mn[2], oz); mx[2] = max(mx[2], oz);
        any = true;
    }
    if (box && __any_sync(0xffffffffu, any)) {
#pragma unroll
        for (int k = 0; k < 3; k++) { mn[k] = __reduce_min_sync(0xffffffffu, mn[k]); mx[k] = __reduce_max_sync(0xffffffffu, mx[k]); }
        if ((threadIdx.x & 31) == 0)
            for (int k = 0; k < 3; k++) { atomicMin(&box[b].mn[k], mn[k]); atomicMax(&box[b].mx[k], mx[k]); }
    }
}

// Sort (keys, vals) of every frame: d_keys[0]/d_vals[0] -> ... ; returns the index (0/1) of the buffer holding the result.
int cvr_radix_sort_pairs(cvr_ctx* c, int B, const int* nPtr, int nOff, int maxN, int passes, const VoxelHdr* hdr, int* result_buf)
{
    const int nRuns = cvr_div_up(maxN, RS_WARP_ELEMS);
    const size_t histStride = (size_t)256 * nRuns + 1;
    const size_t kvStride = (size_t)c->maxN;
    const dim3 g(cvr_frame_blocks(c, B, (size_t)nRuns, RS_BLOCK_WARPS, 8), B);
    int cur = 0;
    for (int pass = 0; pass < passes; pass++) {
        const int shift = pass * 8;
        k_rs_hist<<<g, RS_BLOCK_WARPS * 32, 0, c->stream>>>(c->d_keys[cur], kvStride, nPtr, nOff, shift, (int*)c->d_hist, histStride, nRuns, hdr);
        LAUNCH_CHECK(c);
        k_rs_scan<<<B, 256, 0, c->stream>>>((int*)c->d_hist, histStride, nPtr, nOff, shift, hdr);
        LAUNCH_CHECK(c);
        k_rs_scatter<<<g, RS_BLOCK_WARPS * 32, 0, c->stream>>>(c->d_keys[cur], c->d_vals[cur], c->d_keys[cur ^ 1], c->d_vals[cur ^ 1], kvStride, nPtr,
                                                                nOff, shift, (const int*)c->d_hist, histStride, nRuns, hdr);
        LAUNCH_CHECK(c);
        cur ^= 1;
    }
    *result_buf = cur;
    return CVR_OK;
}

int cvr_launch_voxel(cvr_ctx* c, int B, const cvr_point* in, size_t in_stride, const int* n_in, bool haveBox)
{
    const int maxN = c->maxN;
    const size_t kvStride = (size_t)maxN;
    const float lx = c->cl.leaf[0], ly = c->cl.leaf[1], lz = c->cl.leaf[2];
    if (!haveBox) {                                                    // stand-alone call: measure the input's bounding box
        int rc0 = cvr_launch_chain_init(c, B, false, true, true);
        if (rc0) return rc0;
        k_bbox_minmax<<<dim3(cvr_frame_blocks(c, B, (size_t)maxN, 256, 8), B), 256, 0, c->stream>>>(in, in_stride, n_in, c->d_boxv);
        LAUNCH_CHECK(c);
    }
    const size_t selStride = cvr_sel_stride(maxN);
    int* sel = c->d_sel + (size_t)c->maxB * selStride;                 // selection 1 of the chain (run heads), then selection 0 again (voxel heads)
    int* sel2 = c->d_sel;
    CVR_CUDA(c, cudaMemsetAsync(sel, 0, (size_t)B * selStride * sizeof(int), c->stream));
    CVR_CUDA(c, cudaMemsetAsync(sel2, 0, (size_t)B * selStride * sizeof(int), c->stream));
    k_vox_setup<<<cvr_div_up(B, 128), 128, 0, c->stream>>>(c->d_boxv, n_in, c->d_vhdr, lx, ly, lz, B);
    LAUNCH_CHECK(c);
    uint32_t* pkeys = (uint32_t*)c->d_cellidx;                         // per-point voxel index (scratch of the outlier filters, idle here)
    int* runStart = (int*)c->d_meand;                                  // first point of every run (ditto)
    const dim3 gp(cvr_frame_blocks(c, B, (size_t)maxN, 256, 8), B);
    k_vox_keys<<<gp, 256, 0, c->stream>>>(in, in_stride, c->d_vhdr, lx, ly, lz, pkeys, kvStride);
    LAUNCH_CHECK(c);
    const int nch = cvr_div_up(maxN, CVR_SEL_BLOCK);
    const dim3 gs(cvr_frame_blocks(c, B, (size_t)nch, 1, 8), B);
    k_vox_runs<<<gs, CVR_SEL_THREADS, 0, c->stream>>>(pkeys, kvStride, c->d_vhdr, sel, selStride, c->d_keys[0], c->d_vals[0], runStart);
    LAUNCH_CHECK(c);
    int buf = 0;
    const int hdrInts = (int)(sizeof(VoxelHdr) / sizeof(int));
    int rc = cvr_radix_sort_pairs(c, B, &c->d_vhdr[0].nruns, hdrInts, maxN, 4, c->d_vhdr, &buf);   // sorts the RUNS; the result buffer is per frame: rs_result_buf
    if (rc) return rc;
    k_seg_select<<<gs, CVR_SEL_THREADS, 0, c->stream>>>(c->d_keys[0], c->d_keys[1], kvStride, c->d_vhdr, sel2, selStride, c->d_segstart, (size_t)maxN + 1);
    LAUNCH_CHECK(c);
    k_vox_centroid<<<dim3(cvr_frame_blocks(c, B, (size_t)maxN, 128, 8), B), 128, 0, c->stream>>>(in, in_stride, c->d_vals[0], c->d_vals[1], kvStride, c->d_vhdr, c->d_segstart,
                                                                                                (size_t)maxN + 1, runStart, c->d_vox, (size_t)maxN, c->d_nvox, c->d_boxs);
    LAUNCH_CHECK(c);
    cvr_stage_mark(c, "voxel");
    return CVR_OK;
}
