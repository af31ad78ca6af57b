// k_sor.cu -- pcl::StatisticalOutlierRemoval<PointXYZRGB> for the call sites
// cv_reconstruct/src/point_cloud/PointCloudPostProcessor.cpp:20-24, 38-41 (k = outlier_min_k, alpha = outlier_std_threshold)
// and cv_reconstruct/src/system/MappingSystem.cpp:65-69.
//   uniform grid over the cloud -> counting sort of points by cell -> exact k-nearest-neighbour search by
//   expanding cubic shells until the k-th distance is provably final -> mean distance per point ->
//   global mean / sample sigma in f64 (fixed-order reduction) -> ordered compaction of the kept points.
// Squared distances are accumulated x,y,z in f32 without FMA (FLANN L2 on a CPU build without -mfma).
#include "cvr_internal.cuh"
#include "cvr_scan.cuh"
#include <float.h>



// the search grid of one frame from its bounding box: about two points per cell on a volume-filling cloud
__device__ __forceinline__ SorGrid sor_make_grid(const VoxBox& box, int n, int cap)
{
    SorGrid g;
    g.n = n;
    float ext[3], emax = 0.f;
    for (int k = 0; k < 3; k++) {
        g.mn[k] = g.n > 0 ? ord2f(box.mn[k]) : 0.f;
        const float mx = g.n > 0 ? ord2f(box.mx[k]) : 0.f;
        ext[k] = mx - g.mn[k];
        emax = fmaxf(emax, ext[k]);
    }
    if (!(emax > 0.f)) emax = 1.f;
    double vol = 1.0;
    for (int k = 0; k < 3; k++) { ext[k] = fmaxf(ext[k], emax * 1e-3f); vol *= (double)ext[k]; }
    // surface-like clouds end up sparser than two points per cell, which only costs time
    float s = (float)cbrt(2.0 * vol / (double)max(g.n, 1));
    s = fmaxf(s, emax * 1e-4f);
    for (int it = 0; it < 8; it++) {
        long long tot = 1;
        for (int k = 0; k < 3; k++) { g.dim[k] = (int)(ext[k] / s) + 1; tot *= g.dim[k]; }
        if (tot <= cap) break;
        s *= 1.3f;
    }
    g.cell = s; g.inv = 1.0f / s;
    g.ncell = g.dim[0] * g.dim[1] * g.dim[2];
    return g;
}
// pcl::RadiusOutlierRemoval's grid: cell edge = the radius (a multiple of it when the grid would get too large)
__device__ __forceinline__ SorGrid ror_make_grid(const VoxBox& box, int n, float radius, int cap)
{
    SorGrid g;
    g.n = n;
    float ext[3];
    for (int k = 0; k < 3; k++) {
        g.mn[k] = g.n > 0 ? ord2f(box.mn[k]) : 0.f;
        const float mx = g.n > 0 ? ord2f(box.mx[k]) : 0.f;
        ext[k] = fmaxf(mx - g.mn[k], 0.f);
    }
    float s = radius;
    for (int it = 0; it < 32; it++) {
        long long tot = 1;
        for (int k = 0; k < 3; k++) { g.dim[k] = (int)(ext[k] / s) + 1; tot *= g.dim[k]; }
        if (tot <= cap) break;
        s *= 1.5f;
    }
    g.cell = s; g.inv = 1.0f / s;
    g.ncell = g.dim[0] * g.dim[1] * g.dim[2];
    return g;
}
// Builds the frame's grid (every block computes it, block 0 stores it) and clears the two cell tables it needs.
__global__ void __launch_bounds__(256) k_sor_grid_zero(const VoxBox* __restrict__ box, const int* __restrict__ nPtr, SorGrid* grid, int* ncellOut, int cap, float radius,
                                                       int* cellA, int* cellB, size_t cellStride)
{
    __shared__ SorGrid sg;
    const int b = blockIdx.y;
    if (threadIdx.x == 0) {
        sg = radius > 0.f ? ror_make_grid(box[b], nPtr[b], radius, cap) : sor_make_grid(box[b], nPtr[b], cap);
        if (blockIdx.x == 0) { grid[b] = sg; ncellOut[b] = sg.ncell; }
    }
    __syncthreads();
    const int len = sg.ncell + 1;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < len; i += gridDim.x * blockDim.x) { cellA[(size_t)b * cellStride + i] = 0; cellB[(size_t)b * cellStride + i] = 0; }
}
__global__ void k_sor_cellcount(const cvr_point* __restrict__ pts, size_t stride, const SorGrid* __restrict__ grid, int* cellcnt, size_t cellStride,
                                int* __restrict__ cellOfPt, size_t ptStride)
{
    const int b = blockIdx.y;
    const SorGrid g = grid[b];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < g.n; i += gridDim.x * blockDim.x) {
        const float4 q = *(const float4*)(pts + (size_t)b * stride + i);
        int cx, cy, cz;
        sor_cell(g, q.x, q.y, q.z, cx, cy, cz);
        const int cid = (cz * g.dim[1] + cy) * g.dim[0] + cx;
        cellOfPt[(size_t)b * ptStride + i] = cid;
        atomicAdd(&cellcnt[(size_t)b * cellStride + cid], 1);
    }
}
__global__ void k_sor_scatter(const cvr_point* __restrict__ pts, size_t stride, const SorGrid* __restrict__ grid, const int* __restrict__ cellstart,
                              int* cellfill, size_t cellStride, const int* __restrict__ cellOfPt, size_t ptStride, float4* __restrict__ spts)
{
    const int b = blockIdx.y, n = grid[b].n;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int cid = cellOfPt[(size_t)b * ptStride + i];
        const int pos = cellstart[(size_t)b * cellStride + cid] + atomicAdd(&cellfill[(size_t)b * cellStride + cid], 1);
        float4 q = *(const float4*)(pts + (size_t)b * stride + i);
        q.w = __int_as_float(i);
        spts[(size_t)b * ptStride + pos] = q;
    }
}

// One thread per query point (in cell-sorted order).  best[] holds the k+1 smallest squared distances seen as a binary
// max-heap; the query itself (distance 0) is one of them, exactly like nearestKSearch(i, k+1).  The heap lives in SHARED
// memory, one column per thread (best[i] of thread t at heap[i * 128 + t]: conflict-free however the threads' heap
// positions diverge); as a per-thread array it was dynamically indexed and ended up in local memory.
// Measured and rejected for the 20 M-point case (k = 50; this kernel: 53 ms): a warp per occupied cell with the candidates shared
// through shared memory and the heap updates batched (98 - 140 ms: the batched sifts still run with 5 of 32 lanes busy), the
// same with a radix select (115 ms), re-fitting the cell size to the occupied cells (51 - 123 ms for 5 .. 1 points per cell).
template <int KCAP, int MINB, bool UNROLL>
__global__ void __launch_bounds__(128, MINB) k_sor_knn(const float4* __restrict__ spts, size_t ptStride, const SorGrid* __restrict__ grid,
                                                 const int* __restrict__ cellstart, size_t cellStride, int k, float* __restrict__ meand)
{
    const int b = blockIdx.y;
    const SorGrid g = grid[b];
    const float4* sp = spts + (size_t)b * ptStride;
    const int* cs = cellstart + (size_t)b * cellStride;
    float* out = meand + (size_t)b * ptStride;
    const int K1 = k + 1;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < g.n; t += gridDim.x * blockDim.x) {
    const float4 q = sp[t];
    const int orig = __float_as_int(q.w);
    if (g.n < K1) { out[orig] = 0.f; continue; }
    extern __shared__ float heapsm[];
    struct HeapCol { float* p; __device__ __forceinline__ float& operator[](int i) const { return p[i * 128]; } };
    const HeapCol best{ heapsm + threadIdx.x };
    int cnt = 0;
    int cx, cy, cz;
    sor_cell(g, q.x, q.y, q.z, cx, cy, cz);
    const int Rmax = max(max(g.dim[0], g.dim[1]), g.dim[2]);
    for (int R = 0; R <= Rmax; R++) {
        const int z0 = max(cz - R, 0), z1 = min(cz + R, g.dim[2] - 1);
        const int y0 = max(cy - R, 0), y1 = min(cy + R, g.dim[1] - 1);
        for (int iz = z0; iz <= z1; iz++) {
            const bool zf = (iz == cz - R) || (iz == cz + R);
            for (int iy = y0; iy <= y1; iy++) {
                const bool face = zf || (iy == cy - R) || (iy == cy + R);
                // on a z/y face of the shell the whole x run belongs to it; elsewhere only the two x end caps
                const int xs = cx - R, xe = cx + R;
                for (int part = 0; part < 2; part++) {
                    int xa, xb;
                    if (face) { if (part) break; xa = max(xs, 0); xb = min(xe, g.dim[0] - 1); }
                    else {
                        const int xx = part ? xe : xs;
                        if (part && R == 0) break;
                        if (xx < 0 || xx >= g.dim[0]) continue;
                        xa = xb = xx;
                    }
                    if (xa > xb) continue;
                    const int rowc = (iz * g.dim[1] + iy) * g.dim[0];
                    const int jb = cs[rowc + xa], je = cs[rowc + xb + 1];     // cells of one x run are contiguous in the sorted order
                    // four candidates per turn: their loads are issued together (with one load per turn the warp waited on each
                    // candidate in turn: a third of all stall samples at 20 M points), then each is offered to the heap
                    auto dist2 = [&](int jj) -> float {
                        const float4 p = sp[jj];
                        const float dx = __fsub_rn(q.x, p.x), dy = __fsub_rn(q.y, p.y), dz = __fsub_rn(q.z, p.z);
                        float d2 = __fmul_rn(dx, dx);
                        d2 = __fadd_rn(d2, __fmul_rn(dy, dy));
                        return __fadd_rn(d2, __fmul_rn(dz, dz));
                    };
                    // binary max-heap of the K1 smallest squared distances seen so far (root = the largest of them)
                    auto offer = [&](const float d2) {
                        if (cnt < K1) {
                            int pz = cnt++;
                            while (pz > 0) { const int par = (pz - 1) >> 1; const float pv = best[par]; if (pv >= d2) break; best[pz] = pv; pz = par; }
                            best[pz] = d2;
                        } else if (d2 < best[0]) {
                            int pz = 0;
                            for (;;) {
                                int ch = 2 * pz + 1;
                                if (ch >= K1) break;
                                float cv = best[ch];
                                if (ch + 1 < K1) { const float c2 = best[ch + 1]; if (c2 > cv) { cv = c2; ch++; } }
                                if (cv <= d2) break;
                                best[pz] = cv; pz = ch;
                            }
                            best[pz] = d2;
                        }
                    };
                    if (UNROLL) {
                        for (int j = jb; j < je; j += 4) {
                            const int last = je - 1;
                            const float e0 = dist2(j), e1 = dist2(min(j + 1, last)), e2 = dist2(min(j + 2, last)), e3 = dist2(min(j + 3, last));
                            offer(e0);
                            if (j + 1 < je) offer(e1);
                            if (j + 2 < je) offer(e2);
                            if (j + 3 < je) offer(e3);
                        }
                    } else {                                           // the small clouds of a batch: x runs of a few points, nothing to overlap
                        for (int j = jb; j < je; j++) offer(dist2(j));
                    }
                }
            }
        }
        if (cnt == K1) {
            // every unvisited point lies outside the cube of half-width R*cell around the query's cell, hence farther
            // than R*cell; the 1e-5 margin covers the f32 rounding of both sides
            const float lim = (float)R * g.cell;
            if (best[0] <= lim * lim * (1.0f - 1e-5f)) break;
        }
    }
    // heap sort in place (ascending), so that the sum below runs in the order PCL's sorted neighbour list gives it
    for (int m = K1 - 1; m > 0; m--) {
        const float v = best[m];
        best[m] = best[0];
        int pz = 0;
        for (;;) {
            int ch = 2 * pz + 1;
            if (ch >= m) break;
            float cv = best[ch];
            if (ch + 1 < m) { const float c2 = best[ch + 1]; if (c2 > cv) { cv = c2; ch++; } }
            if (cv <= v) break;
            best[pz] = cv; pz = ch;
        }
        best[pz] = v;
    }
    double s = 0.0;
    for (int j = 1; j < K1; j++) s = __dadd_rn(s, (double)__fsqrt_rn(best[j]));
    out[orig] = (float)__ddiv_rn(s, (double)k);
    }
}

// mean / sample variance over all mean distances, f64, fixed reduction order (deterministic).
__global__ void __launch_bounds__(1024) k_sor_stats(const float* __restrict__ meand, size_t ptStride, const int* __restrict__ nPtr, int k, double alpha, SorStats* st)
{
    __shared__ double ss[1024], sq[1024];
    const int b = blockIdx.x, n = nPtr[b];
    const float* d = meand + (size_t)b * ptStride;
    double s = 0.0, q = 0.0;
    for (int i = threadIdx.x; i < n; i += 1024) { const float v = d[i]; s = __dadd_rn(s, (double)v); q = __dadd_rn(q, (double)__fmul_rn(v, v)); }
    ss[threadIdx.x] = s; sq[threadIdx.x] = q;
    __syncthreads();
    for (int o = 512; o > 0; o >>= 1) {
        if (threadIdx.x < o) { ss[threadIdx.x] = __dadd_rn(ss[threadIdx.x], ss[threadIdx.x + o]); sq[threadIdx.x] = __dadd_rn(sq[threadIdx.x], sq[threadIdx.x + o]); }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const int valid = (n >= k + 1) ? n : 0;
        const double sum = ss[0], sqs = sq[0];
        const double mean = __ddiv_rn(sum, (double)valid);
        const double var = __ddiv_rn(__dsub_rn(sqs, __ddiv_rn(__dmul_rn(sum, sum), (double)valid)), __dsub_rn((double)valid, 1.0));
        const double thr = __dadd_rn(mean, __dmul_rn(alpha, sqrt(var)));
        st[b].sum = sum; st[b].sq = sqs; st[b].thr = thr; st[b].valid = valid;
    }
}
// kept points (mean distance not above the threshold) in input order, by the single-pass ordered selection
__global__ void __launch_bounds__(CVR_SEL_THREADS) k_sor_select(const cvr_point* __restrict__ pts, size_t stride, const float* __restrict__ meand, size_t ptStride,
                                                                const int* __restrict__ nPtr, SorStats* st, int* sel, size_t selStride,
                                                                cvr_point* __restrict__ out, size_t outStride, int* __restrict__ nOut)
{
    __shared__ int ws[33];
    __shared__ int sh;
    const int b = blockIdx.y, n = nPtr[b], nch = (n + CVR_SEL_BLOCK - 1) / CVR_SEL_BLOCK;
    const double thr = st[b].thr;
    unsigned* status = (unsigned*)sel + (size_t)b * selStride;
    int* ticket = (int*)status + (selStride - 1);
    if (nch == 0) { if (blockIdx.x == 0 && threadIdx.x == 0) { nOut[b] = 0; st[b].kept = 0; } return; }
    const float* md = meand + (size_t)b * ptStride;
    const cvr_point* pin = pts + (size_t)b * stride;
    for (;;) {
        const int chunk = sel_next_chunk(ticket, &sh);
        if (chunk >= nch) break;
        const int i0 = chunk * CVR_SEL_BLOCK + threadIdx.x * CVR_SEL_ITEMS;
        int f[CVR_SEL_ITEMS], cnt = 0;
#pragma unroll
        for (int j = 0; j < CVR_SEL_ITEMS; j++) { f[j] = (i0 + j < n) && !((double)md[i0 + j] > thr); cnt += f[j]; }
        int tot;
        const int e = block_excl_scan(cnt, ws, &tot);
        const int base = sel_chain_prefix(status, chunk, tot, &sh);
        cvr_point* o = out + (size_t)b * outStride + base + e;
#pragma unroll
        for (int j = 0; j < CVR_SEL_ITEMS; j++) if (f[j]) { *(uint4*)o = *(const uint4*)(pin + i0 + j); o++; }
        if (chunk == nch - 1 && threadIdx.x == 0) { nOut[b] = base + tot; st[b].kept = base + tot; }
    }
}

// exclusive scan of the per-frame cell counts: one CTA per frame while the table is short, two levels when it is long
static int scan_cells(cvr_ctx* c, int B, int* cellstart, size_t cellStride, int cap, const int* ncell)
{
    if (cap <= (1 << 20)) {                                          // the single-CTA scan only walks the cells in use (lenPtr)
        k_scan_excl<<<B, 1024, 0, c->stream>>>(cellstart, cellStride, cap, ncell, 1);
        LAUNCH_CHECK(c);
        return CVR_OK;
    }
    const int nChunks = cvr_div_up(cap, 1024) + 1;                    // + 1: the chunk that holds index len when len == cap
    const size_t auxStride = (size_t)nChunks + 1;
    k_scan_chunks<<<dim3(nChunks, B), 1024, 0, c->stream>>>(cellstart, cellStride, cap, ncell, c->d_scanaux, auxStride);
    LAUNCH_CHECK(c);
    k_scan_excl<<<B, 1024, 0, c->stream>>>(c->d_scanaux, auxStride, nChunks, nullptr, 1);
    LAUNCH_CHECK(c);
    k_scan_addoff<<<dim3(nChunks, B), 1024, 0, c->stream>>>(cellstart, cellStride, cap, ncell, c->d_scanaux, auxStride, nChunks);
    LAUNCH_CHECK(c);
    return CVR_OK;
}

// shared front half of the two outlier filters: bounding box (unless the producer accumulated it), grid, points sorted by cell
int cvr_build_point_grid(cvr_ctx* c, int B, const cvr_point* in, size_t in_stride, const int* n_in, bool haveBox, float radius, SorGrid** gridOut, int** cellstartOut)
{
    const int maxN = c->maxN;
    const size_t ptStride = (size_t)maxN;
    const int cap = c->sor_grid_cap;
    const size_t cellStride = (size_t)cap + 2;
    SorGrid* grid = (SorGrid*)(c->d_flags + (size_t)c->maxB * 16);
    int* ncell = c->d_flags + (size_t)c->maxB * 32;
    int* cellstart = c->d_cellstart;
    int* cellfill = c->d_cellstart + (size_t)c->maxB * cellStride;
    if (!haveBox) {
        const int rc0 = cvr_launch_chain_init(c, B, false, false, true);
        if (rc0) return rc0;
        k_bbox_minmax<<<dim3(cvr_frame_blocks(c, B, (size_t)maxN, 256, 8), B), 256, 0, c->stream>>>(in, in_stride, n_in, c->d_boxs);
        LAUNCH_CHECK(c);
    }
    const dim3 gp(cvr_frame_blocks(c, B, (size_t)maxN, 256, 8), B);     // strides over the frame's points / cells
    k_sor_grid_zero<<<dim3(cvr_frame_blocks(c, B, (size_t)cap + 1, 256, 8), B), 256, 0, c->stream>>>(c->d_boxs, n_in, grid, ncell, cap, radius, cellstart, cellfill, cellStride);
    LAUNCH_CHECK(c);
    k_sor_cellcount<<<gp, 256, 0, c->stream>>>(in, in_stride, grid, cellstart, cellStride, c->d_cellidx, ptStride);
    LAUNCH_CHECK(c);
    { const int rcs = scan_cells(c, B, cellstart, cellStride, cap, ncell); if (rcs) return rcs; }
    k_sor_scatter<<<gp, 256, 0, c->stream>>>(in, in_stride, grid, cellstart, cellfill, cellStride, c->d_cellidx, ptStride, c->d_spts);
    LAUNCH_CHECK(c);
    *gridOut = grid; *cellstartOut = cellstart;
    return CVR_OK;
}
static int sor_select(cvr_ctx* c, int B, const cvr_point* in, size_t in_stride, const int* n_in)
{
    const int maxN = c->maxN;
    const size_t selStride = cvr_sel_stride(maxN);
    int* sel = c->d_sel + (size_t)2 * c->maxB * selStride;             // selection 2 of the chain
    CVR_CUDA(c, cudaMemsetAsync(sel, 0, (size_t)B * selStride * sizeof(int), c->stream));
    const int nch = cvr_div_up(maxN, CVR_SEL_BLOCK);
    k_sor_select<<<dim3(cvr_frame_blocks(c, B, (size_t)nch, 1, 8), B), CVR_SEL_THREADS, 0, c->stream>>>(in, in_stride, c->d_meand, (size_t)maxN, n_in, c->d_sstats, sel, selStride,
                                                                                                  c->d_sorout, (size_t)maxN, c->d_nout);
    LAUNCH_CHECK(c);
    return CVR_OK;
}

int cvr_launch_sor(cvr_ctx* c, int B, const cvr_point* in, size_t in_stride, const int* n_in, bool haveBox)
{
    const int maxN = c->maxN;
    const size_t ptStride = (size_t)maxN;
    const size_t cellStride = (size_t)c->sor_grid_cap + 2;
    const int k = c->cl.sor_k;
    if (k < 1 || k + 1 > 320) { c->err = "sor_mean_k must be in [1, 319]"; return CVR_ERR_UNSUPPORTED; }
    SorGrid* grid; int* cellstart;
    int rc = cvr_build_point_grid(c, B, in, in_stride, n_in, haveBox, 0.f, &grid, &cellstart);
    if (rc) return rc;
    const float4* sp = c->d_spts;
    const size_t heapBytes = (size_t)(k + 1) * 128 * sizeof(float);   // one heap column per thread
    // a grid-stride kernel: exactly as many CTAs as are resident at once (a part-filled second wave cost a quarter of the time at 20 M points)
    // two forms: 8 CTAs per SM (62 registers) and four candidate loads in flight for one large cloud; 10 CTAs per SM and the plain
    // candidate loop for the many small clouds of a batch
    const bool big = B == 1 && maxN >= (1 << 18);
    const int regCap = big ? 8 : 12;                                   // (12: the batch form as measured best, a little more than is resident at once)
    int perSM = (int)((size_t)c->smem_sm / (heapBytes + 1024));
    perSM = perSM > regCap ? regCap : (perSM < 1 ? 1 : perSM);
    const dim3 g(cvr_frame_blocks(c, B, (size_t)maxN, 128, perSM), B);
    if (big) {
        if (heapBytes > 48 * 1024) CVR_CUDA(c, cudaFuncSetAttribute(k_sor_knn<320, 8, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)heapBytes));
        k_sor_knn<320, 8, true><<<g, 128, heapBytes, c->stream>>>(sp, ptStride, grid, cellstart, cellStride, k, c->d_meand);
    } else {
        if (heapBytes > 48 * 1024) CVR_CUDA(c, cudaFuncSetAttribute(k_sor_knn<320, 10, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)heapBytes));
        k_sor_knn<320, 10, false><<<g, 128, heapBytes, c->stream>>>(sp, ptStride, grid, cellstart, cellStride, k, c->d_meand);
    }
    LAUNCH_CHECK(c);
    k_sor_stats<<<B, 1024, 0, c->stream>>>(c->d_meand, ptStride, n_in, k, c->cl.sor_alpha, c->d_sstats);
    LAUNCH_CHECK(c);
    rc = sor_select(c, B, in, in_stride, n_in);
    if (rc) return rc;
    cvr_stage_mark(c, "sor");
    return CVR_OK;
}

// ------------------------------------------------------------------------------------------------
// SURVEY 8(f) row n3 -- what MappingSystem does with the cloud right after the path.
// pcl::RadiusOutlierRemoval (cv_reconstruct/src/system/MappingSystem.cpp:303-307): keep a point iff at least min_pts + 1 points
// (itself included) lie within `radius` (squared distance <= r^2).  Same uniform grid as the SOR search, cell = radius.
__global__ void __launch_bounds__(128) k_ror_count(const float4* __restrict__ spts, size_t ptStride, const SorGrid* __restrict__ grid,
                                                   const int* __restrict__ cellstart, size_t cellStride, float radius, int min_pts, float* __restrict__ flag)
{
    const int b = blockIdx.y;
    const SorGrid g = grid[b];
    const float4* sp = spts + (size_t)b * ptStride;
    const int* cs = cellstart + (size_t)b * cellStride;
    const float r2 = __fmul_rn(radius, radius);
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < g.n; t += gridDim.x * blockDim.x) {
    const float4 q = sp[t];
    int cx, cy, cz;
    sor_cell(g, q.x, q.y, q.z, cx, cy, cz);
    const int R = (int)ceilf(radius * g.inv);           // rings of cells that can hold a point within the radius (1 unless the grid was coarsened)
    int cnt = 0;
    for (int iz = max(cz - R, 0); iz <= min(cz + R, g.dim[2] - 1) && cnt <= min_pts; iz++)
        for (int iy = max(cy - R, 0); iy <= min(cy + R, g.dim[1] - 1) && cnt <= min_pts; iy++) {
            const int rowc = (iz * g.dim[1] + iy) * g.dim[0];
            const int jb = cs[rowc + max(cx - R, 0)], je = cs[rowc + min(cx + R, g.dim[0] - 1) + 1];
            for (int j = jb; j < je && cnt <= min_pts; j++) {
                const float4 p = sp[j];
                const float dx = __fsub_rn(q.x, p.x), dy = __fsub_rn(q.y, p.y), dz = __fsub_rn(q.z, p.z);
                float d2 = __fmul_rn(dx, dx);
                d2 = __fadd_rn(d2, __fmul_rn(dy, dy));
                d2 = __fadd_rn(d2, __fmul_rn(dz, dz));
                cnt += d2 <= r2;
            }
        }
    // reuse the SOR compaction: "mean distance" 0 keeps, 1 drops against a threshold of 0.5
    flag[(size_t)b * ptStride + __float_as_int(q.w)] = cnt > min_pts ? 0.f : 1.f;
    }
}
__global__ void k_ror_thr(SorStats* st, int B) { int b = blockIdx.x * blockDim.x + threadIdx.x; if (b < B) st[b].thr = 0.5; }

int cvr_launch_ror(cvr_ctx* c, int B, const cvr_point* in, size_t in_stride, const int* n_in, float radius, int min_pts)
{
    const int maxN = c->maxN;
    const size_t ptStride = (size_t)maxN;
    const size_t cellStride = (size_t)c->sor_grid_cap + 2;
    SorGrid* grid; int* cellstart;
    int rc = cvr_build_point_grid(c, B, in, in_stride, n_in, false, radius, &grid, &cellstart);
    if (rc) return rc;
    k_ror_count<<<dim3(cvr_frame_blocks(c, B, (size_t)maxN, 128, 12), B), 128, 0, c->stream>>>(c->d_spts, ptStride, grid, cellstart, cellStride, radius, min_pts, c->d_meand);
    LAUNCH_CHECK(c);
    k_ror_thr<<<cvr_div_up(B, 128), 128, 0, c->stream>>>(c->d_sstats, B);
    LAUNCH_CHECK(c);
    rc = sor_select(c, B, in, in_stride, n_in);
    if (rc) return rc;
    cvr_stage_mark(c, "ror");
    return CVR_OK;
}

// pcl::transformPointCloud (MappingSystem.cpp:318): p' = T * (x, y, z, 1) in f32, ((col0*x + col1*y) + col2*z) + col3, no FMA.
struct Mat4 { float m[16]; };
__global__ void k_transform_cloud(const cvr_point* __restrict__ in, cvr_point* __restrict__ out, int n, Mat4 T)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    cvr_point p = in[i];
    float v[3];
#pragma unroll
    for (int r = 0; r < 3; r++)
        v[r] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T.m[r * 4 + 0], p.x), __fmul_rn(T.m[r * 4 + 1], p.y)), __fmul_rn(T.m[r * 4 + 2], p.z)), T.m[r * 4 + 3]);
    p.x = v[0]; p.y = v[1]; p.z = v[2];
    out[i] = p;
}
int cvr_launch_transform(cvr_ctx* c, const cvr_point* in, cvr_point* out, int n, const float* T)
{
    Mat4 m;
    for (int i = 0; i < 16; i++) m.m[i] = T[i];
    k_transform_cloud<<<cvr_div_up(n, 256), 256, 0, c->stream>>>(in, out, n, m);
    LAUNCH_CHECK(c);
    return CVR_OK;
}
