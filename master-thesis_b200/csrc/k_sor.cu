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


struct SorGrid {
    float mn[3];
    float ext[3];        // extent of the bounding box (floored at a thousandth of the largest)
    float cell, inv;
    int dim[3];
    int ncell;
    int n;
    int r0;              // cube radius (in cells) the warp-per-cell search starts with
};

// cell edge s (grown until the grid fits `cap` cells) -> dims of the grid over the frame's bounding box
__device__ __forceinline__ void sor_fit_cells(SorGrid& g, float s, int cap)
{
    float ext[3], emax = 0.f;
    for (int k = 0; k < 3; k++) { ext[k] = g.ext[k]; emax = fmaxf(emax, ext[k]); }
    for (int it = 0; it < 16; it++) {
        long long tot = 1;
        for (int k = 0; k < 3; k++) { g.dim[k] = (int)(ext[k] / s) + 1; tot *= g.dim[k]; }
        if (tot <= cap) break;
        s *= 1.15f;
    }
    g.cell = s; g.inv = 1.0f / s;
    g.ncell = g.dim[0] * g.dim[1] * g.dim[2];
}
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
    for (int k = 0; k < 3; k++) { ext[k] = fmaxf(ext[k], emax * 1e-3f); vol *= (double)ext[k]; g.ext[k] = ext[k]; }
    // surface-like clouds end up sparser than two points per cell, which only costs time; clouds that fill a fraction of
    // their bounding box end up denser, which k_sor_regrid corrects
    float s = (float)cbrt(2.0 * vol / (double)max(g.n, 1));
    s = fmaxf(s, emax * 1e-4f);
    sor_fit_cells(g, s, cap);
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
        g.ext[k] = ext[k];
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
// Second look at the grid: a cloud that fills only part of its bounding box (a dense scene plus far outliers) has many more
// points per occupied cell than the two the bounding-box estimate aimed at, and every neighbour search would wade through
// them.  k_sor_occupancy counts the occupied cells; k_sor_regrid shrinks the cell edge of such a frame so that OCCUPIED cells
// hold about two points, and flags it; flagged frames get their cell tables cleared and counted again.
__global__ void __launch_bounds__(256) k_sor_occupancy(const int* __restrict__ cellcnt, size_t cellStride, const int* __restrict__ ncell, int* occ)
{
    const int b = blockIdx.y, len = ncell[b];
    int c = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < len; i += gridDim.x * blockDim.x) c += cellcnt[(size_t)b * cellStride + i] > 0;
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&occ[b], c);
}
__global__ void k_sor_regrid(SorGrid* grid, int* ncellOut, int* occ, int* flag, int cap, int B, int k)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    SorGrid g = grid[b];
    const int used = occ[b];
    occ[b] = 0;                                                       // ready for the next call
    const float per = used > 0 ? (float)g.n / (float)used : 0.f;
    int f = 0;
    float perNew = per;
    if (per > 3.0f && g.n > 4096) {
        const float old = g.cell;
        sor_fit_cells(g, g.cell * cbrtf(2.0f / per), cap);
        f = g.ncell > grid[b].ncell ? 1 : 0;                          // (the cap may have pushed the edge back up)
        if (f) perNew = per * (g.cell / old) * (g.cell / old) * (g.cell / old); else g = grid[b];
    }
    // radius of the sphere that holds k + 1 points at this density, in cells: where the cube search starts
    const float rk = cbrtf(3.0f * (float)(k + 1) / (4.0f * 3.14159265f * fmaxf(perNew, 0.05f)));
    g.r0 = min(max((int)ceilf(rk - 0.2f), 1), 3);
    grid[b] = g; ncellOut[b] = g.ncell;
    flag[b] = f;
}
__global__ void __launch_bounds__(256) k_sor_rezero(const int* __restrict__ flag, const int* __restrict__ ncell, int* cellA, int* cellB, size_t cellStride)
{
    const int b = blockIdx.y;
    if (!flag[b]) return;
    const int len = ncell[b] + 1;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < len; i += gridDim.x * blockDim.x) { cellA[(size_t)b * cellStride + i] = 0; cellB[(size_t)b * cellStride + i] = 0; }
}
__device__ __forceinline__ void sor_cell(const SorGrid& g, float x, float y, float z, int& cx, int& cy, int& cz)
{
    cx = min(max((int)((x - g.mn[0]) * g.inv), 0), g.dim[0] - 1);
    cy = min(max((int)((y - g.mn[1]) * g.inv), 0), g.dim[1] - 1);
    cz = min(max((int)((z - g.mn[2]) * g.inv), 0), g.dim[2] - 1);
}
__global__ void k_sor_cellcount(const cvr_point* __restrict__ pts, size_t stride, const SorGrid* __restrict__ grid, int* cellcnt, size_t cellStride,
                                int* __restrict__ cellOfPt, size_t ptStride, const int* __restrict__ onlyFlagged)
{
    const int b = blockIdx.y;
    if (onlyFlagged && !onlyFlagged[b]) return;
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
template <int KCAP>
__global__ void __launch_bounds__(128) k_sor_knn(const float4* __restrict__ spts, size_t ptStride, const SorGrid* __restrict__ grid,
                                                 const int* __restrict__ cellstart, size_t cellStride, int k, float* meand, int onlyMarked)
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
    if (onlyMarked && out[orig] != -1.f) continue;                    // already answered by k_sor_knn_warp
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
                    for (int j = jb; j < je; j++) {
                        const float4 p = sp[j];
                        const float dx = __fsub_rn(q.x, p.x), dy = __fsub_rn(q.y, p.y), dz = __fsub_rn(q.z, p.z);
                        float d2 = __fmul_rn(dx, dx);
                        d2 = __fadd_rn(d2, __fmul_rn(dy, dy));
                        d2 = __fadd_rn(d2, __fmul_rn(dz, dz));
                        // binary max-heap of the K1 smallest squared distances seen so far (root = the largest of them)
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

// ------------------------------------------------------------------------------------------------
// The same search, one WARP per occupied grid cell.  All queries of a cell see exactly the same candidates (the cube of
// cells around it), so the warp stages them once in shared memory (coalesced reads of the cell-sorted points) and then, per
// query, computes the squared distances with the lanes striding over the candidates and SELECTS the (k+1)-th smallest by a
// four-pass radix select on the float bits (a 256-bin histogram per pass) -- no per-thread heaps, no divergence between the
// queries of a warp.  The mean distance is the f64 sum of the square roots of everything below that value plus as many
// copies of it as are still missing (ties).  The cube starts at SorGrid::r0 cells and grows while the (k+1)-th distance is
// not yet provably final (same stop rule as k_sor_knn).  Queries whose cube holds more than KNW_CMAX candidates, or that
// are still open at KNW_RMAX, are marked (-1) and answered by k_sor_knn afterwards.
#define KNW_CMAX 512
#define KNW_WARPS 8
#define KNW_RMAX 3
#define KNW_RUNS 49           // (2 * KNW_RMAX + 1)^2 runs of cells along x
struct KnwWarp { float x[KNW_CMAX], y[KNW_CMAX], z[KNW_CMAX]; unsigned d2[KNW_CMAX]; int hist[256]; int runBeg[KNW_RUNS + 1], runOff[KNW_RUNS + 1]; };

__global__ void __launch_bounds__(KNW_WARPS * 32) k_sor_knn_warp(const float4* __restrict__ spts, size_t ptStride, const SorGrid* __restrict__ grid,
                                                                 const int* __restrict__ cellstart, size_t cellStride, int k, float* __restrict__ meand)
{
    extern __shared__ __align__(16) unsigned char knwsm[];
    KnwWarp& S = ((KnwWarp*)knwsm)[threadIdx.x >> 5];
    const int lane = threadIdx.x & 31, b = blockIdx.y;
    const SorGrid g = grid[b];
    const float4* sp = spts + (size_t)b * ptStride;
    const int* cs = cellstart + (size_t)b * cellStride;
    float* out = meand + (size_t)b * ptStride;
    const int K1 = k + 1;
    const int nWarps = gridDim.x * KNW_WARPS, wg = blockIdx.x * KNW_WARPS + (threadIdx.x >> 5);
    if (g.n < K1) {                                                    // nothing can be judged: every mean distance is 0 (PCL: search fails)
        for (int i = wg * 32 + lane; i < g.n; i += nWarps * 32) out[__float_as_int(sp[i].w)] = 0.f;
        return;
    }
    for (int cbase = wg * 32; cbase < g.ncell; cbase += nWarps * 32) {
        const int cell = cbase + lane;
        const int myBeg = cell < g.ncell ? cs[cell] : 0;
        const int myCnt = cell < g.ncell ? cs[cell + 1] - myBeg : 0;
        unsigned occ = __ballot_sync(0xffffffffu, myCnt > 0);
        while (occ) {
            const int l = __ffs(occ) - 1;
            occ &= occ - 1;
            const int c = cbase + l;
            const int qb = __shfl_sync(0xffffffffu, myBeg, l), qn = __shfl_sync(0xffffffffu, myCnt, l);
            const int cz = c / (g.dim[0] * g.dim[1]), cy = (c - cz * g.dim[0] * g.dim[1]) / g.dim[0], cx = c - (cz * g.dim[1] + cy) * g.dim[0];
            for (int q0 = 0; q0 < qn; q0 += 32) {                      // the cell's queries, 32 at a time
                unsigned pending = (qn - q0 >= 32) ? 0xffffffffu : ((1u << (qn - q0)) - 1u);
                for (int R = g.r0; pending; R++) {
                    if (R > KNW_RMAX) {                                // leave the rest to k_sor_knn
                        if ((pending >> lane) & 1u) out[__float_as_int(sp[qb + q0 + lane].w)] = -1.f;
                        break;
                    }
                    // ---- the cube's runs of cells along x: (2R+1)^2 of them, each a contiguous range of the sorted points ----
                    const int side = 2 * R + 1, nruns = side * side;
                    const int xa = max(cx - R, 0), xb = min(cx + R, g.dim[0] - 1);
                    int nc = 0;
                    for (int p0 = 0; p0 < nruns; p0 += 32) {
                        const int p = p0 + lane;
                        int jb = 0, len = 0;
                        if (p < nruns) {
                            const int iz = cz - R + p / side, iy = cy - R + p % side;
                            if (iz >= 0 && iz < g.dim[2] && iy >= 0 && iy < g.dim[1]) {
                                const int rowc = (iz * g.dim[1] + iy) * g.dim[0];
                                jb = cs[rowc + xa];
                                len = cs[rowc + xb + 1] - jb;
                            }
                        }
                        int inc = len;                                 // inclusive scan of the run lengths
#pragma unroll
                        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
                        if (p < nruns) { S.runBeg[p] = jb; S.runOff[p] = nc + inc - len; }
                        nc += __shfl_sync(0xffffffffu, inc, 31);
                    }
                    if (lane == 0) S.runOff[nruns] = nc;
                    const bool covered = cx - R <= 0 && cx + R >= g.dim[0] - 1 && cy - R <= 0 && cy + R >= g.dim[1] - 1 && cz - R <= 0 && cz + R >= g.dim[2] - 1;
                    if (nc > KNW_CMAX) {
                        if ((pending >> lane) & 1u) out[__float_as_int(sp[qb + q0 + lane].w)] = -1.f;
                        break;
                    }
                    if (nc < K1 && !covered) continue;                 // too few candidates to decide anything: a larger cube
                    __syncwarp();
                    // ---- stage the candidates ----
                    for (int r = 0; r < nruns; r++) {
                        const int jb = S.runBeg[r], off = S.runOff[r], len = S.runOff[r + 1] - off;
                        for (int t = lane; t < len; t += 32) { const float4 pt = sp[jb + t]; S.x[off + t] = pt.x; S.y[off + t] = pt.y; S.z[off + t] = pt.z; }
                    }
                    __syncwarp();
                    // ---- the pending queries, one after the other ----
                    const float lim = (float)R * g.cell, lim2 = lim * lim * (1.0f - 1e-5f);
                    unsigned todo = pending;
                    while (todo) {
                        const int qi = __ffs(todo) - 1;
                        todo &= todo - 1;
                        const float4 q = sp[qb + q0 + qi];
                        for (int t = lane; t < nc; t += 32) {
                            const float dx = __fsub_rn(q.x, S.x[t]), dy = __fsub_rn(q.y, S.y[t]), dz = __fsub_rn(q.z, S.z[t]);
                            float d = __fmul_rn(dx, dx);
                            d = __fadd_rn(d, __fmul_rn(dy, dy));
                            d = __fadd_rn(d, __fmul_rn(dz, dz));
                            S.d2[t] = __float_as_uint(d);                // non-negative floats order like their bit patterns
                        }
                        // radix select of the K1-th smallest: 8 bits per pass, most significant first
                        unsigned prefix = 0;
                        int need = K1;
#pragma unroll 1
                        for (int pass = 0; pass < 4; pass++) {
                            const int shift = 24 - 8 * pass;
#pragma unroll
                            for (int i = 0; i < 8; i++) S.hist[lane * 8 + i] = 0;
                            __syncwarp();
                            for (int t = lane; t < nc; t += 32) {
                                const unsigned v = S.d2[t];
                                if (pass == 0 || (v >> (shift + 8)) == prefix) atomicAdd(&S.hist[(v >> shift) & 255u], 1);
                            }
                            __syncwarp();
                            int h[8], sum = 0;
#pragma unroll
                            for (int i = 0; i < 8; i++) { h[i] = S.hist[lane * 8 + i]; sum += h[i]; }
                            int inc = sum;
#pragma unroll
                            for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
                            const int ex = inc - sum;
                            const unsigned hit = __ballot_sync(0xffffffffu, ex < need && need <= inc);
                            const int hl = __ffs(hit) - 1;             // the lane whose eight bins hold the K1-th smallest
                            int bin = 0, below = ex;
#pragma unroll
                            for (int i = 0; i < 8; i++) { if (below + h[i] < need) { below += h[i]; bin = i + 1; } else break; }
                            bin = __shfl_sync(0xffffffffu, lane * 8 + min(bin, 7), hl);
                            below = __shfl_sync(0xffffffffu, below, hl);
                            need -= below;
                            prefix = (prefix << 8) | (unsigned)bin;
                            __syncwarp();
                        }
                        const unsigned tau = prefix;                   // the K1-th smallest squared distance; `need` copies of it are still to be taken
                        if (!covered && !(__uint_as_float(tau) <= lim2)) continue;   // not provably final in this cube: stays pending
                        double s = 0.0;
                        for (int t = lane; t < nc; t += 32) { const unsigned v = S.d2[t]; if (v < tau) s = __dadd_rn(s, (double)__fsqrt_rn(__uint_as_float(v))); }
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) s = __dadd_rn(s, __shfl_xor_sync(0xffffffffu, s, o));
                        s = __dadd_rn(s, __dmul_rn((double)need, (double)__fsqrt_rn(__uint_as_float(tau))));
                        if (lane == 0) out[__float_as_int(q.w)] = (float)__ddiv_rn(s, (double)k);
                        pending &= ~(1u << qi);
                    }
                    __syncwarp();                                      // the staging area is refilled by the next cube / cell
                }
            }
        }
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
static int sor_build_grid(cvr_ctx* c, int B, const cvr_point* in, size_t in_stride, const int* n_in, bool haveBox, float radius, SorGrid** gridOut, int** cellstartOut)
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
    k_sor_cellcount<<<gp, 256, 0, c->stream>>>(in, in_stride, grid, cellstart, cellStride, c->d_cellidx, ptStride, nullptr);
    LAUNCH_CHECK(c);
    if (radius <= 0.f) {                                               // kNN grid: re-fit to the occupied cells where the cloud is much denser than its box
        int* occ = c->d_flags + (size_t)c->maxB * 33;
        int* flag = c->d_flags + (size_t)c->maxB * 34;
        const dim3 gc(cvr_frame_blocks(c, B, (size_t)cap + 1, 256, 8), B);
        k_sor_occupancy<<<gc, 256, 0, c->stream>>>(cellstart, cellStride, ncell, occ);
        LAUNCH_CHECK(c);
        k_sor_regrid<<<cvr_div_up(B, 128), 128, 0, c->stream>>>(grid, ncell, occ, flag, cap, B, c->cl.sor_k);
        LAUNCH_CHECK(c);
        k_sor_rezero<<<gc, 256, 0, c->stream>>>(flag, ncell, cellstart, cellfill, cellStride);
        LAUNCH_CHECK(c);
        k_sor_cellcount<<<gp, 256, 0, c->stream>>>(in, in_stride, grid, cellstart, cellStride, c->d_cellidx, ptStride, flag);
        LAUNCH_CHECK(c);
    }
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
    int rc = sor_build_grid(c, B, in, in_stride, n_in, haveBox, 0.f, &grid, &cellstart);
    if (rc) return rc;
    const float4* sp = c->d_spts;
    const dim3 g(cvr_frame_blocks(c, B, (size_t)maxN, 128, 12), B);
    const size_t heapBytes = (size_t)(k + 1) * 128 * sizeof(float);   // one heap column per thread
    if (heapBytes > 48 * 1024) CVR_CUDA(c, cudaFuncSetAttribute(k_sor_knn<320>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)heapBytes));
    static const bool warpSearch = !(getenv("CVR_SOR_WARP") && atoi(getenv("CVR_SOR_WARP")) == 0);   // test hook: 0 = per-thread heaps only
    if (warpSearch && k + 1 <= KNW_CMAX / 2) {
        const size_t wsm = sizeof(KnwWarp) * KNW_WARPS;
        CVR_CUDA(c, cudaFuncSetAttribute(k_sor_knn_warp, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsm));
        const int cells = c->sor_grid_cap;
        k_sor_knn_warp<<<dim3(cvr_frame_blocks(c, B, (size_t)cells / 32 + 1, KNW_WARPS, 2), B), KNW_WARPS * 32, wsm, c->stream>>>(sp, ptStride, grid, cellstart, cellStride, k, c->d_meand);
        LAUNCH_CHECK(c);
        k_sor_knn<320><<<g, 128, heapBytes, c->stream>>>(sp, ptStride, grid, cellstart, cellStride, k, c->d_meand, 1);   // whatever the warp search left open
    } else {
        k_sor_knn<320><<<g, 128, heapBytes, c->stream>>>(sp, ptStride, grid, cellstart, cellStride, k, c->d_meand, 0);
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
    int rc = sor_build_grid(c, B, in, in_stride, n_in, false, radius, &grid, &cellstart);
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
