// cvr_scan.cuh -- small per-frame scan / ordered-compaction building blocks (hand-written; no CUB).
// Every batch kernel uses blockIdx.y (or .z) as the frame index and reads its element count from device
// memory, so a whole batch is processed without a host round trip.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define CVR_SEL_BLOCK 1024   // elements per chunk of the ordered selections
#define CVR_SEL_THREADS 256  // threads of a selection block ...
#define CVR_SEL_ITEMS 4      // ... each owning this many consecutive elements of the chunk

// Block-wide exclusive scan of one int per thread (blockDim.x multiple of 32, <= 1024).  Returns the exclusive
// prefix; *total receives the block sum.  `ws` is >= 33 ints of shared memory.
__device__ __forceinline__ int block_excl_scan(int v, int* ws, int* total)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) ws[w] = inc;
    __syncthreads();
    if (w == 0) {
        int s = lane < nw ? ws[lane] : 0;
        int si = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, si, o); if (lane >= o) si += t; }
        ws[lane] = si - s;            // exclusive warp offsets
        if (lane == 31) ws[32] = si;  // block total
    }
    __syncthreads();
    const int r = ws[w] + inc - v;
    if (total) *total = ws[32];
    __syncthreads();
    return r;
}

// Exclusive scan of data[b][0..len) in place, one CTA per frame (grid.x = B); data[b][len] receives the total.
// lenPtr (optional, device) gives a per-frame length <= maxLen; entries beyond it are treated as 0.
static __global__ void k_scan_excl(int* data, size_t frameStride, int maxLen, const int* lenPtr, int lenDiv)
{
    __shared__ int ws[33];
    const int b = blockIdx.x;
    int* d = data + (size_t)b * frameStride;
    int len = maxLen;
    if (lenPtr) len = min(maxLen, (lenPtr[b] + lenDiv - 1) / lenDiv);
    int carry = 0;
    for (int base = 0; base < len; base += blockDim.x) {
        const int i = base + threadIdx.x;
        const int v = i < len ? d[i] : 0;
        int tot;
        const int e = block_excl_scan(v, ws, &tot);
        if (i < len) d[i] = carry + e;
        carry += tot;
    }
    if (threadIdx.x == 0) d[len] = carry;
}

// Two-level form of k_scan_excl for long arrays (the cell table of a multi-million-point cloud): k_scan_chunks scans
// chunks of 1024 in place and writes the chunk totals to aux[b][chunk]; k_scan_excl turns those into chunk offsets (and the
// grand total in aux[b][nChunks]); k_scan_addoff adds them and stores the total at data[b][len].  grid = (nChunks, B).
static __global__ void __launch_bounds__(1024) k_scan_chunks(int* data, size_t frameStride, int maxLen, const int* lenPtr, int* aux, size_t auxStride)
{
    __shared__ int ws[33];
    const int b = blockIdx.y, i = blockIdx.x * 1024 + threadIdx.x;
    int* d = data + (size_t)b * frameStride;
    const int len = lenPtr ? min(maxLen, lenPtr[b]) : maxLen;
    const int v = i < len ? d[i] : 0;
    int tot;
    const int e = block_excl_scan(v, ws, &tot);
    if (i < len) d[i] = e;
    if (threadIdx.x == 0) aux[(size_t)b * auxStride + blockIdx.x] = tot;
}
static __global__ void __launch_bounds__(1024) k_scan_addoff(int* data, size_t frameStride, int maxLen, const int* lenPtr, const int* aux, size_t auxStride,
                                                             int nChunks)
{
    const int b = blockIdx.y, i = blockIdx.x * 1024 + threadIdx.x;
    int* d = data + (size_t)b * frameStride;
    const int len = lenPtr ? min(maxLen, lenPtr[b]) : maxLen;
    if (i < len) d[i] += aux[(size_t)b * auxStride + blockIdx.x];
    if (i == len) d[len] = aux[(size_t)b * auxStride + nChunks];
}

// ---- per-frame bounding box of a point array (float min/max through an order-preserving uint mapping) ----
__device__ __forceinline__ unsigned f2ord(float f) { unsigned u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__device__ __forceinline__ float ord2f(unsigned u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u); }

static __global__ void k_bbox_init(VoxBox* box, int B)
{
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    for (int k = 0; k < 3; k++) { box[b].mn[k] = 0xffffffffu; box[b].mx[k] = 0u; }
}
// points are 16-byte records whose first 12 bytes are x,y,z (cvr_point)
static __global__ void k_bbox_minmax(const void* __restrict__ pts, size_t stride, const int* __restrict__ nPtr, VoxBox* box)
{
    const int b = blockIdx.y, n = nPtr[b];
    const float4* p = (const float4*)pts + (size_t)b * stride;
    unsigned mn[3] = { 0xffffffffu, 0xffffffffu, 0xffffffffu }, mx[3] = { 0, 0, 0 };
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 q = p[i];
        unsigned a = f2ord(q.x), c = f2ord(q.y), d = f2ord(q.z);
        mn[0] = min(mn[0], a); mx[0] = max(mx[0], a); mn[1] = min(mn[1], c); mx[1] = max(mx[1], c); mn[2] = min(mn[2], d); mx[2] = max(mx[2], d);
    }
#pragma unroll
    for (int k = 0; k < 3; k++) { mn[k] = __reduce_min_sync(0xffffffffu, mn[k]); mx[k] = __reduce_max_sync(0xffffffffu, mx[k]); }
    if ((threadIdx.x & 31) == 0 && n > 0)
        for (int k = 0; k < 3; k++) { atomicMin(&box[b].mn[k], mn[k]); atomicMax(&box[b].mx[k], mx[k]); }
}

// ---- single-pass ordered selection ("chained scan") ----------------------------------------------------------------
// An ordered compaction in ONE pass over its input instead of count -> scan -> write: the blocks of a frame draw chunk
// numbers from a ticket counter (so a chunk's predecessors have always been started by a running block: no deadlock
// whatever the number of resident blocks), each publishes its chunk's total and looks back along the status words of the
// earlier chunks until it meets one that already carries an inclusive prefix.
// Per frame: sel[0 .. nchMax] status words, sel[nchMax + 1] the ticket counter; all zero before the launch
// (cvr_sel_stride ints per frame).  Status word: bit 31 = inclusive prefix, bit 30 = chunk aggregate, low 30 bits = value.
#define CVR_SEL_INCL 0x80000000u
#define CVR_SEL_AGG 0x40000000u
static_assert(CVR_SEL_BLOCK == 1024, "cvr_sel_stride (cvr_internal.cuh) assumes 1024-element chunks");

__device__ __forceinline__ int sel_next_chunk(int* ticket, int* sh)
{
    if (threadIdx.x == 0) *sh = atomicAdd(ticket, 1);
    __syncthreads();
    const int c = *sh;
    __syncthreads();
    return c;
}
// exclusive prefix of chunk `chunk` over the frame, given the chunk's own total (all threads call; sh = one shared int).
// The look-back is done by the first warp, 32 predecessors at a time: every lane reads one status word, the nearest
// inclusive prefix ends the search and the aggregates in front of it are summed with one warp reduction.
__device__ __forceinline__ int sel_chain_prefix(unsigned* status, int chunk, int total, int* sh)
{
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        volatile unsigned* st = status;
        if (lane == 0 && chunk > 0) st[chunk] = CVR_SEL_AGG | (unsigned)total;
        int run = 0;
        for (int hi = chunk - 1; hi >= 0;) {                            // window [hi - 31, hi], lane l looks at hi - l
            const int j = hi - lane;
            unsigned v;
            do { v = j >= 0 ? st[j] : CVR_SEL_INCL; } while (__any_sync(0xffffffffu, v == 0u));   // (entries before the frame count as an inclusive 0)
            const unsigned incl = __ballot_sync(0xffffffffu, (v & CVR_SEL_INCL) != 0u);
            const int stop = incl ? __ffs(incl) - 1 : 32;               // nearest lane (= nearest predecessor) that carries an inclusive prefix
            const int add = lane <= stop ? (int)(v & 0x3fffffffu) : 0;
            run += __reduce_add_sync(0xffffffffu, add);
            if (incl) break;
            hi -= 32;
        }
        if (lane == 0) { st[chunk] = CVR_SEL_INCL | (unsigned)(run + total); *sh = run; }
    }
    __syncthreads();
    const int r = *sh;
    __syncthreads();
    return r;
}
