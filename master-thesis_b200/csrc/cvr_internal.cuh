// cvr_internal.cuh -- shared declarations of libcvrecon_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>

#include "../../include/cvrecon_b200.h"

#define CVR_MAX_COST 32767
#define CVR_DISP_SCALE 16

// Normalised StereoSGBM parameters, passed by value to kernels.
struct SgbmDev {
    int W, H;            // image size
    int Wp;              // gray pitch in bytes
    int minD, D, Dp;     // Dp = D padded to a multiple of 64 (DPL = Dp/32 values per lane)
    int bs, SW2, SH2;
    int P1s;             // (short)P1
    int P2;              // full int P2 (delta = (short)(P2 + minLr))
    int P2s;             // (short)P2 : bias of the cost volume
    int uniq, d12, ftzero;
    int minX1, maxX1, W1;
    int INVALID;
    int mode;            // CVR_MODE_*
    int ring;            // pow2 >= bs + 1 (pixel-cost ring buffer of the horizontal box sum)
};

struct BmDev {
    int W, H, Wp;
    int minD, ndisp, Dp;
    int wsz, wsz2, cap, textureThreshold, uniq;
    int lofs, rofs, width1;
    int FILTERED;
    int ring;            // pow2 >= wsz + 1
};

struct CloudDev {
    float Q[16];
    float f, b, cx, cy;
    double sigma_factor;
    float leaf[3];
    int sor_k;
    double sor_alpha;
};

// Per-frame voxel-grid header computed on device.
struct VoxelHdr {
    float mn[3], mx[3];
    int minb[3];
    int div[3];
    int passthrough;     // int32 index overflow: PCL copies input through
    int n;               // number of input points
    int m;               // number of voxels (output points)
    int maxkey;
    int nruns;           // number of runs of consecutive points with equal voxel index (what is actually sorted)
    int pad;
};

// per-frame bounding box of a point array: floats through an order-preserving uint mapping (f2ord, cvr_scan.cuh)
struct VoxBox { unsigned mn[3], mx[3]; };

struct PruneStats {
    int mn, mx;
    unsigned long long sum, sq;
    float a, b;
    int ithr;
    int pad;
};

struct SorStats {
    double sum, sq;      // over mean distances
    double thr;
    int valid;
    int kept;
};

// uniform search grid over one frame's cloud (k_sor.cu builds it; the outlier filters and the ICP search walk it)
struct SorGrid {
    float mn[3];
    float cell, inv;
    int dim[3];
    int ncell;
    int n;
};
#ifdef __CUDACC__
__device__ __forceinline__ void sor_cell(const SorGrid& g, float x, float y, float z, int& cx, int& cy, int& cz)
{
    cx = min(max((int)((x - g.mn[0]) * g.inv), 0), g.dim[0] - 1);
    cy = min(max((int)((y - g.mn[1]) * g.inv), 0), g.dim[1] - 1);
    cz = min(max((int)((z - g.mn[2]) * g.inv), 0), g.dim[2] - 1);
}
#endif

struct StageTimer {
    const char* name;
    cudaEvent_t ev;
};

// One in-flight batch of the pipelined host path (cvr_submit_batch / cvr_wait_batch).
struct PipeSlot {
    uint8_t* d_in[2];        // staged BGR/gray input of this slot
    cvr_point* d_out;        // filtered clouds of the batch, packed back to back (capacity B * maxN)
    int* d_meta;             // offsets / gather header / overflow flag of the packed payload (layout: k_pack_clouds)
    int* h_meta;             // pinned copy of d_meta, valid once evDone has fired
    int16_t* d_outdisp;      // [B][H][W]
    cudaEvent_t evH2D, evDone;   // inputs uploaded / results parked, slot reusable
    bool busy, used;
    cvr_batch io;
    int nframes;
    bool hasCloud;
};

struct cvr_ctx {
    cvr_params prm;
    SgbmDev sg;
    BmDev bm;
    CloudDev cl;
    int device;
    int maxB;
    int maxN;                // point capacity per frame (>= W*H unless filter_only)
    int sm_count;
    cudaStream_t stream;
    std::string err;
    long long launches;
    bool profiling;
    std::vector<StageTimer> timers[3];   // event timers of the synchronous calls [0] and of the two pipeline slots [1], [2]
    int timers_used[3];
    int tset;                            // the set cvr_stage_mark appends to
    std::vector<std::string> t_names;
    std::vector<float> t_ms;

    // ---- device workspace (all sized for maxB frames) ----
    uint8_t* d_in[2];        // staged input images (host path), tight: [B][H][W*ch]
    size_t in_bytes_per_frame;
    uint8_t* d_gray;         // [B][2][H][Wp]
    uint8_t* d_bmpre;        // StereoBM: x-Sobel prefiltered images [B][2][H][Wp]
    int16_t* d_volA;         // hsum, later S : [B][H][W1][Dp]
    int16_t* d_volC;         // C             : [B][H][W1][Dp]
    int16_t* d_lr;           // vertical path state [B][2][W1+2][3][Dp]
    int* d_minlr;            // [B][2][W1+2][4]
    int16_t* d_raw;          // WTA output [B][H][W]
    int2* d_aux;             // WTA result per computed pixel: {minS<<16 | best, S[best-1]<<16 | S[best+1]&0xffff}, x = 0x7fffffff if rejected [B][H][W1]
    int16_t* d_disp;         // after LR check  [B][H][W]
    int16_t* d_med;          // after median / speckle [B][H][W]   (final disparity)
    int* d_label;            // speckle CCL [B][H][W]
    int* d_size;             // speckle component sizes [B][H][W]
    int16_t* d_pruned;       // pruned disparity [B][H][W]
    float* d_dispf;          // [B][H][W]
    uint8_t* d_mask;         // [B][H][W]
    PruneStats* d_pstats;    // [B]
    int* d_sel;              // status words + ticket of the ordered selections (cvr_scan.cuh): [3 uses][B][cvr_sel_stride(maxN)]
    VoxBox* d_boxv;         // per-frame bounding box of the cloud entering the voxel grid [B]
    VoxBox* d_boxs;         // ... of the cloud entering the outlier filters [B]
    cvr_point* d_cloud;      // dense cloud [B][H*W]
    int* d_ncloud;           // [B]
    // voxel
    VoxelHdr* d_vhdr;        // [B]
    uint32_t* d_keys[2];     // [B][H*W]
    uint32_t* d_vals[2];     // [B][H*W]
    uint32_t* d_hist;        // radix histograms [B][256][nTiles]
    int sort_tiles;          // tiles per frame
    int* d_segstart;         // voxel segment starts [B][H*W + 1]
    cvr_point* d_vox;        // voxel output [B][H*W]
    // sor
    int* d_cellstart;        // grid cell starts [B][cells+1]
    int* d_scanaux;          // chunk totals of the two-level cell scan [B][cells/1024 + 3]
    int* d_cellidx;          // point -> cell id
    float* d_meand;          // mean kNN distance [B][H*W]
    SorStats* d_sstats;      // [B]
    cvr_point* d_sorout;     // [B][H*W]
    int* d_nout;             // [B] final count
    int* d_nvox;             // [B] voxel count
    short2* d_rmap1[2];      // rectification: CV_16SC2 map per camera [H][W] (cvr_set_rectification)
    uint16_t* d_rmap2[2];    // rectification: CV_16UC1 fractional map per camera [H][W]
    uint8_t* d_rect[2];      // rectified left / right images of the current batch [B][H][W*4]
    bool rect_ready;
    float4* d_spts;          // sor: points sorted by grid cell (w = original index) [B][maxN]
    uint32_t* d_sortidx;     // sor: point order by cell
    int sor_grid_cap;        // max cells per frame
    int* h_counts;           // pinned [B]
    int* d_flags;            // misc flags [B][4]
    PipeSlot slot[2];
    bool pipe_ready;
    cudaStream_t copy_stream, d2h_stream;
    // pipelined batches (cvr_submit_batch), optional: everything after the path kernels of a batch ("tail": LR check .. pack) is
    // issued on tail_stream (high priority), so that the tail kernels of batch i run beside the matcher of batch i+1
    cudaStream_t tail_stream; cudaStream_t path_stream;   // path_stream: c->stream while a tail is being enqueued on tail_stream
    cudaEvent_t evPath, evAuxFree, evTailLast;
    bool split_tail;         // set by cvr_submit_batch around its enqueue
    bool aux_pending;        // evAuxFree is recorded: the next batch's path kernels wait for it before they overwrite d_aux
    bool tail_dirty;         // work is (or may be) outstanding on tail_stream: other entry points make c->stream wait for evTailLast
    int tail_mode;           // CVR_TAIL_STREAM: 1 = on, 0 = off (default; see cvr_submit_batch)
    long long n_submit, n_wait;
    int* d_meta0; int* h_meta0;   // as PipeSlot::d_meta / h_meta, for the synchronous cvr_process_batch
    // the packed payload of the batch completed last (cvr_process_batch / cvr_wait_batch): what cvr_gather_clouds sends
    bool lp_valid; const cvr_point* lp_pts; const int* lp_dhdr; int lp_nframes; int lp_total;
    // cloud gather (gather.cu): NCCL communicator bound at run time, its stream and staging
    void* comm; bool comm_owned; int comm_nranks, comm_rank;
    cudaStream_t comm_stream;
    int* d_hdr_all; int* h_hdr_all;           // [nranks][2 + maxB] gathered headers (device / pinned)
    cvr_point* d_gather; size_t gather_cap;   // root: receive buffer when the caller wants the clouds on the host
    uint8_t *png_blob, *png_z, *png_raw, *png_tab, *png_out;   // cvr_decode_png_batch: files, zlib streams, scanlines, tables, BGR staging (grown on demand)
    size_t png_blob_cap, png_z_cap, png_raw_cap, png_tab_cap, png_out_cap;
    cudaStream_t png_stream[4]; cudaEvent_t png_ev[5];   // large calls: four chunks of files, each uploaded and decoded on its own stream
    double* d_icp; double* h_icp;   // cvr_icp_align: block partials + totals of its 17 sums (device), the totals (pinned)
    int* d_sgflags;          // per-frame SGBM regime flags (bit 0: negative C seen, bit 1: negative path cost seen) [B]
    int smem_optin;          // max dynamic shared memory per block (opt-in)
    int smem_sm;             // shared memory per SM
    bool force_generic;      // CVR_PATH_KERNELS=generic: use the int32 path kernels (test hook)
    bool keep_S;             // CVR_KEEP_S=1: the last aggregation pass also writes the final S volume (cvr_debug_fetch(1); test hook)
    bool force_no_nn;        // CVR_NO_NN=1: never take the no-neighbour shortcut (test hook)
    int pathv_max_nt;        // CVR_PATHV_NT: upper bound on threads per CTA of the vertical path kernel (tuning hook)
    int pathv_active[3][17]; // concurrent clusters of the vertical path kernel per (CTA size, cluster size): 0 = not asked yet, -1 = cannot run
    int path_h_fwd, path_h_rev;   // horizontal path kernel per direction: 3 = several rows per warp, 2 = one row per warp (CVR_PATH_H_FWD / _REV)
    int pathv4_active[17];   // the same for k_path_v4 (512-thread CTAs) per cluster size
    int path_v_variant;      // 4 = several columns per warp (default), 3 = one column per warp (CVR_PATH_KERNELS=v3), 2 = per-row cluster barrier (=v2)
};

#define CVR_CUDA(ctx, call)                                                                       \
    do {                                                                                          \
        cudaError_t e__ = (call);                                                                 \
        if (e__ != cudaSuccess) {                                                                 \
            (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e__);                     \
            return CVR_ERR_CUDA;                                                                  \
        }                                                                                         \
    } while (0)

// after a kernel launch: report a failed launch (and only count the ones that were accepted)
#define LAUNCH_CHECK(c)                                                                  \
    do {                                                                                 \
        cudaError_t e__ = cudaGetLastError();                                            \
        if (e__ != cudaSuccess) { (c)->err = std::string("kernel launch: ") + cudaGetErrorString(e__) + " at " + __FILE__ + ":" + std::to_string(__LINE__); return CVR_ERR_CUDA; } \
        (c)->launches++;                                                                 \
    } while (0)

// ints in a packed payload's meta record (k_pack_clouds, api.cu)
#define CVR_META_INTS(maxB) (2 * (size_t)(maxB) + 4)

struct cvr_ctx;
static inline int cvr_div_up(int a, int b) { return (a + b - 1) / b; }
static inline size_t cvr_div_up_sz(size_t a, size_t b) { return (a + b - 1) / b; }

// ints per frame of an ordered selection's status record (cvr_scan.cuh): one word per 1024-element chunk, spare, ticket
static inline size_t cvr_sel_stride(int maxItems) { return ((size_t)maxItems + 1023) / 1024 + 2; }

// Blocks per frame (grid.x of a (x, B) grid) for a kernel that strides over a per-frame item count known only on the
// device: enough to cover `maxItems`, but no more than this frame's share of `blocksPerSM` resident blocks per SM --
// a batch of small clouds must not launch W*H-sized grids of idle blocks.
static inline int cvr_frame_blocks(const cvr_ctx* c, int B, size_t maxItems, int itemsPerBlock, int blocksPerSM)
{
    size_t need = (maxItems + (size_t)itemsPerBlock - 1) / (size_t)itemsPerBlock;
    size_t share = ((size_t)c->sm_count * (size_t)blocksPerSM) / (size_t)B;      // rounded down: the whole grid is resident at once
    if (need < 1) need = 1;
    if (share < 1) share = 1;
    return (int)(need < share ? need : share);
}

// ---- stage launchers (each returns cudaError_t of the launch; all on ctx->stream) ----
// k_sgbm.cu
int cvr_launch_gray(cvr_ctx* c, int B, const uint8_t* left, size_t lstep, size_t lfs, const uint8_t* right, size_t rstep,
                    size_t rfs, int channels);
int cvr_launch_sgbm(cvr_ctx* c, int B);     // gray -> d_med (final disparity)
int cvr_sgbm_suggest_batch(cvr_ctx* c, int around);   // batch size near `around` that fills the vertical kernel's cluster waves
// k_rectify.cu: initUndistortRectifyMap (CV_16SC2) for one camera into d_rmap1/2[cam]; remap of B stereo pairs
int cvr_launch_rectify_maps(cvr_ctx* c, int cam, const double* K, const double* D, int nD, const double* R, const double* P, int pcols);
int cvr_launch_remap(cvr_ctx* c, int B, int ch, const uint8_t* dl, size_t lstep, size_t lfs, const uint8_t* dr, size_t rstep, size_t rfs,
                     uint8_t* outL, uint8_t* outR, size_t ostep, size_t ofs);
int cvr_launch_bm(cvr_ctx* c, int B);       // k_bm.cu: gray -> d_med with StereoBM semantics
int cvr_launch_speckle(cvr_ctx* c, int B, int newVal, int maxSize, int maxDiff);   // in place on d_med
// k_cloud.cu
int cvr_launch_prune(cvr_ctx* c, int B, const int16_t* disp_in /*[B][H][W] tight*/);
int cvr_launch_chain_init(cvr_ctx* c, int B, bool stats, bool boxA, bool boxB);   // d_pstats / d_boxv / d_boxs to their neutral values
int cvr_launch_prune_stats(cvr_ctx* c, int B, const int16_t* disp_in);           // batch path: statistics only (the cloud kernel prunes on the fly)
int cvr_launch_cloud(cvr_ctx* c, int B, int mode, const int16_t* disp16, const float* dispf, const uint8_t* mask,
                     const uint8_t* bgr, size_t bstep, size_t bfs, int channels, bool fused);
// k_voxel.cu / k_sor.cu.  haveBox: the producer of `in` already accumulated its bounding box (d_boxv for the voxel grid,
// d_boxs for the outlier filters); otherwise the launcher measures it first.
int cvr_launch_voxel(cvr_ctx* c, int B, const cvr_point* in, size_t in_stride, const int* n_in, bool haveBox);
int cvr_launch_sor(cvr_ctx* c, int B, const cvr_point* in, size_t in_stride, const int* n_in /*device*/, bool haveBox);
int cvr_launch_ror(cvr_ctx* c, int B, const cvr_point* in, size_t in_stride, const int* n_in, float radius, int min_pts);
int cvr_launch_transform(cvr_ctx* c, const cvr_point* in, cvr_point* out, int n, const float* T);
// k_sor.cu: bounding box (unless haveBox), grid (radius > 0: cell = radius, else ~2 points per cell), points sorted by cell into d_spts
int cvr_build_point_grid(cvr_ctx* c, int B, const cvr_point* in, size_t in_stride, const int* n_in, bool haveBox, float radius, SorGrid** gridOut, int** cellstartOut);
void cvr_stage_mark(cvr_ctx* c, const char* name);
void cvr_timers_begin(cvr_ctx* c, int set);   // stage timers of a call: begin, marks, end (after the stream is idle) -> cvr_last_timings
void cvr_timers_end(cvr_ctx* c, int set);
int cvr_join_tail(cvr_ctx* c);         // entry points that work on c->stream alone: outstanding tail-stream work comes first
int cvr_tail_split(cvr_ctx* c);        // matcher launchers: the path kernels of this batch are enqueued, the tail follows
int cvr_tail_aux_consumed(cvr_ctx* c); // ... the tail no longer reads the WTA records (d_aux)
int cvr_path_wait_aux(cvr_ctx* c);     // ... about to enqueue kernels that overwrite d_aux
// message slot read by cvr_last_error(NULL): used by entry points that have no context
void cvr_set_thread_error(const char* msg);
// gather.cu
void cvr_comm_reset(cvr_ctx* c);      // fields to their empty state (context creation)
void cvr_comm_release(cvr_ctx* c);    // destroys an owned communicator and the gather staging

