/*
 * cvrecon_b200.h -- C ABI of libcvrecon_b200.so: the dense per-frame stereo hot path of
 * asad-ahmed-pk/master-thesis (cv_reconstruct) as hand-written sm_100a CUDA.
 *
 * The reference has no FFI/plugin interface; its boundary for this path is the C++ class
 * Reconstruct::Reconstruct3D (cv_reconstruct/include/reconstruct/Reconstruct3D.hpp:25-103),
 * TrackingFrame::SetupFrame (cv_reconstruct/src/system/TrackingFrame.cpp:25-67) and the PCL
 * filter call sites.  Each entry point below names the reference interface it replaces.  A
 * header-only C++ mirror of the reference class on top of this ABI is in
 * master-thesis_b200/host/Reconstruct3D_b200.hpp; the binding a maintainer of the reference
 * would add is shown in INTEGRATION.md.
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success or a
 * negative cvr_status; no exception crosses the ABI; cvr_last_error() returns a message for the
 * last failure on this context.  A context is NOT re-entrant (like the reference's matcher,
 * whose compute() mutates internal buffers): one caller thread per context.  There is no CPU
 * fallback: cvr_create fails when no sm_100-class device is usable.
 *
 * Image layout = cv::Mat: `data`, rows, cols, `step` bytes per row.  Disparity = CV_16SC1,
 * value 16*d, invalid = (minDisparity-1)*16.  Points are 16-byte cvr_point (x,y,z f32 + b,g,r,a);
 * cvr_points_to_pcl() expands to the 32-byte pcl::PointXYZRGB layout.
 */
#ifndef CVRECON_B200_H
#define CVRECON_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cvr_ctx cvr_ctx;

typedef enum cvr_status {
    CVR_OK = 0,
    CVR_ERR_INVALID_ARG = -1,
    CVR_ERR_UNSUPPORTED = -2,     /* parameter combination outside what the kernels implement */
    CVR_ERR_CUDA = -3,
    CVR_ERR_NO_DEVICE = -4,
    CVR_ERR_CAPACITY = -5         /* caller's output buffer too small / batch larger than max_batch */
} cvr_status;

enum { CVR_MODE_SGBM = 0, CVR_MODE_HH = 1 };
enum { CVR_MEM_HOST = 0, CVR_MEM_DEVICE = 1 };
enum { CVR_CLOUD_TRIANGULATE = 0, CVR_CLOUD_REPROJECT = 1 };
enum { CVR_MATCHER_SGBM = 0, CVR_MATCHER_BM = 1 };

/* cv::StereoSGBM parameters exactly as Reconstruct3D::ConfigureSteoreoMatcher sets them
 * (cv_reconstruct/src/reconstruct/Reconstruct3D.cpp:50-64; JSON keys in
 * cv_reconstruct/resources/config/config_default.json:13-21).  The reference computes
 * P1 = 8*3*bs^2 and P2 = 32*3*bs^2 and leaves disp12MaxDiff = 0 and mode = MODE_SGBM. */
typedef struct cvr_sgbm_params {
    int min_disparity;
    int num_disparities;        /* multiple of 16, <= 256 */
    int block_size;             /* odd, >= 1 */
    int P1;
    int P2;
    int disp12_max_diff;        /* 0 => OpenCV uses 1 */
    int pre_filter_cap;         /* <= 63 */
    int uniqueness_ratio;
    int speckle_window_size;    /* 0 => speckle filter off */
    int speckle_range;
    int mode;                   /* CVR_MODE_SGBM | CVR_MODE_HH */
} cvr_sgbm_params;

typedef struct cvr_params {
    int width, height;          /* frame size; fixed per context */
    cvr_sgbm_params sgbm;
    float Q[16];                /* row-major 4x4 for GeneratePointCloud (Reconstruct3D.cpp:31-35) */
    float tri_f, tri_b, tri_cx, tri_cy;   /* Triangulate3D constants (Reconstruct3D.cpp:167-172) */
    double prune_sigma_factor;  /* 1.3 (a double literal) in TrackingFrame.cpp:55 */
    float voxel_leaf[3];        /* MappingSystem.cpp:299: 0.2 */
    int   sor_mean_k;           /* config_default.json:24 outlier_min_k = 30 */
    double sor_stddev_mul;      /* config_default.json:25 outlier_std_threshold = 0.5 */
    int   max_batch;            /* frames in flight per cvr_process_batch call (workspace is sized for it) */
    int   cloud_mode;           /* CVR_CLOUD_TRIANGULATE (live path) | CVR_CLOUD_REPROJECT */
    int   max_points;           /* point capacity per frame for the cloud filters; 0 => width*height */
    int   filter_only;          /* 1 => no stereo workspace: only cvr_voxel_grid / cvr_sor are usable */
    /* "block_matcher" switch of config.json (Reconstruct3D.cpp:41-68): CVR_MATCHER_SGBM | CVR_MATCHER_BM.  The BM branch is
     * cv::StereoBM::create(16, 21) + setBlockSize(SBM.window_size) + setNumDisparities(SBM.num_disparities)
     * (Reconstruct3D.cpp:43-48, :378-380); the remaining fields are OpenCV's StereoBM defaults. */
    int   block_matcher;
    int   bm_window_size;       /* odd, 5..255 (config SBM.window_size = 27) */
    int   bm_num_disparities;   /* multiple of 16, <= 256 (config SBM.num_disparities = 128) */
    int   bm_min_disparity;     /* 0 (values > 0 unsupported) */
    int   bm_pre_filter_cap;    /* 31 */
    int   bm_texture_threshold; /* 10 */
    int   bm_uniqueness_ratio;  /* 15 */
    int   bm_speckle_window_size; /* 0 = off */
    int   bm_speckle_range;     /* 0 */
    /* config.json "requires_rectification" (ConfigParser.cpp:39; ReconstructionSystem.cpp:42-44): 1 => every batch is
     * rectified with the maps of cvr_set_rectification before the matcher, and the rectified left image colours the cloud. */
    int   requires_rectification;
} cvr_params;

typedef struct cvr_point {      /* 16 bytes */
    float x, y, z;
    uint8_t b, g, r, a;
} cvr_point;

/* Fills *p with the reference's defaults: config_default.json SGBM section + P1/P2 rule,
 * Q = diag(1,1,0.005*f,1), KITTI calib (resources/calib/kitti/calib_kitti_rectified.json),
 * leaf 0.2, SOR k=30 alpha=0.5, prune factor 1.3, max_batch 1. */
void cvr_default_params(cvr_params* p, int width, int height);

/* Replaces: Reconstruct3D::Reconstruct3D(stereoSetup, config) (Reconstruct3D.cpp:26-36) --
 * allocates the device workspace once. */
int cvr_create(const cvr_params* p, int device, cvr_ctx** out);
int cvr_destroy(cvr_ctx* ctx);
const char* cvr_last_error(const cvr_ctx* ctx);
const char* cvr_version(void);

/* Replaces: Reconstruct3D::GenerateDisparityMap (Reconstruct3D.cpp:72-84):
 * cvtColor(BGR2GRAY) x2 + StereoSGBM::compute.  channels = 3 or 4 (BGR/BGRA, converted like
 * cv::cvtColor) or 1 (already gray).  Host pointers. */
int cvr_disparity(cvr_ctx* ctx, const uint8_t* left, size_t step_left, const uint8_t* right, size_t step_right,
                  int channels, int16_t* disp, size_t step_disp);

/* Replaces: the two cv::initUndistortRectifyMap(K, D, R, P, size, CV_16SC2, map1, map2) calls of
 * Reconstruct3D::RectifyImages (Reconstruct3D.cpp:351-356), done once per calibration instead of once per frame.
 * K: 3x3 row-major camera matrix; D: n_dist (0, 4, 5, 8 or 12) distortion coefficients k1 k2 p1 p2 k3 k4 k5 k6 s1 s2 s3 s4
 * (the tilt terms of 14-coefficient vectors are not supported); R: 3x3 rectification rotation; P: 3 x p_cols (3 or 4)
 * row-major projection matrix.  Camera 1 = left, 2 = right.  Builds the four maps on the device. */
int cvr_set_rectification(cvr_ctx* ctx, const double* K1, const double* D1, int n_dist1, const double* R1, const double* P1, int p1_cols,
                          const double* K2, const double* D2, int n_dist2, const double* R2, const double* P2, int p2_cols);

/* Replaces: Reconstruct3D::RectifyImages (Reconstruct3D.cpp:330-361): cv::remap(image, rectified, map1, map2,
 * INTER_LINEAR) for both images of one stereo pair (constant border 0).  Host pointers; channels = 1, 3 or 4. */
int cvr_rectify(cvr_ctx* ctx, int channels, const uint8_t* left, size_t step_left, const uint8_t* right, size_t step_right,
                uint8_t* rect_left, size_t step_rect_left, uint8_t* rect_right, size_t step_rect_right);

/* Test hook: copies the maps of one camera (0 = left, 1 = right) to the host: map1 = H*W*2 int16, map2 = H*W uint16. */
int cvr_get_rectification_maps(cvr_ctx* ctx, int camera, int16_t* map1, uint16_t* map2);

/* SURVEY 8(f) row n1 -- the step before the path.  Replaces: the two cv::imdecode(message.LeftImageData / RightImageData,
 * cv::IMREAD_COLOR) calls of MessageConverter::ConvertStereoMessage (cv_reconstruct/src/server/MessageConverter.cpp:65-66) for a
 * batch of n PNG files (what the streamers produce with cv::imencode(".png", ...)): zlib inflate and the PNG scanline filters on
 * the device, output = 8UC3 frames in B, G, R order exactly as cv::imdecode returns them (image i at bgr + i * image_stride, rows
 * `step` bytes apart), in host or device memory according to mem_out -- device output can be handed to cvr_submit_batch as a
 * CVR_MEM_DEVICE input without ever touching the host.  png[i] / png_bytes[i] are host pointers / sizes of the encoded files.
 * Every image must have the context's width x height.  Decoded on the device: 8-bit, non-interlaced, colour types gray, RGB and
 * RGBA (alpha dropped); other PNG flavours return CVR_ERR_UNSUPPORTED, a damaged stream CVR_ERR_INVALID_ARG (the Adler-32 of
 * every zlib stream is verified; the CRCs of the PNG chunks are not).
 * A DEFLATE stream is decoded by one thread, so the rate is set by the number of files per call (a B200 keeps 4736 streams in
 * flight): hand over several batches of frames at once where latency allows, and keep the files in pinned host memory. */
int cvr_decode_png_batch(cvr_ctx* ctx, int n, const uint8_t* const* png, const size_t* png_bytes, int mem_out, uint8_t* bgr, size_t step,
                         size_t image_stride);

/* Replaces: TrackingFrame::SetupFrame + PruneDisparityImage (TrackingFrame.cpp:25-67).
 * disp is pruned in place; disp_f32 (CV_32F, tight W*4 step) and mask (CV_8U, tight) are outputs
 * and may be NULL. */
int cvr_prune(cvr_ctx* ctx, int16_t* disp_inout, size_t step_disp, float* disp_f32, uint8_t* mask);

/* Replaces: Reconstruct3D::GeneratePointCloud (Reconstruct3D.cpp:106-153):
 * reprojectImageTo3D(disp, Q, true), skip Z == 10000, recentre, flip, BGR->RGB, x negated.
 * Output order = row-major over surviving pixels.  *n_out receives the count. */
int cvr_reproject_cloud(cvr_ctx* ctx, const int16_t* disp, size_t step_disp, const uint8_t* bgr, size_t step_bgr,
                        cvr_point* out, int capacity, int* n_out);

/* Replaces: Reconstruct3D::Triangulate3D (Reconstruct3D.cpp:156-216); mask may be NULL. */
int cvr_triangulate(cvr_ctx* ctx, const float* disp_f32, size_t step_disp, const uint8_t* bgr, size_t step_bgr,
                    const uint8_t* mask, size_t step_mask, cvr_point* out, int capacity, int* n_out);

/* Replaces: pcl::VoxelGrid<PointXYZRGB>::filter at MappingSystem.cpp:297-300 (leaf from cvr_params).
 * Output in ascending voxel index.  *passthrough = 1 when PCL would warn "leaf size too small"
 * (int32 index overflow) and copy the input through. */
int cvr_voxel_grid(cvr_ctx* ctx, const cvr_point* in, int n, cvr_point* out, int capacity, int* m_out, int* passthrough);

/* Replaces: PointCloudPostProcessor::RemoveOutliers (PointCloudPostProcessor.cpp:38-41) =
 * pcl::StatisticalOutlierRemoval with meanK / stddevMulThresh from cvr_params.  in == out allowed. */
int cvr_sor(cvr_ctx* ctx, const cvr_point* in, int n, cvr_point* out, int capacity, int* kept_out);

/* The two filters above on one cloud that may already live in DEVICE memory (mem = CVR_MEM_HOST | CVR_MEM_DEVICE for `in` and
 * `out` alike): stages = 4 (VoxelGrid), 8 (StatisticalOutlierRemoval) or 12 (VoxelGrid then SOR, the order of MappingSystem.cpp:297-300
 * followed by PointCloudPostProcessor::RemoveOutliers).  Lets a cloud the path produced be filtered again (e.g. a merged map
 * block) without a round trip through the host. */
int cvr_filter_cloud(cvr_ctx* ctx, int mem, const cvr_point* in, int n, int stages, cvr_point* out, int capacity, int* n_out);

/* SURVEY 8(f) row n3 -- the consumer-side steps MappingSystem applies to the cloud right after the path.
 * Replaces: pcl::RadiusOutlierRemoval::filter at MappingSystem.cpp:303-307 (setRadiusSearch(1.0), setMinNeighborsInRadius(300)):
 * a point is kept iff at least min_neighbors + 1 points (itself included) lie within `radius`.  Input order preserved. */
int cvr_radius_outlier_removal(cvr_ctx* ctx, const cvr_point* in, int n, float radius, int min_neighbors, cvr_point* out, int capacity,
                               int* kept_out);
/* Replaces: pcl::transformPointCloud(cloud, cloud, pose) at MappingSystem.cpp:318 (T: row-major 4x4, f32; colours untouched). */
int cvr_transform_cloud(cvr_ctx* ctx, const cvr_point* in, int n, const float* T, cvr_point* out);

/* SURVEY 8(f) row n4 -- the alignment MappingSystem sets up between a key frame's cloud and the previous map block.
 * Replaces: pcl::IterativeClosestPoint<PointXYZRGB, PointXYZRGB>::align at MappingSystem.cpp:321-341 (the same settings at
 * :104-127; PointCloudPostProcessor.cpp:26-31 configures a second instance): point-to-point ICP with PCL's defaults -- nearest
 * neighbour correspondences within max_correspondence_distance, transformation by SVD, the three stop rules of PCL's default
 * convergence criteria (iterations, transformation epsilon, absolute / relative change of the mean squared correspondence
 * distance).  cvr_icp_default_params fills in the reference's numbers (5, 50, 1e-8, 1).  source / target / aligned are host or
 * device arrays according to `mem`; aligned (optional) receives the source cloud moved by T; T is row-major, applied to column
 * vectors like Eigen::Matrix4f (p' = T * p), and maps the source onto the target (getFinalTransformation()).
 * state: 0 not converged (fewer than three correspondences), 1 iterations, 2 transformation epsilon, 3 absolute, 4 relative MSE. */
typedef struct cvr_icp_params {
    double max_correspondence_distance;
    int max_iterations;
    double transformation_epsilon;
    double euclidean_fitness_epsilon;
} cvr_icp_params;
typedef struct cvr_icp_result {
    float T[16];
    int iterations, converged, state, correspondences;   /* correspondences of the last iteration */
    double fitness;                                       /* getFitnessScore(): mean squared distance to the nearest target point */
} cvr_icp_result;
void cvr_icp_default_params(cvr_icp_params* p);
int cvr_icp_align(cvr_ctx* ctx, int mem, const cvr_point* source, int n_source, const cvr_point* target, int n_target, const cvr_icp_params* params,
                  cvr_point* aligned, cvr_icp_result* result);

/* Frame-parallel fused path: for each of nframes frames
 *   GenerateDisparityMap -> SetupFrame(prune) -> Triangulate3D|GeneratePointCloud -> VoxelGrid -> SOR
 * (the per-frame chain of ReconstructionSystem::ProcessStereoFrame, ReconstructionSystem.cpp:35-61,
 * plus the filter call sites).  Frame i of a buffer is at base + i*frame_stride bytes.  mem selects
 * host or device pointers for ALL buffers of the call.  Optional outputs may be NULL. */
typedef struct cvr_batch {
    int mem;                        /* CVR_MEM_HOST | CVR_MEM_DEVICE */
    int channels;                   /* 1, 3 or 4 */
    const uint8_t* left;  size_t left_step,  left_frame_stride;
    const uint8_t* right; size_t right_step, right_frame_stride;
    int16_t* disp;        size_t disp_step,  disp_frame_stride;     /* optional: final disparity (pre-prune) */
    cvr_point* cloud;     size_t cloud_capacity;                    /* per frame, in points; frame i at cloud + i*capacity */
    int32_t* cloud_count;                                           /* [nframes] */
    int stages;                     /* bit 0 disparity, 1 prune+cloud, 2 voxel, 3 sor; 0 => all */
    /* Packed form of the cloud output; when cloud_packed is set it is used INSTEAD of `cloud`: the clouds of all frames back
     * to back, frame i = cloud_packed[cloud_offsets[i] .. cloud_offsets[i+1]) -- the layout MapDataBase::InsertBlock
     * (cv_reconstruct/src/system/MapDataBase.cpp:28-54) appends block by block.  One device->host copy per batch instead of
     * one per frame.  cloud_offsets has nframes + 1 entries; cloud_count is still filled when given. */
    cvr_point* cloud_packed; size_t cloud_packed_capacity;          /* in points, for the whole batch */
    int32_t* cloud_offsets;
} cvr_batch;
int cvr_process_batch(cvr_ctx* ctx, int nframes, const cvr_batch* io);

/* No counterpart in the reference (which processes one frame per call).  The vertical path kernel gives every frame of a
 * batch to one thread-block cluster and the GPU holds a fixed number of those at once (found by an occupancy query: e.g. 26
 * clusters of 5 CTAs on a 148-SM B200 at KITTI shape), so a batch that is a multiple of that number has no part-filled
 * last wave.  Returns the batch size within a quarter of `around` (and <= max_batch) with the lowest cost per frame, or a
 * negative cvr_status. */
int cvr_suggest_batch(cvr_ctx* ctx, int around);

/* Pipelined form of cvr_process_batch: up to two batches in flight, so that the upload of batch i+1 overlaps the kernels
 * of batch i and the download of batch i-1.  This is how the frame loop
 * of the reference's server thread (cv_reconstruct/src/server/ReconstructionServer.cpp:134-153: dequeue a stereo message,
 * ProcessStereoFrame, repeat) is meant to drive the library at full rate: submit the frames just dequeued, then collect
 * the previous batch.
 *   cvr_submit_batch  enqueues upload + the whole chain and returns at once; the io struct is copied, the buffers it
 *                     points to (inputs AND outputs, host or device) must stay valid and untouched until the matching
 *                     cvr_wait_batch.  Host batches are staged through the slot; device batches are read and written in place.
 *   cvr_wait_batch    blocks until the OLDEST submitted batch is done and its outputs are in the caller's buffers
 * At most two submits may be outstanding (a third returns CVR_ERR_CAPACITY).  The first submit allocates the two
 * pipeline slots (inputs + parked outputs).  Mixing with the synchronous entry points is allowed once every
 * submitted batch has been waited for. */
int cvr_submit_batch(cvr_ctx* ctx, int nframes, const cvr_batch* io);
int cvr_wait_batch(cvr_ctx* ctx);
/* Error behaviour of the batched calls: a cloud that does not fit cloud_capacity / cloud_packed_capacity makes the call (for
 * submitted batches: the matching cvr_wait_batch) return CVR_ERR_CAPACITY -- for host AND device buffers; device buffers then
 * hold the points that fitted and the true counts.  A failed cvr_wait_batch still releases its pipeline slot. */

/* ---- multi-GPU: gather of the filtered clouds to the mapping rank ------------------------------------------
 * Replaces: the hand-over of every processed frame's cloud to the one map, MappingSystem::AddFrame -> MapDataBase::InsertBlock
 * (cv_reconstruct/src/system/MappingSystem.cpp:344-345, MapDataBase.cpp:28-54), for frames sharded over one process per GPU.
 * Frames are independent, so this is the only exchange on the path: per-frame counts by ncclAllGather, the packed clouds by
 * grouped ncclSend / ncclRecv straight from the device buffer the filters wrote, on a stream of its own (it overlaps the
 * kernels of the next submitted batch).  libnccl.so.2 is bound at run time (dlopen; the copy already loaded into the process
 * is preferred), nothing links against it.
 *
 *   cvr_comm_get_unique_id  rank 0 makes an id (CVR_COMM_ID_BYTES bytes) and hands it to the other processes by any means
 *   cvr_comm_init           every rank: ncclCommInitRank on the context's device
 *   cvr_comm_attach         alternative: use an ncclComm_t the host already owns (not destroyed by the library)
 *   cvr_gather_clouds       collective over all ranks: sends the clouds of the batch this rank completed last
 *                           (cvr_process_batch, or the most recent cvr_wait_batch) to `root`.  On root, `out` (host or device
 *                           memory according to `mem`) receives the points rank by rank, frames in batch order;
 *                           frame_offsets[0 .. total_frames] are the frame boundaries in `out`, frames_of_rank[r] the number
 *                           of frames rank r contributed.  Other ranks pass NULL / 0 for the outputs.  Blocks until this
 *                           rank's part is done.  Call it before submitting the batch after next (which reuses the buffer
 *                           the payload is parked in).  Every rank must have been created with the same max_batch.
 *                           Without a communicator the call is the one-rank case (a local copy).
 *   cvr_gather_layout       the host arithmetic of the gather, exported for tests: headers = nranks records of header_ints
 *                           int32 {nframes, total points, per-frame counts...}; writes rank_offset[0..nranks] (points) and,
 *                           if wanted, the global frame_offsets; returns the total number of frames or a negative status. */
#define CVR_COMM_ID_BYTES 128
int cvr_comm_get_unique_id(void* id);
int cvr_comm_init(cvr_ctx* ctx, int nranks, int rank, const void* id);
int cvr_comm_attach(cvr_ctx* ctx, void* nccl_comm);
int cvr_comm_destroy(cvr_ctx* ctx);
int cvr_gather_clouds(cvr_ctx* ctx, int root, int mem, cvr_point* out, size_t out_capacity, int32_t* frame_offsets, int frame_offsets_capacity,
                      int32_t* frames_of_rank);
int cvr_gather_layout(const int32_t* headers, int nranks, int header_ints, int64_t* rank_offset, int32_t* frame_offsets, int frame_offsets_capacity,
                      int32_t* frames_of_rank);

/* No counterpart in the reference: TrackingFrame::SetupFrame / cvr_prune for callers that are not allowed to
 * allocate GPU memory.  Think of a frame loop that must keep a fixed latency (cudaMalloc/cudaFree stall the whole
 * device while they run), a plug-in host or safety-minded integrator that forbids device allocation once start-up
 * is over, or a process whose memory pool belongs to another library.  Such a caller cannot create a cvr_ctx
 * (cvr_create allocates the workspace) -- this call needs none.  It takes DEVICE pointers to tight width*height
 * arrays: d_disp in, and any of d_pruned (int16), d_disp_f32 (float), d_mask (uint8) out (NULL = not wanted), and
 * enqueues the work on `stream` (a cudaStream_t, NULL = default stream) of the CURRENT device without waiting for
 * it.  It calls no cudaMalloc/cudaFree nor any allocating library routine: the little scratch it needs (a 256-byte
 * statistics record and one byte per pixel) sits in a fixed array built into the library, which every GPU that
 * loads the library keeps until the process exits (4 MiB + 256 B).
 * Limit: width*height <= CVR_NOALLOC_MAX_PIXELS, larger frames are refused with CVR_ERR_CAPACITY and a message in
 * cvr_last_error(NULL).  The first call on each device looks up the array's device address; make that first call
 * before starting any stream capture.  All calls on a device share the one array and it keeps whatever the last
 * call left there (the call resets what it relies on), so two calls whose GPU work can overlap -- different
 * streams or threads -- race with each other: keep them ordered.  d_pruned may be the same buffer as d_disp (pruning in
 * place, as TrackingFrame::PruneDisparityImage does).  The looked-up address does not survive cudaDeviceReset. */
#define CVR_NOALLOC_MAX_PIXELS (4u << 20)
int cvr_prune_noalloc(const int16_t* d_disp, int width, int height, double sigma_factor,
                      int16_t* d_pruned, float* d_disp_f32, uint8_t* d_mask, void* stream);

/* Expands 16-byte points to pcl::PointXYZRGB (32 B: x,y,z,1.0f, b,g,r,a, 12 B pad).  Host side. */
void cvr_points_to_pcl(const cvr_point* in, int n, void* pcl_points_32B);

/* Replaces: pcl::io::savePCDFileBinary(path, cloud) as the reference dumps its clouds (ReconstructionServer.cpp:165,
 * MappingSystem.cpp:61, 84; test/reconstruct_test_kitti_rectified.cpp:77): a PCD v0.7 file, DATA binary, FIELDS x y z rgb
 * (4 x 4 bytes per point, rgb = the bytes b, g, r, a read as one 32-bit word -- PCL's packed colour), unorganised
 * (WIDTH n, HEIGHT 1).  Host memory in, host side only; pcl::io::loadPCDFile / pcl_viewer / CloudCompare read it. */
int cvr_save_pcd(const char* path, const cvr_point* points, int n);

/* Test/inspection hooks (no reference counterpart): copy an intermediate of the most recent
 * cvr_disparity call to host.  which: 0 = C volume, 1 = S volume (both int16 [H][W1][Dp]),
 * 2 = raw disparity before median/speckle (int16 [H][W]), 3 = after the LR check.  *dp receives the padded D pitch.
 * Of the most recent cvr_sor call: 4 = mean k-NN distance per input point (bytes/4 floats, input order),
 * 5 = {double sum, sum of squares, threshold; int valid, kept} (32 bytes). */
int cvr_debug_fetch(cvr_ctx* ctx, int which, void* host_dst, size_t bytes, int* w1, int* dp);

/* Number of kernel launches issued by this context since creation (bench.py's gpu_launches). */
long long cvr_launch_count(const cvr_ctx* ctx);
/* Device time (ms, CUDA events on the context's stream) of each stage of the last call, and the
 * launch count of that call; names[i] are static strings.  Returns the number of stages written. */
int cvr_last_timings(const cvr_ctx* ctx, const char** names, float* ms, int max_stages);
/* Enable/disable per-stage event timing (adds event records between stages). */
int cvr_set_profiling(cvr_ctx* ctx, int enable);

#ifdef __cplusplus
}
#endif
#endif /* CVRECON_B200_H */
