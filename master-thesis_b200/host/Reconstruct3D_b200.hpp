// Reconstruct3D_b200.hpp -- header-only C++14 mirror of the reference's reconstruction-stage interface on top of the
// C ABI (include/cvrecon_b200.h).  Same method names, argument meaning and error behaviour as
//   Reconstruct::Reconstruct3D            cv_reconstruct/include/reconstruct/Reconstruct3D.hpp:25-103
//   System::TrackingFrame::SetupFrame     cv_reconstruct/src/system/TrackingFrame.cpp:25-44
//   PointCloudPostProcessor::RemoveOutliers  cv_reconstruct/include/point_cloud/PointCloudPostProcessor.hpp:38
// OpenCV / PCL / Eigen headers are not available in this image, so images are passed as cvr::Mat views (the fields
// of cv::Mat that matter: data, rows, cols, step) and clouds as std::vector<cvr_point>; INTEGRATION.md shows the
// two-line adapters for cv::Mat and pcl::PointCloud<pcl::PointXYZRGB>.
// Errors: the reference lets cv::Exception propagate; this mirror throws std::runtime_error with cvr_last_error().
#pragma once
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/cvrecon_b200.h"

namespace cvr {

struct Mat {            // non-owning view with cv::Mat's layout
    void* data;
    int rows, cols;
    std::size_t step;   // bytes per row
    int channels;
};

struct StereoCalib {    // the fields of Camera::Calib::StereoCalib the path reads (CameraCalib.hpp:44-65)
    float fxL, fyL, cxL, cyL;   // LeftCameraCalib.K(0,0), K(1,1), K(0,2), K(1,2)
    float fxR, fyR, cxR, cyR;   // RightCameraCalib
    float T0;                   // T(0): baseline
};

struct RectificationCalib {   // what RectifyImages reads from m_StereoCameraSetup (Reconstruct3D.cpp:333-349), row-major doubles
    double K1[9], K2[9];       // LeftCameraCalib.K / RightCameraCalib.K
    std::vector<double> D1, D2;   // distortion coefficients (0, 4, 5, 8 or 12)
    double R1[9], R2[9];       // Rectification.RL / RR
    double P1[12], P2[12];     // Rectification.PL / PR (3x4)
};

enum StereoBlockMatcherType { STEREO_BLOCK_MATCHER = 0, STEREO_SEMI_GLOBAL_BLOCK_MATCHER = 1 };   // Reconstruct3DTypes.hpp:12-15

struct SGBMConfig {     // Config::Reconstruction (Config.hpp:25-45): BlockMatcherType, SBM, SGBM
    int NumDisparities = 16, BlockSize = 128, UniquenessRatio = 2, SpeckleRange = 2, SpeckleWindowSize = 100, PreFilterCap = 10,
        MinDisparity = 0;
    StereoBlockMatcherType BlockMatcherType = STEREO_SEMI_GLOBAL_BLOCK_MATCHER;
    int SBMNumDisparities = 16, SBMWindowSize = 21;                          // Config::Reconstruction::SBM (Config.hpp:31-34)
    int OutlierMinK = 30;            // point_cloud_post_processing.outlier_min_k (config_default.json:24)
    double OutlierStdDevThreshold = 0.5;
};

class Reconstruct3D {
public:
    // Reconstruct3D::Reconstruct3D(stereoSetup, config) (Reconstruct3D.cpp:26-36) + ConfigureSteoreoMatcher (:39-69)
    Reconstruct3D(const StereoCalib& calib, const SGBMConfig& cfg, int width, int height, int device = 0, int maxBatch = 1,
                  int mode = CVR_MODE_SGBM)
        : m_calib(calib)
    {
        cvr_params p;
        cvr_default_params(&p, width, height);
        p.sgbm.min_disparity = cfg.MinDisparity;
        p.sgbm.pre_filter_cap = cfg.PreFilterCap;
        p.sgbm.block_size = cfg.BlockSize;
        p.sgbm.num_disparities = cfg.NumDisparities;
        p.sgbm.speckle_range = cfg.SpeckleRange;
        p.sgbm.speckle_window_size = cfg.SpeckleWindowSize;
        p.sgbm.uniqueness_ratio = cfg.UniquenessRatio;
        p.sgbm.P1 = 8 * 3 * cfg.BlockSize * cfg.BlockSize;      // Reconstruct3D.cpp:63
        p.sgbm.P2 = 32 * 3 * cfg.BlockSize * cfg.BlockSize;     // Reconstruct3D.cpp:64
        p.sgbm.mode = mode;
        // Reconstruct3D.cpp:41-48: the STEREO_BLOCK_MATCHER branch = StereoBM::create(16, 21) + block size + disparities
        p.block_matcher = cfg.BlockMatcherType == STEREO_BLOCK_MATCHER ? CVR_MATCHER_BM : CVR_MATCHER_SGBM;
        p.bm_window_size = cfg.SBMWindowSize; p.bm_num_disparities = cfg.SBMNumDisparities;
        for (int i = 0; i < 16; i++) p.Q[i] = (i % 5 == 0) ? 1.0f : 0.0f;
        const float f = (calib.fxL + calib.fxR) / 2.0f;         // Reconstruct3D.cpp:33
        p.Q[10] = static_cast<float>(0.005 * f);                // Reconstruct3D.cpp:34
        p.tri_f = (calib.fxL + calib.fyL) / 2.0f;               // Reconstruct3D.cpp:168
        p.tri_b = calib.T0; p.tri_cx = calib.cxL; p.tri_cy = calib.cyL;
        p.sor_mean_k = cfg.OutlierMinK; p.sor_stddev_mul = cfg.OutlierStdDevThreshold;
        p.max_batch = maxBatch;
        m_width = width; m_height = height; m_params = p; m_device = device;
        if (cvr_create(&p, device, &m_ctx) != CVR_OK) throw std::runtime_error(std::string("Reconstruct3D: ") + cvr_last_error(nullptr));
    }
    ~Reconstruct3D() { cvr_destroy(m_ctx); }
    Reconstruct3D(const Reconstruct3D&) = delete;
    Reconstruct3D& operator=(const Reconstruct3D&) = delete;

    // The calibration half of RectifyImages: the reference rebuilds both maps on every call (Reconstruct3D.cpp:351-356); here once.
    void SetRectification(const RectificationCalib& r)
    {
        check(cvr_set_rectification(m_ctx, r.K1, r.D1.data(), static_cast<int>(r.D1.size()), r.R1, r.P1, 4, r.K2, r.D2.data(),
                                    static_cast<int>(r.D2.size()), r.R2, r.P2, 4));
    }
    // void RectifyImages(const cv::Mat& leftImage, const cv::Mat& rightImage, cv::Mat& rectLeftImage, cv::Mat& rectRightImage) const
    void RectifyImages(const Mat& leftImage, const Mat& rightImage, Mat& rectLeftImage, Mat& rectRightImage) const
    {
        check(cvr_rectify(m_ctx, leftImage.channels, static_cast<const uint8_t*>(leftImage.data), leftImage.step,
                          static_cast<const uint8_t*>(rightImage.data), rightImage.step, static_cast<uint8_t*>(rectLeftImage.data),
                          rectLeftImage.step, static_cast<uint8_t*>(rectRightImage.data), rectRightImage.step));
    }
    // cv::Mat GenerateDisparityMap(const cv::Mat& leftImage, const cv::Mat& rightImage) const  -> CV_16SC1 rows x cols
    std::vector<int16_t> GenerateDisparityMap(const Mat& leftImage, const Mat& rightImage) const
    {
        std::vector<int16_t> disp(static_cast<std::size_t>(m_width) * m_height);
        check(cvr_disparity(m_ctx, static_cast<const uint8_t*>(leftImage.data), leftImage.step, static_cast<const uint8_t*>(rightImage.data),
                            rightImage.step, leftImage.channels, disp.data(), static_cast<std::size_t>(m_width) * 2));
        return disp;
    }
    // TrackingFrame::SetupFrame: prunes `disparity` in place, fills the CV_32F disparity and the CV_8U mask
    void SetupFrame(std::vector<int16_t>& disparity, std::vector<float>& disparityF32, std::vector<uint8_t>& mask) const
    {
        disparityF32.resize(disparity.size()); mask.resize(disparity.size());
        check(cvr_prune(m_ctx, disparity.data(), static_cast<std::size_t>(m_width) * 2, disparityF32.data(), mask.data()));
    }
    // pcl::PointCloud<pcl::PointXYZRGB> GeneratePointCloud(const cv::Mat& disparity, const cv::Mat& cameraImage) const
    std::vector<cvr_point> GeneratePointCloud(const std::vector<int16_t>& disparity, const Mat& cameraImage) const
    {
        std::vector<cvr_point> out(static_cast<std::size_t>(m_width) * m_height);
        int n = 0;
        check(cvr_reproject_cloud(m_ctx, disparity.data(), static_cast<std::size_t>(m_width) * 2, static_cast<const uint8_t*>(cameraImage.data),
                                  cameraImage.step, out.data(), static_cast<int>(out.size()), &n));
        out.resize(n);
        return out;
    }
    // ... Triangulate3D(const cv::Mat& disparity, const cv::Mat& cameraImage, cv::InputArray& mask = cv::noArray()) const
    std::vector<cvr_point> Triangulate3D(const std::vector<float>& disparity, const Mat& cameraImage, const std::vector<uint8_t>* mask = nullptr) const
    {
        std::vector<cvr_point> out(static_cast<std::size_t>(m_width) * m_height);
        int n = 0;
        check(cvr_triangulate(m_ctx, disparity.data(), static_cast<std::size_t>(m_width) * 4, static_cast<const uint8_t*>(cameraImage.data), cameraImage.step,
                              mask ? mask->data() : nullptr, static_cast<std::size_t>(m_width), out.data(), static_cast<int>(out.size()), &n));
        out.resize(n);
        return out;
    }
    // pcl::VoxelGrid<pcl::PointXYZRGB>::filter with setLeafSize(0.2, 0.2, 0.2) (MappingSystem.cpp:297-300)
    std::vector<cvr_point> VoxelGridFilter(const std::vector<cvr_point>& cloud) const
    {
        std::vector<cvr_point> out(cloud.size() ? cloud.size() : 1);
        int m = 0, pass = 0;
        check(cvr_voxel_grid(m_ctx, cloud.data(), static_cast<int>(cloud.size()), out.data(), static_cast<int>(out.size()), &m, &pass));
        out.resize(m);
        return out;
    }
    // void PointCloudPostProcessor::RemoveOutliers(input, output) -- called with input == output in the reference
    void RemoveOutliers(const std::vector<cvr_point>& input, std::vector<cvr_point>& output) const
    {
        std::vector<cvr_point> out(input.size() ? input.size() : 1);
        int k = 0;
        check(cvr_sor(m_ctx, input.data(), static_cast<int>(input.size()), out.data(), static_cast<int>(out.size()), &k));
        out.resize(k);
        output.swap(out);
    }
    // void GetCameraParameters(float& fx, float& fy, float& cx, float& cy, int camNumber = 0) const (Reconstruct3D.cpp:87-103)
    void GetCameraParameters(float& fx, float& fy, float& cx, float& cy, int camNumber = 0) const
    {
        if (camNumber == 1) { fx = m_calib.fxR; fy = m_calib.fyR; cx = m_calib.cxR; cy = m_calib.cyR; }
        else { fx = m_calib.fxL; fy = m_calib.fyL; cx = m_calib.cxL; cy = m_calib.cyL; }
    }
    // pcl::PointXYZRGB TriangulateUV(float u, float v, float disparity, const cv::Vec3b& color) const (Reconstruct3D.cpp:305-328):
    // the single-point form of Triangulate3D the sparse tracker calls; host arithmetic, same operation order.
    cvr_point TriangulateUV(float u, float v, float disparity, const uint8_t colorBGR[3]) const
    {
        const float b = m_calib.T0, f = (m_calib.fxL + m_calib.fyL) / 2.0f;
        cvr_point pt;
        pt.z = f * b / disparity;
        pt.x = (u - m_calib.cxL) * (pt.z / f);
        pt.y = -(v - m_calib.cyL) * (pt.z / f);
        pt.r = colorBGR[2]; pt.g = colorBGR[1]; pt.b = colorBGR[0]; pt.a = 255;
        return pt;
    }
    // pcl::PointXYZ BackProjectPoint(float x, float y) const (Reconstruct3D.cpp:283-302): the ray through pixel (x, y) at Z = 1
    void BackProjectPoint(float x, float y, float xyz[3]) const
    {
        const float Z = 1.0f;
        xyz[0] = (x - m_calib.cxL) * Z / m_calib.fxL; xyz[1] = (y - m_calib.cyL) * Z / m_calib.fyL; xyz[2] = Z;
    }
    // The three setters (Reconstruct3D.cpp:364-388) act on the cv::StereoMatcher the reference holds; here the workspace
    // depends on these parameters, so the context is rebuilt (rectification maps, if any, must be set again).
    void SetStereoBMWindowSize(int size)        // m_StereoMatcher->setBlockSize(size): whichever matcher is active
    {
        if (m_params.block_matcher == CVR_MATCHER_BM) m_params.bm_window_size = size; else m_params.sgbm.block_size = size;
        rebuild();
    }
    void SetStereoBMNumDisparities(int num)     // m_StereoMatcher->setNumDisparities(num)
    {
        if (m_params.block_matcher == CVR_MATCHER_BM) m_params.bm_num_disparities = num; else m_params.sgbm.num_disparities = num;
        rebuild();
    }
    void SetBlockMatcherType(StereoBlockMatcherType type)   // a NEW matcher with OpenCV's defaults, as the reference creates it
    {
        if (type == STEREO_BLOCK_MATCHER) {                  // cv::StereoBM::create(16, 21)
            m_params.block_matcher = CVR_MATCHER_BM; m_params.bm_num_disparities = 16; m_params.bm_window_size = 21;
            m_params.bm_min_disparity = 0; m_params.bm_pre_filter_cap = 31; m_params.bm_texture_threshold = 10; m_params.bm_uniqueness_ratio = 15;
            m_params.bm_speckle_window_size = 0; m_params.bm_speckle_range = 0;
        } else {                                             // cv::StereoSGBM::create(0, 16, 3): every other parameter 0, MODE_SGBM
            m_params.block_matcher = CVR_MATCHER_SGBM;
            m_params.sgbm.min_disparity = 0; m_params.sgbm.num_disparities = 16; m_params.sgbm.block_size = 3; m_params.sgbm.P1 = 0; m_params.sgbm.P2 = 0;
            m_params.sgbm.disp12_max_diff = 0; m_params.sgbm.pre_filter_cap = 0; m_params.sgbm.uniqueness_ratio = 0;
            m_params.sgbm.speckle_window_size = 0; m_params.sgbm.speckle_range = 0; m_params.sgbm.mode = CVR_MODE_SGBM;
        }
        rebuild();
    }
    cvr_ctx* context() const { return m_ctx; }

    // ---- batched / multi-GPU side (no counterpart class in the reference: it processes one frame per call on one machine) ----
    // What MapDataBase::InsertBlock (MapDataBase.cpp:28-54) receives per processed frame, for a whole batch: the filtered
    // clouds back to back, frame i = points[offsets[i] .. offsets[i + 1]).
    struct MapBlocks {
        std::vector<cvr_point> points;
        std::vector<int32_t> offsets;          // frames + 1 entries
        std::vector<int32_t> framesOfRank;     // GatherMapBlocks on the root: frames contributed by each rank
        std::size_t frames() const { return offsets.empty() ? 0 : offsets.size() - 1; }
    };
    // ReconstructionSystem::ProcessStereoFrame (ReconstructionSystem.cpp:35-61) + the filter call sites for `n` frames at once
    // (n <= maxBatch of the constructor): left[i] / right[i] are the frames, all of the constructor's size.
    MapBlocks ProcessStereoFrames(const Mat* left, const Mat* right, int n, std::size_t pointCapacity = 0) const
    {
        const std::size_t rowBytes = static_cast<std::size_t>(m_width) * left[0].channels, frameBytes = rowBytes * m_height;
        std::vector<uint8_t> l(frameBytes * n), r(frameBytes * n);                                // one dense staging copy per camera
        for (int i = 0; i < n; i++)
            for (int y = 0; y < m_height; y++) {
                std::memcpy(&l[i * frameBytes + y * rowBytes], static_cast<const uint8_t*>(left[i].data) + y * left[i].step, rowBytes);
                std::memcpy(&r[i * frameBytes + y * rowBytes], static_cast<const uint8_t*>(right[i].data) + y * right[i].step, rowBytes);
            }
        MapBlocks out;
        out.points.resize(pointCapacity ? pointCapacity : static_cast<std::size_t>(m_width) * m_height * n);
        out.offsets.assign(n + 1, 0);
        cvr_batch io;
        std::memset(&io, 0, sizeof(io));
        io.mem = CVR_MEM_HOST; io.channels = left[0].channels; io.stages = 15;
        io.left = l.data(); io.left_step = rowBytes; io.left_frame_stride = frameBytes;
        io.right = r.data(); io.right_step = rowBytes; io.right_frame_stride = frameBytes;
        io.cloud_packed = out.points.data(); io.cloud_packed_capacity = out.points.size(); io.cloud_offsets = out.offsets.data();
        check(cvr_process_batch(m_ctx, n, &io));
        out.points.resize(static_cast<std::size_t>(out.offsets[n]));
        return out;
    }
    // One process per GPU: every rank joins the communicator once (the 128-byte id comes from rank 0's NewCommId by whatever
    // channel the host has), then GatherMapBlocks after each batch brings all ranks' map blocks to the mapping rank
    // (MappingSystem.cpp:344-345: AddFrame -> InsertBlock) -- a collective call; only the root's return value is filled.
    // pcl::io::savePCDFileBinary(path, cloud) (ReconstructionServer.cpp:165, MappingSystem.cpp:61, 84)
    static void SavePCD(const std::string& path, const std::vector<cvr_point>& cloud)
    {
        if (cvr_save_pcd(path.c_str(), cloud.data(), static_cast<int>(cloud.size())) != CVR_OK)
            throw std::runtime_error(std::string("cvrecon_b200: ") + cvr_last_error(nullptr));
    }
    static std::vector<uint8_t> NewCommId()
    {
        std::vector<uint8_t> id(CVR_COMM_ID_BYTES);
        if (cvr_comm_get_unique_id(id.data()) != CVR_OK) throw std::runtime_error(std::string("cvrecon_b200: ") + cvr_last_error(nullptr));
        return id;
    }
    void JoinComm(int nranks, int rank, const std::vector<uint8_t>& id) { check(cvr_comm_init(m_ctx, nranks, rank, id.data())); m_nranks = nranks; m_rank = rank; }
    MapBlocks GatherMapBlocks(int root, std::size_t pointCapacity, int maxFramesPerRank) const
    {
        MapBlocks out;
        if (m_rank != root) { check(cvr_gather_clouds(m_ctx, root, CVR_MEM_HOST, nullptr, 0, nullptr, 0, nullptr)); return out; }
        out.points.resize(pointCapacity);
        out.offsets.assign(static_cast<std::size_t>(m_nranks) * maxFramesPerRank + 1, 0);
        out.framesOfRank.assign(m_nranks, 0);
        check(cvr_gather_clouds(m_ctx, root, CVR_MEM_HOST, out.points.data(), out.points.size(), out.offsets.data(),
                                static_cast<int>(out.offsets.size()), out.framesOfRank.data()));
        int frames = 0;
        for (int v : out.framesOfRank) frames += v;
        out.offsets.resize(frames + 1);
        out.points.resize(static_cast<std::size_t>(out.offsets[frames]));
        return out;
    }

private:
    void check(int rc) const { if (rc != CVR_OK) throw std::runtime_error(std::string("cvrecon_b200: ") + cvr_last_error(m_ctx)); }
    void rebuild()
    {
        cvr_ctx* fresh = nullptr;
        if (cvr_create(&m_params, m_device, &fresh) != CVR_OK) throw std::runtime_error(std::string("Reconstruct3D: ") + cvr_last_error(nullptr));
        cvr_destroy(m_ctx);
        m_ctx = fresh;
    }
    cvr_params m_params;
    int m_device = 0;
    cvr_ctx* m_ctx = nullptr;
    StereoCalib m_calib;
    int m_width = 0, m_height = 0;
    int m_nranks = 1, m_rank = 0;
};

}  // namespace cvr
