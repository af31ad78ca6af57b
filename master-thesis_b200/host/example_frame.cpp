// example_frame.cpp -- one frame through the C++ mirror, the way ReconstructionSystem::ProcessStereoFrame
// (cv_reconstruct/src/system/ReconstructionSystem.cpp:35-61) drives the reference class.  Reads two raw BGR frames
// (width*height*3 bytes each) and writes disparity (int16) and the filtered cloud (16-byte points) as raw files.
// Build: g++ -std=c++14 -I include master-thesis_b200/host/example_frame.cpp -L master-thesis_b200/lib -lcvrecon_b200
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>

#include "Reconstruct3D_b200.hpp"

static std::vector<uint8_t> slurp(const char* path, std::size_t n)
{
    std::vector<uint8_t> b(n);
    std::ifstream f(path, std::ios::binary);
    if (!f.read(reinterpret_cast<char*>(b.data()), static_cast<std::streamsize>(n))) { std::cerr << "cannot read " << path << "\n"; std::exit(2); }
    return b;
}

int main(int argc, char** argv)
{
    if (argc < 7) { std::cerr << "usage: example_frame W H left.bgr right.bgr out_disp.i16 out_cloud.bin [block_size num_disparities [out_bm_disp.i16]]\n"; return 2; }
    const int W = std::atoi(argv[1]), H = std::atoi(argv[2]);
    cvr::StereoCalib calib{959.791f, 956.9251f, 696.0217f, 224.1806f, 903.7596f, 901.9653f, 695.7519f, 224.2509f, 0.532671f};
    cvr::SGBMConfig cfg;
    cfg.BlockSize = argc > 7 ? std::atoi(argv[7]) : 27; cfg.NumDisparities = argc > 8 ? std::atoi(argv[8]) : 128;
    cfg.UniquenessRatio = 10; cfg.SpeckleRange = 2; cfg.SpeckleWindowSize = 100; cfg.PreFilterCap = 10; cfg.MinDisparity = 0;
    try {
        cvr::Reconstruct3D recon(calib, cfg, W, H);
        std::vector<uint8_t> l = slurp(argv[3], static_cast<std::size_t>(W) * H * 3), r = slurp(argv[4], static_cast<std::size_t>(W) * H * 3);
        cvr::Mat left{l.data(), H, W, static_cast<std::size_t>(W) * 3, 3}, right{r.data(), H, W, static_cast<std::size_t>(W) * 3, 3};
        std::vector<int16_t> disparity = recon.GenerateDisparityMap(left, right);
        std::ofstream(argv[5], std::ios::binary).write(reinterpret_cast<const char*>(disparity.data()), static_cast<std::streamsize>(disparity.size() * 2));
        std::vector<float> dispF; std::vector<uint8_t> mask;
        recon.SetupFrame(disparity, dispF, mask);
        std::vector<cvr_point> cloud = recon.Triangulate3D(dispF, left, &mask);
        cloud = recon.VoxelGridFilter(cloud);
        recon.RemoveOutliers(cloud, cloud);
        std::ofstream(argv[6], std::ios::binary).write(reinterpret_cast<const char*>(cloud.data()), static_cast<std::streamsize>(cloud.size() * sizeof(cvr_point)));
        std::cout << "points " << cloud.size() << "\n";
        if (argc > 9) {      // the reference's setters: a new StereoBM with OpenCV's defaults, then window size and disparity range
            recon.SetBlockMatcherType(cvr::STEREO_BLOCK_MATCHER);
            recon.SetStereoBMNumDisparities(cfg.NumDisparities);
            recon.SetStereoBMWindowSize(cfg.BlockSize);
            std::vector<int16_t> bm = recon.GenerateDisparityMap(left, right);
            std::ofstream(argv[9], std::ios::binary).write(reinterpret_cast<const char*>(bm.data()), static_cast<std::streamsize>(bm.size() * 2));
            const uint8_t white[3] = {255, 255, 255};
            const cvr_point q = recon.TriangulateUV(100.f, 50.f, 16.f, white);
            std::cout << "uv " << q.x << " " << q.y << " " << q.z << "\n";
        }
    } catch (const std::exception& e) {
        std::cerr << e.what() << "\n";
        return 3;
    }
    return 0;
}
