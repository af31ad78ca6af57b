// example_gather.cpp -- a C++ host driving the batched path and the cloud gather: what one process per GPU does with a shard of
// the frame sequence (ReconstructionServer.cpp:134-153 dequeues frames, ReconstructionSystem::ProcessStereoFrame processes
// them, MappingSystem.cpp:344-345 hands each cloud to the one map).  Every rank processes `n` frames of its own, then
// GatherMapBlocks brings all clouds to rank 0, which checks them against what it computes itself for the other ranks' frames.
//   example_gather W H n frames.bgr                      one process (the gather degenerates to a local copy)
//   RANK=r WORLD_SIZE=w LOCAL_RANK=g CVR_ID_FILE=path example_gather W H n frames.bgr      one process per GPU; rank 0 leaves
//       the NCCL id in CVR_ID_FILE, the others pick it up there (any channel of the host would do)
// frames.bgr holds w * n stereo pairs: for pair p the left then the right frame, W*H*3 bytes each; rank r owns pairs r*n .. r*n+n-1.
// Build: g++ -std=c++14 -I include master-thesis_b200/host/example_gather.cpp -L master-thesis_b200/lib -lcvrecon_b200
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <thread>

#include "Reconstruct3D_b200.hpp"

int main(int argc, char** argv)
{
    if (argc < 5) { std::cerr << "usage: example_gather W H n frames.bgr\n"; return 2; }
    const int W = std::atoi(argv[1]), H = std::atoi(argv[2]), n = std::atoi(argv[3]);
    const int rank = std::getenv("RANK") ? std::atoi(std::getenv("RANK")) : 0, world = std::getenv("WORLD_SIZE") ? std::atoi(std::getenv("WORLD_SIZE")) : 1;
    const int dev = std::getenv("LOCAL_RANK") ? std::atoi(std::getenv("LOCAL_RANK")) : 0;
    const std::size_t fb = static_cast<std::size_t>(W) * H * 3;
    std::vector<uint8_t> all(fb * 2 * n * world);
    { std::ifstream f(argv[4], std::ios::binary); if (!f.read(reinterpret_cast<char*>(all.data()), static_cast<std::streamsize>(all.size()))) { std::cerr << "cannot read frames\n"; return 2; } }
    cvr::StereoCalib calib{959.791f, 956.9251f, 696.0217f, 224.1806f, 903.7596f, 901.9653f, 695.7519f, 224.2509f, 0.532671f};
    cvr::SGBMConfig cfg;
    cfg.BlockSize = 9; cfg.NumDisparities = 64; cfg.UniquenessRatio = 10; cfg.SpeckleRange = 2; cfg.SpeckleWindowSize = 100; cfg.PreFilterCap = 10;
    cfg.OutlierMinK = 8;
    try {
        cvr::Reconstruct3D recon(calib, cfg, W, H, dev, n);
        if (world > 1) {
            const char* idf = std::getenv("CVR_ID_FILE");
            if (!idf) { std::cerr << "CVR_ID_FILE not set\n"; return 2; }
            std::vector<uint8_t> id(CVR_COMM_ID_BYTES);
            if (rank == 0) {
                id = cvr::Reconstruct3D::NewCommId();
                std::ofstream((std::string(idf) + ".tmp").c_str(), std::ios::binary).write(reinterpret_cast<const char*>(id.data()), CVR_COMM_ID_BYTES);
                std::rename((std::string(idf) + ".tmp").c_str(), idf);
            } else {
                for (int tries = 0; tries < 600; tries++) {
                    std::ifstream f(idf, std::ios::binary);
                    if (f.read(reinterpret_cast<char*>(id.data()), CVR_COMM_ID_BYTES)) break;
                    std::this_thread::sleep_for(std::chrono::milliseconds(100));
                }
            }
            recon.JoinComm(world, rank, id);
        }
        auto shard = [&](int r) {
            std::vector<cvr::Mat> L, R;
            for (int i = 0; i < n; i++) {
                uint8_t* base = all.data() + fb * 2 * (static_cast<std::size_t>(r) * n + i);
                L.push_back(cvr::Mat{base, H, W, static_cast<std::size_t>(W) * 3, 3});
                R.push_back(cvr::Mat{base + fb, H, W, static_cast<std::size_t>(W) * 3, 3});
            }
            return recon.ProcessStereoFrames(L.data(), R.data(), n);
        };
        std::vector<cvr::Reconstruct3D::MapBlocks> want;
        if (rank == 0) for (int r = 1; r < world; r++) want.push_back(shard(r));      // what the other ranks are going to send
        const cvr::Reconstruct3D::MapBlocks mine = shard(rank);                        // the batch completed last is the one gathered
        const cvr::Reconstruct3D::MapBlocks got = recon.GatherMapBlocks(0, static_cast<std::size_t>(W) * H * n * world, n);
        if (rank == 0) {
            bool ok = got.frames() == static_cast<std::size_t>(n) * world;
            std::size_t f = 0;
            for (int r = 0; r < world && ok; r++) {
                const cvr::Reconstruct3D::MapBlocks& w = r == 0 ? mine : want[r - 1];
                for (int i = 0; i < n && ok; i++, f++) {
                    const std::size_t len = w.offsets[i + 1] - w.offsets[i];
                    ok = static_cast<std::size_t>(got.offsets[f + 1] - got.offsets[f]) == len &&
                         std::memcmp(&got.points[got.offsets[f]], &w.points[w.offsets[i]], len * sizeof(cvr_point)) == 0;
                }
            }
            std::cout << (ok ? "gather ok" : "gather MISMATCH") << " ranks " << world << " frames " << got.frames() << " points " << got.points.size() << "\n";
            return ok ? 0 : 1;
        }
    } catch (const std::exception& e) {
        std::cerr << e.what() << "\n";
        return 3;
    }
    return 0;
}
