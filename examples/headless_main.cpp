// headless_main.cpp -- RunNrcHpm (reference src/main.cpp:109-277) without the window: same construction order, same
// per-frame sequence (camera update, ResetStats, Render, wait, stats every 25 frames, VolumeData::Update), PPM / PFM files of the
// last frame instead of a swapchain.  Every literal of the reference's main is an option (en::NrcConfig, include/nrc_host.hpp):
//   headless_main --frames=100 --width=800 --height=800 --out=frame.ppm [--help]
// Build: see INTEGRATION.md.
#include <chrono>
#include <iostream>
#include <memory>
#include <thread>
#include "nrc_b200.hpp"
#include "nrc_host.hpp"

// Handle exchange of the data-parallel ranks through a directory every rank can see (one process per GPU; any transport would do:
// MPI_Allgather, a socket, torch.distributed in the Python host).
static std::vector<unsigned char> ExchangeHandles(const std::string& dir, const std::vector<unsigned char>& mine, int rank, int world)
{
    {
        std::ofstream f(dir + "/handle." + std::to_string(rank) + ".tmp", std::ios::binary);
        f.write(reinterpret_cast<const char*>(mine.data()), (std::streamsize)mine.size());
    }
    std::rename((dir + "/handle." + std::to_string(rank) + ".tmp").c_str(), (dir + "/handle." + std::to_string(rank)).c_str());
    std::vector<unsigned char> all;
    for (int r = 0; r < world; r++) {
        std::vector<unsigned char> h;
        for (int tries = 0; tries < 1200 && h.size() != mine.size(); tries++) {
            std::ifstream f(dir + "/handle." + std::to_string(r), std::ios::binary);
            if (f) h.assign((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
            if (h.size() != mine.size()) std::this_thread::sleep_for(std::chrono::milliseconds(50));
        }
        if (h.size() != mine.size()) throw std::runtime_error("rank " + std::to_string(r) + " never published its exchange handle in " + dir);
        all.insert(all.end(), h.begin(), h.end());
    }
    return all;
}

int main(int argc, char** argv)
{
    try {
        const en::NrcConfig cfg = en::ParseNrcConfig(argc, argv);
        const uint32_t width = cfg.width, height = cfg.height;
        std::cout << nrc_version() << std::endl;
        en::Context ctx(cfg.device);                                          // VulkanAPI::Init, main.cpp:120
        ctx.SetPrecision(cfg.precision == "fp32" ? NRC_MLP_FP32 : NRC_MLP_BF16_TC);
        std::vector<float> density = cfg.density.empty() ? en::MakeStandInDensity(cfg.gridX, cfg.gridY, cfg.gridZ)
                                                         : en::ReadFileDensity3D(cfg.density, cfg.gridX, cfg.gridY, cfg.gridZ);   // main.cpp:123
        en::VolumeData volumeData(ctx, density.data(), cfg.gridX, cfg.gridY, cfg.gridZ);   // main.cpp:124-129
        int hdrW = 1024, hdrH = 512;
        std::vector<float> hdr4f = cfg.envmap.empty() ? en::MakeStandInSky((uint32_t)hdrW, (uint32_t)hdrH)
                                                      : en::ReadFileHdr4f(cfg.envmap, hdrW, hdrH);   // main.cpp:131-132
        auto hdrCdf = en::Hdr4fToCdf(ctx, hdr4f, (uint32_t)hdrW, (uint32_t)hdrH);         // main.cpp:133
        en::HdrEnvMap hdrEnvMap(ctx, (uint32_t)hdrW, (uint32_t)hdrH, hdr4f, hdrCdf[0], hdrCdf[1]);   // main.cpp:134-139
        hdrEnvMap.SetStrengths(cfg.envDirect, cfg.envHpm);
        en::Camera camera({cfg.camPos[0], cfg.camPos[1], cfg.camPos[2]}, {cfg.camDir[0], cfg.camDir[1], cfg.camDir[2]}, {0.0f, 1.0f, 0.0f},
                          (float)width / (float)height, cfg.fovDeg * 3.14159265358979323846f / 180.0f, cfg.nearPlane, cfg.farPlane);   // main.cpp:142-149
        en::DirLight dirLight(cfg.dirZenith, cfg.dirAzimuth, {1.0f, 1.0f, 1.0f}, cfg.dirStrength);   // main.cpp:151 (strength 0 there: target == 0)
        en::PointLight pointLight({cfg.pointPos[0], cfg.pointPos[1], cfg.pointPos[2]}, {cfg.pointColor[0], cfg.pointColor[1], cfg.pointColor[2]},
                                  cfg.pointStrength);                          // main.cpp:152
        en::NeuralRadianceCache nrc(ctx, cfg.lr, cfg.weightDecay, cfg.beta1);  // main.cpp:156
        uint32_t res[16];
        en::Check(nrc_mrhe_resolutions(cfg.minRes, cfg.maxRes, cfg.levels < 2 ? 2 : cfg.levels,
                                       cfg.resolutions == "intended" ? NRC_RES_INTENDED : NRC_RES_AS_SHIPPED_STD140, res), "nrc_mrhe_resolutions");
        en::MRHE mrhe(ctx, cfg.mrheLr, cfg.mrheWeightDecay, cfg.levels, cfg.hashTableSize, 2, cfg.minRes, cfg.maxRes, res);   // main.cpp:157
        if (cfg.fixedPreset) mrhe.SetFixedMode(true);
        if (cfg.adam) { nrc.SetAdam(true); mrhe.SetAdam(true); }
        nrc_volume_params& vp = volumeData.Params();
        vp.useNN = cfg.useNN; vp.showNonNN = cfg.showNonNN; vp.densityFactor = cfg.densityFactor; vp.g = cfg.g;
        vp.noNnSpp = cfg.noNnSpp; vp.withNnSpp = cfg.withNnSpp;
        volumeData.SetSeed(cfg.seed);
        en::NrcHpmRenderer renderer(ctx, width, height, cfg.trainWidth, cfg.trainHeight, camera, volumeData, dirLight, pointLight, hdrEnvMap, nrc,
                                    mrhe);                                     // main.cpp:159-166
        if (!cfg.loadCheckpoint.empty()) en::LoadCheckpoint(cfg.loadCheckpoint, nrc, mrhe);   // resume a trained cache (not in the reference)
        // --gpus=N --rank=r --rendezvous=<dir>: one process per GPU; this rank renders rows r, r+N, ... and traces train invocations
        // r, r+N, ...; the gradients are summed over NVLink inside the fused step kernel.
        // NRC_SPLIT_FRAME=1: the same three-call frame with a world of one rank.
        const bool split = cfg.gpus > 1 || std::getenv("NRC_SPLIT_FRAME") != nullptr;
        std::unique_ptr<en::DataParallel> dp;
        if (split) {
            dp.reset(new en::DataParallel(nrc, mrhe, cfg.rank, cfg.gpus));
            if (cfg.gpus > 1) {
                const char* dir = std::getenv("NRC_RENDEZVOUS_DIR");
                if (!dir) throw std::runtime_error("--gpus > 1 needs NRC_RENDEZVOUS_DIR=<directory shared by the ranks> for the handle exchange");
                dp->Connect(ExchangeHandles(dir, dp->Handle(), cfg.rank, cfg.gpus));
                renderer.SetPartitionStrided((uint32_t)cfg.rank, (uint32_t)cfg.gpus, (uint32_t)cfg.rank, (uint32_t)cfg.gpus);
            } else {
                dp->Connect(dp->Handle());
            }
        }
        for (int frame = 0; frame < cfg.frames; frame++) {                    // main.cpp:181-248
            if (cfg.weightsEvery > 0 && frame % cfg.weightsEvery == 0) nrc.PrintWeights();   // main.cpp:183-187
            if (cfg.orbit != 0.0f) camera.RotateAroundOrigin({0.0f, 1.0f, 0.0f}, cfg.orbit);   // "camera training" orbit, main.cpp:200-203
            camera.SetAspectRatio(width, height);
            camera.UpdateUniformBuffer();                                     // main.cpp:212-213
            nrc.ResetStats();                                                 // main.cpp:215
            if (split) { renderer.FrameTrain(); dp->ReduceStep(); renderer.FrameRender(); }
            else renderer.Render();                                           // main.cpp:216
            ctx.WaitIdle();                                                   // main.cpp:217
            if (cfg.statsEvery > 0 && frame % cfg.statsEvery == 0)
                std::cout << "frame " << frame << " NRC MSE Loss: " << nrc.GetStats().mseLoss << std::endl;   // main.cpp:220-224
            volumeData.Update(true);                                          // main.cpp:230
        }
        if (!cfg.out.empty()) {
            std::vector<unsigned char> bgra(renderer.GetImageDataSize());
            renderer.ReadImage(bgra.data());
            en::WritePpm(cfg.out, bgra.data(), width, height);
            std::cout << "wrote " << cfg.out << std::endl;
        }
        if (!cfg.outPfm.empty()) {
            std::vector<float> rgba((size_t)width * height * 4);
            renderer.ReadImage(rgba.data(), NRC_FRAME_RGBA32F);
            en::WritePfm(cfg.outPfm, rgba.data(), width, height);
            std::cout << "wrote " << cfg.outPfm << std::endl;
        }
        if (!cfg.saveCheckpoint.empty()) en::SaveCheckpoint(cfg.saveCheckpoint, nrc, mrhe);
        if (dp) { if (dp->TimedOut()) std::cerr << "exchange handshake timed out" << std::endl; dp->Destroy(); }
        renderer.Destroy(); mrhe.Destroy(); nrc.Destroy(); hdrEnvMap.Destroy(); volumeData.Destroy();   // main.cpp:253-274
    } catch (const std::exception& e) {
        std::cerr << "error: " << e.what() << std::endl;
        return 1;
    }
    return 0;
}
