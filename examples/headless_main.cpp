// headless_main.cpp -- RunNrcHpm (reference src/main.cpp:109-277) without the window: same construction order, same
// per-frame sequence (camera update, ResetStats, Render, wait, stats every 25 frames, VolumeData::Update), a PPM of the
// last frame instead of a swapchain.  Build: see INTEGRATION.md.
//   headless_main [frames=100] [width=800] [height=800] [out.ppm]
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <memory>
#include "nrc_b200.hpp"

// deterministic stand-ins for the blobs the reference repo does not ship (.MISSING_LARGE_BLOBS)
static std::vector<float> MakeDensity(uint32_t nx, uint32_t ny, uint32_t nz)
{
    std::vector<float> d((size_t)nx * ny * nz);
    for (uint32_t x = 0; x < nx; x++)
        for (uint32_t y = 0; y < ny; y++)
            for (uint32_t z = 0; z < nz; z++) {
                float fx = ((x + 0.5f) / nx - 0.5f) / 0.42f, fy = ((y + 0.5f) / ny - 0.45f) / 0.36f, fz = ((z + 0.5f) / nz - 0.5f) / 0.44f;
                float r = std::sqrt(fx * fx + fy * fy + fz * fz);
                float n = 0.5f + 0.5f * std::sin(0.35f * x) * std::sin(0.29f * y + 1.0f) * std::sin(0.23f * z + 2.0f);
                float v = (1.0f - r + 0.5f * (n - 0.5f)) * 2.5f;
                d[((size_t)x * ny + y) * nz + z] = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
            }
    return d;
}

static std::vector<float> MakeSky(uint32_t w, uint32_t h)
{
    std::vector<float> e((size_t)w * h * 4);
    for (uint32_t y = 0; y < h; y++)
        for (uint32_t x = 0; x < w; x++) {
            float t = (y + 0.5f) / h;
            float* p = &e[((size_t)y * w + x) * 4];
            p[0] = 0.25f + 0.35f * (1 - t); p[1] = 0.35f + 0.35f * (1 - t); p[2] = 0.55f + 0.4f * t; p[3] = 1.0f;
        }
    return e;
}

int main(int argc, char** argv)
{
    int frames = argc > 1 ? std::atoi(argv[1]) : 100;
    uint32_t width = argc > 2 ? (uint32_t)std::atoi(argv[2]) : 800, height = argc > 3 ? (uint32_t)std::atoi(argv[3]) : 800;
    const char* out = argc > 4 ? argv[4] : "frame.ppm";
    try {
        std::cout << nrc_version() << std::endl;
        en::Context ctx(0);                                                   // VulkanAPI::Init, main.cpp:120
        auto density = MakeDensity(125, 85, 153);                             // ReadFileDensity3D, main.cpp:123
        en::VolumeData volumeData(ctx, density.data(), 125, 85, 153);         // main.cpp:124-129
        uint32_t hdrW = 1024, hdrH = 512;
        auto hdr4f = MakeSky(hdrW, hdrH);                                     // ReadFileHdr4f, main.cpp:131-132
        auto hdrCdf = en::Hdr4fToCdf(ctx, hdr4f, hdrW, hdrH);                 // main.cpp:133
        en::HdrEnvMap hdrEnvMap(ctx, hdrW, hdrH, hdr4f, hdrCdf[0], hdrCdf[1]);   // main.cpp:134-139
        en::Camera camera({0.0f, 0.0f, -64.0f}, {0.0f, 0.0f, 1.0f}, {0.0f, 1.0f, 0.0f}, (float)width / (float)height, 1.0471975512f, 0.1f,
                          100.0f);                                            // main.cpp:142-149
        en::DirLight dirLight(-1.57f, 0.0f, {1.0f, 1.0f, 1.0f}, 1.0f);       // main.cpp:151 (strength 0 there: target == 0)
        en::PointLight pointLight({0.0f, 0.0f, 0.0f}, {1.0f, 1.0f, 1.0f}, 1.0f);   // main.cpp:152
        en::NeuralRadianceCache nrc(ctx, 0.001f, 0.0f, 0.5f);                 // main.cpp:156
        en::MRHE mrhe(ctx, 0.01f, 0.0f);                                      // main.cpp:157
        volumeData.Params().useNN = 1;
        en::NrcHpmRenderer renderer(ctx, width, height, 100, 100, camera, volumeData, dirLight, pointLight, hdrEnvMap, nrc, mrhe);   // main.cpp:159-166
        if (const char* ck = std::getenv("NRC_LOAD_CHECKPOINT")) en::LoadCheckpoint(ck, nrc, mrhe);   // resume a trained cache (not in the reference)
        // NRC_SPLIT_FRAME=1: the frame in its data-parallel form (train | exchange+steps | render) with a world of one rank;
        // a multi-GPU host constructs DataParallel(nrc, mrhe, rank, world), gathers Handle() over its own transport and
        // gives every rank its rows with renderer.SetPartition.
        const bool split = std::getenv("NRC_SPLIT_FRAME") != nullptr;
        std::unique_ptr<en::DataParallel> dp;
        if (split) {
            dp.reset(new en::DataParallel(nrc, mrhe, 0, 1));
            dp->Connect(dp->Handle());
        }
        for (int frame = 0; frame < frames; frame++) {                        // main.cpp:181-248
            if (frame % 100 == 0 && std::getenv("NRC_PRINT_WEIGHTS")) nrc.PrintWeights();   // main.cpp:183-187
            camera.RotateAroundOrigin({0.0f, 1.0f, 0.0f}, 0.01f);             // "camera training" orbit, main.cpp:200-203
            camera.SetAspectRatio(width, height);
            camera.UpdateUniformBuffer();                                     // main.cpp:212-213
            nrc.ResetStats();                                                 // main.cpp:215
            if (split) { renderer.FrameTrain(); dp->ReduceStep(); renderer.FrameRender(); }
            else renderer.Render();                                           // main.cpp:216
            ctx.WaitIdle();                                                   // main.cpp:217
            if (frame % 25 == 0) std::cout << "frame " << frame << " NRC MSE Loss: " << nrc.GetStats().mseLoss << std::endl;   // main.cpp:220-224
            volumeData.Update(true);                                          // main.cpp:230
        }
        std::vector<unsigned char> bgra(renderer.GetImageDataSize());
        renderer.ReadImage(bgra.data());
        std::ofstream f(out, std::ios::binary);
        f << "P6\n" << width << " " << height << "\n255\n";
        for (size_t i = 0; i < (size_t)width * height; i++) { char rgb[3] = {(char)bgra[4 * i + 2], (char)bgra[4 * i + 1], (char)bgra[4 * i]}; f.write(rgb, 3); }
        if (const char* ck = std::getenv("NRC_SAVE_CHECKPOINT")) en::SaveCheckpoint(ck, nrc, mrhe);
        if (dp) { if (dp->TimedOut()) std::cerr << "exchange handshake timed out" << std::endl; dp->Destroy(); }
        renderer.Destroy(); mrhe.Destroy(); nrc.Destroy(); hdrEnvMap.Destroy(); volumeData.Destroy();   // main.cpp:253-274
        std::cout << "wrote " << out << std::endl;
    } catch (const std::exception& e) {
        std::cerr << "error: " << e.what() << std::endl;
        return 1;
    }
    return 0;
}
