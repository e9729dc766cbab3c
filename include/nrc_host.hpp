// nrc_host.hpp -- the host side either side of the hot path, in C++ like the reference's (north_star: "host code stays C++"):
//
//   en::ReadFileDensity3D   src/read_file.cpp:50-78    raw float32 grid, index = x*Y*Z + y*Z + z
//   en::ReadFileHdr4f       src/read_file.cpp:80-121   Radiance RGBE (.hdr) -> float RGBA, flipped vertically.  The reference calls
//                                                      stb_image (stbi_set_flip_vertically_on_load(true); stbi_loadf(..., 4)); stb is
//                                                      an un-vendored vcpkg dependency (CMakeLists.txt:32-34), so its published
//                                                      algorithm is restated here (stb_image.h: stbi__hdr_load / stbi__hdr_convert:
//                                                      value = mantissa byte * 2^(e - 136), alpha = 1, new-style RLE scanlines and flat
//                                                      files) -- dependency-free.  Cross-checked against OpenCV on the reference's own
//                                                      data/image/photostudio.hdr (tests/test_cpp_host.py).
//   en::WritePpm / WritePfm frame writers (the reference presents to a swapchain instead, src/main.cpp:245)
//   en::NrcConfig           ONE struct for every literal of RunNrcHpm (src/main.cpp:112-166) and the ctor defaults behind them
//                           (src/MRHE.cpp:83-87, src/VolumeData.cpp:82-89, src/HdrEnvMap.cpp:108), with a --key=value parser
//                           (the reference has no CLI or config file at all; everything is an ImGui slider or a literal).
//
// Header-only, no CUDA, no third-party code: it compiles with any C++17 compiler and is tested on the CPU.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace en {

// ---- ReadFileDensity3D (src/read_file.cpp:50-78) --------------------------------------------------------------------------------
// The reference returns vector<vector<vector<float>>>[x][y][z]; the flat array in the same order (x slowest, z fastest) is what
// nrc_volume_create takes.  Throws like the reference's Log::Error(..., true) when the file is missing or short.
inline std::vector<float> ReadFileDensity3D(const std::string& fileName, size_t xSize, size_t ySize, size_t zSize)
{
    std::ifstream f(fileName, std::ios::binary);
    if (!f) throw std::runtime_error("Failed to open file " + fileName);
    const size_t total = xSize * ySize * zSize;
    std::vector<float> d(total);
    f.read(reinterpret_cast<char*>(d.data()), (std::streamsize)(total * sizeof(float)));
    if ((size_t)f.gcount() != total * sizeof(float)) throw std::runtime_error("Failed to read " + std::to_string(total) + " floats from " + fileName);
    return d;
}

inline void WriteFileDensity3D(const std::string& fileName, const std::vector<float>& density)
{
    std::ofstream f(fileName, std::ios::binary);
    if (!f) throw std::runtime_error("Failed to open file " + fileName);
    f.write(reinterpret_cast<const char*>(density.data()), (std::streamsize)(density.size() * sizeof(float)));
}

// ---- Radiance .hdr (RGBE) ----------------------------------------------------------------------------------------------------------
// bytes of a file -> width, height, RGBE bytes [h][w][4] in file (top-down) order
inline std::vector<unsigned char> DecodeHdr(const std::vector<unsigned char>& data, int& width, int& height)
{
    size_t pos = 0;
    auto line = [&]() -> std::string {
        size_t end = pos;
        while (end < data.size() && data[end] != '\n') end++;
        if (end >= data.size()) throw std::runtime_error("not a Radiance HDR file (truncated header)");
        std::string s(data.begin() + (long)pos, data.begin() + (long)end);
        pos = end + 1;
        return s;
    };
    const std::string first = line();
    if (first != "#?RADIANCE" && first != "#?RGBE") throw std::runtime_error("not a Radiance HDR file");
    bool valid = false;
    for (;;) {
        const std::string s = line();
        if (s.empty()) break;
        if (s == "FORMAT=32-bit_rle_rgbe") valid = true;
    }
    if (!valid) throw std::runtime_error("unsupported HDR format");
    {
        const std::string s = line();
        char a[8] = {0}, b[8] = {0};
        int h = 0, w = 0;
        if (std::sscanf(s.c_str(), "%3s %d %3s %d", a, &h, b, &w) != 4 || std::strcmp(a, "-Y") != 0 || std::strcmp(b, "+X") != 0 || w <= 0 || h <= 0)
            throw std::runtime_error("unsupported HDR data layout");
        width = w; height = h;
    }
    const size_t W = (size_t)width, H = (size_t)height;
    std::vector<unsigned char> img(W * H * 4);
    auto flat = [&]() {
        if (pos + W * H * 4 > data.size()) throw std::runtime_error("corrupt HDR: truncated pixel data");
        std::memcpy(img.data(), data.data() + pos, W * H * 4);
    };
    if (width < 8 || width >= 32768) { flat(); return img; }   // flat data
    for (size_t y = 0; y < H; y++) {
        if (pos + 4 > data.size()) throw std::runtime_error("corrupt HDR: truncated scanline");
        const int c1 = data[pos], c2 = data[pos + 1], len = data[pos + 2];
        if (c1 != 2 || c2 != 2 || (len & 0x80)) {
            // not run-length encoded: the whole remaining image is flat (stb's fallback on the first scanline)
            if (y != 0) throw std::runtime_error("corrupt HDR: mixed flat and RLE scanlines");
            flat();
            return img;
        }
        if (((len << 8) | data[pos + 3]) != width) throw std::runtime_error("corrupt HDR: invalid decoded scanline length");
        pos += 4;
        for (int k = 0; k < 4; k++) {
            size_t i = 0;
            while (i < W) {
                if (pos >= data.size()) throw std::runtime_error("corrupt HDR: truncated run");
                size_t count = data[pos++];
                if (count > 128) {   // a run of one value
                    count -= 128;
                    if (count == 0 || i + count > W || pos >= data.size()) throw std::runtime_error("corrupt HDR run");
                    const unsigned char v = data[pos++];
                    for (size_t j = 0; j < count; j++) img[(y * W + i + j) * 4 + (size_t)k] = v;
                } else {             // literal bytes
                    if (count == 0 || i + count > W || pos + count > data.size()) throw std::runtime_error("corrupt HDR run");
                    for (size_t j = 0; j < count; j++) img[(y * W + i + j) * 4 + (size_t)k] = data[pos + j];
                    pos += count;
                }
                i += count;
            }
        }
    }
    return img;
}

// ReadFileHdr4f: float RGBA [h][w][4], row 0 = BOTTOM row of the picture (the reference flips on load, src/read_file.cpp:83), alpha 1.
inline std::vector<float> ReadFileHdr4f(const std::string& fileName, int& width, int& height)
{
    std::ifstream f(fileName, std::ios::binary);
    if (!f) throw std::runtime_error("Failed to open file " + fileName);
    std::vector<unsigned char> data((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    const std::vector<unsigned char> rgbe = DecodeHdr(data, width, height);
    const size_t W = (size_t)width, H = (size_t)height;
    std::vector<float> out(W * H * 4);
    for (size_t y = 0; y < H; y++)
        for (size_t x = 0; x < W; x++) {
            const unsigned char* p = &rgbe[((H - 1 - y) * W + x) * 4];
            float* o = &out[(y * W + x) * 4];
            const float scale = p[3] ? std::ldexp(1.0f, (int)p[3] - 136) : 0.0f;   // stbi__hdr_convert
            o[0] = (float)p[0] * scale; o[1] = (float)p[1] * scale; o[2] = (float)p[2] * scale; o[3] = 1.0f;
        }
    return out;
}

// ---- frame writers -------------------------------------------------------------------------------------------------------------------
// BGRA8 frame (VK_FORMAT_B8G8R8A8_UNORM, what NrcHpmRenderer::ReadImage returns) -> binary PPM
inline void WritePpm(const std::string& fileName, const unsigned char* bgra, uint32_t width, uint32_t height)
{
    std::ofstream f(fileName, std::ios::binary);
    if (!f) throw std::runtime_error("Failed to open file " + fileName);
    f << "P6\n" << width << " " << height << "\n255\n";
    std::vector<char> row((size_t)width * 3);
    for (uint32_t y = 0; y < height; y++) {
        for (uint32_t x = 0; x < width; x++) {
            const unsigned char* p = bgra + ((size_t)y * width + x) * 4;
            row[3 * x] = (char)p[2]; row[3 * x + 1] = (char)p[1]; row[3 * x + 2] = (char)p[0];
        }
        f.write(row.data(), (std::streamsize)row.size());
    }
}

// RGBA32F frame (vec4 outColor of nrc-forward.frag) -> colour PFM: little-endian floats, rows bottom-up as the format wants
inline void WritePfm(const std::string& fileName, const float* rgba, uint32_t width, uint32_t height)
{
    std::ofstream f(fileName, std::ios::binary);
    if (!f) throw std::runtime_error("Failed to open file " + fileName);
    f << "PF\n" << width << " " << height << "\n-1.0\n";
    std::vector<float> row((size_t)width * 3);
    for (uint32_t y = 0; y < height; y++) {
        const float* src = rgba + (size_t)(height - 1 - y) * width * 4;
        for (uint32_t x = 0; x < width; x++) { row[3 * x] = src[4 * x]; row[3 * x + 1] = src[4 * x + 1]; row[3 * x + 2] = src[4 * x + 2]; }
        f.write(reinterpret_cast<const char*>(row.data()), (std::streamsize)(row.size() * sizeof(float)));
    }
}

// ---- NrcConfig: every literal of RunNrcHpm in one place -------------------------------------------------------------------------------
struct NrcConfig {
    // frame and train grid (src/main.cpp:112-113, 161; the dispatch hard-codes 100 x 100, NrcHpmRenderer.cpp:604)
    uint32_t width = 800, height = 800, trainWidth = 100, trainHeight = 100;
    int frames = 100;
    // data (src/main.cpp:123, 132): empty path = the deterministic stand-in (the blobs are missing from the reference, .MISSING_LARGE_BLOBS)
    std::string density;                               // raw float32 grid
    uint32_t gridX = 125, gridY = 85, gridZ = 153;
    std::string envmap;                                // Radiance .hdr
    // volume (src/VolumeData.cpp:82-89)
    uint32_t useNN = 1, showNonNN = 0;
    float densityFactor = 0.4f, g = 0.7f;
    int noNnSpp = 1, withNnSpp = 1;
    uint32_t seed = 1;                                 // VolumeData::Update's per-frame `random` vec4: a seeded LCG instead of glm::linearRand
    // lights (src/main.cpp:151-152 build both with strength 0, which makes the training target identically zero: 1 here)
    float dirZenith = -1.57f, dirAzimuth = 0.0f, dirStrength = 1.0f;
    float pointPos[3] = {0.0f, 0.0f, 0.0f}, pointColor[3] = {1.0f, 1.0f, 1.0f}, pointStrength = 1.0f;
    float envDirect = 1.0f, envHpm = 8.0f;             // src/HdrEnvMap.cpp:108
    // camera (src/main.cpp:142-149) and the "camera training" orbit (main.cpp:200-203)
    float camPos[3] = {0.0f, 0.0f, -64.0f}, camDir[3] = {0.0f, 0.0f, 1.0f}, fovDeg = 60.0f, nearPlane = 0.1f, farPlane = 100.0f;
    float orbit = 0.01f;                               // radians per frame about +y; 0 = fixed camera
    // NeuralRadianceCache(lr, wd, beta1) and MRHE(lr, wd) (src/main.cpp:156-157; src/MRHE.cpp:83-87)
    float lr = 0.001f, weightDecay = 0.0f, beta1 = 0.5f;
    float mrheLr = 0.01f, mrheWeightDecay = 0.0f;
    uint32_t levels = 16, hashTableSize = 16384, minRes = 16, maxRes = 512;
    std::string resolutions = "intended";              // "intended" | "as_shipped_std140" 
    std::string precision = "bf16";                    // "bf16" (tcgen05) | "fp32" (1e-3 parity path)
    bool fixedPreset = false, adam = false;            // opt-in, not the shipped behaviour
    // host
    int device = 0, gpus = 1;                          // gpus > 1: rows / samples of rank `rank` (one process per GPU)
    int rank = 0;
    int statsEvery = 25, weightsEvery = 0;             // src/main.cpp:220-224, 183-187
    std::string out = "frame.ppm", outPfm, saveCheckpoint, loadCheckpoint;
};

namespace detail {
inline bool ParseFloats(const std::string& v, float* dst, int n)
{
    std::stringstream ss(v);
    std::string tok;
    for (int i = 0; i < n; i++) {
        if (!std::getline(ss, tok, ',')) return false;
        char* end = nullptr;
        dst[i] = std::strtof(tok.c_str(), &end);
        if (end == tok.c_str() || *end != '\0') return false;
    }
    return !std::getline(ss, tok, ',');
}
template <class T> inline bool ParseNumber(const std::string& v, T& dst)
{
    char* end = nullptr;
    const double x = std::strtod(v.c_str(), &end);
    if (end == v.c_str() || *end != '\0') return false;
    dst = (T)x;
    return true;
}
inline bool ParseBool(const std::string& v, bool& dst)
{
    if (v == "1" || v == "true" || v == "on" || v.empty()) { dst = true; return true; }
    if (v == "0" || v == "false" || v == "off") { dst = false; return true; }
    return false;
}
}   // namespace detail

inline const char* NrcConfigUsage()
{
    return "options (--key=value; every default is the reference's literal unless noted):\n"
           "  --width --height --train-width --train-height --frames\n"
           "  --density=<raw float32 file> --grid=X,Y,Z   --envmap=<file.hdr>\n"
           "  --use-nn --show-non-nn --density-factor --g --spp (sets both spp values) --no-nn-spp --with-nn-spp --seed\n"
           "  --dir-zenith --dir-azimuth --dir-strength --point-pos=x,y,z --point-color=r,g,b --point-strength --env-direct --env-hpm\n"
           "  --cam-pos=x,y,z --cam-dir=x,y,z --fov (degrees) --near --far --orbit (radians per frame)\n"
           "  --lr --weight-decay --beta1 --mrhe-lr --mrhe-weight-decay --levels --table-size (T) --min-res --max-res\n"
           "  --resolutions=intended|as_shipped_std140 --precision=bf16|fp32 --fixed --adam\n"
           "  --device --gpus --rank --stats-every --weights-every --out=<ppm> --out-pfm=<pfm> --save-checkpoint --load-checkpoint\n";
}

// Parses --key=value (and bare --flag) arguments into cfg; throws std::runtime_error naming the offending argument.
inline NrcConfig ParseNrcConfig(int argc, const char* const* argv, NrcConfig cfg = NrcConfig())
{
    using namespace detail;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        if (a.rfind("--", 0) != 0) throw std::runtime_error("unexpected argument '" + a + "'\n" + NrcConfigUsage());
        const size_t eq = a.find('=');
        const std::string k = a.substr(2, eq == std::string::npos ? std::string::npos : eq - 2);
        const std::string v = eq == std::string::npos ? std::string() : a.substr(eq + 1);
        bool ok = true, flag = false;
#define NRC_NUM(name, field) else if (k == name) ok = ParseNumber(v, cfg.field)
#define NRC_STR(name, field) else if (k == name) cfg.field = v
        if (k == "help") throw std::runtime_error(NrcConfigUsage());
        NRC_NUM("width", width); NRC_NUM("height", height); NRC_NUM("train-width", trainWidth); NRC_NUM("train-height", trainHeight);
        NRC_NUM("frames", frames);
        NRC_STR("density", density); NRC_STR("envmap", envmap);
        else if (k == "grid") { float t[3]; ok = ParseFloats(v, t, 3); if (ok) { cfg.gridX = (uint32_t)t[0]; cfg.gridY = (uint32_t)t[1]; cfg.gridZ = (uint32_t)t[2]; } }
        NRC_NUM("use-nn", useNN); NRC_NUM("show-non-nn", showNonNN); NRC_NUM("density-factor", densityFactor); NRC_NUM("g", g);
        NRC_NUM("no-nn-spp", noNnSpp); NRC_NUM("with-nn-spp", withNnSpp); NRC_NUM("seed", seed);
        else if (k == "spp") { ok = ParseNumber(v, cfg.noNnSpp); cfg.withNnSpp = cfg.noNnSpp; }
        NRC_NUM("dir-zenith", dirZenith); NRC_NUM("dir-azimuth", dirAzimuth); NRC_NUM("dir-strength", dirStrength);
        else if (k == "point-pos") ok = ParseFloats(v, cfg.pointPos, 3);
        else if (k == "point-color") ok = ParseFloats(v, cfg.pointColor, 3);
        NRC_NUM("point-strength", pointStrength); NRC_NUM("env-direct", envDirect); NRC_NUM("env-hpm", envHpm);
        else if (k == "cam-pos") ok = ParseFloats(v, cfg.camPos, 3);
        else if (k == "cam-dir") ok = ParseFloats(v, cfg.camDir, 3);
        NRC_NUM("fov", fovDeg); NRC_NUM("near", nearPlane); NRC_NUM("far", farPlane); NRC_NUM("orbit", orbit);
        NRC_NUM("lr", lr); NRC_NUM("weight-decay", weightDecay); NRC_NUM("beta1", beta1);
        NRC_NUM("mrhe-lr", mrheLr); NRC_NUM("mrhe-weight-decay", mrheWeightDecay);
        NRC_NUM("levels", levels); NRC_NUM("table-size", hashTableSize); NRC_NUM("min-res", minRes); NRC_NUM("max-res", maxRes);
        NRC_STR("resolutions", resolutions); NRC_STR("precision", precision);
        else if (k == "fixed") { flag = true; ok = ParseBool(v, cfg.fixedPreset); }
        else if (k == "adam") { flag = true; ok = ParseBool(v, cfg.adam); }
        NRC_NUM("device", device); NRC_NUM("gpus", gpus); NRC_NUM("rank", rank);
        NRC_NUM("stats-every", statsEvery); NRC_NUM("weights-every", weightsEvery);
        NRC_STR("out", out); NRC_STR("out-pfm", outPfm); NRC_STR("save-checkpoint", saveCheckpoint); NRC_STR("load-checkpoint", loadCheckpoint);
        else throw std::runtime_error("unknown option '--" + k + "'\n" + NrcConfigUsage());
#undef NRC_NUM
#undef NRC_STR
        if (!ok || (eq == std::string::npos && !flag)) throw std::runtime_error("bad value for '--" + k + "': '" + v + "'");
    }
    if (cfg.width == 0 || cfg.height == 0 || cfg.trainWidth == 0 || cfg.trainHeight == 0) throw std::runtime_error("empty frame or train grid");
    if (cfg.levels < 1 || cfg.levels > 16) throw std::runtime_error("--levels must be 1..16");
    if (cfg.resolutions != "intended" && cfg.resolutions != "as_shipped_std140") throw std::runtime_error("--resolutions must be intended or as_shipped_std140");
    if (cfg.precision != "bf16" && cfg.precision != "fp32") throw std::runtime_error("--precision must be bf16 or fp32");
    if (cfg.gpus < 1 || cfg.rank < 0 || cfg.rank >= cfg.gpus) throw std::runtime_error("--rank must be in [0, --gpus)");
    return cfg;
}

// Deterministic stand-ins for the blobs the reference does not ship (.MISSING_LARGE_BLOBS): a lumpy ellipsoid and a gradient sky.
inline std::vector<float> MakeStandInDensity(uint32_t nx, uint32_t ny, uint32_t nz)
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

inline std::vector<float> MakeStandInSky(uint32_t w, uint32_t h)
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

}   // namespace en
