// nrc_b200.hpp -- header-only C++ facade over the C-ABI (nrc_b200.h) that keeps the reference's class surface
// (namespace en, same constructor arguments and method names, Vulkan handle parameters dropped), so a headless
// main.cpp with the body shape of RunNrcHpm (src/main.cpp:109-277 of the reference) compiles against it:
//
//   en::NeuralRadianceCache  include/engine/graphics/NeuralRadianceCache.hpp:16-30
//   en::MRHE                 include/engine/graphics/MRHE.hpp:14-22
//   en::NrcHpmRenderer       include/engine/graphics/renderer/NrcHpmRenderer.hpp:20-40
//   en::Camera               include/engine/graphics/Camera.hpp:24   (glm::perspective/lookAt/inverse restated, Camera.cpp:164-179)
//   en::DirLight/PointLight  include/engine/graphics/DirLight.hpp:26, PointLight.hpp:16
//   en::VolumeData           include/engine/objects/VolumeData.hpp:22  (+ vk::Texture3D, src/Texture3D.cpp:10-43)
//   en::HdrEnvMap            include/engine/graphics/HdrEnvMap.hpp:15-20
//
// Errors: the reference's ASSERT_VULKAN throws std::runtime_error (src/Log.cpp:22-27); so does every failing call here.
#pragma once
#include <array>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>
#include "nrc_b200.h"

namespace en {

inline void Check(nrc_status s, const char* what)
{
    if (s != NRC_OK) throw std::runtime_error(std::string(what) + ": " + nrc_last_error());
}

// replaces VulkanAPI::Init/Shutdown (src/VulkanAPI.cpp:35-65)
class Context {
public:
    explicit Context(int device = 0) { Check(nrc_ctx_create(device, &m_Handle), "nrc_ctx_create"); }
    ~Context() { nrc_ctx_destroy(m_Handle); }
    Context(const Context&) = delete;
    Context& operator=(const Context&) = delete;
    nrc_ctx* Get() const { return m_Handle; }
    void WaitIdle() const { Check(nrc_sync(m_Handle), "nrc_sync"); }   // vkQueueWaitIdle, src/main.cpp:217
    void SetPrecision(nrc_mlp_precision p) const { Check(nrc_ctx_set_precision(m_Handle, p), "nrc_ctx_set_precision"); }
private:
    nrc_ctx* m_Handle = nullptr;
};

// ---- minimal column-major math (what the reference takes from glm) ---------------------------------------------
struct vec3 { float x, y, z; };
inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline vec3 cross(vec3 a, vec3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline vec3 normalize(vec3 a) { return a * (1.0f / std::sqrt(dot(a, a))); }

struct mat4 {
    float m[16];   // m[col*4 + row]
    static mat4 identity() { mat4 r{}; r.m[0] = r.m[5] = r.m[10] = r.m[15] = 1.0f; return r; }
};
inline mat4 operator*(const mat4& a, const mat4& b)
{
    mat4 r{};
    for (int c = 0; c < 4; c++)
        for (int row = 0; row < 4; row++) {
            float s = 0.0f;
            for (int k = 0; k < 4; k++) s += a.m[k * 4 + row] * b.m[c * 4 + k];
            r.m[c * 4 + row] = s;
        }
    return r;
}
inline mat4 perspective(float fovy, float aspect, float zNear, float zFar)   // glm::perspective, RH, clip z in [-1,1]
{
    float t = std::tan(fovy / 2.0f);
    mat4 r{};
    r.m[0] = 1.0f / (aspect * t);
    r.m[5] = 1.0f / t;
    r.m[10] = -(zFar + zNear) / (zFar - zNear);
    r.m[11] = -1.0f;
    r.m[14] = -(2.0f * zFar * zNear) / (zFar - zNear);
    return r;
}
inline mat4 lookAt(vec3 eye, vec3 center, vec3 up)   // glm::lookAt, RH
{
    vec3 f = normalize(center - eye), s = normalize(cross(f, up)), u = cross(s, f);
    mat4 r = mat4::identity();
    r.m[0] = s.x; r.m[4] = s.y; r.m[8] = s.z;
    r.m[1] = u.x; r.m[5] = u.y; r.m[9] = u.z;
    r.m[2] = -f.x; r.m[6] = -f.y; r.m[10] = -f.z;
    r.m[12] = -dot(s, eye); r.m[13] = -dot(u, eye); r.m[14] = dot(f, eye);
    return r;
}
inline mat4 rotate(float angle, vec3 axis)   // glm::rotate(angle, axis)
{
    float c = std::cos(angle), s = std::sin(angle);
    vec3 a = normalize(axis), t = a * (1.0f - c);
    mat4 r = mat4::identity();
    r.m[0] = c + t.x * a.x; r.m[1] = t.x * a.y + s * a.z; r.m[2] = t.x * a.z - s * a.y;
    r.m[4] = t.y * a.x - s * a.z; r.m[5] = c + t.y * a.y; r.m[6] = t.y * a.z + s * a.x;
    r.m[8] = t.z * a.x + s * a.y; r.m[9] = t.z * a.y - s * a.x; r.m[10] = c + t.z * a.z;
    return r;
}
inline mat4 inverse(const mat4& a)   // Gauss-Jordan in double, rounded to float (glm::inverse to ~1 ulp)
{
    double w[4][8];
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) { w[r][c] = a.m[c * 4 + r]; w[r][4 + c] = (r == c) ? 1.0 : 0.0; }
    for (int i = 0; i < 4; i++) {
        int p = i;
        for (int r = i + 1; r < 4; r++) if (std::fabs(w[r][i]) > std::fabs(w[p][i])) p = r;
        for (int c = 0; c < 8; c++) std::swap(w[i][c], w[p][c]);
        double d = w[i][i];
        for (int c = 0; c < 8; c++) w[i][c] /= d;
        for (int r = 0; r < 4; r++)
            if (r != i) { double f = w[r][i]; for (int c = 0; c < 8; c++) w[r][c] -= f * w[i][c]; }
    }
    mat4 out{};
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) out.m[c * 4 + r] = (float)w[r][4 + c];
    return out;
}
inline vec3 transformPoint(const mat4& m, vec3 v)
{
    return {m.m[0] * v.x + m.m[4] * v.y + m.m[8] * v.z + m.m[12], m.m[1] * v.x + m.m[5] * v.y + m.m[9] * v.z + m.m[13],
            m.m[2] * v.x + m.m[6] * v.y + m.m[10] * v.z + m.m[14]};
}

// ---- scene objects ---------------------------------------------------------------------------------------------
class Camera {   // en::Camera(pos, viewDir, up, aspectRatio, fov, nearPlane, farPlane), Camera.hpp:24
public:
    Camera(vec3 pos, vec3 viewDir, vec3 up, float aspectRatio, float fov, float nearPlane, float farPlane)
        : m_Pos(pos), m_ViewDir(viewDir), m_Up(up), m_AspectRatio(aspectRatio), m_Fov(fov), m_Near(nearPlane), m_Far(farPlane)
    {
        UpdateUniformBuffer();
    }
    void UpdateUniformBuffer()   // Camera.cpp:164-173
    {
        mat4 pv = perspective(m_Fov, m_AspectRatio, m_Near, m_Far) * lookAt(m_Pos, m_Pos + m_ViewDir, m_Up);
        mat4 inv = inverse(pv);
        for (int i = 0; i < 16; i++) { m_Data.projView[i] = pv.m[i]; m_Data.invProjView[i] = inv.m[i]; }
        m_Data.pos[0] = m_Pos.x; m_Data.pos[1] = m_Pos.y; m_Data.pos[2] = m_Pos.z;
    }
    void RotateAroundOrigin(vec3 axis, float angle)   // Camera.cpp:175-179 ("camera training" orbit, main.cpp:200-203)
    {
        m_Pos = transformPoint(rotate(angle, axis), m_Pos);
        m_ViewDir = normalize(m_Pos) * -1.0f;
    }
    void SetAspectRatio(uint32_t width, uint32_t height) { m_AspectRatio = (float)width / (float)height; }
    vec3 GetPos() const { return m_Pos; }
    const nrc_camera& GetData() const { return m_Data; }
private:
    vec3 m_Pos, m_ViewDir, m_Up;
    float m_AspectRatio, m_Fov, m_Near, m_Far;
    nrc_camera m_Data{};
};

class DirLight {   // en::DirLight(zenith, azimuth, color, strength), DirLight.hpp:26
public:
    DirLight(float zenith, float azimuth, vec3 color, float strength)
    {
        m_Data.zenith = zenith; m_Data.azimuth = azimuth; m_Data.strength = strength;
        m_Data.color[0] = color.x; m_Data.color[1] = color.y; m_Data.color[2] = color.z;
        UpdateDir();
    }
    void SetZenith(float z) { m_Data.zenith = z; UpdateDir(); }
    void SetAzimuth(float a) { m_Data.azimuth = a; UpdateDir(); }
    void SetStrength(float s) { m_Data.strength = s; }
    const nrc_dir_light& GetData() const { return m_Data; }
private:
    void UpdateDir()   // VecFromAngles, src/DirLight.cpp:5-14
    {
        vec3 d = transformPoint(rotate(m_Data.azimuth, {0, 1, 0}) * rotate(m_Data.zenith, {1, 0, 0}), {0, 1, 0});
        m_Data.dir[0] = d.x; m_Data.dir[1] = d.y; m_Data.dir[2] = d.z;
    }
    nrc_dir_light m_Data{};
};

class PointLight {   // en::PointLight(pos, color, strength), PointLight.hpp:16
public:
    PointLight(vec3 pos, vec3 color, float strength)
    {
        m_Data.pos[0] = pos.x; m_Data.pos[1] = pos.y; m_Data.pos[2] = pos.z;
        m_Data.color[0] = color.x; m_Data.color[1] = color.y; m_Data.color[2] = color.z;
        m_Data.strength = strength;
    }
    void SetStrength(float s) { m_Data.strength = s; }
    const nrc_point_light& GetData() const { return m_Data; }
private:
    nrc_point_light m_Data{};
};

// vk::Texture3D(density3D, LINEAR, CLAMP_TO_BORDER, OPAQUE_BLACK) + en::VolumeData(&tex), src/main.cpp:123-129.
// density3D[x][y][z] as ReadFileDensity3D returns it (src/read_file.cpp:50-78).
class VolumeData {
public:
    VolumeData(const Context& ctx, const std::vector<std::vector<std::vector<float>>>& density3D)
    {
        uint32_t nx = (uint32_t)density3D.size(), ny = (uint32_t)density3D[0].size(), nz = (uint32_t)density3D[0][0].size();
        std::vector<float> flat((size_t)nx * ny * nz);
        for (uint32_t x = 0; x < nx; x++)
            for (uint32_t y = 0; y < ny; y++)
                for (uint32_t z = 0; z < nz; z++) flat[((size_t)x * ny + y) * nz + z] = density3D[x][y][z];
        Create(ctx, flat.data(), nx, ny, nz);
    }
    VolumeData(const Context& ctx, const float* densityXYZ, uint32_t nx, uint32_t ny, uint32_t nz) { Create(ctx, densityXYZ, nx, ny, nz); }
    // VolumeData::Update (VolumeData.cpp:105-109): a new `random` vec4 per frame; the reference draws it from an unseeded
    // glm::linearRand, here it is an explicit input or a seeded LCG
    void Update(bool /*cameraChanged*/)
    {
        for (int i = 0; i < 4; i++) {
            m_Lcg = m_Lcg * 1664525u + 1013904223u;
            m_Params.random[i] = (float)(m_Lcg >> 8) / 16777216.0f;
        }
        Push();
    }
    void Update(const float random4[4]) { for (int i = 0; i < 4; i++) m_Params.random[i] = random4[i]; Push(); }
    void SetSeed(uint32_t seed) { m_Lcg = 0x2545F491u ^ (seed * 2654435761u); }   // of the LCG behind Update(bool)
    nrc_volume_params& Params() { return m_Params; }   // useNN, showNonNN, densityFactor, g, spp: the ImGui sliders (VolumeData.cpp:116-128)
    void Push() const { Check(nrc_volume_set_params(m_Handle, &m_Params), "nrc_volume_set_params"); }
    nrc_volume* Get() const { return m_Handle; }
    void Destroy() { nrc_volume_destroy(m_Handle); m_Handle = nullptr; }
private:
    void Create(const Context& ctx, const float* d, uint32_t nx, uint32_t ny, uint32_t nz)
    {
        Check(nrc_volume_create(ctx.Get(), d, nx, ny, nz, &m_Handle), "nrc_volume_create");
        m_Params = nrc_volume_params{{0, 0, 0, 0}, 0u, 0u, 0.4f, 0.7f, 1, 1};   // VolumeData.cpp:82-89
    }
    nrc_volume* m_Handle = nullptr;
    nrc_volume_params m_Params{};
    uint32_t m_Lcg = 0x2545F491u;
};

class HdrEnvMap {   // en::HdrEnvMap(width, height, hdr4f, cdfX, cdfY), HdrEnvMap.hpp:15-20
public:
    HdrEnvMap(const Context& ctx, uint32_t width, uint32_t height, const std::vector<float>& hdr4f, const std::vector<float>& hdrCdfX,
              const std::vector<float>& hdrCdfY)
    {
        Check(nrc_envmap_create(ctx.Get(), width, height, hdr4f.data(), hdrCdfX.empty() ? nullptr : hdrCdfX.data(),
                                hdrCdfY.empty() ? nullptr : hdrCdfY.data(), &m_Handle), "nrc_envmap_create");
    }
    void SetStrengths(float directStrength, float hpmStrength) { Check(nrc_envmap_set_strengths(m_Handle, directStrength, hpmStrength), "nrc_envmap_set_strengths"); }
    void SetImportanceSampling(bool on) { Check(nrc_envmap_set_importance_sampling(m_Handle, on ? 1 : 0), "nrc_envmap_set_importance_sampling"); }   // opt-in, fwd.frag:715
    nrc_envmap* Get() const { return m_Handle; }
    void Destroy() { nrc_envmap_destroy(m_Handle); m_Handle = nullptr; }
private:
    nrc_envmap* m_Handle = nullptr;
};

// en::Hdr4fToCdf (src/read_file.cpp:141-206): {invCdfX (h*w), invCdfY (h)}
inline std::array<std::vector<float>, 2> Hdr4fToCdf(const Context& ctx, const std::vector<float>& hdr4f, uint32_t width, uint32_t height)
{
    std::array<std::vector<float>, 2> out{std::vector<float>((size_t)width * height), std::vector<float>(height)};
    Check(nrc_hdr4f_to_cdf(ctx.Get(), hdr4f.data(), width, height, out[0].data(), out[1].data()), "nrc_hdr4f_to_cdf");
    return out;
}

// ---- NN state ----------------------------------------------------------------------------------------------------
class NeuralRadianceCache {   // NeuralRadianceCache.hpp:8-61
public:
    struct StatsData { float mseLoss; };   // NeuralRadianceCache.hpp:11-14
    // seed replaces std::random_device (src/NeuralRadianceCache.cpp:629): W ~ N(0,1)*0.01, biases 0
    NeuralRadianceCache(const Context& ctx, float learningRate, float weightDecay, float beta1, uint64_t seed = 1)
    {
        Check(nrc_mlp_create(ctx.Get(), learningRate, weightDecay, beta1, seed, &m_Handle), "nrc_mlp_create");
    }
    // opt-in fixed preset (no reference counterpart): second moment + bias correction in nrc-step
    void SetAdam(bool on, float beta2 = 0.999f, float eps = 1e-8f) const { Check(nrc_mlp_set_adam(m_Handle, on ? 1 : 0, beta2, eps), "nrc_mlp_set_adam"); }
    void Destroy() { nrc_mlp_destroy(m_Handle); m_Handle = nullptr; }
    void ResetStats() { Check(nrc_mlp_reset_stats(m_Handle), "nrc_mlp_reset_stats"); }
    const StatsData& GetStats() { Check(nrc_mlp_get_stats(m_Handle, &m_Stats.mseLoss), "nrc_mlp_get_stats"); return m_Stats; }
    void PrintWeights() const   // NeuralRadianceCache.cpp:558-594
    {
        std::vector<float> p(NRC_MLP_PARAM_COUNT);
        Check(nrc_mlp_get_params(m_Handle, p.data(), nullptr, nullptr), "nrc_mlp_get_params");
        for (int l = 0; l < 6; l++) {
            std::printf("Weights %d: [", l);
            int n = l < 5 ? 4096 : 192;
            for (int i = 0; i < n; i++) std::printf("%s%f", i ? ", " : "", p[(size_t)l * 4096 + i]);
            std::printf("]\n");
        }
    }
    // checkpoint surface (the reference has none)
    void SetParams(const std::vector<float>& params) { Check(nrc_mlp_set_params(m_Handle, params.data()), "nrc_mlp_set_params"); }
    std::vector<float> GetParams() const
    {
        std::vector<float> p(NRC_MLP_PARAM_COUNT);
        Check(nrc_mlp_get_params(m_Handle, p.data(), nullptr, nullptr), "nrc_mlp_get_params");
        return p;
    }
    nrc_mlp* Get() const { return m_Handle; }
private:
    nrc_mlp* m_Handle = nullptr;
    StatsData m_Stats{0.0f};
};

class MRHE {   // MRHE.hpp:7-47; the constants of src/MRHE.cpp:83-87 are defaulted arguments here
public:
    MRHE(const Context& ctx, float learningRate, float weightDecay, uint32_t levelCount = 16, uint32_t hashTableSize = 16384,
         uint32_t featureCount = 2, uint32_t minRes = 16, uint32_t maxRes = 512, const uint32_t* resolutions = nullptr, uint64_t seed = 2)
    {
        Check(nrc_mrhe_create(ctx.Get(), learningRate, weightDecay, levelCount, hashTableSize, featureCount, minRes, maxRes, resolutions, seed,
                              &m_Handle), "nrc_mrhe_create");
        m_Shape[0] = levelCount; m_Shape[1] = hashTableSize; m_Shape[2] = featureCount;
    }
    const uint32_t* GetShape() const { return m_Shape; }   // {levels, entries per level, features}
    // opt-in fixed preset (no reference counterpart): corrected interpolation / gradient scatter, Adam in mrhe-step
    void SetFixedMode(bool on) const { Check(nrc_mrhe_set_fixed_mode(m_Handle, on ? 1 : 0), "nrc_mrhe_set_fixed_mode"); }
    void SetAdam(bool on, float beta1 = 0.9f, float beta2 = 0.999f, float eps = 1e-8f) const
    {
        Check(nrc_mrhe_set_adam(m_Handle, on ? 1 : 0, beta1, beta2, eps), "nrc_mrhe_set_adam");
    }
    void Destroy() { nrc_mrhe_destroy(m_Handle); m_Handle = nullptr; }
    size_t GetHashTableSize() const { return nrc_mrhe_table_bytes(m_Handle); }   // bytes, MRHE.cpp:267-270
    void PrintHashTables() const   // MRHE.cpp:234-260
    {
        std::vector<float> t(GetHashTableSize() / sizeof(float));
        Check(nrc_mrhe_get_tables(m_Handle, t.data(), nullptr), "nrc_mrhe_get_tables");
        std::printf("HashTables: [");
        for (size_t i = 0; i < t.size() && i < 256; i++) std::printf("%s%f", i ? ", " : "", t[i]);
        std::printf("]\n");
    }
    nrc_mrhe* Get() const { return m_Handle; }
private:
    nrc_mrhe* m_Handle = nullptr;
    uint32_t m_Shape[3] = {0, 0, 0};
};

// ---- NrcHpmRenderer (NrcHpmRenderer.hpp:17-96) --------------------------------------------------------------------------
class NrcHpmRenderer {
public:
    NrcHpmRenderer(const Context& ctx, uint32_t width, uint32_t height, uint32_t trainWidth, uint32_t trainHeight, const Camera& camera,
                   const VolumeData& volumeData, const DirLight& dirLight, const PointLight& pointLight, const HdrEnvMap& hdrEnvMap,
                   const NeuralRadianceCache& nrc, const MRHE& mrhe)
        : m_Width(width), m_Height(height), m_Camera(camera), m_VolumeData(volumeData), m_DirLight(dirLight), m_PointLight(pointLight),
          m_HdrEnvMap(hdrEnvMap), m_Nrc(nrc), m_Mrhe(mrhe)
    {
        Check(nrc_renderer_create(ctx.Get(), width, height, trainWidth, trainHeight, volumeData.Get(), hdrEnvMap.Get(), nrc.Get(), mrhe.Get(),
                                  &m_Handle), "nrc_renderer_create");
    }
    // Render(queue): one submit of train -> nrc-step -> mrhe-step -> render (NrcHpmRenderer.cpp:57-72); asynchronous.
    // Like the reference's descriptor sets, the referenced objects' CURRENT state is what the frame sees.
    void Render() const
    {
        Check(nrc_renderer_set_camera(m_Handle, &m_Camera.GetData()), "nrc_renderer_set_camera");
        Check(nrc_renderer_set_dir_light(m_Handle, &m_DirLight.GetData()), "nrc_renderer_set_dir_light");
        Check(nrc_renderer_set_point_light(m_Handle, &m_PointLight.GetData()), "nrc_renderer_set_point_light");
        m_VolumeData.Push();
        Check(nrc_render(m_Handle), "nrc_render");
    }
    // The same frame in its two halves, for data-parallel ranks: FrameTrain(); DataParallel::ReduceStep(); FrameRender().
    // SetPartition gives this rank rows [y0,y1) of the frame and train invocations [i0,i1).
    void SetPartition(uint32_t y0, uint32_t y1, uint32_t i0, uint32_t i1) { Check(nrc_renderer_set_partition(m_Handle, y0, y1, i0, i1), "nrc_renderer_set_partition"); }
    // interleaved partition (rows yFirst, yFirst + yStride, ...; train invocations likewise): the load-balanced split data-parallel hosts use
    void SetPartitionStrided(uint32_t yFirst, uint32_t yStride, uint32_t iFirst, uint32_t iStride)
    {
        Check(nrc_renderer_set_partition_strided(m_Handle, yFirst, yStride, iFirst, iStride), "nrc_renderer_set_partition_strided");
    }
    void FrameTrain() const
    {
        Check(nrc_renderer_set_camera(m_Handle, &m_Camera.GetData()), "nrc_renderer_set_camera");
        Check(nrc_renderer_set_dir_light(m_Handle, &m_DirLight.GetData()), "nrc_renderer_set_dir_light");
        Check(nrc_renderer_set_point_light(m_Handle, &m_PointLight.GetData()), "nrc_renderer_set_point_light");
        m_VolumeData.Push();
        Check(nrc_frame_train(m_Handle), "nrc_frame_train");
    }
    void FrameRender() const { Check(nrc_frame_render(m_Handle), "nrc_frame_render"); }
    void ResizeFrame(uint32_t width, uint32_t height)   // NrcHpmRenderer.cpp:103-126
    {
        Check(nrc_resize(m_Handle, width, height), "nrc_resize");
        m_Width = width; m_Height = height;
    }
    void* GetImage() const { return nrc_frame_ptr(m_Handle, NRC_FRAME_BGRA8); }            // device address (stands in for the VkImage)
    size_t GetImageDataSize() const { return nrc_frame_bytes(m_Handle, NRC_FRAME_BGRA8); }  // width*height*4
    void ReadImage(void* dst, nrc_frame_format fmt = NRC_FRAME_BGRA8) const { Check(nrc_read_frame(m_Handle, dst, fmt), "nrc_read_frame"); }
    // Presenting frame k while frame k+1 is recorded (src/main.cpp:197-230): queue the copy of this rank's rows into page-locked
    // host memory, keep calling Render(), collect with WaitReadback(ticket).
    int ReadImageAsync(void* dst, nrc_frame_format fmt = NRC_FRAME_BGRA8) const
    {
        int ticket = -1;
        Check(nrc_read_frame_partition_async(m_Handle, dst, fmt, &ticket), "nrc_read_frame_partition_async");
        return ticket;
    }
    void WaitReadback(int ticket) const { Check(nrc_readback_wait(m_Handle, ticket), "nrc_readback_wait"); }
    // initial capacity of the render vertex pool (0 = default); an overflowing frame is traced again before it is handed out
    void SetTrainEnvLight(bool on) { Check(nrc_renderer_set_train_env_light(m_Handle, on ? 1 : 0), "nrc_renderer_set_train_env_light"); }   // opt-in, train.comp:1118
    void SetPoolCapacity(uint64_t vertices) { Check(nrc_renderer_set_pool_capacity(m_Handle, vertices), "nrc_renderer_set_pool_capacity"); }
    uint64_t GetPoolCapacity() const { uint64_t v = 0; Check(nrc_renderer_get_pool_capacity(m_Handle, &v), "nrc_renderer_get_pool_capacity"); return v; }
    void Destroy() { nrc_renderer_destroy(m_Handle); m_Handle = nullptr; }
    nrc_renderer* Get() const { return m_Handle; }
private:
    nrc_renderer* m_Handle = nullptr;
    uint32_t m_Width, m_Height;
    const Camera& m_Camera;               // the reference keeps const references too (NrcHpmRenderer.hpp:49-55)
    const VolumeData& m_VolumeData;
    const DirLight& m_DirLight;
    const PointLight& m_PointLight;
    const HdrEnvMap& m_HdrEnvMap;
    const NeuralRadianceCache& m_Nrc;
    const MRHE& m_Mrhe;
};

// ---- data-parallel exchange fused with the optimiser steps (nrc_dp_*; the reference is single-GPU) ---------------------------------
// One process per GPU.  Construct on every rank, exchange Handle() bytes in rank order by any transport (MPI_Allgather, a file ...),
// Connect(), then one ReduceStep() per frame between the renderer's FrameTrain() and FrameRender().
class DataParallel {
public:
    DataParallel(NeuralRadianceCache& nrc, MRHE& mrhe, int rank, int world)
    {
        Check(nrc_dp_create(nrc.Get(), mrhe.Get(), rank, world, &m_Handle), "nrc_dp_create");
        m_Bytes.resize(nrc_dp_handle_bytes());
        Check(nrc_dp_export(m_Handle, m_Bytes.data()), "nrc_dp_export");
    }
    const std::vector<unsigned char>& Handle() const { return m_Bytes; }
    void Connect(const std::vector<unsigned char>& allHandlesInRankOrder)
    {
        Check(nrc_dp_connect(m_Handle, allHandlesInRankOrder.data(), 0), "nrc_dp_connect");
    }
    void ReduceStep() { Check(nrc_dp_reduce_step(m_Handle), "nrc_dp_reduce_step"); }
    bool TimedOut() const { int v = 0; Check(nrc_dp_status(m_Handle, &v), "nrc_dp_status"); return v != 0; }
    void Destroy() { nrc_dp_destroy(m_Handle); m_Handle = nullptr; }

private:
    nrc_dp* m_Handle = nullptr;
    std::vector<unsigned char> m_Bytes;
};

// ---- checkpoint of the learnt state (the reference has none).  Same file layout as
// neuralradiancecaching_b200/checkpoint.py: "NRCB200\0", u32 {version, nParams, nTables, L, T, F}, then float32 params, momentum, tables.
// Version 2 (written only when Adam of the fixed preset is on) appends u32 flags {1 MLP, 2 tables}, then per set bit u32 steps and
// the moments (MLP: second; tables: first, second).  LoadCheckpoint restores them into objects that have Adam switched on.
inline void SaveCheckpoint(const std::string& path, const NeuralRadianceCache& nrc, const MRHE& mrhe)
{
    const uint32_t* sh = mrhe.GetShape();
    std::vector<float> p(NRC_MLP_PARAM_COUNT), m(NRC_MLP_PARAM_COUNT), t((size_t)sh[0] * sh[1] * sh[2]);
    Check(nrc_mlp_get_params(nrc.Get(), p.data(), nullptr, m.data()), "nrc_mlp_get_params");
    Check(nrc_mrhe_get_tables(mrhe.Get(), t.data(), nullptr), "nrc_mrhe_get_tables");
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("SaveCheckpoint: cannot open " + path);
    uint32_t stepsMlp = 0, stepsTab = 0;
    const uint32_t flags = (nrc_mlp_get_adam_state(nrc.Get(), nullptr, &stepsMlp) == NRC_OK ? 1u : 0u) |
                           (nrc_mrhe_get_adam_state(mrhe.Get(), nullptr, nullptr, &stepsTab) == NRC_OK ? 2u : 0u);
    std::vector<float> second, m1, m2;
    if (flags & 1u) { second.resize(p.size()); Check(nrc_mlp_get_adam_state(nrc.Get(), second.data(), nullptr), "nrc_mlp_get_adam_state"); }
    if (flags & 2u) {
        m1.resize(t.size()); m2.resize(t.size());
        Check(nrc_mrhe_get_adam_state(mrhe.Get(), m1.data(), m2.data(), nullptr), "nrc_mrhe_get_adam_state");
    }
    const uint32_t head[6] = {flags ? 2u : 1u, (uint32_t)p.size(), (uint32_t)t.size(), sh[0], sh[1], sh[2]};
    bool ok = std::fwrite("NRCB200\0", 1, 8, f) == 8 && std::fwrite(head, 4, 6, f) == 6 && std::fwrite(p.data(), 4, p.size(), f) == p.size() &&
              std::fwrite(m.data(), 4, m.size(), f) == m.size() && std::fwrite(t.data(), 4, t.size(), f) == t.size();
    if (flags) ok = ok && std::fwrite(&flags, 4, 1, f) == 1;
    if (flags & 1u) ok = ok && std::fwrite(&stepsMlp, 4, 1, f) == 1 && std::fwrite(second.data(), 4, second.size(), f) == second.size();
    if (flags & 2u)
        ok = ok && std::fwrite(&stepsTab, 4, 1, f) == 1 && std::fwrite(m1.data(), 4, m1.size(), f) == m1.size() &&
             std::fwrite(m2.data(), 4, m2.size(), f) == m2.size();
    ok = (std::fclose(f) == 0) && ok;
    if (!ok) throw std::runtime_error("SaveCheckpoint: short write to " + path);
}

inline void LoadCheckpoint(const std::string& path, NeuralRadianceCache& nrc, MRHE& mrhe)
{
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("LoadCheckpoint: cannot open " + path);
    char magic[8];
    uint32_t head[6];
    bool ok = std::fread(magic, 1, 8, f) == 8 && std::memcmp(magic, "NRCB200\0", 8) == 0 && std::fread(head, 4, 6, f) == 6 && (head[0] == 1u || head[0] == 2u);
    const uint32_t* sh = mrhe.GetShape();
    ok = ok && head[1] == (uint32_t)NRC_MLP_PARAM_COUNT && head[3] == sh[0] && head[4] == sh[1] && head[5] == sh[2] &&
         head[2] == sh[0] * sh[1] * sh[2];
    std::vector<float> p, m, t;
    if (ok) {
        p.resize(head[1]); m.resize(head[1]); t.resize(head[2]);
        ok = std::fread(p.data(), 4, p.size(), f) == p.size() && std::fread(m.data(), 4, m.size(), f) == m.size() &&
             std::fread(t.data(), 4, t.size(), f) == t.size();
    }
    uint32_t flags = 0, stepsMlp = 0, stepsTab = 0;
    std::vector<float> second, m1, m2;
    if (ok && head[0] == 2u) {
        ok = std::fread(&flags, 4, 1, f) == 1;
        if (ok && (flags & 1u)) { second.resize(p.size()); ok = std::fread(&stepsMlp, 4, 1, f) == 1 && std::fread(second.data(), 4, second.size(), f) == second.size(); }
        if (ok && (flags & 2u)) {
            m1.resize(t.size()); m2.resize(t.size());
            ok = std::fread(&stepsTab, 4, 1, f) == 1 && std::fread(m1.data(), 4, m1.size(), f) == m1.size() &&
                 std::fread(m2.data(), 4, m2.size(), f) == m2.size();
        }
    }
    std::fclose(f);
    if (!ok) throw std::runtime_error("LoadCheckpoint: " + path + " is not a checkpoint of this cache shape");
    std::vector<float> zp(p.size(), 0.0f), zt(t.size(), 0.0f);
    Check(nrc_mlp_set_params(nrc.Get(), p.data()), "nrc_mlp_set_params");
    Check(nrc_mlp_set_state(nrc.Get(), zp.data(), m.data()), "nrc_mlp_set_state");
    Check(nrc_mrhe_set_tables(mrhe.Get(), t.data(), zt.data()), "nrc_mrhe_set_tables");
    // Adam state goes into objects that have Adam switched on (SetAdam before LoadCheckpoint); otherwise it is skipped
    if ((flags & 1u) && nrc_mlp_get_adam_state(nrc.Get(), nullptr, nullptr) == NRC_OK)
        Check(nrc_mlp_set_adam_state(nrc.Get(), second.data(), stepsMlp), "nrc_mlp_set_adam_state");
    if ((flags & 2u) && nrc_mrhe_get_adam_state(mrhe.Get(), nullptr, nullptr, nullptr) == NRC_OK)
        Check(nrc_mrhe_set_adam_state(mrhe.Get(), m1.data(), m2.data(), stepsTab), "nrc_mrhe_set_adam_state");
}

}   // namespace en
