// C++ host side above the libvolrt C ABI: the reference's own objects for the hot path, same names and
// argument meaning, so that its main() (source/Etna/TestMain.cpp) reads the same with the Vulkan draw
// replaced by the CUDA march.  Header-only; links against libvolrt.so only (include/volrt.h).
//
//   reference (source/Etna/Core/Vulkan/...)                         here
//   vkc::Renderer::Init/Shutdown/AddRenderPass/EnqueueRenderPass/   vkc::Renderer (VulkanRenderer.h:58-100)
//        BeginFrame/EndFrame/GetSwapchainExtent/GetCurrentFrame
//   vkc::Texture3D(unsigned char*, VkExtent3D)                      vkc::Texture3D (VulkanTexture.h:69-74)
//   vkc::UniformBuffer<T>(count).Update(&data, frame)               vkc::UniformBuffer<T> (VulkanUniformBuffer.h:12-61)
//   DescriptorSetWriter.WriteBuffer(0|1,...).WriteImage(2,...)      vkc::DescriptorSet (TestMain.cpp:147-153)
//   vkCmdDrawIndexed inside the RenderPassDelegate                  RenderPassContext::DrawIndexed (TestMain.cpp:216)
//   Error(...) macro -> throws Exception (Core/Utils.h:22-29)       vkc::Exception (same behaviour: throws)
//
// Left out: window, swapchain, ImGui, validation — no graphics stack exists on the build machines.  The
// colour target is a headless device frame; EndFrame() reads it back into page-locked host memory, one
// buffer per frame in flight (MAX_FRAMES_IN_FLIGHT = 2, VulkanRenderer.cpp:13).
#pragma once
#include <cstdint>
#include <cstring>
#include <functional>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "../include/volrt.h"

namespace vkc {

struct Extent2D { uint32_t width, height; };
struct Extent3D { uint32_t width, height, depth; };
struct Offset2D { int32_t x, y; };
struct Rect2D { Offset2D offset; Extent2D extent; };

class Exception : public std::runtime_error {
public:
    Exception(int line, const char* file, const std::string& msg)
        : std::runtime_error(std::string(file) + ":" + std::to_string(line) + ": " + msg) {}
};
#define VOLRT_HOST_ERROR(msg) throw ::vkc::Exception(__LINE__, __FILE__, (msg))

// the reference keeps its device in a global singleton (VulkanContext.cpp:11-33); so does this mirror
class Context {
public:
    static vr_context*& Get() { static vr_context* ctx = nullptr; return ctx; }
    static void Check(vr_status st, const char* what) {
        if (st != VR_OK) VOLRT_HOST_ERROR(std::string(what) + ": " + vr_last_error(Get()));
    }
};

class Texture3D {
public:
    // VulkanTexture.cpp:111-156: RGBA8 UNORM, tightly packed, x fastest; sampler = linear, mirrored repeat
    Texture3D(unsigned char* data, Extent3D extent) : Extent(extent) {
        vr_volume_desc d = {sizeof(vr_volume_desc), extent.width, extent.height, extent.depth,
                            VR_VOL_RGBA8_UNORM, VR_ADDRESS_MIRRORED_REPEAT, VR_LAYOUT_TEXTURE};
        Context::Check(vr_upload_volume(Context::Get(), &d, data), "Texture3D");
    }
    // the start-up bake of TestMain.cpp:43-92 done on the device instead of by FastNoise2 on the CPU
    static Texture3D Bake(const vr_noise_desc* desc = nullptr) {
        vr_noise_desc d;
        if (desc) d = *desc; else vr_noise_desc_default(&d);
        Context::Check(vr_generate_volume(Context::Get(), &d, nullptr), "Texture3D::Bake");
        return Texture3D({d.size, d.size, d.size});
    }
    Extent3D Extent;
private:
    explicit Texture3D(Extent3D e) : Extent(e) {}
};

template <class T>
class UniformBuffer {
public:
    explicit UniformBuffer(uint32_t count = 1) : Storage(count), HostMappings(count) {
        for (uint32_t i = 0; i < count; ++i) HostMappings[i] = &Storage[i];
    }
    UniformBuffer(const UniformBuffer&) = delete;
    UniformBuffer& operator=(const UniformBuffer&) = delete;
    void Update(const T* data, uint32_t index = 0) { std::memcpy(HostMappings[index], data, sizeof(T)); }   // VulkanUniformBuffer.h:58-61
    std::vector<T> Storage;
    std::vector<void*> HostMappings;
};

// bindings 0,1 = uniform blocks, 2 = the sampler3D (TestMain.cpp:132-153); one set per frame in flight
struct DescriptorSet {
    const void* Ubo[2] = {nullptr, nullptr};
    uint32_t UboSize[2] = {0, 0};
    DescriptorSet& WriteBuffer(uint32_t binding, const void* mapping, uint32_t range) {
        if (binding > 1) VOLRT_HOST_ERROR("uniform bindings are 0 (ObjectShaderData) and 1 (GlobalShaderData)");
        Ubo[binding] = mapping; UboSize[binding] = range;
        return *this;
    }
    DescriptorSet& WriteImage(uint32_t binding, const Texture3D&) {
        if (binding != 2) VOLRT_HOST_ERROR("the sampler3D is binding 2");
        return *this;
    }
};

struct RenderPassCreateInfo {          // the fields of the reference's struct that reach the hot path
    vr_output_format Format = VR_OUT_RGBA8_UNORM;   // reference: swapchain format (VulkanSwapchain.cpp:181-191)
    vr_params Params;                               // shader literals (frag.glsl:29-32,42,63,66-69) + extensions
    RenderPassCreateInfo() { vr_params_default(&Params); }
};

class Renderer;
struct RenderPassContext {
    Renderer* Owner;
    uint32_t FrameIndex;               // index of the frame in flight
    const DescriptorSet* Bound = nullptr;
    void BindDescriptorSet(const DescriptorSet& s) { Bound = &s; }
    inline void DrawIndexed(uint32_t indexCount);   // the cube's 36 indices (TestMain.cpp:94-112,216)
};
using RenderPassDelegate = std::function<void(RenderPassContext&&)>;

class Renderer {
public:
    static constexpr uint32_t MaxFramesInFlight = 2;            // VulkanRenderer.cpp:13

    // gpus > 1: CUDA devices 0 .. gpus-1 behind the one context (vr_create_info.device_count); the frame lives on device 0
    void Init(uint32_t width = 1280, uint32_t height = 720, vr_output_format fmt = VR_OUT_RGBA8_UNORM, int device = 0, int gpus = 1) {
        vr_create_info ci = {};
        ci.struct_size = sizeof(vr_create_info);
        ci.device = device; ci.width = width; ci.height = height; ci.format = (uint32_t)fmt; ci.max_views = 1;
        if (gpus > 1) {
            ci.device_count = (uint32_t)(gpus > VR_MAX_DEVICES ? VR_MAX_DEVICES : gpus);
            for (uint32_t k = 0; k < ci.device_count; ++k) ci.devices[k] = (int32_t)k;
        }
        vr_context* ctx = nullptr;
        vr_status st = vr_create(&ci, &ctx);
        if (st != VR_OK) VOLRT_HOST_ERROR(std::string("Renderer::Init: ") + vr_last_error(nullptr));
        Context::Get() = ctx;
        Extent = {width, height};
        PixelBytes = fmt == VR_OUT_RGBA16F ? 8 : fmt == VR_OUT_RGBA32F ? 16 : 4;
        for (auto& b : Readback) {
            b = vr_host_alloc((size_t)width * height * PixelBytes);
            if (!b) VOLRT_HOST_ERROR("Renderer::Init: vr_host_alloc failed");
        }
    }
    void Shutdown() {
        if (!Context::Get()) return;
        vr_wait(Context::Get());
        for (auto& b : Readback) { vr_host_free(b); b = nullptr; }
        vr_destroy(Context::Get());
        Context::Get() = nullptr;
    }
    void AddRenderPass(const std::string& name, const RenderPassCreateInfo& initInfo) { Passes[name] = initInfo; }
    void EnqueueRenderPass(const std::string& name, Rect2D area, const std::vector<std::string>& /*dependencies*/ = {},
                           RenderPassDelegate&& delegate = [](RenderPassContext&&) {}) {
        if (!Passes.count(name)) VOLRT_HOST_ERROR("unknown render pass " + name);
        Queue.push_back({name, area, std::move(delegate)});
    }
    // wait for the slot's previous frame (the reference waits on the frame fence, VulkanRenderer.cpp:142-159)
    void BeginFrame() { Context::Check(vr_wait(Context::Get()), "BeginFrame"); }
    // record + submit every queued pass (VulkanRenderer.cpp:161-230); returns immediately (VR_RENDER_ASYNC)
    void EndFrame() {
        for (auto& q : Queue) {
            Current = &Passes[q.Name];
            Area = q.Area;
            q.Delegate(RenderPassContext{this, CurrentFrame, nullptr});
        }
        Queue.clear();
        LastFrame = CurrentFrame;
        CurrentFrame = (CurrentFrame + 1) % MaxFramesInFlight;
    }
    [[nodiscard]] Extent2D GetSwapchainExtent() const { return Extent; }
    [[nodiscard]] uint32_t GetFramesCount() const { return MaxFramesInFlight; }
    [[nodiscard]] uint32_t GetCurrentFrame() const { return CurrentFrame; }
    // headless stand-in for presentation: the finished pixels of the most recently submitted frame
    const void* WaitAndMapLastFrame() {
        Context::Check(vr_wait(Context::Get()), "WaitAndMapLastFrame");
        return Readback[LastFrame];
    }
    float LastGpuMs() {
        vr_stats s; s.struct_size = sizeof s;
        Context::Check(vr_get_stats(Context::Get(), &s), "vr_get_stats");
        return s.gpu_ms;
    }
    uint32_t PixelBytes = 4;

private:
    friend struct RenderPassContext;
    void Draw(const DescriptorSet& set, uint32_t frame) {
        if (!set.Ubo[0] || !set.Ubo[1]) VOLRT_HOST_ERROR("DrawIndexed: descriptor set has no uniform buffers");
        if (Area.offset.x != 0 || Area.offset.y != 0 || Area.extent.width != Extent.width || Area.extent.height != Extent.height)
            VOLRT_HOST_ERROR("DrawIndexed: the render area must be the whole target (TestMain.cpp:192)");
        vr_params p = Current->Params;
        p.ubo0 = set.Ubo[0]; p.ubo0_size = set.UboSize[0];
        p.ubo1 = set.Ubo[1]; p.ubo1_size = set.UboSize[1];
        Context::Check(vr_set_params(Context::Get(), &p), "vr_set_params");
        vr_target t;
        std::memset(&t, 0, sizeof t);
        t.struct_size = sizeof t;
        t.memory_kind = VR_MEM_HOST;
        t.dst = Readback[frame];
        t.flags = VR_RENDER_ASYNC;
        Context::Check(vr_render(Context::Get(), &t), "vr_render");
    }
    struct Queued { std::string Name; Rect2D Area; RenderPassDelegate Delegate; };
    std::map<std::string, RenderPassCreateInfo> Passes;
    std::vector<Queued> Queue;
    RenderPassCreateInfo* Current = nullptr;
    Rect2D Area{};
    Extent2D Extent{0, 0};
    uint32_t CurrentFrame = 0, LastFrame = 0;
    void* Readback[MaxFramesInFlight] = {nullptr, nullptr};
};

inline void RenderPassContext::DrawIndexed(uint32_t indexCount) {
    if (indexCount != 36) VOLRT_HOST_ERROR("DrawIndexed: the path draws the 36-index cube (TestMain.cpp:94-112)");
    if (!Bound) VOLRT_HOST_ERROR("DrawIndexed: no descriptor set bound");
    Owner->Draw(*Bound, FrameIndex);
}

}  // namespace vkc
