// Headless restatement of the reference's main() (source/Etna/TestMain.cpp:41-263) over the vkc-shaped
// host mirror: same start-up order, same structs, same per-frame updates, the A/D/W/S/Q/E keys replaced by
// a scripted key trace, the window by PPM files.  SURVEY.md §8(f) rank 1 ("interactive host loop
// without a window").  Everything heavy happens in libvolrt (CUDA); this file is host logic only.
//
//   volrt_headless [--frames N] [--keys "D*60,W*30,A*20"] [--out DIR] [--every K] [--size WxH]
//                  [--gpus N]      CUDA devices 0..N-1 behind the one context (row bands, peer stores into device 0's frame)
//                  [--std140]      send GlobalShaderData in the shader's std140 layout (quirk Q2 "fixed": the medium scrolls)
//                  [--dt SECONDS]  frameTime advances by this much per frame instead of following the wall clock
//                  [--scroll-yzw]  also write -frameTime into MediaScroll[0].y/.z/.w — the only element the reference writes,
//                                  [0].x, is read by a commented-out line of the shader (frag.glsl:65), so its medium never moves
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "glm_lite.hpp"
#include "volrt_host.hpp"

struct ObjectShaderData {          // TestMain.cpp:27-32
    glm::mat4 Model;
    glm::mat4 View;
    glm::mat4 Projection;
};

#pragma pack(push, 4)
struct GlobalShaderData {          // TestMain.cpp:34-39 — glm::vec3 is 12 bytes and glm::mat4 4-byte aligned, so
    glm::mat4 WorldToLocal;        // MediaScroll starts at byte 76, NOT at the std140 offset 80
    glm::vec3 CameraPosition;
    glm::mat4 MediaScroll;
};
#pragma pack(pop)
// What the SHADER declares (frag.glsl:9-14, std140: a vec3 is padded to 16 bytes, MediaScroll at byte 80).  The reference
// never sends this layout — with its 140-byte struct the -frameTime it writes lands in padding and the medium does not
// scroll.  --std140 sends this one instead: the animation the shader was written for.
struct GlobalShaderDataStd140 {
    glm::mat4 WorldToLocal;
    glm::vec3 CameraPosition;
    float pad0;
    glm::mat4 MediaScroll;
};
static_assert(sizeof(GlobalShaderDataStd140) == 144, "std140 image of GlobalShaderData");
static_assert(sizeof(ObjectShaderData) == 192, "ObjectShaderData");
static_assert(sizeof(GlobalShaderData) == 140, "GlobalShaderData is 140 bytes in the reference's C++ layout");

static std::vector<char> parse_keys(const std::string& spec, int frames) {
    // "D*60,W*30" -> one key (or '-') per frame; the last entry repeats
    std::vector<char> out;
    size_t i = 0;
    while (i < spec.size()) {
        char k = spec[i++];
        int n = 1;
        if (i < spec.size() && spec[i] == '*') { n = std::atoi(spec.c_str() + i + 1); while (i < spec.size() && spec[i] != ',') ++i; }
        if (i < spec.size() && spec[i] == ',') ++i;
        for (int j = 0; j < n; ++j) out.push_back(k);
    }
    if (out.empty()) out.push_back('-');
    while ((int)out.size() < frames) out.push_back(out.back());
    return out;
}

static void write_ppm(const std::string& path, const unsigned char* rgba, uint32_t w, uint32_t h) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return;
    std::fprintf(f, "P6\n%u %u\n255\n", w, h);
    std::vector<unsigned char> row(w * 3);
    for (uint32_t y = 0; y < h; ++y) {
        for (uint32_t x = 0; x < w; ++x) for (int c = 0; c < 3; ++c) row[x * 3 + c] = rgba[((size_t)y * w + x) * 4 + c];
        std::fwrite(row.data(), 1, row.size(), f);
    }
    std::fclose(f);
}

int main(int argc, char** argv) {
    int frames = 120, every = 0, gpus = 1;
    bool std140 = false, scroll_yzw = false;
    float scripted_dt = -1.0f;
    uint32_t W = 1280, H = 720;                         // VulkanContext.cpp:24
    std::string keys = "D*60,W*30,A*30", out_dir;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "--frames" && i + 1 < argc) frames = std::atoi(argv[++i]);
        else if (a == "--keys" && i + 1 < argc) keys = argv[++i];
        else if (a == "--out" && i + 1 < argc) out_dir = argv[++i];
        else if (a == "--every" && i + 1 < argc) every = std::atoi(argv[++i]);
        else if (a == "--std140") std140 = true;
        else if (a == "--scroll-yzw") scroll_yzw = true;
        else if (a == "--dt" && i + 1 < argc) scripted_dt = (float)std::atof(argv[++i]);
        else if (a == "--size" && i + 1 < argc) { if (std::sscanf(argv[++i], "%ux%u", &W, &H) != 2) return 2; }
        else if (a == "--gpus" && i + 1 < argc) gpus = std::atoi(argv[++i]);
        else { std::fprintf(stderr, "unknown argument %s\n", a.c_str()); return 2; }
    }
    try {
        vkc::Renderer renderer;
        renderer.Init(W, H, VR_OUT_RGBA8_UNORM, 0, gpus);                         // TestMain.cpp:121; --gpus N: one context, N devices
        // TestMain.cpp:43-92 + :128 — the 128^3 RGBA8 noise volume, baked on the device (incl. quirk Q1)
        vkc::Texture3D texture = vkc::Texture3D::Bake();

        vkc::UniformBuffer<ObjectShaderData> objectUniformBuffer(renderer.GetFramesCount());    // :130
        vkc::UniformBuffer<GlobalShaderData> globalUniformBuffer(renderer.GetFramesCount());
        vkc::UniformBuffer<GlobalShaderDataStd140> globalUniformBufferStd140(renderer.GetFramesCount());
        std::vector<vkc::DescriptorSet> perFrameSets(renderer.GetFramesCount());                 // :138-154
        for (uint32_t i = 0; i < renderer.GetFramesCount(); ++i)
            perFrameSets[i].WriteBuffer(0, objectUniformBuffer.HostMappings[i], sizeof(ObjectShaderData))
                           .WriteBuffer(1, std140 ? globalUniformBufferStd140.HostMappings[i] : globalUniformBuffer.HostMappings[i],
                                        std140 ? sizeof(GlobalShaderDataStd140) : sizeof(GlobalShaderData))
                           .WriteImage(2, texture);

        vkc::RenderPassCreateInfo createInfo;                                                    // :156-166
        renderer.AddRenderPass("BasePass", createInfo);

        const uint32_t indicesCount = 36;                                                        // :94-112
        float cubePhi = 0, cubeTheta = 0, rotationSpeed = 100.f, controlledFrameTime = 0;
        const std::vector<char> trace = parse_keys(keys, frames);
        auto startTime = std::chrono::high_resolution_clock::now();
        double gpu_ms_sum = 0.0;
        unsigned long long checksum = 0;

        for (int frame = 0; frame < frames; ++frame) {                                           // :173
            const float deltaTime = 0.016f;
            switch (trace[frame]) {                                                              // :177-188
                case 'A': cubePhi -= rotationSpeed * deltaTime; break;
                case 'D': cubePhi += rotationSpeed * deltaTime; break;
                case 'W': cubeTheta -= rotationSpeed * deltaTime; break;
                case 'S': cubeTheta += rotationSpeed * deltaTime; break;
                case 'Q': controlledFrameTime += deltaTime * 0.2f; break;
                case 'E': controlledFrameTime -= deltaTime * 0.2f; break;
                default: break;
            }
            vkc::Rect2D rect = {{0, 0}, renderer.GetSwapchainExtent()};                          // :192
            renderer.EnqueueRenderPass("BasePass", rect, {},
                [&perFrameSets, indicesCount](vkc::RenderPassContext&& rpc) {                   // :194-217
                    rpc.BindDescriptorSet(perFrameSets[rpc.FrameIndex]);
                    rpc.DrawIndexed(indicesCount);
                });

            auto rot = glm::rotate(glm::mat4(1.0f), glm::radians(cubePhi), glm::vec3(0.0f, 0.0f, 1.0f));   // :222
            ObjectShaderData osd;
            osd.Model = glm::rotate(rot, glm::radians(cubeTheta), glm::vec3(0.0f, 1.0f, 0.0f));
            osd.View = glm::lookAt(glm::vec3(3.0f, 3.0f, 3.0f), glm::vec3(0.0f, 0.0f, 0.0f), glm::vec3(0.0f, 0.0f, 1.0f));
            osd.Projection = glm::perspective(glm::radians(45.0f), (float)1280 / (float)720, 0.1f, 10.0f);
            osd.Projection[1][1] *= -1;                                                          // :228

            glm::mat4 worldToLocal = glm::inverse(osd.Model);                                    // :230
            auto now = std::chrono::high_resolution_clock::now();
            float frameTime = std::chrono::duration<float>(now - startTime).count();            // :232
            if (scripted_dt >= 0.0f) frameTime = scripted_dt * (float)frame;
            glm::mat4 mediaScroll;
            mediaScroll[0][0] = -frameTime;                                                      // :233-238
            if (scroll_yzw) mediaScroll[0][1] = mediaScroll[0][2] = mediaScroll[0][3] = -frameTime;
            GlobalShaderData gsd;
            gsd.WorldToLocal = worldToLocal;
            gsd.CameraPosition = glm::vec3(3.0f, 3.0f, 3.0f);                                   // :242
            gsd.MediaScroll = mediaScroll;

            objectUniformBuffer.Update(&osd, renderer.GetCurrentFrame());                       // :248
            globalUniformBuffer.Update(&gsd, renderer.GetCurrentFrame());                       // :249
            if (std140) {
                GlobalShaderDataStd140 g2;
                g2.WorldToLocal = gsd.WorldToLocal; g2.CameraPosition = gsd.CameraPosition; g2.pad0 = 0.0f; g2.MediaScroll = gsd.MediaScroll;
                globalUniformBufferStd140.Update(&g2, renderer.GetCurrentFrame());
            }

            renderer.BeginFrame();                                                               // :251
            renderer.EndFrame();                                                                 // :255

            if ((every > 0 && frame % every == 0) || frame == frames - 1) {
                const unsigned char* px = (const unsigned char*)renderer.WaitAndMapLastFrame();
                unsigned long long s = 0;
                for (size_t i = 0; i < (size_t)W * H * 4; i += 4) s += px[i];
                checksum = checksum * 1315423911ull + s;
                gpu_ms_sum += renderer.LastGpuMs();
                if (!out_dir.empty()) {
                    char name[64];
                    std::snprintf(name, sizeof name, "/frame_%04d.ppm", frame);
                    write_ppm(out_dir + name, px, W, H);
                }
                std::printf("frame %4d key %c phi %7.2f theta %7.2f sum(R) %llu gpu %.3f ms\n", frame, trace[frame], cubePhi, cubeTheta, s, renderer.LastGpuMs());
            }
        }
        renderer.WaitAndMapLastFrame();
        double wall = std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - startTime).count();
        std::printf("frames %d  wall %.3f s  %.1f frames/s (2 frames in flight, frame read back to host every frame)  checksum %llx\n",
                    frames, wall, frames / wall, checksum);
        (void)controlledFrameTime;      // unused in the reference as well (TestMain.cpp:172,185-188)
        renderer.Shutdown();                                                                     // :260
    } catch (const std::exception& e) {
        std::fprintf(stderr, "fatal: %s\n", e.what());
        return 1;
    }
    return 0;
}
