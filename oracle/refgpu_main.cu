// TEST / MEASUREMENT INFRASTRUCTURE -- NOT PRODUCT CODE.
// The reference's OWN device path on this GPU: src/pathtracer.h (Pathtracer::Init / Render / Finish and every
// __global__ kernel of src/kernels.h) compiled UNMODIFIED where it lies, for sm_100a, driven like src/main.cpp:273-319
// drives it.  The only substitution: the four CUDA<->OpenGL interop calls of InteropTex (pathtracer.h:15-43) are
// renamed to local functions that hand out a plain cudaArray (surface load/store) instead of a mapped GL texture --
// "GL interop becomes a headless device buffer".  No graphics API is touched.
// Uses: (1) parity of the whole frame loop (shade, connect, radiance-cache kernels have no host build) against the
// reference's real device kernels; (2) the reference's throughput on the same B200 next to the product's.
//   refgpu_WxH <scene.txt>        (see parseScene below; writes <out>.pfm and <out>.bin float4 sums)
#include <map>
#include <fstream>
#include <sstream>
#include <cstring>
#include <cassert>
#include <string>
#include <vector>
#include <chrono>
#include "types.h"
#include "bvhBuilder.h"
#include "scene.h"

// ---- headless replacements for the GL interop used by InteropTex -------------------------------------------------
static std::vector<cudaArray_t> g_images;
static cudaError_t stubRegisterImage(cudaGraphicsResource** res, GLuint, GLenum, unsigned int) {
    cudaChannelFormatDesc d = cudaCreateChannelDesc(32, 32, 32, 32, cudaChannelFormatKindFloat);
    cudaArray_t a = nullptr;
    cudaError_t e = cudaMallocArray(&a, &d, WINDOW_WIDTH, WINDOW_HEIGHT, cudaArraySurfaceLoadStore);
    g_images.push_back(a);
    *res = (cudaGraphicsResource*)(uintptr_t)g_images.size();
    return e;
}
static cudaError_t stubMap(int, cudaGraphicsResource**, cudaStream_t = 0) { return cudaSuccess; }
static cudaError_t stubUnmap(int, cudaGraphicsResource**, cudaStream_t = 0) { return cudaSuccess; }
static cudaError_t stubMappedArray(cudaArray_t* arr, cudaGraphicsResource* res, unsigned int, unsigned int) {
    *arr = g_images[(uintptr_t)res - 1];
    return cudaSuccess;
}
#define cudaGraphicsGLRegisterImage stubRegisterImage
#define cudaGraphicsMapResources stubMap
#define cudaGraphicsUnmapResources stubUnmap
#define cudaGraphicsSubResourceGetMappedArray stubMappedArray
#include "pathtracer.h"

static void readImage(int idx, std::vector<float>& out) {
    out.resize((size_t)WINDOW_WIDTH * WINDOW_HEIGHT * 4);
    cudaSafe(cudaMemcpy2DFromArray(out.data(), WINDOW_WIDTH * 16, g_images[idx], 0, 0, WINDOW_WIDTH * 16, WINDOW_HEIGHT, cudaMemcpyDeviceToHost));
}

int main(int argc, char** argv) {
    if (argc < 2) { fprintf(stderr, "usage: %s scene.txt\n", argv[0]); return 2; }
    std::ifstream in(argv[1]);
    if (!in) { fprintf(stderr, "cannot open %s\n", argv[1]); return 2; }
    Scene scene;
    std::vector<uint> models;
    float3 eye = make_float3(0, 2, -6), dir = make_float3(0, 0, 1);
    float cd = 1.5f, focal = 5.0f, aperture = 0.01f;
    int clearFrames = 1, convergeFrames = 1, nee = 1, cache = 1, timedFrames = 0;
    std::string out = "refgpu_out", line, key;
    while (std::getline(in, line)) {
        std::istringstream s(line);
        if (!(s >> key) || key[0] == '#') continue;
        if (key == "material") {            // diffuse3 emission3 reflect glossy transmit ior absorption3
            float v[13]; for (float& x : v) s >> x;
            Material m = Material::DIFFUSE(make_float3(v[0], v[1], v[2]));
            m.emission = make_float3(v[3], v[4], v[5]); m.reflect = v[6]; m.glossy = v[7]; m.transmit = v[8]; m.refractive_index = v[9];
            m.absorption = make_float3(v[10], v[11], v[12]);
            scene.addMaterial(m);
        } else if (key == "model") {        // path scale rot3 off3 material useMtl
            std::string path; float sc, r[3], o[3]; uint mat; int useMtl;
            s >> path >> sc >> r[0] >> r[1] >> r[2] >> o[0] >> o[1] >> o[2] >> mat >> useMtl;
            models.push_back(scene.addModel(path, sc, make_float3(r[0], r[1], r[2]), make_float3(o[0], o[1], o[2]), mat, useMtl != 0));
        } else if (key == "object") {       // model material(-1 = per triangle) pos3 rot3 scale3
            int model, mat; float p[3], r[3], sc[3];
            s >> model >> mat >> p[0] >> p[1] >> p[2] >> r[0] >> r[1] >> r[2] >> sc[0] >> sc[1] >> sc[2];
            GameObject g(models[model], mat < 0 ? 0xffffffffu : (uint)mat);
            g.position = make_float3(p[0], p[1], p[2]); g.rotation = make_float3(r[0], r[1], r[2]); g.scale = make_float3(sc[0], sc[1], sc[2]);
            scene.addObject(g);
        } else if (key == "sphere") { float p[3], r; uint mat; s >> p[0] >> p[1] >> p[2] >> r >> mat; scene.addSphere(Sphere(make_float3(p[0], p[1], p[2]), r, mat)); }
        else if (key == "plane") { float n[3], d; uint mat; s >> n[0] >> n[1] >> n[2] >> d >> mat; scene.addPlane(Plane(make_float3(n[0], n[1], n[2]), d, mat)); }
        else if (key == "camera") s >> eye.x >> eye.y >> eye.z >> dir.x >> dir.y >> dir.z >> cd >> focal >> aperture;
        else if (key == "render") s >> clearFrames >> convergeFrames >> nee >> cache >> timedFrames;
        else if (key == "interactive_depth") s >> scene.interactive_depth;
        else if (key == "out") s >> out;
    }
    scene.finalize();
    for (size_t i = 0; i < scene.objects.size(); i++) scene.instances[i] = ConvertToInstance(scene.objects[i]);     // data half of Scene::update, scene.h:396-400
    BuildTopLevelBVH(scene.topLevelBVH.data(), scene.instances, scene.models.data(), scene.objects.size());
    Camera camera(eye, dir, cd, focal, aperture);
    camera.update(nullptr);

    Pathtracer app(scene, 1, 2);
    app.Init();                                      // loads ./bluenoise.png and ./cave.hdr from the working directory
    HNEE = nee != 0; HCACHE = cache != 0;            // main.cpp toggles them from the keyboard; synced on the next clearing frame
    for (int i = 0; i < clearFrames; i++) { app.Render(camera, 0.0f, 0.0f, true); app.Finish(); }
    for (int i = 0; i < convergeFrames; i++) { app.Render(camera, 0.0f, 0.0f, false); app.Finish(); }
    std::vector<float> lum, alb;
    readImage(0, lum); readImage(1, alb);
    { std::ofstream f(out + ".bin", std::ios::binary); f.write((const char*)lum.data(), lum.size() * 4); f.write((const char*)alb.data(), alb.size() * 4); }
    if (timedFrames > 0) {                           // throughput of further converge samples, one Finish per frame as in main.cpp:308-319
        cudaSafe(cudaDeviceSynchronize());
        auto t0 = std::chrono::steady_clock::now();
        for (int i = 0; i < timedFrames; i++) { app.Render(camera, 0.0f, 0.0f, false); app.Finish(); }
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        printf("REFGPU timed_frames %d total_ms %.3f ms_per_sample %.4f\n", timedFrames, ms, ms / timedFrames);
    }
    printf("REFGPU ok triangles %zu instances %zu size %dx%d depth %d\n", scene.allVertices.size(), scene.objects.size(), WINDOW_WIDTH, WINDOW_HEIGHT, MAX_RAY_DEPTH);
    return 0;
}
