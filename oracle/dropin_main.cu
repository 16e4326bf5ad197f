// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.
// Drop-in check: the reference's OWN host code (Scene::addModel with its vendored tinyobjloader, createBVHBinned,
// BuildTopLevelBVH, Camera -- compiled where it lies under /root/reference/src, nothing copied) drives the
// product through include/ptb200_pathtracer.h, the `class Pathtracer : public Application` adapter that replaces
// src/pathtracer.h.  This is what a reference maintainer's main.cpp does minus the GLFW window
// (src/main.cpp:273-319): Init, Render(shouldClear), Render x N, Finish -- then a PFM instead of a swap.
//   dropin_WxH <model.obj> <out.pfm> <converge samples> [eye xyz] [emissive 0|1] [device top-level BVH 0|1]
#include <map>
#include <fstream>
#include <sstream>
#include <cstring>
#include <cassert>
#include <string>
#include "types.h"
#include "bvhBuilder.h"
#include "scene.h"
#include "ptb200_pathtracer.h"

int main(int argc, char** argv) {
    if (argc < 4) { fprintf(stderr, "usage: %s model.obj out.pfm samples [ex ey ez] [emissive]\n", argv[0]); return 2; }
    const int samples = atoi(argv[3]);
    float3 eye = make_float3(0, 2, -6);
    if (argc >= 7) eye = make_float3(atof(argv[4]), atof(argv[5]), atof(argv[6]));
    const bool emissive = argc >= 8 && atoi(argv[7]) != 0;
    const bool deviceTlas = argc >= 9 && atoi(argv[8]) != 0;

    Scene scene;
    Material white = Material::DIFFUSE(make_float3(0.7f, 0.6f, 0.5f));
    const MATERIAL_ID whiteId = scene.addMaterial(white);
    Material lamp = Material::DIFFUSE(make_float3(1));
    lamp.emission = make_float3(10);
    const MATERIAL_ID lampId = scene.addMaterial(lamp);
    const uint model = scene.addModel(argv[1], 1.0f, make_float3(0), make_float3(0), whiteId, false);
    scene.addObject(GameObject(model));
    if (emissive) {                                   // a second, emissive instance of the same mesh above the first: NEE + guiding get work
        GameObject l(model, lampId);
        l.position = make_float3(0, 6, 0); l.scale = make_float3(0.3f);
        scene.addObject(l);
    }
    scene.addPlane(Plane(make_float3(0, -1, 0), -3, whiteId));
    scene.finalize();
    // the data half of Scene::update (scene.h:396-400); the keyboard half needs a GLFW window
    for (size_t i = 0; i < scene.objects.size(); i++) scene.instances[i] = ConvertToInstance(scene.objects[i]);
    BuildTopLevelBVH(scene.topLevelBVH.data(), scene.instances, scene.models.data(), scene.objects.size());
    // with the device top-level build the host tree is only needed for the first upload: scramble the boxes afterwards so a
    // render that still used them would not match
    auto scramble = [&]() { for (auto& n : scene.topLevelBVH) { n.box.vmin = make_float3(1e6f); n.box.vmax = make_float3(1e6f + 1); } };

    Camera camera(eye, normalize(make_float3(0, 1, 0) - eye), 1.5f, 5.0f, 0.01f);          // initialCamera() intrinsics, stateLoader.h:30-33
    camera.update(nullptr);                                                                  // recalculate(), types.h:590-600

    Pathtracer app(scene, 0, 0);
    app.Init();
    app.deviceTopLevelBVH = deviceTlas;
    if (deviceTlas) scramble();
    app.SetSkyConstant(1, 1, 1);
    app.Render(camera, 0.0f, 0.0f, true);
    for (int i = 0; i < samples; i++) app.Render(camera, 0.0f, 0.0f, false);
    app.Finish();
    app.WritePFM(argv[2]);
    printf("dropin ok: %zu triangles, %zu instances, %d converge samples -> %s\n", scene.allVertices.size(), scene.objects.size(), samples, argv[2]);
    return 0;
}
