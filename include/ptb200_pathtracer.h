/* ptb200_pathtracer.h -- drop-in for the reference's src/pathtracer.h.
 *
 * `class Pathtracer : public Application` with the reference's exact method signatures
 * (src/application.h:7-20, src/pathtracer.h:46-310), forwarding to the C ABI of libptb200.so.
 * Include it INSTEAD of pathtracer.h in the reference's translation unit (after application.h's
 * dependencies are visible, as main.cpp does) and link with -lptb200.  Every POD of the ABI is
 * layout-compatible with the reference struct it mirrors, so the reference's own Scene vectors and
 * Camera are passed as they are.  Errors print and exit like the reference's cudaSafe
 * (src/use_cuda.h:32-40).  The GL textures are not touched: read the image with ReadLuminance /
 * ReadAlbedo / WritePFM (headless).
 */
#ifndef PTB200_PATHTRACER_H
#define PTB200_PATHTRACER_H
#include <cstdio>
#include <cstdlib>
#include "application.h"
#include "globals.h"
#include "ptb200.h"

class Pathtracer : public Application
{
private:
    pt_ctx* ctx = nullptr;
    static void die(pt_ctx* c, const char* what) { fprintf(stderr, "%s: %s\n", what, pt_last_error(c)); exit(1); }

public:
    /* true: the top-level BVH is rebuilt on the device from scene.instances on every clearing frame
     * (pt_update_instances_device_tlas), so the host loop may drop its BuildTopLevelBVH call (scene.h:396-400) */
    bool deviceTopLevelBVH = false;

    Pathtracer(Scene& scene, GLuint luminanceTexture, GLuint albedoTexture) : Application(scene, luminanceTexture, albedoTexture) {}
    ~Pathtracer() { pt_destroy(ctx); }

    virtual void Init() override
    {
        static_assert(sizeof(TriangleV) == sizeof(pt_triangle_v) && sizeof(TriangleD) == sizeof(pt_triangle_d), "triangle layouts");
        static_assert(sizeof(Material) == sizeof(pt_material) && sizeof(Model) == sizeof(pt_model), "material/model layouts");
        static_assert(sizeof(Instance) == sizeof(pt_instance) && sizeof(TopLevelBVH) == sizeof(pt_tlas_node), "instance/TLAS layouts");
        static_assert(sizeof(Sphere) == sizeof(pt_sphere) && sizeof(Plane) == sizeof(pt_plane) && sizeof(Camera) == sizeof(pt_camera), "primitive/camera layouts");
        pt_config cfg; pt_default_config(&cfg);
        cfg.width = WINDOW_WIDTH; cfg.height = WINDOW_HEIGHT; cfg.max_ray_depth = MAX_RAY_DEPTH;   // constants.h:4-13
        cfg.interactive_depth = scene.interactive_depth;                                          // scene.h:137
        if (pt_create(&ctx, &cfg)) die(nullptr, "pt_create");
        pt_scene_desc d = {};
        d.vertices = (const pt_triangle_v*)scene.allVertices.data();   d.num_triangles = (uint32_t)scene.allVertices.size();
        d.vertexData = (const pt_triangle_d*)scene.allVertexData.data();
        d.models = (const pt_model*)scene.models.data();               d.num_models = (uint32_t)scene.models.size();      // Model::bvh = host BVH2
        d.materials = (const pt_material*)scene.materials.data();      d.num_materials = (uint32_t)scene.materials.size();
        d.instances = (const pt_instance*)scene.instances;             d.num_instances = (uint32_t)scene.objects.size();
        d.topBvh = (const pt_tlas_node*)scene.topLevelBVH.data();      d.num_tlas_nodes = (uint32_t)scene.topLevelBVH.size();
        d.spheres = (const pt_sphere*)scene.spheres.data();            d.num_spheres = (uint32_t)scene.spheres.size();
        d.planes = (const pt_plane*)scene.planes.data();               d.num_planes = (uint32_t)scene.planes.size();
        if (pt_upload_scene(ctx, &d) < 0) die(ctx, "pt_upload_scene");
        HNEE = true; HCACHE = true;                                                               // pathtracer.h:213-214
    }

    virtual void Render(const Camera& camera, float currentTime, float frameTime, bool shouldClear) override
    {
        (void)currentTime; (void)frameTime;
        if (shouldClear) {
            pt_set_toggles(ctx, HNEE, HCACHE);                                                    // pathtracer.h:242-243
            if (deviceTopLevelBVH ? pt_update_instances_device_tlas(ctx, (const pt_instance*)scene.instances, (uint32_t)scene.objects.size())
                                  : pt_update_instances(ctx, (const pt_instance*)scene.instances, (uint32_t)scene.objects.size(),
                                                        (const pt_tlas_node*)scene.topLevelBVH.data(), (uint32_t)scene.topLevelBVH.size()))   // :245-246
                die(ctx, "pt_update_instances");
        }
        if (pt_render(ctx, (const pt_camera*)&camera, shouldClear)) die(ctx, "pt_render");
    }

    virtual void Finish() override { if (pt_finish(ctx)) die(ctx, "pt_finish"); }

    /* headless replacements for the two GL textures (src/main.cpp:214-260) */
    void SetSkyConstant(float r, float g, float b) { pt_set_sky_constant(ctx, r, g, b); }
    void ReadLuminance(float* rgbw) { if (pt_read_image(ctx, rgbw, 0)) die(ctx, "pt_read_image"); }
    void ReadAlbedo(float* rgbw) { if (pt_read_image(ctx, rgbw, 1)) die(ctx, "pt_read_image"); }
    /* streaming read-back into pinned host memory while the next samples render; WaitRead(ticket) or Finish() before use */
    uint64_t ReadLuminanceAsync(float* rgbw) { uint64_t t = 0; if (pt_read_image_async(ctx, rgbw, 0, &t)) die(ctx, "pt_read_image_async"); return t; }
    void WaitRead(uint64_t ticket) { if (pt_wait_read(ctx, ticket)) die(ctx, "pt_wait_read"); }
    void WritePFM(const char* path) { if (pt_write_pfm(ctx, path, 0)) die(ctx, "pt_write_pfm"); }
    pt_ctx* Context() { return ctx; }
};

#endif
