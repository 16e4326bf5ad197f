// render_pfm -- headless render-to-PFM through the C ABI only (no Python): the replacement for the reference's GLFW
// window loop (src/main.cpp:273-376).  Loads an OBJ (+MTL) with the library's host scene code, adds a ground plane and
// an optional emissive copy of the mesh as a light, renders N converge samples and writes the display value rgb / w
// (optionally through the reference's blur / gamma / vignette pass) as a colour PFM.
//   render_pfm model.obj out.pfm [width height samples depth light(0|1) post(0|1)]
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include "ptb200.h"
#include "ptb200_scene.h"

static void die(const char* what, const char* msg) { fprintf(stderr, "%s: %s\n", what, msg ? msg : ""); exit(1); }

int main(int argc, char** argv) {
    if (argc < 3) { fprintf(stderr, "usage: %s model.obj out.pfm [width height samples depth light post]\n", argv[0]); return 2; }
    const uint32_t W = argc > 3 ? atoi(argv[3]) : 640, H = argc > 4 ? atoi(argv[4]) : 480;
    const int samples = argc > 5 ? atoi(argv[5]) : 64, depth = argc > 6 ? atoi(argv[6]) : 8;
    const bool light = argc > 7 && atoi(argv[7]) != 0, post = argc > 8 && atoi(argv[8]) != 0;

    pts_scene* sc = pts_scene_new();
    pt_material white = {}; white.diffuse_color = {0.7f, 0.7f, 0.7f}; white.refractive_index = 1.0f;
    pt_material lamp = white; lamp.diffuse_color = {1, 1, 1}; lamp.emission = {12, 12, 12};
    const uint32_t whiteId = pts_add_material(sc, &white), lampId = pts_add_material(sc, &lamp);
    const float zero[3] = {0, 0, 0}, one[3] = {1, 1, 1};
    const int model = pts_add_model_obj(sc, argv[1], 1.0f, zero, zero, whiteId, 0);
    if (model < 0) die("pts_add_model_obj", pts_last_error(sc));
    pts_add_object(sc, (uint32_t)model, 0xffffffffu, zero, zero, one);
    if (light) { const float up[3] = {0, 7, 0}, small[3] = {0.3f, 0.3f, 0.3f}; pts_add_object(sc, (uint32_t)model, lampId, up, zero, small); }
    const float n[3] = {0, -1, 0};
    pts_add_plane(sc, n, -3.0f, whiteId);
    if (pts_finalize(sc)) die("pts_finalize", pts_last_error(sc));
    pt_scene_desc desc;
    if (pts_get_desc(sc, &desc)) die("pts_get_desc", pts_last_error(sc));

    pt_config cfg; pt_default_config(&cfg);
    cfg.width = W; cfg.height = H; cfg.max_ray_depth = (uint32_t)depth;
    pt_ctx* ctx = nullptr;
    if (pt_create(&ctx, &cfg)) die("pt_create", pt_last_error(nullptr));
    if (pt_upload_scene(ctx, &desc) < 0) die("pt_upload_scene", pt_last_error(ctx));

    const float eye[3] = {1.0f, 3.0f, -8.0f};
    float d[3] = {0.0f - eye[0], 1.0f - eye[1], 0.0f - eye[2]};
    const float len = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    const float cam9[9] = {eye[0], eye[1], eye[2], d[0] / len, d[1] / len, d[2] / len, 1.5f, 5.0f, 0.01f};   // initialCamera() intrinsics, stateLoader.h:30-33
    pt_camera cam; pts_camera_make(cam9, W, H, &cam);

    if (pt_render(ctx, &cam, 1)) die("pt_render", pt_last_error(ctx));                 // clearing frame
    for (int i = 0; i < samples; i++) if (pt_render(ctx, &cam, 0)) die("pt_render", pt_last_error(ctx));
    if (pt_finish(ctx)) die("pt_finish", pt_last_error(ctx));
    if (!post) { if (pt_write_pfm(ctx, argv[2], 0)) die("pt_write_pfm", pt_last_error(ctx)); }
    else {                                                                             // the reference's display pass, then PFM
        std::vector<float> rgba((size_t)W * H * 4);
        if (pt_postprocess(ctx, nullptr, nullptr, (float)(samples + 1), 1, rgba.data())) die("pt_postprocess", pt_last_error(ctx));
        FILE* f = fopen(argv[2], "wb"); if (!f) die("fopen", argv[2]);
        fprintf(f, "PF\n%u %u\n-1.0\n", W, H);
        for (uint32_t y = 0; y < H; y++) for (uint32_t x = 0; x < W; x++) fwrite(&rgba[((size_t)(H - 1 - y) * W + x) * 4], sizeof(float), 3, f);
        fclose(f);
    }
    pt_stats st; pt_get_stats(ctx, &st);
    printf("render_pfm ok: %u triangles, %llu samples, %llu rays -> %s\n", desc.num_triangles, (unsigned long long)st.samples,
           (unsigned long long)(st.extend_rays + st.shadow_rays), argv[2]);
    pt_destroy(ctx); pts_scene_free(sc);
    return 0;
}
