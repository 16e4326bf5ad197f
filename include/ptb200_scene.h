/* ptb200_scene.h -- host-side scene surface of libptb200.so (plain C++ inside, no GPU needed).
 *
 * Mirrors the reference's `Scene` (src/scene.h:120-402): addMaterial / addModel / addObject /
 * addSphere / addPlane / finalize / update, with the binned-SAH BVH2 builder (src/bvhBuilder.h:35-274),
 * the agglomerative top-level BVH (src/scene.h:81-118), instance matrices (src/scene.h:9-25) and the
 * camera frame (src/types.h:590-600).  pts_get_desc() yields the pt_scene_desc that pt_upload_scene takes.
 */
#ifndef PTB200_SCENE_H
#define PTB200_SCENE_H
#include "ptb200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pts_scene pts_scene;

PT_API pts_scene* pts_scene_new(void);
PT_API void pts_scene_free(pts_scene*);
PT_API const char* pts_last_error(const pts_scene*);
PT_API double pts_last_build_ms(const pts_scene*);                /* wall time of the last BVH2 build */
/* Where models added from now on get their BVH2 built: device >= 0 -> pt_build_bvh on that GPU (same tree),
 * -1 (default) -> the host builder.  A failed device build fails the add_model call; it never falls back silently. */
PT_API void pts_set_build_device(pts_scene*, int device);

PT_API uint32_t pts_add_material(pts_scene*, const pt_material*);  /* Scene::addMaterial, scene.h:140-144 */
PT_API int pts_set_material(pts_scene*, uint32_t idx, const pt_material*);
/* Scene::addModel (scene.h:159-347) on an OBJ file; returns the model id or a negative pt_status. */
PT_API int pts_add_model_obj(pts_scene*, const char* path, float scale, const float* rot3, const float* off3, uint32_t material, int use_mtl);
/* Same tail (BVH build, Model record) for in-memory triangles: v = n*9 floats, nrm = n*3 or NULL
 * (=> normalize(cross(e1,e2)), scene.h:297-299), uv = n*6 or NULL, mats = n or NULL (=> `material`). */
PT_API int pts_add_model_soup(pts_scene*, const float* v, const float* nrm, const float* uv, const uint32_t* mats, uint32_t n, uint32_t material);
/* File names the MTL gave for map_Kd (normal_map = 0) / norm (normal_map = 1); "" when none.  The caller
 * decodes the image and stores a pt_texture_create handle in the material (pts_set_material). */
PT_API const char* pts_material_texture_name(const pts_scene*, uint32_t material, int normal_map);

PT_API uint32_t pts_add_object(pts_scene*, uint32_t model, uint32_t material, const float* pos3, const float* rot3, const float* scale3);  /* GameObject, types.h:416-429 */
PT_API int pts_set_object(pts_scene*, uint32_t idx, const float* pos3, const float* rot3, const float* scale3);
PT_API void pts_add_sphere(pts_scene*, const float* pos3, float radius, uint32_t material);
PT_API void pts_add_plane(pts_scene*, const float* normal3, float d, uint32_t material);
PT_API int pts_finalize(pts_scene*);   /* Scene::finalize (scene.h:354-365) + first update */
PT_API int pts_update(pts_scene*);     /* instances + top-level BVH (scene.h:396-400) */
PT_API int pts_get_desc(pts_scene*, pt_scene_desc* out);   /* pointers stay valid until the scene changes */

/* Camera ctor + update() with no key pressed (types.h:609-637).  in9 = eye3, viewDir3, d, focalLength, aperture */
PT_API void pts_camera_make(const float* in9, uint32_t width, uint32_t height, pt_camera* out);

#ifdef __cplusplus
}
#endif
#endif
