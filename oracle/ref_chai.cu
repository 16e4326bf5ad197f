// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.
// oracle/_ref/libptref_chai.so: the reference's OWN scene-script binding, getScriptedScene (src/sceneBuilder.h:271-311), with
// ChaiScript 6.1.0 as vendored under /root/reference/chaiscript, compiled UNMODIFIED where it lies.  Everything of
// ref_harness.cu is part of this library too, so the resulting Scene is read back through the same ref_counts / ref_ptr.
// tests/test_cpu_scene_script.py runs the reference's example_scene.chai / sponza.chai through this binding and through the
// product's interpreter (cuda_pathtracer_b200/scene_script.py) and compares the arrays.
#define CHAI
#include "ref_harness.cu"
#include "sceneBuilder.h"

// Runs the script in the current working directory (model paths in scripts are relative, as for the reference binary).
// Returns a RefScene* (finalized, instances + TLAS updated) or nullptr with the ChaiScript error on stderr.
REF_API void* ref_script_scene(const char* filename) {
    try {
        RefScene* s = new RefScene{getScriptedScene(filename), true};
        ref_update(s);
        return s;
    } catch (const std::exception& e) {
        fprintf(stderr, "ref_script_scene: %s\n", e.what());
        return nullptr;
    }
}
