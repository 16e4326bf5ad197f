// TEST INFRASTRUCTURE (only tests/ may load this).  The reference's image decoder -- stb_image v2.22, vendored at
// /root/reference/src/stb_image.h and used by loadTexture / loadTextureL / loadTextureHDR (src/use_cuda.h:108-263) --
// compiled where it lies, so that tests/test_cpu_textures.py can compare cuda_pathtracer_b200/textures.py with the bytes
// and floats the reference's loaders start from.  Built by `make -C oracle ref` into oracle/_ref/libstbref.so.
#define STB_IMAGE_IMPLEMENTATION
#include "stb_image.h"

extern "C" {
unsigned char* ref_stbi_load(const char* file, int* w, int* h, int* channels, int req) { return stbi_load(file, w, h, channels, req); }
// stbi_ldr_to_hdr_gamma(gamma) + stbi_loadf, the pair loadTextureL / loadTextureHDR call (use_cuda.h:169-170, 205-206)
float* ref_stbi_loadf(const char* file, int* w, int* h, int* channels, int req, float gamma) {
    stbi_ldr_to_hdr_gamma(gamma);
    return stbi_loadf(file, w, h, channels, req);
}
void ref_stbi_free(void* p) { stbi_image_free(p); }
}
