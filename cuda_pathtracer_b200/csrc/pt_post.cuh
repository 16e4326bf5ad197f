// pt_post.cuh -- the reference's display post-process as CUDA kernels.  The reference runs
// these as GLSL compute / fragment shaders on the two RGBA32F textures (src/main.cpp:46-171, 316-376); the math is
// plain arithmetic, restated here statement by statement (graphics API: unavailable on this box):
//   k_post_horz     gauss_horz  (main.cpp:111-141): demodulate (luminance / max(albedo, 0.001) * nrSamples), 7-tap
//   k_post_vert     gauss_vert  (main.cpp:143-171): 8-tap (-4..3, as written), same kernel weights
//   k_post_resolve  quad_fs / quad_fs_blurred (main.cpp:46-109): divide by w, remodulate by albedo when blurred,
//                   gamma 2.0, vignette 1 - |uv - 0.5|^2 on all four components
// The hard-coded 640 / 480 bounds of the shaders (main.cpp:130,161) are the image width / height here.
#pragma once
#include "pt_device.cuh"

namespace ptb {

__device__ __forceinline__ float gaussWeight(int i, float spread) { return expf(-(float)(i * i) * 0.5f * spread) / sqrtf(2.0f * 3.1415926f); }

__global__ void __launch_bounds__(256) k_post_horz(const float4* __restrict__ lum, const float4* __restrict__ albedo, float4* dst, int W, int H, float nrSamples) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= W || y >= H) return;
    const float spread = fmaxf(0.7f, nrSamples / 200.0f);
    float sum = 0.0f; float3 colorSum = f3(0.0f);
    for (int i = -3; i < 4; i++) {
        if (x + i < 0 || x + i >= W) continue;
        const float4 l = lum[(size_t)y * W + x + i], a = albedo[(size_t)y * W + x + i];
        const float3 c = f3(l.x / fmaxf(a.x, 0.001f), l.y / fmaxf(a.y, 0.001f), l.z / fmaxf(a.z, 0.001f)) * nrSamples;
        const float g = gaussWeight(i, spread);
        colorSum += c * g; sum += g;
    }
    dst[(size_t)y * W + x] = make_float4(colorSum.x / sum, colorSum.y / sum, colorSum.z / sum, nrSamples);
}

__global__ void __launch_bounds__(256) k_post_vert(const float4* __restrict__ src, float4* dst, int W, int H, float nrSamples) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= W || y >= H) return;
    const float spread = fmaxf(0.7f, nrSamples / 200.0f);
    float sum = 0.0f; float3 colorSum = f3(0.0f);
    for (int i = -4; i < 4; i++) {
        if (y + i < 0 || y + i >= H) continue;
        const float4 l = src[(size_t)(y + i) * W + x];
        const float g = gaussWeight(i, spread);
        colorSum += f3(l) * g; sum += g;
    }
    dst[(size_t)y * W + x] = make_float4(colorSum.x / sum, colorSum.y / sum, colorSum.z / sum, nrSamples);
}

__global__ void __launch_bounds__(256) k_post_resolve(const float4* __restrict__ lum, const float4* __restrict__ albedo, float4* dst, int W, int H, int blurred) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= W || y >= H) return;
    const float u = ((float)x + 0.5f) / (float)W - 0.5f, v = ((float)y + 0.5f) / (float)H - 0.5f;     // uv of the texel centre
    const float4 l = lum[(size_t)y * W + x];
    float3 c = f3(l.x / l.w, l.y / l.w, l.z / l.w);
    if (blurred) { const float4 a = albedo[(size_t)y * W + x]; c = c * f3(a.x / a.w, a.y / a.w, a.z / a.w); }
    const float invGamma = 1.0f / 2.0f;
    const float vig = 1.0f - (u * u + v * v);
    dst[(size_t)y * W + x] = make_float4(powf(c.x, invGamma) * vig, powf(c.y, invGamma) * vig, powf(c.z, invGamma) * vig, vig);
}

// ---- L2 gather ceiling: random 16-byte or 32-byte reads over an L2-resident array, the access
// pattern of BVH-node / triangle fetches.  Every thread issues `iters` independent hashed loads (8 in flight).
template <int VEC>   // float4 loads per access: 1 = 16 B, 2 = 32 B
__global__ void __launch_bounds__(256) k_gather_probe(const float4* __restrict__ data, uint nElems /* power of two, in accesses */, uint iters, float* sink) {
    uint idx = wang_hash(blockIdx.x * blockDim.x + threadIdx.x);
    float acc = 0.0f;
    for (uint i = 0; i < iters; i += 8) {
        float4 v[8 * VEC];
#pragma unroll
        for (int k = 0; k < 8; k++) {
            idx = idx * 1664525u + 1013904223u;
            const uint e = (idx >> 7) & (nElems - 1u);
#pragma unroll
            for (int j = 0; j < VEC; j++) v[k * VEC + j] = __ldg(data + (size_t)e * VEC + j);
        }
#pragma unroll
        for (int k = 0; k < 8 * VEC; k++) acc += v[k].x + v[k].w;
    }
    if (acc == 123.456f) *sink = acc;
}

}  // namespace ptb
