// Backward of the GaussianModel getters between the decoder and the rasteriser:
//   scaling = exp(_scaling), rotation = F.normalize(_rotation), opacity = sigmoid(_opacity)
// (gaussian_splatting/scene/gaussian_model.py:99-121; in the reference this is three autograd nodes
// and their elementwise kernels).  One pass over the P Gaussians, in place: the gradients with respect to
// the ACTIVATED values, as ggd_raster_backward produced them, become gradients with respect to the raw
// decoder outputs, which is what ggd_decode_backward takes.  HBM-bound: 64 B read + 32 B written per Gaussian.
#include "common.cuh"

namespace ggd {
namespace {

__global__ void __launch_bounds__(256)
activation_backward_kernel(int P, const float* __restrict__ act_scale, const float4* __restrict__ rotation,
                           const float* __restrict__ act_opacity, float* __restrict__ g_scale,
                           float4* __restrict__ g_rot, float* __restrict__ g_opacity)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    // d exp(s) / ds = exp(s)
#pragma unroll
    for (int k = 0; k < 3; ++k) g_scale[3 * i + k] *= act_scale[3 * i + k];
    // d sigmoid(o) / do = sigma (1 - sigma)
    const float sg = act_opacity[i];
    g_opacity[i] *= sg * (1.0f - sg);
    // y = r / max(|r|, eps), eps = 1e-12 (torch.nn.functional.normalize): dr = (g - y (y . g)) / |r| above eps
    const float4 r = rotation[i];
    const float4 g = g_rot[i];
    const float n = sqrtf(r.x * r.x + r.y * r.y + r.z * r.z + r.w * r.w);
    float4 d;
    if (n > 1e-12f) {
        const float inv = 1.0f / n;
        const float4 y = make_float4(r.x * inv, r.y * inv, r.z * inv, r.w * inv);
        const float dot = y.x * g.x + y.y * g.y + y.z * g.z + y.w * g.w;
        d = make_float4((g.x - y.x * dot) * inv, (g.y - y.y * dot) * inv, (g.z - y.z * dot) * inv, (g.w - y.w * dot) * inv);
    } else {
        d = make_float4(g.x * 1e12f, g.y * 1e12f, g.z * 1e12f, g.w * 1e12f);
    }
    g_rot[i] = d;
}

}  // namespace
}  // namespace ggd

using namespace ggd;

extern "C" int ggd_activation_backward(void* stream, int P, const float* act_scale, const float* rotation,
                                       const float* act_opacity, float* g_scale, float* g_rotation, float* g_opacity)
{
    if (P <= 0 || !act_scale || !rotation || !act_opacity || !g_scale || !g_rotation || !g_opacity) {
        set_error("ggd_activation_backward: bad arguments");
        return GGD_E_INVALID;
    }
    if ((reinterpret_cast<uintptr_t>(rotation) & 15) || (reinterpret_cast<uintptr_t>(g_rotation) & 15)) {
        set_error("ggd_activation_backward: rotation arrays must be 16-byte aligned");
        return GGD_E_INVALID;
    }
    activation_backward_kernel<<<(P + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        P, act_scale, reinterpret_cast<const float4*>(rotation), act_opacity, g_scale,
        reinterpret_cast<float4*>(g_rotation), g_opacity);
    GGD_LAUNCH_CHECK("activation_backward_kernel");
    return GGD_OK;
}
