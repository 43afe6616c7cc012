// extern "C" entry points of the decoder half (include/ggd.h).
#include "decode.cuh"

namespace ggd {

int launch_pack_weights(cudaStream_t s, const DecoderLayout& L, const float* flat, float* packed);
int launch_planes_channels_last(cudaStream_t s, const float* src, void* dst, int D, int H, int W, bool half_out);
}  // namespace ggd
#include "decode_args.cuh"
#include "decode_bwd.cuh"

namespace ggd { int debug_decode_profile(unsigned long long out[12]); }
using namespace ggd;

extern "C" int ggd_debug_decode_profile(uint64_t out[12])
{
    if (!out) { set_error("ggd_debug_decode_profile: null pointer"); return GGD_E_INVALID; }
    return debug_decode_profile(reinterpret_cast<unsigned long long*>(out));
}

namespace {
int check_cfg(const ggd_decoder_config* c, DecoderLayout* L)
{
    if (!c) { set_error("null decoder config"); return GGD_E_INVALID; }
    if (c->precision != 0 && c->precision != 1) { set_error("precision must be 0 (fp32) or 1 (fp16 tensor core)"); return GGD_E_INVALID; }
    if (c->hidden != kHidden) { set_error("hidden_dim %d unsupported: kernels are specialised for %d", c->hidden, kHidden); return GGD_E_INVALID; }
    if (c->arch != 0 && c->arch != 1) { set_error("arch must be 0 (eg3d) or 1 (panohead)"); return GGD_E_INVALID; }
    if (c->plane_depth < 1 || (c->arch == 0 && c->plane_depth != 1)) { set_error("bad plane_depth %d for arch %d", c->plane_depth, c->arch); return GGD_E_INVALID; }
    if (c->plane_h < 1 || c->plane_w < 1 || !(c->box_warp > 0.f)) { set_error("bad plane size / box_warp"); return GGD_E_INVALID; }
    if (!make_decoder_layout(c->variant, c->pos_freqs, L)) { set_error("unsupported decoder variant %d / pos_freqs %d", c->variant, c->pos_freqs); return GGD_E_INVALID; }
    return GGD_OK;
}
}  // namespace

extern "C" {

int ggd_decode_param_counts(const ggd_decoder_config* cfg, size_t* flat_floats, size_t* packed_floats)
{
    DecoderLayout L;
    if (int e = check_cfg(cfg, &L)) return e;
    if (flat_floats) *flat_floats = L.flat_count;
    if (packed_floats) *packed_floats = cfg->precision == 1 ? (tc_packed_bytes(L) + 3) / 4 : L.pk_count;
    return GGD_OK;
}

int ggd_decode_param_offset(const ggd_decoder_config* cfg, int head, int layer, int is_bias,
                            size_t* offset, int* fan_in, int* fan_out)
{
    DecoderLayout L;
    if (int e = check_cfg(cfg, &L)) return e;
    if (head < 0 || head >= kHeads || layer < 0 || layer >= kLayers || !offset) { set_error("bad head/layer"); return GGD_E_INVALID; }
    *offset = is_bias ? L.flat_b[head][layer] : L.flat_w[head][layer];
    if (fan_in) *fan_in = layer == 0 ? L.in_dim[head] : kHidden;
    if (fan_out) *fan_out = layer == kLayers - 1 ? head_out(head) : kHidden;
    return GGD_OK;
}

int ggd_decode_pack_weights(void* stream, const ggd_decoder_config* cfg, const float* flat_params, float* packed)
{
    DecoderLayout L;
    if (int e = check_cfg(cfg, &L)) return e;
    if (!flat_params || !packed) { set_error("ggd_decode_pack_weights: null pointer"); return GGD_E_INVALID; }
    if (cfg->precision == 1) return launch_pack_weights_tc(static_cast<cudaStream_t>(stream), L, flat_params, packed);
    return launch_pack_weights(static_cast<cudaStream_t>(stream), L, flat_params, packed);
}

int ggd_planes_channels_last(void* stream, const ggd_decoder_config* cfg, const float* planes_nchw, void* planes_cl)
{
    DecoderLayout L;
    if (int e = check_cfg(cfg, &L)) return e;
    if (!planes_nchw || !planes_cl) { set_error("ggd_planes_channels_last: null pointer"); return GGD_E_INVALID; }
    StageScope t(static_cast<cudaStream_t>(stream), kStPlanes);
    return launch_planes_channels_last(static_cast<cudaStream_t>(stream), planes_nchw, planes_cl,
                                       cfg->plane_depth, cfg->plane_h, cfg->plane_w, cfg->precision == 1);
}

int ggd_decode_forward(void* stream, const ggd_decoder_config* cfg, const void* planes_cl,
                       const float* xyz, int P, const float* packed,
                       float* out_xyz, float* out_scale, float* out_rotation,
                       float* out_opacity, float* out_color,
                       float* act_scale, float* act_rotation, float* act_opacity)
{
    DecodeFwdArgs a;
    if (int e = check_cfg(cfg, &a.L)) return e;
    if (P <= 0) { set_error("P must be positive (got %d)", P); return GGD_E_INVALID; }
    if (!planes_cl || !xyz || !packed || !out_xyz || !out_scale || !out_rotation || !out_opacity || !out_color) { set_error("ggd_decode_forward: null required pointer"); return GGD_E_INVALID; }
    if ((reinterpret_cast<uintptr_t>(planes_cl) & 15) || (reinterpret_cast<uintptr_t>(packed) & 15)) { set_error("planes_cl and packed must be 16-byte aligned"); return GGD_E_INVALID; }
    a.planes = static_cast<const float*>(planes_cl); a.arch = cfg->arch; a.D = cfg->plane_depth; a.H = cfg->plane_h; a.W = cfg->plane_w;
    a.coord_scale = 2.0f / cfg->box_warp; a.pos_freqs = cfg->pos_freqs; a.scale_bias = cfg->scale_bias;
    a.xyz = xyz; a.P = P; a.packed = packed;
    a.out[kXyz] = out_xyz; a.out[kScale] = out_scale; a.out[kRot] = out_rotation;
    a.out[kOpacity] = out_opacity; a.out[kColor] = out_color;
    a.act_scale = act_scale; a.act_rot = act_rotation; a.act_opacity = act_opacity;
    int dev = 0, sms = 0;
    GGD_CUDA_CHECK(cudaGetDevice(&dev));
    GGD_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    StageScope t(static_cast<cudaStream_t>(stream), kStDecode);
    if (cfg->precision == 1) return launch_decode_tc(static_cast<cudaStream_t>(stream), a, sms);
    return launch_decode_fwd(static_cast<cudaStream_t>(stream), a, sms);
}

size_t ggd_decode_backward_scratch_bytes(const ggd_decoder_config* cfg, int P)
{
    DecoderLayout L;
    if (check_cfg(cfg, &L) || P <= 0) return 0;
    const size_t simt = bwd_stage(L.pos_dim, (size_t)P).bytes;
    const size_t tc = align_up(bwd_stage_tc((size_t)P).bytes, 256) + align_up(tc_packed_bytes(L), 256) + align_up(bwd_tc_packed_bytes(), 256);
    return simt > tc ? simt : tc;
}

int ggd_decode_backward(void* stream, const ggd_decoder_config* cfg, const float* planes_cl_f32, const void* planes_cl_f16,
                        const float* xyz, int P, const float* flat_params, const float* packed_f32,
                        const float* out_xyz, const float* out_scale, const float* out_rotation,
                        const float* out_opacity, const float* out_color,
                        const float* dL_dxyz, const float* dL_dscale, const float* dL_drotation,
                        const float* dL_dopacity, const float* dL_dcolor,
                        void* scratch, size_t scratch_bytes, float* dflat, float* dplanes)
{
    DecodeBwdArgs a;
    if (int e = check_cfg(cfg, &a.L)) return e;
    if (P <= 0) { set_error("P must be positive (got %d)", P); return GGD_E_INVALID; }
    // the tensor-core chain samples the fp16 planes when they are given and then never touches the fp32 ones
    const bool f32_planes_unused = cfg && cfg->precision == 1 && cfg->pos_freqs == 0 && planes_cl_f16 != nullptr;
    if ((!planes_cl_f32 && !f32_planes_unused) || !xyz || !flat_params || !packed_f32 || !scratch || !dflat) { set_error("ggd_decode_backward: null required pointer"); return GGD_E_INVALID; }
    if (a.L.variant != 0 && (!out_xyz || !out_scale || !out_rotation || !out_opacity || !out_color)) { set_error("ggd_decode_backward: chained variants need the forward results"); return GGD_E_INVALID; }
    const size_t need = ggd_decode_backward_scratch_bytes(cfg, P);
    if (scratch_bytes < need) { set_error("ggd_decode_backward: scratch too small (%zu < %zu)", scratch_bytes, need); return GGD_E_CAPACITY; }
    if (reinterpret_cast<uintptr_t>(scratch) & 127) { set_error("scratch must be 128-byte aligned"); return GGD_E_INVALID; }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    a.planes = planes_cl_f32; a.arch = cfg->arch; a.D = cfg->plane_depth; a.H = cfg->plane_h; a.W = cfg->plane_w;
    a.coord_scale = 2.0f / cfg->box_warp; a.pos_freqs = cfg->pos_freqs; a.xyz = xyz; a.P = P;
    a.packed = packed_f32; a.flat = flat_params;
    a.out[kXyz] = out_xyz; a.out[kScale] = out_scale; a.out[kRot] = out_rotation; a.out[kOpacity] = out_opacity; a.out[kColor] = out_color;
    a.dout[kXyz] = dL_dxyz; a.dout[kScale] = dL_dscale; a.dout[kRot] = dL_drotation; a.dout[kOpacity] = dL_dopacity; a.dout[kColor] = dL_dcolor;
    a.stage = static_cast<uint8_t*>(scratch); a.dflat = dflat; a.dplanes = dplanes;
    GGD_CUDA_CHECK(cudaMemsetAsync(dflat, 0, sizeof(float) * a.L.flat_count, s));
    if (dplanes) GGD_CUDA_CHECK(cudaMemsetAsync(dplanes, 0, sizeof(float) * 3 * (size_t)kFeat * cfg->plane_depth * cfg->plane_h * cfg->plane_w, s));
    int dev = 0, sms = 0;
    GGD_CUDA_CHECK(cudaGetDevice(&dev));
    GGD_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (cfg->precision == 1 && cfg->pos_freqs == 0) {
        // tensor-core chain: stage | forward fp16 pack | transposed fp16 pack, all inside the scratch buffer
        DecodeBwdTcArgs t;
        t.b = a;
        uint8_t* pf = a.stage + align_up(bwd_stage_tc((size_t)P).bytes, 256);
        uint8_t* pb = pf + align_up(tc_packed_bytes(a.L), 256);
        if (int e = launch_pack_weights_tc(s, a.L, flat_params, pf)) return e;
        if (int e = launch_pack_weights_bwd_tc(s, a.L, flat_params, pb)) return e;
        t.packed_fwd = pf; t.packed_bwd = pb;
        t.planes16 = static_cast<const __half*>(planes_cl_f16);
        if (int e = launch_decode_bwd_chain_tc(s, t, sms)) return e;
        return launch_decode_bwd_dw_mn(s, a.L, a.stage, (size_t)P, dflat, sms);
    }
    if (int e = launch_decode_bwd_chain(s, a, sms)) return e;
    return launch_decode_bwd_dw(s, a.L, a.stage, (size_t)P, dflat, sms);
}

}  // extern "C"
