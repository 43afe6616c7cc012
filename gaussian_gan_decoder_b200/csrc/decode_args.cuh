// Argument block of the fused decode forward kernel.
#pragma once
#include "decode.cuh"

namespace ggd {

struct DecodeFwdArgs {
    DecoderLayout L;
    const float* planes;   // channels-last [3][D][H][W][32]
    int arch;              // 0 eg3d (bilinear planes), 1 panohead (trilinear grids)
    int D, H, W;
    float coord_scale;     // 2 / box_warp
    int pos_freqs;
    float scale_bias;      // -2 (parallel, sequential) or -2.5 (reverse)
    const float* xyz;      // [P][3]
    int P;
    const float* packed;   // pack_weights_kernel output
    float* out[kHeads];    // raw attributes, reference shapes [P][3|3|4|1|3]
    float* act_scale;      // optional: exp(scale)      [P][3]
    float* act_rot;        // optional: normalize(rot)  [P][4]
    float* act_opacity;    // optional: sigmoid(opac)   [P][1]
};

int launch_decode_fwd(cudaStream_t s, const DecodeFwdArgs& a, int num_sms);
// tensor-core (fp16 tcgen05) flavour: decode_tc.cu
int launch_decode_tc(cudaStream_t s, const DecodeFwdArgs& a, int num_sms);
int launch_pack_weights_tc(cudaStream_t s, const DecoderLayout& L, const float* flat, void* packed);
size_t tc_packed_bytes(const DecoderLayout& L);

}  // namespace ggd
