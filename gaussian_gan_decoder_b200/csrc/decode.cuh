// Shared description of the five-head decoder (host + device).
#pragma once
#include "common.cuh"

namespace ggd {

// canonical head ids == order of the reference's get_params_custom()
// (main/decoder_models/parallel_decoder.py:82-92): xyz, scale, rotation, opacity, color
enum Head { kXyz = 0, kScale = 1, kRot = 2, kOpacity = 3, kColor = 4, kHeads = 5 };
#ifdef __CUDACC__
__host__ __device__
#endif
constexpr int head_out(int h) { return h == 2 ? 4 : (h == 3 ? 1 : 3); }
constexpr int kFeat = 32;     // tri-plane channels per sample
constexpr int kHidden = 128;  // hidden width the kernels are specialised for
constexpr int kLayers = 4;
constexpr int kMaxIn = 128;   // layer-0 fan-in after padding (32 + 63 + 11 = 106 max)

struct DecoderLayout {
    int variant;          // 0 parallel, 1 sequential, 2 sequential_reverse
    int pos_dim;          // 3, or 63 with the 10-frequency embedding
    int order[kHeads];    // execution order (chained variants feed earlier outputs forward)
    int in_dim[kHeads];   // layer-0 fan-in of each head (indexed by head id)
    int in_pad[kHeads];   // rounded up to a multiple of 32
    // flat nn.Linear-layout parameter buffer (what Adam / the all-reduce see)
    size_t flat_w[kHeads][kLayers], flat_b[kHeads][kLayers], flat_count;
    // packed fp32 buffer of the SIMT kernel: W^T as [K_pad][N_pad] + bias[N_pad]
    size_t pk_w[kHeads][kLayers], pk_b[kHeads][kLayers], pk_count;
};

// Fills the layout; returns false for an unsupported configuration.
bool make_decoder_layout(int variant, int pos_freqs, DecoderLayout* L);

}  // namespace ggd
