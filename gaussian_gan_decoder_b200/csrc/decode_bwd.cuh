// Decoder backward: shared declarations (decode_bwd.cu = chain kernel, decode_bwd_dw.cu = weight-gradient GEMMs).
#pragma once
#include "decode.cuh"
#include <cuda_fp16.h>

namespace ggd {

constexpr int kBwdPts = 64;   // points per tile of the backward chain kernel == K of one staged operand tile

// byte offset of element (row, col) in a K-major no-swizzle fp16 operand with K_pad columns
__host__ __device__ __forceinline__ uint32_t canon_off(uint32_t r, uint32_t k, uint32_t K_pad)
{
    return (r >> 3) * (K_pad * 16) + (k >> 3) * 128 + (r & 7) * 16 + (k & 7) * 2;
}

// Staging area between the two kernels.  For every head h, hidden layer l in {0,1,2} and 64-point
// tile t the chain kernel writes two fp16 MMA operands, both K-major with K = the 64 points:
//   A[h][l][t] : dZ_l^T  [128 rows (output unit n)] x [64 points]
//   B[h][l][t] : H_l^T   [rows(l) (input unit k, incl. the constant-one row)] x [64 points]
// so that dW_l[n][k] = sum over tiles of A * B^T is a plain K-major x K-major tensor-core GEMM.
struct BwdStage {
    int k0;                       // rows of B for layer 0 (48 or 112); 144 for layers 1, 2
    size_t tiles;
    size_t a_tile, b_tile[3];     // bytes per tile
    size_t a_off[kHeads][3], b_off[kHeads][3];   // byte offsets of the [tile] arrays
    size_t bytes;
};
__host__ __device__ inline BwdStage bwd_stage(int pos_dim, size_t P)
{
    BwdStage s;
    s.k0 = pos_dim == 3 ? 48 : 112;
    s.tiles = (P + kBwdPts - 1) / kBwdPts;
    s.a_tile = (size_t)kHidden * kBwdPts * 2;
    s.b_tile[0] = (size_t)s.k0 * kBwdPts * 2;
    s.b_tile[1] = s.b_tile[2] = (size_t)(kHidden + 16) * kBwdPts * 2;
    size_t o = 0;
    for (int h = 0; h < kHeads; ++h)
        for (int l = 0; l < 3; ++l) {
            s.a_off[h][l] = o; o += s.a_tile * s.tiles;
            s.b_off[h][l] = o; o += s.b_tile[l] * s.tiles;
        }
    s.bytes = o;
    return s;
}

struct DecodeBwdArgs {
    DecoderLayout L;
    const float* planes;    // fp32 channels-last [3][D][H][W][32]
    int arch, D, H, W;
    float coord_scale;
    int pos_freqs;
    const float* xyz;       // [P][3]
    int P;
    const float* packed;    // fp32 forward pack (W^T, zero padded)
    const float* flat;      // nn.Linear-layout parameters
    const float* out[kHeads];    // saved forward results (chained inputs of later heads)
    const float* dout[kHeads];   // dL/d result, any may be null
    uint8_t* stage;         // BwdStage area
    float* dflat;           // [flat_count], last-layer weights/biases accumulated here by the chain kernel
    float* dplanes;         // optional: dL/d feature_planes in the generator's NCHW layout [3][32*D][H][W]
};

int launch_decode_bwd_chain(cudaStream_t s, const DecodeBwdArgs& a, int num_sms);
int launch_decode_bwd_dw(cudaStream_t s, const DecoderLayout& L, const uint8_t* stage, size_t P, float* dflat, int num_sms);

}  // namespace ggd
