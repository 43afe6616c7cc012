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

// tensor-core chain (decode_bwd_tc.cu): 128-point tiles, [point][feature] operands
struct BwdStageTc {
    size_t tiles, x0_tile, h_tile, dz_tile;
    size_t x0_off, h_off[kHeads][2], dz_off[kHeads][3];   // H_1, H_2 and dZ_0..2 per head; X_0 once
    size_t bytes;
};
__host__ __device__ inline BwdStageTc bwd_stage_tc(size_t P)
{
    BwdStageTc s;
    s.tiles = (P + 127) / 128;
    s.x0_tile = 128 * 48 * 2;
    s.h_tile = 128 * 144 * 2;
    s.dz_tile = 128 * 128 * 2;
    size_t o = 0;
    s.x0_off = o; o += s.x0_tile * s.tiles;
    for (int h = 0; h < kHeads; ++h) {
        for (int l = 0; l < 2; ++l) { s.h_off[h][l] = o; o += s.h_tile * s.tiles; }
        for (int l = 0; l < 3; ++l) { s.dz_off[h][l] = o; o += s.dz_tile * s.tiles; }
    }
    s.bytes = o;
    return s;
}
size_t bwd_tc_packed_bytes();
int launch_pack_weights_bwd_tc(cudaStream_t s, const DecoderLayout& L, const float* flat, void* packed);
struct DecodeBwdTcArgs {
    DecodeBwdArgs b;              // planes (fp32 channels-last), xyz, outs, douts, stage, dflat, dplanes
    const uint8_t* packed_fwd;    // decode_tc.cu forward pack (fp16, bias columns folded in)
    const uint8_t* packed_bwd;    // transposed fp16 pack (launch_pack_weights_bwd_tc)
    const __half* planes16;       // optional: the forward's fp16 channels-last planes; the recompute then samples what the
                                  // forward sampled (and half the bytes: the fp32 planes do not stay in L2 next to the staging traffic)
};
int launch_decode_bwd_chain_tc(cudaStream_t s, const DecodeBwdTcArgs& a, int num_sms);
int launch_decode_bwd_dw_mn(cudaStream_t s, const DecoderLayout& L, const uint8_t* stage, size_t P, float* dflat, int num_sms);

int launch_decode_bwd_chain(cudaStream_t s, const DecodeBwdArgs& a, int num_sms);
int launch_decode_bwd_dw(cudaStream_t s, const DecoderLayout& L, const uint8_t* stage, size_t P, float* dflat, int num_sms);

}  // namespace ggd
