// Body of the decoder backward chain kernel (included by decode_bwd.cu).
namespace ggd {
namespace {

__global__ void __launch_bounds__(kThreads, 1)
decode_bwd_chain_kernel(const DecodeBwdArgs a)
{
    extern __shared__ __align__(16) float smem[];
    float* base = smem;                         // [128][64]  feat | pos | chained inputs | ... | 1
    float* z[3] = {base + kMaxIn * kPts, base + (kMaxIn + kHidden) * kPts, base + (kMaxIn + 2 * kHidden) * kPts};
    float* hA = z[2] + kHidden * kPts;          // [128][64]
    float* hB = hA + kHidden * kPts;            // [128][64]
    float* wbuf = hB + kHidden * kPts;          // [32][128]
    float* dfeat = wbuf + kKC * kHidden;        // [32][64]
    float* dchain = dfeat + kFeat * kPts;       // [16][64]
    float* dybuf = dchain + 16 * kPts;          // [4][64]
    const DecoderLayout& L = a.L;
    const BwdStage st = bwd_stage(L.pos_dim, (size_t)a.P);
    const int tid = threadIdx.x, pt = tid & (kPts - 1), q = tid >> 6;
    const int tiles = (int)st.tiles;
    const size_t plane_stride = (size_t)a.D * a.H * a.W * kFeat;
    const int chain0 = kFeat + L.pos_dim;
    const int k0 = st.k0;

    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int p0 = tile * kPts;
        __syncthreads();
        for (int i = tid; i < kMaxIn * kPts; i += kThreads) base[i] = 0.f;
        for (int i = tid; i < (kFeat + 16) * kPts; i += kThreads) dfeat[i] = 0.f;   // dfeat + dchain
        __syncthreads();
        // ---- sampling (fp32), 8 lanes per point ----
        {
            const int sub = tid & 7, grp = tid >> 3;
            for (int it = 0; it < kPts / 32; ++it) {
                const int lp = it * 32 + grp, gp = p0 + lp;
                if (gp < a.P) {
                    const float x = a.coord_scale * __ldg(a.xyz + 3 * (size_t)gp);
                    const float y = a.coord_scale * __ldg(a.xyz + 3 * (size_t)gp + 1);
                    const float zc = a.coord_scale * __ldg(a.xyz + 3 * (size_t)gp + 2);
                    float4 s0, s1, s2;
                    if (a.arch == 0) {
                        s0 = sample4(a.planes, 0, 1, a.H, a.W, x, y, 0.f, sub);
                        s1 = sample4(a.planes + plane_stride, 0, 1, a.H, a.W, x, zc, 0.f, sub);
                        s2 = sample4(a.planes + 2 * plane_stride, 0, 1, a.H, a.W, zc, x, 0.f, sub);
                    } else {
                        s0 = sample4(a.planes, 1, a.D, a.H, a.W, x, y, zc, sub);
                        s1 = sample4(a.planes + plane_stride, 1, a.D, a.H, a.W, x, zc, y, sub);
                        s2 = sample4(a.planes + 2 * plane_stride, 1, a.D, a.H, a.W, y, zc, x, sub);
                    }
                    base[(4 * sub + 0) * kPts + lp] = (s0.x + s1.x + s2.x) / 3.f;
                    base[(4 * sub + 1) * kPts + lp] = (s0.y + s1.y + s2.y) / 3.f;
                    base[(4 * sub + 2) * kPts + lp] = (s0.z + s1.z + s2.z) / 3.f;
                    base[(4 * sub + 3) * kPts + lp] = (s0.w + s1.w + s2.w) / 3.f;
                }
            }
        }
        // ---- position rows, chained inputs from the saved forward results, constant one ----
        if (tid < kPts) {
            const int gp = p0 + tid;
            if (gp < a.P) {
                float px[3];
                for (int j = 0; j < 3; ++j) { px[j] = __ldg(a.xyz + 3 * (size_t)gp + j); base[(kFeat + j) * kPts + tid] = px[j]; }
                for (int f = 0; f < a.pos_freqs; ++f) {
                    const float freq = 1.0f + (float)f * (511.0f / 9.0f);
                    for (int j = 0; j < 3; ++j) {
                        base[(kFeat + 3 + 6 * f + j) * kPts + tid] = sinf(px[j] * freq);
                        base[(kFeat + 6 + 6 * f + j) * kPts + tid] = cosf(px[j] * freq);
                    }
                }
                if (L.variant != 0) {
                    int col = chain0;
                    for (int hi = 0; hi < kHeads - 1; ++hi) {
                        const int h = L.order[hi], n = head_out(h);
                        for (int j = 0; j < n; ++j) base[(col + j) * kPts + tid] = __ldg(a.out[h] + (size_t)gp * n + j);
                        col += n;
                    }
                }
            }
            base[(k0 - 1) * kPts + tid] = 1.0f;     // the constant-one row of the layer-0 operand (bias gradient)
        }
        __syncthreads();

        // ---- heads in reverse chaining order ----
        for (int hi = kHeads - 1; hi >= 0; --hi) {
            const int h = L.order[hi], n_out = head_out(h);
            int my_chain = chain0;                       // chain column of this head's first output
            for (int e = 0; e < hi; ++e) my_chain += head_out(L.order[e]);
            const float* pk = a.packed;
            uint8_t* stB[3] = {a.stage + st.b_off[h][0] + (size_t)tile * st.b_tile[0],
                               a.stage + st.b_off[h][1] + (size_t)tile * st.b_tile[1],
                               a.stage + st.b_off[h][2] + (size_t)tile * st.b_tile[2]};
            // forward recompute: z_l kept; H_l (the input of layer l, with its constant-one row) staged
            store_tile(stB[0], base, k0, k0, -1);
            dense_fwd(base, L.in_pad[h], pk + L.pk_w[h][0], pk + L.pk_b[h][0], z[0], hA, wbuf);
            __syncthreads();
            store_tile(stB[1], hA, kHidden, kHidden + 16, kHidden);
            dense_fwd(hA, kHidden, pk + L.pk_w[h][1], pk + L.pk_b[h][1], z[1], hB, wbuf);
            __syncthreads();
            store_tile(stB[2], hB, kHidden, kHidden + 16, kHidden);
            dense_fwd(hB, kHidden, pk + L.pk_w[h][2], pk + L.pk_b[h][2], z[2], hA, wbuf);
            __syncthreads();
            // last layer (recomputed for the activation derivative) and dL/dy
            for (int i = tid; i < kHidden * 4; i += kThreads) wbuf[i] = __ldg(pk + L.pk_w[h][3] + i);
            __syncthreads();
            if (tid < kPts) {
                const int gp = p0 + tid;
                float dy[4] = {0.f, 0.f, 0.f, 0.f};
                if (gp < a.P) {
                    float y[4];
                    const float* b3 = pk + L.pk_b[h][3];
                    for (int j = 0; j < 4; ++j) y[j] = __ldg(b3 + j);
                    for (int k = 0; k < kHidden; ++k) {
                        const float x = hA[k * kPts + tid];
                        const float4 w = reinterpret_cast<const float4*>(wbuf)[k];
                        y[0] = fmaf(x, w.x, y[0]); y[1] = fmaf(x, w.y, y[1]); y[2] = fmaf(x, w.z, y[2]); y[3] = fmaf(x, w.w, y[3]);
                    }
                    for (int j = 0; j < n_out; ++j) {
                        float gsum = a.dout[h] ? __ldg(a.dout[h] + (size_t)gp * n_out + j) : 0.f;
                        if (L.variant != 0 && hi < kHeads - 1) gsum += dchain[(my_chain - chain0 + j) * kPts + tid];
                        float d = 1.f;
                        if (h == kXyz) d = 0.01f;
                        if (h == kScale) d = -1.0f / (1.0f + expf(-(y[j] + 5.0f)));   // d/dy of -softplus(y+5)+c
                        dy[j] = gsum * d;
                    }
                }
                for (int j = 0; j < 4; ++j) dybuf[j * kPts + tid] = dy[j];
            }
            __syncthreads();
            // last layer weight / bias gradient: reduced over the tile here (4 x 129 values)
            if (tid < kHidden) {
                for (int j = 0; j < n_out; ++j) {
                    float sacc = 0.f;
                    for (int p = 0; p < kPts; ++p) sacc = fmaf(dybuf[j * kPts + p], hA[tid * kPts + p], sacc);
                    atomicAdd(a.dflat + L.flat_w[h][3] + (size_t)j * kHidden + tid, sacc);
                }
            } else if (tid < kHidden + n_out) {
                const int j = tid - kHidden;
                float sacc = 0.f;
                for (int p = 0; p < kPts; ++p) sacc += dybuf[j * kPts + p];
                atomicAdd(a.dflat + L.flat_b[h][3] + j, sacc);
            }
            // dZ2 = (W3^T dy) * gelu'(z2), in place
            {
                const float* W3 = a.flat + L.flat_w[h][3];
                for (int j = 0; j < kOutPerThread; ++j) {
                    const int k = q * kOutPerThread + j;
                    float dh = 0.f;
                    for (int o = 0; o < n_out; ++o) dh = fmaf(__ldg(W3 + o * kHidden + k), dybuf[o * kPts + pt], dh);
                    z[2][k * kPts + pt] = dh * gelu_grad(z[2][k * kPts + pt]);
                }
            }
            __syncthreads();
            float acc[kOutPerThread];
            for (int l = 2; l >= 0; --l) {
                store_tile(a.stage + st.a_off[h][l] + (size_t)tile * st.a_tile, z[l], kHidden, kHidden, -1);
                const int in = l == 0 ? L.in_dim[h] : kHidden;
                dense_bwd(z[l], a.flat + L.flat_w[h][l], in, acc, wbuf);
                if (l > 0) {
#pragma unroll
                    for (int j = 0; j < kOutPerThread; ++j) {
                        const int k = q * kOutPerThread + j;
                        z[l - 1][k * kPts + pt] = acc[j] * gelu_grad(z[l - 1][k * kPts + pt]);
                    }
                } else {
                    // dX0: features accumulate over heads; chained columns go to earlier heads' outputs
#pragma unroll
                    for (int j = 0; j < kOutPerThread; ++j) {
                        const int k = q * kOutPerThread + j;
                        if (k < kFeat) dfeat[k * kPts + pt] += acc[j];
                        else if (k >= chain0 && k < in) dchain[(k - chain0) * kPts + pt] += acc[j];
                    }
                }
                __syncthreads();
            }
        }

        // ---- feature gradient -> plane gradient (mean over 3 planes, bilinear / trilinear scatter) ----
        if (a.dplanes) {
            const int sub = tid & 7, grp = tid >> 3;
            const size_t gplane = (size_t)kFeat * a.D * a.H * a.W;      // one plane in the generator layout
            for (int it = 0; it < kPts / 32; ++it) {
                const int lp = it * 32 + grp, gp = p0 + lp;
                if (gp >= a.P) continue;
                const float x = a.coord_scale * __ldg(a.xyz + 3 * (size_t)gp);
                const float y = a.coord_scale * __ldg(a.xyz + 3 * (size_t)gp + 1);
                const float zc = a.coord_scale * __ldg(a.xyz + 3 * (size_t)gp + 2);
                const float4 g = make_float4(dfeat[(4 * sub) * kPts + lp] / 3.f, dfeat[(4 * sub + 1) * kPts + lp] / 3.f,
                                             dfeat[(4 * sub + 2) * kPts + lp] / 3.f, dfeat[(4 * sub + 3) * kPts + lp] / 3.f);
                if (a.arch == 0) {
                    scatter4(a.dplanes, 0, 1, a.H, a.W, x, y, 0.f, sub, g);
                    scatter4(a.dplanes + gplane, 0, 1, a.H, a.W, x, zc, 0.f, sub, g);
                    scatter4(a.dplanes + 2 * gplane, 0, 1, a.H, a.W, zc, x, 0.f, sub, g);
                } else {
                    scatter4(a.dplanes, 1, a.D, a.H, a.W, x, y, zc, sub, g);
                    scatter4(a.dplanes + gplane, 1, a.D, a.H, a.W, x, zc, y, sub, g);
                    scatter4(a.dplanes + 2 * gplane, 1, a.D, a.H, a.W, y, zc, x, sub, g);
                }
            }
        }
    }
}

}  // namespace

size_t decode_bwd_smem_bytes()
{
    return sizeof(float) * (size_t)((kMaxIn + 5 * kHidden) * kPts + kKC * kHidden + (kFeat + 16 + 4) * kPts);
}

int launch_decode_bwd_chain(cudaStream_t s, const DecodeBwdArgs& a, int num_sms)
{
    const size_t smem = decode_bwd_smem_bytes();
    GGD_CUDA_CHECK(cudaFuncSetAttribute(decode_bwd_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int tiles = (a.P + kPts - 1) / kPts;
    decode_bwd_chain_kernel<<<tiles < num_sms ? tiles : num_sms, kThreads, smem, s>>>(a);
    GGD_LAUNCH_CHECK("decode_bwd_chain_kernel");
    return GGD_OK;
}

}  // namespace ggd
