    // (included inside decode_bwd_chain_tc_kernel, after the epilogue role)
    else if (warp == 16) {
        // ================================= MMA issuer (+ staging stores) =================================
        if (lane == 0) {
            const uint32_t sb = ptx::smem_u32(smem);
            const uint32_t id128 = ptx::idesc_f16(kM, 128), id16 = ptx::idesc_f16(kM, 16), id48 = ptx::idesc_f16(kM, kK0);
            const uint32_t id_dw3 = ptx::idesc_f16(128, 16) | (1u << 15) | (1u << 16);   // both operands MN-major (K = points)
            uint32_t cnt = 0, n_ready = 0;
            // descriptors built once per call, K steps unrolled at compile time (ptx::mma_f16_ss_steps)
            auto mma_k = [&](uint32_t d, uint32_t a_addr, uint32_t a_K, uint32_t b_addr, uint32_t b_K, uint32_t idesc, auto ksteps, bool acc0) {
                ptx::mma_f16_ss_steps<decltype(ksteps)::value>(d, ptx::smem_desc(a_addr, 128, a_K * 16), ptx::smem_desc(b_addr, 128, b_K * 16),
                                                                16, 16, idesc, acc0);
            };
            using std::integral_constant;
            for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
                for (int hi = kHeads - 1; hi >= 0; --hi) {
                    const int h = L.order[hi];
                    for (int s = 0; s < 8; ++s, ++cnt) {
                        const uint32_t slot = cnt % kWSlots;
                        const uint32_t w_addr = sb + kOffW + slot * kSlotBytes;
                        PROF_T(t0);
                        ptx::mbar_wait(&bars[BW_FULL + slot], (cnt / kWSlots) & 1);
                        PROF_ADD(3, t0);
                        PROF_T(t1);
                        ptx::mbar_wait(&bars[B_READY], n_ready & 1);
                        PROF_ADD(4, t1);
                        PROF_T(t2);
                        ++n_ready;
                        ptx::tc_fence_after();
                        bool stored = false;
                        if (s == 0 && hi == kHeads - 1) { bulk_s2g(a.stage + st.x0_off + (size_t)tile * st.x0_tile, smem + kOffX0, (uint32_t)st.x0_tile); stored = true; }
                        if (s == 1 || s == 2) { bulk_s2g(a.stage + st.h_off[h][s - 1] + (size_t)tile * st.h_tile, smem + kOffA, (uint32_t)st.h_tile); stored = true; }
                        if (s >= 5) { const int l = 7 - s; bulk_s2g(a.stage + st.dz_off[h][l] + (size_t)tile * st.dz_tile, smem + kOffDZ, (uint32_t)st.dz_tile); stored = true; }
                        if (stored) bulk_commit();
                        const uint32_t dacc = tmem_base + kColAcc;
                        switch (s) {
                            case 0: mma_k(dacc, sb + kOffX0, kK0, w_addr, kK0, id128, integral_constant<int, kK0 / 16>{}, false); break;
                            case 1: case 2: mma_k(dacc, sb + kOffA, kKH, w_addr, kKH, id128, integral_constant<int, kKH / 16>{}, false); break;
                            case 3: mma_k(tmem_base + kColY, sb + kOffA, kKH, w_addr, kKH, id16, integral_constant<int, kKH / 16>{}, false); break;
                            case 4:
                                mma_k(dacc, sb + kOffDY, 16, w_addr, 16, id128, integral_constant<int, 1>{}, false);
                                // dW3[k][j] += sum over the tile's points H3[p][k] dY[p][j]: H3 is still in the hidden operand
                                // buffer ([point][unit], 144 columns), dY is [point][16]; contraction over the strided index
                                ptx::mma_f16_ss_steps<kM / 16>(tmem_base + col_dw3(hi), ptx::smem_desc(sb + kOffA, kKH * 16, 128),
                                                               ptx::smem_desc(sb + kOffDY, 256, 128), 2 * (kKH * 16) / 16, 2 * 256 / 16, id_dw3,
                                                               tile != (int)blockIdx.x);
                                break;
                            case 5: case 6: mma_k(dacc, sb + kOffDZ, kHidden, w_addr, kHidden, id128, integral_constant<int, kHidden / 16>{}, false); break;
                            default: mma_k(tmem_base + kColDX, sb + kOffDZ, kHidden, w_addr, kHidden, id48, integral_constant<int, kHidden / 16>{}, hi != kHeads - 1); break;
                        }
                        PROF_T(t3);
                        if (stored) bulk_wait_read();      // measured: committing first and letting the epilogue wait for the
                        PROF_ADD(5, t3);                   // drain right before it overwrites the source changes nothing
                        ptx::mma_commit(&bars[B_ACC]);
                        ptx::mma_commit(&bars[BW_EMPTY + slot]);
                        PROF_ADD(6, t2);
                    }
                }
            }
            bulk_wait_all();
        }
    } else {
        // ================================ weight loader ================================
        if (lane == 0) {
            uint32_t cnt = 0;
            for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
                for (int hi = kHeads - 1; hi >= 0; --hi) {
                    const int h = L.order[hi];
                    for (int s = 0; s < 8; ++s, ++cnt) {
                        const uint32_t slot = cnt % kWSlots;
                        if (cnt >= (uint32_t)kWSlots) ptx::mbar_wait(&bars[BW_EMPTY + slot], ((cnt / kWSlots) - 1) & 1);
                        const StepDesc d = step_weights(s);
                        const uint8_t* src = s < 4 ? args.packed_fwd + (size_t)h * kFwdPackHead + d.src_off
                                                   : args.packed_bwd + (size_t)h * kBwdPackHead + d.src_off;
                        ptx::mbar_arrive_expect_tx(&bars[BW_FULL + slot], d.bytes);
                        ptx::bulk_g2s(smem + kOffW + slot * kSlotBytes, src, d.bytes, &bars[BW_FULL + slot]);
                    }
                }
            }
        }
    }
