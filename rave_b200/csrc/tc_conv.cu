// tc_conv.cu -- the tcgen05 implicit-GEMM convolution kernel of the bf16 engine.
//
// One persistent CTA per SM, warp-specialised:
//   warps 0-3  epilogue    TMEM -> registers (tcgen05.ld) -> bias / residual / activation -> HBM
//                          (fused DilatedUnit: also writes act(conv3 + b) back to smem as the A
//                          operand of the 1x1 conv -- the intermediate never leaves the SM)
//   warps 4-7  A loaders   channels-last rows HBM -> registers -> pre-activation (LeakyReLU /
//                          Snake / AdaIN affine, zero padding, stream edges) -> smem in the
//                          UMMA K-major no-swizzle core-matrix layout
//   warp  8    MMA issuer  one thread: tcgen05.mma (M=128, N=N_TILE, K=16) per tap and K step;
//                          a dilated / shifted tap is only a different start row of the same
//                          smem tile (start address + 16 B per row), accumulators in TMEM
//   warp  9    W producer  weight blocks HBM/L2 -> smem with cp.async.bulk on an mbarrier ring,
//                          or once for the whole kernel when the layer's weights fit
// Stages hand over through mbarriers only; tcgen05.commit releases smem stages and publishes
// accumulators.  smem A tile:  addr(row, c8) = c8*(R_ALLOC*16) + row*16   (c8 = 8-channel group)
//        weight block [N][KC]: addr(n,   c8) = c8*(N*16)       + n*16
#include "common.cuh"
#include "tc_conv.cuh"
#include "tc_ptx.cuh"

namespace rb {
namespace tc {

namespace {

enum { BAR_A_FULL = 0, BAR_A_EMPTY = 4, BAR_W_FULL = 8, BAR_W_EMPTY = 12, BAR_D_FULL = 16,
       BAR_D_EMPTY = 18, BAR_H_FULL = 20, BAR_H_EMPTY = 22, BAR_D2_FULL = 24, BAR_D2_EMPTY = 26,
       BAR_COUNT = 28 };

__device__ __forceinline__ float act_f(float v, int kind, float alpha) {
  if (kind == RB_ACT_LEAKY) return v > 0.f ? v : 0.2f * v;
  if (kind == RB_ACT_SNAKE) {
    const float s = sinf(alpha * v);
    return v + (1.0f / (alpha + 1e-9f)) * (s * s);
  }
  return v;
}

__device__ __forceinline__ void unpack8(const uint4& q, float* f) {
  const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  return make_uint4(pack2(f[0], f[1]), pack2(f[2], f[3]), pack2(f[4], f[5]), pack2(f[6], f[7]));
}

struct Work {
  int nt, b, t0;
};
__device__ __forceinline__ Work decode(const TcParams& p, int w) {
  const int n_rt = p.B * p.tiles_per_stream;
  Work k;
  k.nt = w / n_rt;
  const int rt = w - k.nt * n_rt;
  k.b = rt / p.tiles_per_stream;
  k.t0 = (rt - k.b * p.tiles_per_stream) * p.MT * 128;
  return k;
}

// ------------------------------------------------------------------------------ A loader
// Rows [t0 - off, t0 - off + R) x channels [kc*KC, kc*KC + KC) of stream b -> smem stage.
__device__ __forceinline__ void load_a_chunk(const TcParams& p, const Work& wk, int kc, uint8_t* stage,
                                             int tid) {
  const int c8n = p.KC >> 3;
  const int units = p.R * c8n;
  const int row0 = wk.t0 - p.off;
  const size_t sbase = ((size_t)wk.b * p.x.ld + p.x.hist) * p.x.C + (size_t)kc * p.KC;
  const bool plain = p.pre_act == RB_ACT_NONE && p.aff_a == nullptr;
  for (int u0 = tid; u0 < units; u0 += 128 * 4) {
    uint4 v[4];
    int row[4], c8[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int u = u0 + i * 128;
      row[i] = u / c8n;
      c8[i] = u - row[i] * c8n;
      const int t = row0 + row[i];
      v[i] = make_uint4(0u, 0u, 0u, 0u);
      if (u < units && t >= p.x.lo && t < p.x.n)
        v[i] = __ldg(reinterpret_cast<const uint4*>(p.x.p + sbase + (ptrdiff_t)t * p.x.C + c8[i] * 8));
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int u = u0 + i * 128;
      if (u >= units) break;
      if (!plain) {
        float f[8];
        unpack8(v[i], f);
        const int ch = kc * p.KC + c8[i] * 8;            // view channel of f[0]
        if (p.aff_a) {
          const int t = row0 + row[i];
          if (t >= p.x.lo && t < p.x.n) {                // padding stays zero (F.pad comes after AdaIN)
            const int c0 = ch % p.aff_C;
            const float* a = p.aff_a + (size_t)wk.b * p.aff_C + c0;
            const float* bb = p.aff_b + (size_t)wk.b * p.aff_C + c0;
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] = fmaf(f[e], __ldg(a + e), __ldg(bb + e));
          }
        }
        if (p.pre_act == RB_ACT_LEAKY) {
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = f[e] > 0.f ? f[e] : 0.2f * f[e];
        } else if (p.pre_act == RB_ACT_SNAKE) {
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = act_f(f[e], RB_ACT_SNAKE, __ldg(p.alpha + ch + e));
        }
        v[i] = pack8(f);
      }
      *reinterpret_cast<uint4*>(stage + (size_t)c8[i] * p.R_ALLOC * 16 + (size_t)row[i] * 16) = v[i];
    }
  }
}

// ----------------------------------------------------------------------------- epilogues
// 16 accumulator columns of one row: + bias, + residual, optional sign mask -> 32 B of bf16.
__device__ __forceinline__ void store16(const TcParams& p, const uint32_t* r, const float* bias_s, int cg,
                                        int b, int t, bool valid) {
  float v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]) + bias_s[cg + i];
  if (!valid) return;
  if (p.res.p) {
    const __nv_bfloat16* rp =
        p.res.p + ((size_t)b * p.res.ld + p.res.hist + (t - p.res_delay)) * p.res.C + cg;
    float f[8];
    unpack8(__ldg(reinterpret_cast<const uint4*>(rp)), f);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] += f[i];
    unpack8(__ldg(reinterpret_cast<const uint4*>(rp + 8)), f);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[8 + i] += f[i];
  }
  if (p.out_mask && !(t & 1)) {
#pragma unroll
    for (int i = 1; i < 16; i += 2) v[i] = -v[i];
  }
  __nv_bfloat16* yp =
      const_cast<__nv_bfloat16*>(p.y.p) + ((size_t)b * p.y.ld + p.y.hist + t) * p.y.C + cg;
  *reinterpret_cast<uint4*>(yp) = pack8(v);
  *reinterpret_cast<uint4*>(yp + 8) = pack8(v + 8);
}

}  // namespace

__global__ void __launch_bounds__(TC_THREADS, 1) tc_conv_kernel(const __grid_constant__ TcParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + BAR_COUNT * 8);
  float* bias_s = reinterpret_cast<float*>(smem + p.off_bias);
  const int N_total = p.N_TILE * p.n_ntiles;
  float* bias2_s = bias_s + N_total;              // fused only: [C]
  float* alpha2_s = bias2_s + p.N_TILE;           // fused + Snake only: [C]
  uint8_t* a_smem = smem + p.off_a;
  uint8_t* w_smem = smem + p.off_w;
  uint8_t* h_smem = smem + p.off_h;
  const uint32_t bar0 = smem_u32(bars);
  auto BAR = [&](int i) { return bar0 + 8u * (uint32_t)i; };

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    for (int i = 0; i < 4; ++i) {
      mbar_init(BAR(BAR_A_FULL + i), 128);
      mbar_init(BAR(BAR_A_EMPTY + i), 1);
      mbar_init(BAR(BAR_W_FULL + i), 1);
      mbar_init(BAR(BAR_W_EMPTY + i), 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(BAR(BAR_D_FULL + i), 1);
      mbar_init(BAR(BAR_D_EMPTY + i), 128);
      mbar_init(BAR(BAR_H_FULL + i), 128);
      mbar_init(BAR(BAR_H_EMPTY + i), 1);
      mbar_init(BAR(BAR_D2_FULL + i), 1);
      mbar_init(BAR(BAR_D2_EMPTY + i), 128);
    }
    fence_mbar_init();
  }
  for (int i = tid; i < N_total; i += TC_THREADS) bias_s[i] = p.bias ? p.bias[i] : 0.f;
  if (p.fused) {
    for (int i = tid; i < p.N_TILE; i += TC_THREADS) {
      bias2_s[i] = p.bias2 ? p.bias2[i] : 0.f;
      alpha2_s[i] = p.alpha2 ? p.alpha2[i] : 1.f;
    }
  }
  if (warp == 8) {
    tmem_alloc(smem_u32(tmem_slot), p.tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  const uint32_t lbo_a = (uint32_t)p.R_ALLOC * 16u;
  const uint32_t lbo_w = (uint32_t)p.N_TILE * 16u;
  const int ksteps = p.KC >> 4;
  const int C2 = p.N_TILE;                        // fused: channels of the unit

  if (warp >= 4 && warp < 8) {
    // ============================================================== A loaders (128 threads)
    const int ltid = tid - 128;
    uint32_t it = 0;
    for (int w = blockIdx.x; w < p.n_work; w += gridDim.x) {
      const Work wk = decode(p, w);
      for (int kc = 0; kc < p.n_kchunks; ++kc) {
        if (__ldg(p.masks + wk.nt * p.n_kchunks + kc) == 0) continue;
        const uint32_t sa = it % p.SA, ph = (it / p.SA) & 1;
        mbar_wait(BAR(BAR_A_EMPTY + sa), ph ^ 1);
        load_a_chunk(p, wk, kc, a_smem + (size_t)sa * p.a_stage_bytes, ltid);
        fence_proxy_async();
        mbar_arrive(BAR(BAR_A_FULL + sa));
        ++it;
      }
    }
  } else if (warp == 9) {
    // =========================================================== W producer (one thread)
    if (lane == 0) {
      if (p.w_resident) {
        const uint32_t total = (uint32_t)p.nt_base[p.n_ntiles] * p.w_blk_bytes +
                               (p.fused ? (uint32_t)(C2 / p.KC) * p.w_blk_bytes : 0u);
        mbar_expect_tx(BAR(BAR_W_FULL), total);
        for (uint32_t o = 0; o < total; o += 32768u) {
          const uint32_t n = total - o < 32768u ? total - o : 32768u;
          bulk_g2s(smem_u32(w_smem + o), reinterpret_cast<const uint8_t*>(p.w) + o, n, BAR(BAR_W_FULL));
        }
      } else {
        uint32_t it = 0;
        for (int w = blockIdx.x; w < p.n_work; w += gridDim.x) {
          const Work wk = decode(p, w);
          int gblk = p.nt_base[wk.nt];
          for (int kc = 0; kc < p.n_kchunks; ++kc) {
            const uint64_t mask = __ldg(p.masks + wk.nt * p.n_kchunks + kc);
            for (int j = 0; j < p.n_taps; ++j) {
              if (!((mask >> j) & 1)) continue;
              const uint32_t sw = it % p.SW, ph = (it / p.SW) & 1;
              mbar_wait(BAR(BAR_W_EMPTY + sw), ph ^ 1);
              mbar_expect_tx(BAR(BAR_W_FULL + sw), p.w_blk_bytes);
              bulk_g2s(smem_u32(w_smem + (size_t)sw * p.w_blk_bytes),
                       reinterpret_cast<const uint8_t*>(p.w) + (size_t)gblk * p.w_blk_bytes, p.w_blk_bytes,
                       BAR(BAR_W_FULL + sw));
              ++gblk;
              ++it;
            }
          }
        }
      }
    }
  } else if (warp == 8) {
    // =========================================================== MMA issuer (one thread)
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(128, p.N_TILE);
      const uint32_t a0 = smem_u32(a_smem), w0 = smem_u32(w_smem), h0 = smem_u32(h_smem);
      const uint32_t d_stage_cols = p.fused ? 2u * C2 : (uint32_t)(p.MT * p.N_TILE);
      uint32_t it_a = 0, it_w = 0;
      if (p.w_resident) {
        mbar_wait(BAR(BAR_W_FULL), 0);
        tc_fence_after();
      }
      // GEMM1 of local work item i (accumulator stage i % SD)
      auto gemm1 = [&](uint32_t i, const Work& wk) {
        const uint32_t sd = i % p.SD, phd = (i / p.SD) & 1;
        mbar_wait(BAR(BAR_D_EMPTY + sd), phd ^ 1);
        tc_fence_after();
        const uint32_t dcol = tmem + sd * d_stage_cols;
        uint32_t acc = 0;
        int blk = 0;
        for (int kc = 0; kc < p.n_kchunks; ++kc) {
          const uint64_t mask = __ldg(p.masks + wk.nt * p.n_kchunks + kc);
          if (mask == 0) continue;
          const uint32_t sa = it_a % p.SA, pha = (it_a / p.SA) & 1;
          mbar_wait(BAR(BAR_A_FULL + sa), pha);
          tc_fence_after();
          const uint32_t abase = a0 + sa * p.a_stage_bytes;
          for (int j = 0; j < p.n_taps; ++j) {
            if (!((mask >> j) & 1)) continue;
            uint32_t wbase, sw = 0;
            if (p.w_resident) {
              wbase = w0 + (uint32_t)blk * p.w_blk_bytes;
            } else {
              sw = it_w % p.SW;
              mbar_wait(BAR(BAR_W_FULL + sw), (it_w / p.SW) & 1);
              tc_fence_after();
              wbase = w0 + sw * p.w_blk_bytes;
            }
            for (int mt = 0; mt < p.MT; ++mt) {
              const uint32_t arow = abase + (uint32_t)(j * p.dil + mt * 128) * 16u;
              for (int kk = 0; kk < ksteps; ++kk) {
                const uint64_t ad = make_desc(arow + (uint32_t)kk * 2u * lbo_a, lbo_a, 128u);
                const uint64_t bd = make_desc(wbase + (uint32_t)kk * 2u * lbo_w, lbo_w, 128u);
                mma_bf16(dcol + (uint32_t)(mt * p.N_TILE), ad, bd, idesc, acc | (uint32_t)kk);
              }
            }
            acc = 1;
            if (!p.w_resident) {
              mma_commit(BAR(BAR_W_EMPTY + sw));
              ++it_w;
            }
            ++blk;
          }
          mma_commit(BAR(BAR_A_EMPTY + sa));
          ++it_a;
        }
        mma_commit(BAR(BAR_D_FULL + sd));
      };
      // GEMM2 (fused): D2 = h * W1^T
      auto gemm2 = [&](uint32_t i) {
        const uint32_t sd = i % p.SD, phd = (i / p.SD) & 1;
        const uint32_t sh = i % p.SH, phh = (i / p.SH) & 1;
        mbar_wait(BAR(BAR_H_FULL + sh), phh);
        mbar_wait(BAR(BAR_D2_EMPTY + sd), phd ^ 1);
        tc_fence_after();
        const uint32_t dcol = tmem + sd * d_stage_cols + (uint32_t)C2;
        const uint32_t hbase = h0 + sh * p.h_stage_bytes;
        const int k16n = C2 >> 4;
        for (int k16 = 0; k16 < k16n; ++k16) {
          const int kc = k16 / ksteps, kk = k16 - kc * ksteps;
          const uint64_t ad = make_desc(hbase + (uint32_t)k16 * 2u * 2048u, 2048u, 128u);
          const uint64_t bd =
              make_desc(w0 + (uint32_t)(p.n_blk1 + kc) * p.w_blk_bytes + (uint32_t)kk * 2u * lbo_w, lbo_w, 128u);
          mma_bf16(dcol, ad, bd, idesc, (uint32_t)k16);
        }
        mma_commit(BAR(BAR_H_EMPTY + sh));
        mma_commit(BAR(BAR_D2_FULL + sd));
      };
      uint32_t i = 0;
      for (int w = blockIdx.x; w < p.n_work; w += gridDim.x, ++i) {
        gemm1(i, decode(p, w));
        if (p.fused && i >= 1) gemm2(i - 1);
      }
      if (p.fused && i >= 1) gemm2(i - 1);
    }
  } else {
    // ============================================================ epilogue (warps 0-3)
    const int rit = warp * 32 + lane;                       // row within a 128-row sub-tile
    const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
    const uint32_t d_stage_cols = p.fused ? 2u * C2 : (uint32_t)(p.MT * p.N_TILE);
    auto epi_plain = [&](uint32_t i, const Work& wk) {
      const uint32_t sd = i % p.SD, phd = (i / p.SD) & 1;
      mbar_wait(BAR(BAR_D_FULL + sd), phd);
      tc_fence_after();
      for (int mt = 0; mt < p.MT; ++mt) {
        const int t = wk.t0 + mt * 128 + rit;
        const bool valid = t < p.y.n;
        for (int c0 = 0; c0 < p.N_TILE; c0 += 16) {
          uint32_t r[16];
          tmem_ld16(tmem + lane_base + sd * d_stage_cols + (uint32_t)(mt * p.N_TILE + c0), r);
          tmem_ld_wait();
          store16(p, r, bias_s, wk.nt * p.N_TILE + c0, wk.b, t, valid);
        }
      }
      tc_fence_before();
      mbar_arrive(BAR(BAR_D_EMPTY + sd));
    };
    // fused E1: h = act2(D1 + bias) -> smem (A operand of the 1x1 conv)
    auto epi_h = [&](uint32_t i) {
      const uint32_t sd = i % p.SD, phd = (i / p.SD) & 1;
      const uint32_t sh = i % p.SH, phh = (i / p.SH) & 1;
      mbar_wait(BAR(BAR_D_FULL + sd), phd);
      mbar_wait(BAR(BAR_H_EMPTY + sh), phh ^ 1);
      tc_fence_after();
      uint8_t* hb = h_smem + (size_t)sh * p.h_stage_bytes + (size_t)rit * 16;
      for (int c0 = 0; c0 < C2; c0 += 16) {
        uint32_t r[16];
        tmem_ld16(tmem + lane_base + sd * d_stage_cols + (uint32_t)c0, r);
        tmem_ld_wait();
        float v[16];
#pragma unroll
        for (int e = 0; e < 16; ++e)
          v[e] = act_f(__uint_as_float(r[e]) + bias_s[c0 + e], p.act2, alpha2_s[c0 + e]);
        *reinterpret_cast<uint4*>(hb + (size_t)(c0 >> 3) * 2048) = pack8(v);
        *reinterpret_cast<uint4*>(hb + (size_t)((c0 >> 3) + 1) * 2048) = pack8(v + 8);
      }
      fence_proxy_async();
      tc_fence_before();
      mbar_arrive(BAR(BAR_H_FULL + sh));
      mbar_arrive(BAR(BAR_D_EMPTY + sd));
    };
    // fused E2: y = D2 + bias2 + x
    auto epi_out = [&](uint32_t i, const Work& wk) {
      const uint32_t sd = i % p.SD, phd = (i / p.SD) & 1;
      mbar_wait(BAR(BAR_D2_FULL + sd), phd);
      tc_fence_after();
      const int t = wk.t0 + rit;
      const bool valid = t < p.y.n;
      for (int c0 = 0; c0 < C2; c0 += 16) {
        uint32_t r[16];
        tmem_ld16(tmem + lane_base + sd * d_stage_cols + (uint32_t)(C2 + c0), r);
        tmem_ld_wait();
        store16(p, r, bias2_s, c0, wk.b, t, valid);
      }
      tc_fence_before();
      mbar_arrive(BAR(BAR_D2_EMPTY + sd));
    };
    uint32_t i = 0;
    Work prev{0, 0, 0};
    for (int w = blockIdx.x; w < p.n_work; w += gridDim.x, ++i) {
      const Work wk = decode(p, w);
      if (!p.fused) {
        epi_plain(i, wk);
      } else {
        epi_h(i);
        if (i >= 1) epi_out(i - 1, prev);
        prev = wk;
      }
    }
    if (p.fused && i >= 1) epi_out(i - 1, prev);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    tc_fence_after();
    tmem_dealloc(tmem, p.tmem_cols);
  }
}

int launch_tc_conv(const TcParams& p, cudaStream_t s) {
  if (p.n_work <= 0) return RB_OK;
  static int n_sm = 0;
  static bool attr_set = false;
  if (!attr_set) {
    int dev = 0;
    RB_CUDA(cudaGetDevice(&dev));
    RB_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
    RB_CUDA(cudaFuncSetAttribute(tc_conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set = true;
  }
  RB_CHECK(p.smem_bytes <= 227u * 1024u, RB_ERR_UNSUPPORTED, "tc conv needs %u B of shared memory",
           p.smem_bytes);
  const int grid = p.n_work < n_sm ? p.n_work : n_sm;
  tc_conv_kernel<<<grid, TC_THREADS, p.smem_bytes, s>>>(p);
  ++g_launches;
  RB_CUDA(cudaGetLastError());
  return RB_OK;
}

}  // namespace tc
}  // namespace rb
