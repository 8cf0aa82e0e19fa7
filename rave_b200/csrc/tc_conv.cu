// tc_conv.cu -- the tcgen05 implicit-GEMM convolution kernel of the bf16 engine.
//
// One persistent CTA per SM, warp-specialised (15 warps):
//   warps 0-7   epilogue    TMEM -> registers (tcgen05.ld) -> bias / residual / sign mask /
//                           output activation -> HBM.  Warp w owns TMEM lanes 32*(w%4).. (hardware
//                           rule) and the 32-column groups of parity w/4.  Fused DilatedUnit: also
//                           writes act(conv3 + b) back to smem as the A operand of the 1x1 conv --
//                           the intermediate never leaves the SM.
//   warps 8-11  act pass    only when the input tensor is stored un-activated (Snake nets, raw
//                           program inputs): pre-activation in place in smem between the TMA
//                           landing and the MMA.  LeakyReLU nets store every tensor already
//                           activated (the producer's epilogue applies it; a residual read inverts
//                           it with min(z, 5z)), so these warps idle and TMA feeds the MMA directly.
//   warp  12    MMA issuer  one elected thread: tcgen05.mma (M=128, N=N_TILE, K=16) per tap and K
//                           step; a dilated / shifted tap is only a different start row of the
//                           same smem tile (descriptor start address + 16 B per row), fp32
//                           accumulators in TMEM, double-buffered against the epilogue
//   warp  13    W producer  weight blocks L2 -> smem with cp.async.bulk on an mbarrier ring,
//                           or once for the whole kernel when the layer's weights fit
//   warp  14    A producer  one thread: TMA tensor-map boxes (rows x channel groups) of the input
//                           view -> smem; strided / phase views, zero padding and stream edges
//                           are all the tensor map's job (out-of-range rows are zero-filled)
// Stages hand over through mbarriers only; tcgen05.commit releases smem stages and publishes
// accumulators.  smem A tile:  addr(row, c8) = c8*(R_ALLOC*16) + row*16   (c8 = 8-channel group)
//        weight block [N][KC]: addr(n,   c8) = c8*(N*16)       + n*16
// The activation kinds are template parameters so that each instantiation stays small.
#include <cuda.h>

#include <cstdlib>
#include <cstring>
#include <type_traits>

#include "common.cuh"
#include "tc_conv.cuh"
#include "tc_ptx.cuh"

namespace rb {
namespace tc {

namespace {

constexpr int MAX_SA = 12;       // A stages: enough bytes in flight per SM to cover HBM latency
constexpr int MAX_SD = 4;        // accumulator / h stages
constexpr int MAX_SW = 8;        // streamed weight ring
enum { BAR_A_FULL = 0, BAR_A_EMPTY = MAX_SA, BAR_RAW_FULL = 2 * MAX_SA, BAR_W_FULL = 3 * MAX_SA,
       BAR_W_EMPTY = BAR_W_FULL + MAX_SW, BAR_D_FULL = BAR_W_EMPTY + MAX_SW, BAR_D_EMPTY = BAR_D_FULL + MAX_SD,
       BAR_H_FULL = BAR_D_EMPTY + MAX_SD, BAR_H_EMPTY = BAR_H_FULL + MAX_SD, BAR_D2_FULL = BAR_H_EMPTY + MAX_SD,
       BAR_D2_EMPTY = BAR_D2_FULL + MAX_SD, BAR_COUNT = BAR_D2_EMPTY + MAX_SD };
static_assert(BAR_COUNT * 8 + 8 <= 640, "barriers + TMEM slot must fit below the tap masks (smem + 640)");
constexpr int N_LOAD = 128;      // loader threads
constexpr int W_LOAD0 = 12;      // activation-pass warps 12-15 (un-activated inputs only)
constexpr int E1_WARPS = 8;      // fused: warps [0, 8) build the h tile, the other output warps store y

template <int ACT>
__device__ __forceinline__ float act_f(float v, float alpha, float inv_alpha) {
  if (ACT == RB_ACT_LEAKY) return fmaxf(v, 0.2f * v);
  if (ACT == RB_ACT_SNAKE) {
    // x + sin^2(a x) / (a + 1e-9)   (blocks.py:845-853); bf16 storage: fast sine is ample.  The reciprocal is a
    // per-channel constant computed at load: one MUFU operation per element instead of two.
    const float s = __sinf(alpha * v);
    return fmaf(inv_alpha, s * s, v);
  }
  return v;
}

// Waveform head tanh(wave * sigmoid(amp)) (blocks.py:700-707) with the hardware tanh: sigmoid(a) = 0.5 +
// 0.5 tanh(a / 2), so an output costs two MUFU operations instead of two exponentials and two reciprocals (the
// last conv's epilogue was MUFU-bound: 4 x 4096 operations per tile at 16 per clock).  tanh.approx.f32 is good
// to 2^-11 relative, a quarter of the bf16 rounding of the stored value.
__device__ __forceinline__ float tanh_hw(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float head_value(float wave, float amp) {
  const float g = fmaf(0.5f, tanh_hw(0.5f * amp), 0.5f);
  return tanh_hw(wave * g);
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

// debug event trace (CTA 0, first 64 work items): see tools_trace.py
// `tmask` (in scope at every use) is 0 for lane 0 of CTA 0 when tracing is on, else all ones: one
// compare per site in production runs
#define TRACE(tile, ev) do { if (((tile) | tmask) < 64u) \
    p.trace[(tile) * 16u + (ev)] = (unsigned long long)clock64(); } while (0)

struct Work {
  int nt, b, t0;
};
// The work items of one persistent CTA: w = blockIdx.x + i * gridDim.x, decoded once with
// divisions and then advanced by the pre-decomposed grid stride (adds and compares only).
// FLAT is a compile-time property (the cta_group::2 variants): its two integer divisions per work() -- inlined at
// every call site -- spread the hot code of the narrow fused units over more instruction-cache lines when the
// choice was a run-time branch (C = 64 units 776 -> 865 kcycles, icc hit rate 99.4 -> 93.7 %).
template <bool FLAT>
struct WorkIterT {
  int nt, b, tile, left;
  // w = (b * tiles_per_stream + tile) * n_ntiles + nt: the n-tiles of one row tile run at the same
  // time on neighbouring CTAs, so the A tile is read from HBM once and from L2 afterwards
  // Pair mode (p.pair = 1): the two CTAs of a cluster walk the same work sequence -- same n-tile,
  // hence the same weight blocks in the same order, which lets each load half of every block and
  // multicast it to both -- and take the even / odd 128*MT-row tile of each slot.
  // Flat mode (p.flat_units = u > 0; the cta_group::2 pairs): the (stream, 128*MT-row unit)
  // pairs are numbered stream-major, u units per stream, and a work slot is 2^pair CONSECUTIVE units of one
  // n-tile -- so a cluster spans neighbouring streams when a stream has fewer units than the cluster has CTAs
  // (the deepest layers: 256 rows per stream).  `tile` holds the slot's linear index, `b` the grid stride.
  __device__ __forceinline__ void init(const TcParams& p) {
    const int w = (int)blockIdx.x >> p.pair, stride = (int)gridDim.x >> p.pair;
    if (FLAT) {
      nt = 0; tile = w; b = stride;
      left = w < p.n_work ? (p.n_work - w + stride - 1) / stride : 0;
      return;
    }
    const int rt = w / p.n_ntiles;
    nt = w - rt * p.n_ntiles;
    b = rt / p.tiles_per_stream;
    tile = rt - b * p.tiles_per_stream;
    left = w < p.n_work ? (p.n_work - w + stride - 1) / stride : 0;
  }
  __device__ __forceinline__ void next(const TcParams& p) {
    if (FLAT) { tile += b; --left; return; }
    nt += p.step_nt;
    if (nt >= p.n_ntiles) { nt -= p.n_ntiles; ++tile; }
    tile += p.step_tile;
    if (tile >= p.tiles_per_stream) { tile -= p.tiles_per_stream; ++b; }
    b += p.step_b;
    --left;
  }
  // two items on in one step (the sets of output warps of a fused unit take alternate items)
  __device__ __forceinline__ void next2(const TcParams& p) {
    if (FLAT) { tile += 2 * b; left -= 2; return; }
    nt += p.step2_nt;
    if (nt >= p.n_ntiles) { nt -= p.n_ntiles; ++tile; }
    tile += p.step2_tile;
    if (tile >= p.tiles_per_stream) { tile -= p.tiles_per_stream; ++b; }
    b += p.step2_b;
    left -= 2;
  }
  __device__ __forceinline__ Work work(const TcParams& p) const {
    // p.reverse: the launch walks its row tiles from the END of the tensor.  Consecutive layers alternate, so a
    // layer starts on the rows its producer wrote last -- the ~100 MB of them that are still in L2.
    if (FLAT) {
      // a CTA past the last unit gets a tile far beyond the end: TMA zero-fills its loads, its epilogue stores nothing
      const int tl_ = p.reverse ? p.n_work - 1 - tile : tile;
      const int rt = tl_ / p.n_ntiles, nt_ = tl_ - rt * p.n_ntiles;
      const int g = (rt << p.pair) + ((int)blockIdx.x & ((1 << p.pair) - 1));
      const int bs = g / p.flat_units, tl = g - bs * p.flat_units;
      return Work{nt_, bs < p.B ? bs : 0, bs < p.B ? tl * p.MT * 128 : (1 << 24)};
    }
    if (p.reverse)
      return Work{nt, p.B - 1 - b, (((p.tiles_per_stream - 1 - tile) << p.pair) + ((int)blockIdx.x & p.pair)) * p.MT * 128};
    return Work{nt, b, ((tile << p.pair) + ((int)blockIdx.x & p.pair)) * p.MT * 128};
  }
};

// Stage index + phase bit of an mbarrier ring, advanced without divisions.
struct Ring {
  uint32_t idx = 0, ph = 0;
  __device__ __forceinline__ void next(int n) {
    if (++idx == (uint32_t)n) { idx = 0; ph ^= 1; }
  }
  // k positions on (any k, n >= 1)
  __device__ __forceinline__ void advance(int k, int n) {
    idx += (uint32_t)k;
    while (idx >= (uint32_t)n) { idx -= (uint32_t)n; ph ^= 1; }
  }
};

// ------------------------------------------------------------------------------ A tile
struct ChunkRef {
  int kc;
  uint32_t sa, ph;
};

// (row, 8-channel group) units of one chunk, strided over the act-pass threads: consecutive
// threads take consecutive rows of one channel group (consecutive 16-byte units: no conflicts).
template <typename F>
__device__ __forceinline__ void for_units(const TcParams& p, int tid, F&& f) {
  const int c8n = p.KC >> 3;
  if ((N_LOAD % c8n) == 0) {                     // c8n in {1,2,4,8,16}: fixed channel group per thread
    const int rthr = N_LOAD / c8n;
    const int c8 = tid / rthr;
#pragma unroll 4
    for (int row = tid % rthr; row < p.R; row += rthr) f(row, c8);
  } else {
    const int units = p.R * c8n;
    for (int u = tid; u < units; u += N_LOAD) {
      const int c8 = u / p.R;
      f(u - c8 * p.R, c8);
    }
  }
}

// One chunk (KC/8 channel groups x R_ALLOC rows starting at view row `row0`) as TMA boxes.
// Tensor-map coordinates: {0, phase, row - lo, channel group, stream}; rows outside [lo, n)
// arrive as zeros (conv zero padding, stream edges), so no thread touches the tile.
// Called by the WHOLE producer warp: lane 0 posts the byte count, then lane i issues copy i (one
// per channel group, at most 16) -- one instruction issues them all.  Issued one after the other
// by a single lane (~100 dependent instructions per tile, the warp shares its scheduler with
// busy output warps) the producer itself set the tile rate of the narrow fused units: ncu showed
// the GEMM1 issuers waiting for A_FULL 40 % of the time while the producer never waited for a free stage.
__device__ __forceinline__ void issue_a_tma(const TcParams& p, const void* tmap, int b, int row0, int kc,
                                            uint32_t stage, uint32_t bar, int lane) {
  if (lane == 0) mbar_expect_tx(bar, p.a_tx_bytes);
  __syncwarp();
  const int c8n = p.KC >> 3;
  const int kg0 = kc * c8n;
  if (p.a_flat && row0 >= p.x.lo && row0 + p.R_ALLOC <= p.x.n) {
    // interior tile of a unit-stride view: the rows of one channel group are one contiguous run
    // in HBM and in the smem tile, so a linear bulk copy per group replaces the tensor box (whose
    // 16-byte inner dimension caps the TMA unit at ~32 B/clk); edge tiles keep the zero-filling box
    // (a unit-stride view has a single phase: group g sits g * sA units into the stream)
    if (lane < c8n) {
      const uint4* src = reinterpret_cast<const uint4*>(p.x.p) +
                         ((long long)b * p.x.sb + (p.x.hist + row0) + (long long)(kg0 + lane) * p.x.sA);
      const uint32_t gbytes = (uint32_t)p.R_ALLOC * 16u;
      bulk_g2s(stage + (uint32_t)lane * gbytes, src, gbytes, bar);
    }
    return;
  }
  const int row = row0 - p.x.lo;
  const int nbox = (c8n / p.a_gb) * p.a_rsplit;            // <= 16
  if (lane < nbox) {
    const int gi = lane / p.a_rsplit, rs = lane - gi * p.a_rsplit;
    const int g = gi * p.a_gb;
    const int vg = kg0 + g;
    int ph, cg;
    if (p.x.g1_shift >= 0) { ph = vg >> p.x.g1_shift; cg = vg & (p.x.G1 - 1); }
    else { ph = vg / p.x.G1; cg = vg - ph * p.x.G1; }
    tma_load_5d(stage + (uint32_t)(g * p.R_ALLOC + rs * p.a_box_rows) * 16u, tmap, 0, ph,
                row + rs * p.a_box_rows, cg, b, bar);
  }
}

// LeakyReLU(0.2) on two packed bf16: max(x, 0.2 x).  0.2 is not a bf16, so the product is
// formed as fma(x, c1, x * c2) with c1 + c2 = 0.2 to ~1e-6: one rounding of 0.2 x like the
// fp32 reference followed by a bf16 store, at 3 packed instructions per pair.
__device__ __forceinline__ uint32_t leaky2(uint32_t w) {
  const __nv_bfloat162 x = *reinterpret_cast<const __nv_bfloat162*>(&w);
  const __nv_bfloat162 c1 = __floats2bfloat162_rn(0.2001953125f, 0.2001953125f);
  const __nv_bfloat162 c2 = __floats2bfloat162_rn(-0.0001953125f, -0.0001953125f);
  const __nv_bfloat162 y = __hmax2(x, __hfma2(x, c1, __hmul2(x, c2)));
  return *reinterpret_cast<const uint32_t*>(&y);
}
template <int PRE>
__device__ __forceinline__ void finish_a_chunk(const TcParams& p, int kc, uint8_t* stage, int tid) {
  if (PRE == RB_ACT_NONE) return;
  const size_t col = (size_t)p.R_ALLOC * 16;
  for_units(p, tid, [&](int row, int c8) {
    uint4* q = reinterpret_cast<uint4*>(stage + (size_t)c8 * col + (size_t)row * 16);
    if (PRE == RB_ACT_LEAKY) {
      uint4 v = *q;
      v.x = leaky2(v.x); v.y = leaky2(v.y); v.z = leaky2(v.z); v.w = leaky2(v.w);
      *q = v;
    } else {
      float f[8];
      unpack8(*q, f);
      const float4* al = reinterpret_cast<const float4*>(p.alpha + kc * p.KC + c8 * 8);
      const float4* il = reinterpret_cast<const float4*>(p.alpha + p.alpha_n + kc * p.KC + c8 * 8);
      const float4 a0 = __ldg(al), a1 = __ldg(al + 1), i0 = __ldg(il), i1 = __ldg(il + 1);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float ia[8] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y, i1.z, i1.w};
#pragma unroll
      for (int e = 0; e < 8; ++e) f[e] = act_f<PRE>(f[e], a[e], ia[e]);
      *q = pack8(f);
    }
  });
}

// ----------------------------------------------------------------------------- epilogues
// NC (16 or 32) accumulator columns of one row from TMEM, + bias (smem, float4 reads issued
// before the TMEM wait), packed fp32 adds.
template <int NC>
__device__ __forceinline__ void load_acc(uint32_t taddr, const float* bias_s, float* v) {
  uint32_t r[NC];
  if (NC == 32) tmem_ld32(taddr, r); else tmem_ld16(taddr, r);
  float4 b[4];
#pragma unroll
  for (int g = 0; g < 4; ++g) b[g] = *reinterpret_cast<const float4*>(bias_s + g * 4);
  tmem_ld_wait();
#pragma unroll
  for (int h = 0; h < NC / 16; ++h) {
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const int e = h * 16 + g * 4;
      v[e + 0] = __uint_as_float(r[e + 0]); v[e + 1] = __uint_as_float(r[e + 1]);
      v[e + 2] = __uint_as_float(r[e + 2]); v[e + 3] = __uint_as_float(r[e + 3]);
      add2(v[e + 0], v[e + 1], b[g].x, b[g].y);
      add2(v[e + 2], v[e + 3], b[g].z, b[g].w);
    }
    if (h + 1 < NC / 16) {
#pragma unroll
      for (int g = 0; g < 4; ++g) b[g] = *reinterpret_cast<const float4*>(bias_s + 16 + g * 4);
    }
  }
}
// LeakyReLU(0.2) on NC fp32 values: packed multiply + max.
template <int NC>
__device__ __forceinline__ void leaky_n(float* v) {
#pragma unroll
  for (int e = 0; e < NC; e += 2) {
    float u0, u1;
    mul2(u0, u1, v[e], v[e + 1], 0.2f, 0.2f);
    v[e] = fmaxf(v[e], u0);
    v[e + 1] = fmaxf(v[e + 1], u1);
  }
}

// 16-byte-unit offset of channel group g inside one stream of a view (32-bit: a stream's
// tensor is far below 2^31 units).
__device__ __forceinline__ int goff32(const TcView& v, int g) {
  int hi, lo;
  if (v.g1_shift >= 0) {
    hi = g >> v.g1_shift;
    lo = g & (v.G1 - 1);
  } else {
    hi = g / v.G1;
    lo = g - hi * v.G1;
  }
  return lo * v.sA + hi * v.sB;
}

// The common case of drain() below, with the address arithmetic hoisted out of the chunk loop:
// NT a multiple of 32, channel groups of a chunk sA apart in y (and in the residual), no waveform
// head, no sign mask.  RES: 0 none, 1 global (L2-prefetched), 2 the A tile in smem (fused unit).
template <int RES, bool SPIN = false, bool SB = false>
__device__ __forceinline__ void drain_fast(const TcParams& p, uint32_t bar_full, uint32_t parity, uint32_t tbase,
                                           const float* bs, const Work& wk, int rit, int part, int nparts, int ch0,
                                           int MTn, int NT, const uint8_t* res_s, uint32_t tr,
                                           uint32_t bar_full1 = 0, uint32_t bar_empty0 = 0) {
  // bar_full1 != 0: split accumulator (gemm1_cg2) -- sub-tile 1 has its own FULL barrier, and sub-tile 0 is given
  // back (bar_empty0) as soon as this warp has drained its share of it
  constexpr uint32_t tmask = 0;
  const int row = wk.t0 + rit;
  const int cstep = nparts * 32;
  bool handed = bar_full1 == 0;
  // per-thread offsets in 16-byte units from the (warp-uniform) tensor bases: 32 bits cover 64 GB
  const uint32_t yoff = (uint32_t)((long long)wk.b * p.y.sb) + (uint32_t)((p.y.hist + row) * p.y.sT);
  const uint32_t roff = RES == 1 ? (uint32_t)((long long)wk.b * p.res.sb) +
                                       (uint32_t)((p.res.hist - p.res_delay + row) * p.res.sT)
                                 : 0u;
  uint4* const ybase = reinterpret_cast<uint4*>(const_cast<__nv_bfloat16*>(p.y.p));
  const uint4* const rbase = reinterpret_cast<const uint4*>(p.res.p);
  const uint4* const srow = RES == 2 ? reinterpret_cast<const uint4*>(res_s) + (rit + p.off) : nullptr;
  uint4 rr[4];
  auto load_rr = [&](int mt, int c0) {
    if (RES == 1) {
      if (row + mt * 128 >= p.y.n) return;
      const uint32_t o = roff + (uint32_t)(mt * 128 * p.res.sT + goff32(p.res, (ch0 + c0) >> 3));
#pragma unroll
      for (int q = 0; q < 4; ++q) rr[q] = __ldg(rbase + (o + (uint32_t)(q * p.res.sA)));
    }
  };
  int mt = 0, c0 = part * 32;
  if (c0 >= NT) mt = MTn;
  if (RES == 1 && (rit & 7) == 0) {
    // pull this warp's residual rows into L2 while the MMAs of this work item run (see drain())
    for (int pm = 0; pm < MTn && row + pm * 128 < p.y.n; ++pm)
      for (int pc = part * 32; pc < NT; pc += cstep) {
        const uint32_t o = roff + (uint32_t)(pm * 128 * p.res.sT + goff32(p.res, (ch0 + pc) >> 3));
#pragma unroll
        for (int q = 0; q < 4; ++q)
          asm volatile("prefetch.global.L2 [%0];" ::"l"(rbase + (o + (uint32_t)(q * p.res.sA))));
      }
  }
  if (RES == 1 && mt < MTn) load_rr(mt, c0);
  mbar_wait_t<SPIN>(bar_full, parity, p.cg2_spin);
  tc_fence_after();
  TRACE(tr, 12);
  while (mt < MTn) {
    int mtn = mt, c0n = c0 + cstep;
    if (c0n >= NT) { c0n = part * 32; ++mtn; }
    if (RES == 2) {                                         // the stage is certainly full by now
      const uint4* sp = srow + ((ch0 + c0) >> 3) * p.R_ALLOC;
#pragma unroll
      for (int q = 0; q < 4; ++q) rr[q] = sp[q * p.R_ALLOC];
    }
    float v[32];
    load_acc<32>(tbase + (uint32_t)(mt * NT + c0), bs + ch0 + c0, v);
    TRACE(tr, 13);
    const bool valid = row + mt * 128 < p.y.n;
    if (SB) {
      // per-stream bias of this row's class (TcParams::sbias); all lanes but the edge rows' read the same addresses
      const int t = row + mt * 128;
      const int cls = t == 0 ? 1 : (t == p.y.n - 1 ? 2 : 0);
      const float4* sp = reinterpret_cast<const float4*>(p.sbias + ((size_t)wk.b * 3 + cls) * p.sbias_n + ch0 + c0);
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float4 b4 = __ldg(sp + q);
        add2(v[4 * q], v[4 * q + 1], b4.x, b4.y);
        add2(v[4 * q + 2], v[4 * q + 3], b4.z, b4.w);
      }
    }
    if (RES != 0 && valid) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint4 r4 = rr[q];
        const uint32_t w[4] = {r4.x, r4.y, r4.z, r4.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float lo = __uint_as_float(w[e] << 16), hi = __uint_as_float(w[e] & 0xffff0000u);
          if (p.res_act) {                                  // stored as leaky(x): x = min(z, 5 z), exact in fp32
            float l5, h5;
            mul2(l5, h5, lo, hi, 5.f, 5.f);
            lo = fminf(lo, l5);
            hi = fminf(hi, h5);
          }
          add2(v[q * 8 + 2 * e], v[q * 8 + 2 * e + 1], lo, hi);
        }
      }
    }
    if (RES == 1 && mtn < MTn) load_rr(mtn, c0n);           // next chunk's residual, in flight during the stores
    if (valid) {
      if (p.out_act) leaky_n<32>(v);
      TRACE(tr, 14);
      const uint32_t o = yoff + (uint32_t)(mt * 128 * p.y.sT + goff32(p.y, (ch0 + c0) >> 3));
#pragma unroll
      for (int q = 0; q < 4; ++q) ybase[o + (uint32_t)(q * p.y.sA)] = pack8(v + q * 8);
    }
    if (!handed && mtn != mt) {
      handed = true;
      tc_fence_before();
      __syncwarp();
      if ((threadIdx.x & 31) == 0) mbar_arrive(bar_empty0);
      if (mtn < MTn) {
        mbar_wait_t<SPIN>(bar_full1, parity, p.cg2_spin);
        tc_fence_after();
      }
    }
    mt = mtn;
    c0 = c0n;
  }
  if (!handed) {                                            // a warp without columns still hands sub-tile 0 back
    __syncwarp();
    if ((threadIdx.x & 31) == 0) mbar_arrive(bar_empty0);
  }
  TRACE(tr, 15);
  tc_fence_before();
}

// Fused unit, residual in smem: the same work as drain_fast<2> as a ROLLED loop over 16-column
// pieces.  The output warps of one scheduler run E1 and E2 code at the same time; unrolled 32-wide
// bodies (~110 + ~250 instructions, plus the poll loops) overflow the ~6 KB L0 instruction cache
// and ncu shows stall_no_inst as the top stall of the narrow fused units (a sample at every 128-byte
// line).  Half the body, executed twice, fits.
__device__ __forceinline__ void drain_smem16(const TcParams& p, uint32_t bar_full, uint32_t parity, uint32_t tbase,
                                             const float* bs, const Work& wk, int rit, int part, int nparts,
                                             int NT, const uint8_t* res_s) {
  // A fused unit's tensors are plain row views (one phase): channel group g sits g * sA units into the stream,
  // so every address below advances by adds -- goff32()'s phase split (a division when G1 is not a power of
  // two, compiled in either way) and the parameter reloads were ~25 of the ~160 instructions of a piece, and
  // the narrow units are bound by issue slots.
  const int row = wk.t0 + rit;
  const uint32_t ysA = (uint32_t)p.y.sA, RA = (uint32_t)p.R_ALLOC;
  uint4* yp = reinterpret_cast<uint4*>(const_cast<__nv_bfloat16*>(p.y.p)) +
              ((uint32_t)((long long)wk.b * p.y.sb) + (uint32_t)((p.y.hist + row) * p.y.sT) + (uint32_t)(part * 4) * ysA);
  const uint4* sp = reinterpret_cast<const uint4*>(res_s) + (rit + p.off) + (uint32_t)(part * 4) * RA;
  const int jump = 16 + (nparts - 1) * 32;                  // from the second piece of a chunk to the next chunk
  const bool valid = row < p.y.n;
  const bool ract = p.res_act != 0, oact = p.out_act != 0;
  mbar_wait(bar_full, parity);
  tc_fence_after();
  // (requesting a piece's accumulator one piece ahead was measured: +8 % instructions, no fewer cycles)
#pragma unroll 1
  for (int c = part * 32; c < NT;) {
    const uint4 r0 = sp[0], r1 = sp[RA];
    float v[16];
    load_acc<16>(tbase + (uint32_t)c, bs + c, v);
    const int d = (c & 16) ? jump : 16;
    const uint32_t w[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      float lo = __uint_as_float(w[e] << 16), hi = __uint_as_float(w[e] & 0xffff0000u);
      if (ract) {                                           // stored as leaky(x): x = min(z, 5 z), exact in fp32
        float l5, h5;
        mul2(l5, h5, lo, hi, 5.f, 5.f);
        lo = fminf(lo, l5);
        hi = fminf(hi, h5);
      }
      add2(v[2 * e], v[2 * e + 1], lo, hi);
    }
    if (oact) leaky_n<16>(v);
    if (valid) {
      yp[0] = pack8(v);
      yp[ysA] = pack8(v + 8);
    }
    yp += (uint32_t)(d >> 3) * ysA;
    sp += (uint32_t)(d >> 3) * RA;
    c += d;
  }
  tc_fence_before();
}

// Fused unit whose residual comes from global memory (C = 96 / 128: the A tiles cannot wait in smem until
// E2): 16-column pieces like drain_smem16.  The residual of the warp's first two pieces is requested BEFORE the
// accumulator barrier and each later piece two pieces ahead of its use (the tensor is this unit's own input, which
// the A producer is pulling through L2 at the same time, so a piece of work covers the latency).  The event trace
// of the first version -- 32-column chunks, one ahead -- showed E2 at 5200-6300 cycles per tile against ~1800 for E1
// on the same amount of data: the second chunk's load was only issued after the first chunk's math.  (A four-deep
// queue filled before the barrier worked as well but cost 24 register moves per piece in this rolled loop.)
__device__ __forceinline__ void drain_res16(const TcParams& p, uint32_t bar_full, uint32_t parity, uint32_t tbase,
                                            const float* bs, const Work& wk, int rit, int part, int nparts, int NT,
                                            uint32_t tr, bool by16 = false) {
  constexpr uint32_t tmask = 0;                             // `tr` already carries the trace mask
  const int row = wk.t0 + rit;
  const uint32_t ysA = (uint32_t)p.y.sA, rsA = (uint32_t)p.res.sA;
  // pieces of a warp: both halves of its 32-column chunks (part * 32, + 16, then the next chunk nparts * 32 on),
  // or -- by16, units with fewer pieces than warps -- one 16-column piece every nparts * 16 columns
  const int c_first = by16 ? part * 16 : part * 32;
  const int jump = by16 ? nparts * 16 : 16 + (nparts - 1) * 32;   // from the second piece of a chunk to the next chunk
  auto step = [&](int c) { return (by16 || (c & 16)) ? jump : 16; };
  uint4* yp = reinterpret_cast<uint4*>(const_cast<__nv_bfloat16*>(p.y.p)) +
              ((uint32_t)((long long)wk.b * p.y.sb) + (uint32_t)((p.y.hist + row) * p.y.sT) + (uint32_t)(c_first >> 3) * ysA);
  const uint4* rp = reinterpret_cast<const uint4*>(p.res.p) +
                    ((uint32_t)((long long)wk.b * p.res.sb) + (uint32_t)((p.res.hist - p.res_delay + row) * p.res.sT) +
                     (uint32_t)(c_first >> 3) * rsA);
  const bool valid = row < p.y.n;
  const bool ract = p.res_act != 0, oact = p.out_act != 0;
  uint4 q0, q1, q2, q3;                                     // residual of the current piece (q0, q1) and the next (q2, q3)
  q0 = q1 = q2 = q3 = make_uint4(0u, 0u, 0u, 0u);
  int cq = c_first;                                         // column of the next piece to request
  if (cq < NT && valid) { q0 = __ldg(rp); q1 = __ldg(rp + rsA); }
  { const int d = step(cq); rp += (uint32_t)(d >> 3) * rsA; cq += d; }
  if (cq < NT && valid) { q2 = __ldg(rp); q3 = __ldg(rp + rsA); }
  { const int d = step(cq); rp += (uint32_t)(d >> 3) * rsA; cq += d; }
  mbar_wait(bar_full, parity);
  tc_fence_after();
  TRACE(tr, 12);
#pragma unroll 1
  for (int c = c_first; c < NT;) {
    float v[16];
    load_acc<16>(tbase + (uint32_t)c, bs + c, v);
    const int d = step(c);
    TRACE(tr, 13);
    const uint32_t w[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
    // the queue moves up and the piece two ahead is requested before this piece's math
    q0 = q2;
    q1 = q3;
    if (cq < NT && valid) { q2 = __ldg(rp); q3 = __ldg(rp + rsA); }
    { const int d = step(cq); rp += (uint32_t)(d >> 3) * rsA; cq += d; }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      float lo = __uint_as_float(w[e] << 16), hi = __uint_as_float(w[e] & 0xffff0000u);
      if (ract) {                                           // stored as leaky(x): x = min(z, 5 z), exact in fp32
        float l5, h5;
        mul2(l5, h5, lo, hi, 5.f, 5.f);
        lo = fminf(lo, l5);
        hi = fminf(hi, h5);
      }
      add2(v[2 * e], v[2 * e + 1], lo, hi);
    }
    if (oact) leaky_n<16>(v);
    TRACE(tr, 14);
    if (valid) {
      yp[0] = pack8(v);
      yp[ysA] = pack8(v + 8);
    }
    yp += (uint32_t)(d >> 3) * ysA;
    c += d;
  }
  TRACE(tr, 15);
  tc_fence_before();
}

// Drains part `part` of `nparts` of the 32-column chunks of `MTn` sub-tiles x `NT` accumulator
// columns (TMEM address `tbase`, lane base included) into y at channel offset ch0: accumulator
// + bias (+ residual) -> sign mask / output activation -> bf16 -> 16-byte stores.  The residual
// of a chunk is requested as soon as the previous chunk has consumed its own (before the
// accumulator barrier for the first chunk), so its latency overlaps the math and the stores.
__device__ __forceinline__ void drain(const TcParams& p, uint32_t bar_full, uint32_t parity, uint32_t tbase,
                                      const float* bs, const Work& wk, int rit, int part, int nparts, int ch0,
                                      int MTn, int NT, const uint8_t* res_s = nullptr, uint32_t tr = 0xffffffffu) {
  constexpr uint32_t tmask = 0;                             // `tr` already carries the trace mask
  const bool has_res = p.res.p != nullptr;
  const uint4* rs = reinterpret_cast<const uint4*>(p.res.p) + (long long)wk.b * p.res.sb;
  uint4* ys = reinterpret_cast<uint4*>(const_cast<__nv_bfloat16*>(p.y.p)) + (long long)wk.b * p.y.sb;
  // four consecutive channel groups of a chunk sit sA apart when they share a phase
  const bool quad_y = (p.y.G1 & 3) == 0, quad_r = (p.res.G1 & 3) == 0;
  const int cstep = nparts * 32;
  const int row = wk.t0 + rit;
  uint4 rr[4];
  auto load_rr = [&](int mt, int c0) {
    const int t = row + mt * 128;
    if (t >= p.y.n) return;
    const int cg = (ch0 + c0) >> 3;
    if (res_s) {                                          // fused unit: rows off.. of the A tile in smem
      const uint4* sp = reinterpret_cast<const uint4*>(res_s) + (cg * p.R_ALLOC + rit + p.off);
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if (q * 8 < NT - c0) rr[q] = sp[q * p.R_ALLOC];
      return;
    }
    const uint4* rp = rs + ((p.res.hist - p.res_delay + t) * p.res.sT + (quad_r ? goff32(p.res, cg) : 0));
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (q * 8 < NT - c0) rr[q] = __ldg(rp + (quad_r ? q * p.res.sA : goff32(p.res, cg + q)));
  };
  int mt = 0, c0 = part * 32;
  if (c0 >= NT) mt = MTn;
  if (has_res && !res_s && (rit & 7) == 0) {
    // The residual tensor was written two kernels ago and has usually left L2: pull this warp's
    // rows of every chunk into L2 now (one prefetch per 128-byte line, no registers), while the
    // MMAs of this work item run; the loads below then see L2 latency instead of HBM latency.
    for (int pm = 0; pm < MTn; ++pm) {
      const int t = row + pm * 128;
      if (t >= p.y.n) break;
      for (int pc = part * 32; pc < NT; pc += cstep) {
        const int cg = (ch0 + pc) >> 3;
        const uint4* rp = rs + ((p.res.hist - p.res_delay + t) * p.res.sT + (quad_r ? goff32(p.res, cg) : 0));
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (q * 8 < NT - pc)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(rp + (quad_r ? q * p.res.sA : goff32(p.res, cg + q))));
      }
    }
  }
  if (has_res && !res_s && mt < MTn) load_rr(mt, c0);
  mbar_wait(bar_full, parity);
  tc_fence_after();
  TRACE(tr, 12);
  if (res_s && mt < MTn) load_rr(mt, c0);                 // the stage is certainly full by now
  while (mt < MTn) {
    int mtn = mt, c0n = c0 + cstep;
    if (c0n >= NT) { c0n = part * 32; ++mtn; }
    const int t = row + mt * 128;
    const bool valid = t < p.y.n;
    const int cg = (ch0 + c0) >> 3;
    const uint32_t taddr = tbase + (uint32_t)(mt * NT + c0);
    auto body = [&](auto nc_tag) {
      constexpr int NC = decltype(nc_tag)::value;
      float v[NC];
      load_acc<NC>(taddr, bs + ch0 + c0, v);
      TRACE(tr, 13);
      if (has_res && valid) {
#pragma unroll
        for (int q = 0; q < NC / 8; ++q) {
          const uint4 r4 = rr[q];
          const uint32_t w[4] = {r4.x, r4.y, r4.z, r4.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float lo = __uint_as_float(w[e] << 16), hi = __uint_as_float(w[e] & 0xffff0000u);
            if (p.res_act) {
              // stored as z = leaky(x): x = min(z, 5 z), exact in fp32 (no second bf16 rounding)
              float l5, h5;
              mul2(l5, h5, lo, hi, 5.f, 5.f);
              lo = fminf(lo, l5);
              hi = fminf(hi, h5);
            }
            add2(v[q * 8 + 2 * e], v[q * 8 + 2 * e + 1], lo, hi);
          }
        }
      }
      if (has_res && mtn < MTn) load_rr(mtn, c0n);        // next chunk's residual, in flight during the stores
      if (!valid) return;
      if (NC == 32 && p.head) {
        // waveform head (blocks.py:700-707, amplitude_modulation): tanh(wave * sigmoid(amp)), then the
        // reverse_half sign of the PQMF synthesis input (pqmf.py:13-17): odd bands of even real rows
        const int sub = (ch0 + c0) >> 5;                  // real row = t * head + sub
        const bool even_row = (((t * p.head) + sub) & 1) == 0;
        float o[16];
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const float th = head_value(v[e], v[16 + e]);
          o[e] = ((e & 1) && even_row) ? -th : th;
        }
        uint4* hp = ys + (p.y.hist + t) * p.y.sT;
        hp[goff32(p.y, 2 * sub)] = pack8(o);
        hp[goff32(p.y, 2 * sub + 1)] = pack8(o + 8);
        return;
      }
      if (p.out_mask && !(t & 1)) {
#pragma unroll
        for (int e = 1; e < NC; e += 2) v[e] = -v[e];
      }
      if (p.out_act) leaky_n<NC>(v);                      // stored activated: LeakyReLU(0.2)
      TRACE(tr, 14);
      uint4* yp = ys + ((p.y.hist + t) * p.y.sT + (quad_y ? goff32(p.y, cg) : 0));
#pragma unroll
      for (int q = 0; q < NC / 8; ++q) yp[quad_y ? q * p.y.sA : goff32(p.y, cg + q)] = pack8(v + q * 8);
    };
    if (NT - c0 >= 32) body(std::integral_constant<int, 32>{});
    else body(std::integral_constant<int, 16>{});
    mt = mtn;
    c0 = c0n;
  }
  TRACE(tr, 15);
  tc_fence_before();
}

// Two 16-column accumulator pieces + bias; both TMEM loads are in flight before the one wait.
__device__ __forceinline__ void load_acc2x16(uint32_t ta, uint32_t tb, const float* ba, const float* bb, float* v) {
  uint32_t r[32];
  tmem_ld16(ta, r);
  tmem_ld16(tb, r + 16);
  tmem_ld_wait();
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const float* bp = h ? bb : ba;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const float4 b4 = *reinterpret_cast<const float4*>(bp + g * 4);
      const int e = h * 16 + g * 4;
      v[e + 0] = __uint_as_float(r[e + 0]); v[e + 1] = __uint_as_float(r[e + 1]);
      v[e + 2] = __uint_as_float(r[e + 2]); v[e + 3] = __uint_as_float(r[e + 3]);
      add2(v[e + 0], v[e + 1], b4.x, b4.y);
      add2(v[e + 2], v[e + 3], b4.z, b4.w);
    }
  }
}

// Layers whose output rows interleave in memory -- row-batched narrow layers, transposed-conv phases, the
// audio rows of the synthesis bank, the waveform head (TcParams::pair_mode).  drain() / drain_fast() give
// every lane one 16-byte unit per store and the lanes of a warp sit 2, 4 or 8 units apart: each instruction
// fills half of every 32-byte sector it touches, and since the L1 -> crossbar path moves sectors the layer
// pays for twice its output (ncu on the first conv: l1tex__m_l1tex2xbar_write_bytes 2.15 GB for 1.07 GB of
// output, request cycles 78 % busy, DRAM 39 %).  Here a lane gathers the two pieces that are neighbours in
// memory -- phase 2i and 2i+1 of one channel group, or two consecutive groups of an audio row -- and writes
// them with one 32-byte store.  Single n-tile, no residual.
template <bool SPIN>
__device__ __forceinline__ void drain_pair(const TcParams& p, uint32_t bar_full, uint32_t parity, uint32_t tbase,
                                           const float* bs, const Work& wk, int rit, int part, int nparts, int MTn,
                                           int NT) {
  const int row = wk.t0 + rit;
  uint4* const ybase = reinterpret_cast<uint4*>(const_cast<__nv_bfloat16*>(p.y.p));
  const uint32_t yoff = (uint32_t)((long long)wk.b * p.y.sb) + (uint32_t)((p.y.hist + row) * p.y.sT);
  const bool oact = p.out_act != 0;
  mbar_wait_t<SPIN>(bar_full, parity, p.cg2_spin);
  tc_fence_after();
  if (p.pair_mode == 3) {
    // waveform head (blocks.py:700-707) with two real rows per view row: columns [32 s, 32 s + 16) wave and
    // [32 s + 16, 32 s + 32) amplitude of real row 2 t + s.  Part g < 2 takes channel group g of both rows:
    // tanh(wave * sigmoid(amp)), reverse_half (pqmf.py:13-17: odd bands of even rows, i.e. s = 0), 32 bytes out.
    // (tcgen05.ld is .sync.aligned: every lane of the warp executes it, rows past the end only skip the store)
    if (part < 2) {
      uint32_t r[32];
      const uint32_t t0 = tbase + (uint32_t)(8 * part);
      tmem_ld8(t0, r);
      tmem_ld8(t0 + 16u, r + 8);
      tmem_ld8(t0 + 32u, r + 16);
      tmem_ld8(t0 + 48u, r + 24);
      tmem_ld_wait();
      float o[16];
#pragma unroll
      for (int sub = 0; sub < 2; ++sub) {
        const float* bw = bs + 32 * sub + 8 * part;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float wv = __uint_as_float(r[16 * sub + e]) + bw[e];
          const float am = __uint_as_float(r[16 * sub + 8 + e]) + bw[16 + e];
          const float th = head_value(wv, am);
          o[8 * sub + e] = ((e & 1) && sub == 0) ? -th : th;
        }
      }
      if (row < p.y.n) st_global_256(ybase + (yoff + (uint32_t)(part * p.y.sA)), pack8(o), pack8(o + 8));
    }
  } else if (p.pair_mode == 2) {
    // consecutive channel groups are consecutive units (audio rows): 32 columns = 64 contiguous bytes
    for (int mt = 0; mt < MTn; ++mt) {
      const bool valid = row + mt * 128 < p.y.n;
      for (int c0 = part * 32; c0 < NT; c0 += nparts * 32) {
        float v[32];
        load_acc<32>(tbase + (uint32_t)(mt * NT + c0), bs + c0, v);
        if (oact) leaky_n<32>(v);
        uint4* yp = ybase + (yoff + (uint32_t)(mt * 128 * p.y.sT + (c0 >> 3)));
        if (valid) {
          st_global_256(yp, pack8(v), pack8(v + 8));
          st_global_256(yp + 2, pack8(v + 16), pack8(v + 24));
        }
      }
    }
  } else {
    // phases: view channel n = ph * cpo + c; unit(group g, phase ph) = g * sA + ph
    // (no divisions here: this preamble runs once per warp and tile, and these launches have one or two pieces
    // per warp and tile -- the two divisions were ~50 of its 134 instructions)
    const int cpo = p.pair_dcol, kper = cpo >> 4;
    const int units = p.pair_units;                           // (NT / (2 cpo)) * kper, from the host
    bs += wk.nt * NT;                                         // this n-tile's bias columns ...
    const uint32_t ph0 = (uint32_t)(wk.nt * p.pair_ph_tile);  // ... and its first phase
    int pp0 = 0, k0 = part;                                   // unit index = part, part + nparts, ...
    while (k0 >= kper) { k0 -= kper; ++pp0; }
    for (int mt = 0; mt < MTn; ++mt) {
      const bool valid = row + mt * 128 < p.y.n;
      int pp = pp0, k = k0;
      for (int u = part; u < units; u += nparts) {
        const int cA = 2 * pp * cpo + 16 * k;
        float v[32];
        load_acc2x16(tbase + (uint32_t)(mt * NT + cA), tbase + (uint32_t)(mt * NT + cA + cpo), bs + cA,
                     bs + cA + cpo, v);
        if (oact) leaky_n<32>(v);
        uint4* yp = ybase + (yoff + (uint32_t)(mt * 128 * p.y.sT + 2 * k * p.y.sA + 2 * pp) + ph0);
        if (valid) {
          st_global_256(yp, pack8(v), pack8(v + 16));
          st_global_256(yp + p.y.sA, pack8(v + 8), pack8(v + 24));
        }
        k += nparts;
        while (k >= kper) { k -= kper; ++pp; }
      }
    }
  }
  tc_fence_before();
}

// Fused E1 chunk: h = act2(D1 + bias) as bf16 into the h tile (core-matrix layout, 2048 B per
// channel group).
template <int ACT2, int NC, bool HT>
__device__ __forceinline__ void h_chunk(uint32_t taddr, const float* bs, const float* alpha2, int ia_off, uint8_t* hdst) {
  float v[NC];
  load_acc<NC>(taddr, bs, v);
  if (ACT2 == RB_ACT_LEAKY) {
    leaky_n<NC>(v);
  } else if (ACT2 == RB_ACT_SNAKE) {
#pragma unroll
    for (int e = 0; e < NC; ++e) v[e] = act_f<ACT2>(v[e], alpha2[e], alpha2[e + ia_off]);
  }
  if (!HT) {
#pragma unroll
    for (int q = 0; q < NC / 8; ++q) *reinterpret_cast<uint4*>(hdst + (size_t)q * 2048) = pack8(v + q * 8);
  } else {
    // h stays in tensor memory: the NC fp32 columns just read become NC / 2 columns of packed bf16 pairs at the
    // same column base -- the A operand of the 1x1 conv's MMAs (K step k = columns [16 k, 16 k + 8))
    static_assert(NC == 16, "h in TMEM: 16-column pieces");
    const uint4 lo = pack8(v), hi = pack8(v + 8);
    const uint32_t r[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
    tmem_st8(taddr, r);
  }
}

}  // namespace

// issued by the elected lane of the MMA warp only
#define MMA(...) do { if (leader) mma_bf16_lohi(__VA_ARGS__); } while (0)
#define COMMIT(bar) do { if (leader) mma_commit(bar); } while (0)


// DR: how accumulators are drained -- 0 generic drain(), 1 drain_fast (global or no residual),
// 2 drain_fast with the residual read from the A tile in smem (fused only).  One drain per
// instantiation keeps every variant inside the 96-register budget without spills.
// TR: event trace compiled in (tools_trace.py); production launches use TR = false.
// NS: sets of 8 output warps of a fused unit without activation pass (2 or 3).  ncu on the narrow
// units: 5 warps per scheduler, 1.1 eligible, 9 cycles per issued instruction -- the serial
// wait / tcgen05.ld / math / store chain of an E1 or E2 chunk is latency-bound, and the two sets
// were busy all the time.  A third set (28 warps, 72 registers) puts a third tile in flight.
// CG2: the two CTAs of a cluster pair run ONE tcgen05.mma.cta_group::2 stream (M = 256: each CTA's own 128-row
// tile, N_TILE = 256 with each CTA holding 128 rows of every weight block).  Per SM and MMA the tensor core then
// reads 4 KB of A + 4 KB of B from shared memory instead of 4 + 8, and only half of every weight block lands in
// each SM: the streamed wide layers were bound by the 128 B/clk shared-memory port (profiles/r01_ubench.txt), not
// by the tensor pipe.  The leader (cluster rank 0) issues; the peer's MMA warp relays its barriers (below).
// PTX wants one cta_group for every tcgen05 instruction of a kernel, hence a template parameter.
// HT: fused unit with h in tensor memory (TcParams::h_tmem) -- a template parameter because the narrow units are
// bound by issue slots: the same choice as a run-time flag cost the C = 64 units 5 % (+16 M instructions).
template <int PRE, int ACT2, bool FUSED, int DR, bool TR, int NS, bool CG2, bool HT>
__global__ void __launch_bounds__(NS == 3 ? 896 : TC_THREADS, 1)
    tc_conv_kernel(const __grid_constant__ TcParams p, const __grid_constant__ TcTmap tmx) {
  static_assert(NS == 2 || (NS == 3 && FUSED && PRE == RB_ACT_NONE), "three sets: fused units without activation pass");
  static_assert(!CG2 || (!FUSED && PRE == RB_ACT_NONE && NS == 2), "cta_group::2: plain streamed layers");
  static_assert(!HT || FUSED, "h in tensor memory: fused units");
  using WorkIter = WorkIterT<CG2>;                          // cta_group::2 launches number their work flat
  constexpr int W_MMA = NS == 3 ? 24 : 16, W_WPROD = W_MMA + 1, W_APROD = W_MMA + 2, W_MMA2 = W_MMA + 3;
  constexpr int N_THREADS = NS == 3 ? 896 : TC_THREADS;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + BAR_COUNT * 8);
  uint64_t* mask_s = reinterpret_cast<uint64_t*>(smem + 640);   // [n_ntiles][n_kchunks] <= TC_MAX_MASKS entries
  float* bias_s = reinterpret_cast<float*>(smem + p.off_bias);
  const int N_total = p.N_TILE * p.n_ntiles;
  float* bias2_s = bias_s + N_total;              // fused only: [C]
  float* alpha2_s = bias2_s + p.N_TILE;           // fused + Snake only: [C]
  uint8_t* a_smem = smem + p.off_a;
  uint8_t* w_smem = smem + p.off_w;
  uint8_t* h_smem = smem + p.off_h;
  const uint32_t bar0 = smem_u32(bars);
  auto BAR = [&](int i) { return bar0 + 8u * (uint32_t)i; };

  // read %tid.x ONCE: ptxas otherwise rematerialises S2R SR_TID.X inside the epilogue loops to save a
  // register (-1 % cycles over the step)
  int tid;
  asm volatile("mov.u32 %0, %%tid.x;" : "=r"(tid));
  const int warp = tid >> 5, lane = tid & 31;
  const uint32_t tmask = (TR && p.trace && blockIdx.x == 0 && lane == 0) ? 0u : 0xffffffffu;
  // Programmatic dependent launch: let the next kernel of the stream start its prologue on SMs
  // this grid has already left; everything that touches activations waits below.
  asm volatile("griddepcontrol.launch_dependents;");
  // Output warps: when no activation pass is needed, warps 12-15 drain accumulators too.
  //   plain:  warps [0, W_OUT_END) share the 32-column chunks round-robin
  //   fused:  warps [0, E1_WARPS) run E1 (h tile), warps [E1_WARPS, W_OUT_END) run E2 (output)
  constexpr int W_OUT_END = PRE == RB_ACT_NONE ? 16 : 12;   // (three sets: 24 output warps, 8 per tile and phase)
  constexpr int W_OUT0 = FUSED ? E1_WARPS : 0;
  constexpr int N_OUT = (W_OUT_END - W_OUT0) * 32;
  constexpr int OUT_PARTS = (W_OUT_END - W_OUT0) / 4;
  constexpr int E1_PARTS = E1_WARPS / 4;

  // cta_group::2: the leader's FULL / D_EMPTY barriers also count one relayed arrival from the peer
  const uint32_t cg2_lead = (CG2 && (cluster_ctarank() & 1u) == 0) ? 1u : 0u;
  if (tid == 0) {
    for (int i = 0; i < MAX_SA; ++i) {
      mbar_init(BAR(BAR_A_FULL + i), (PRE == RB_ACT_NONE ? 1 : N_LOAD / 32) + cg2_lead);
      mbar_init(BAR(BAR_A_EMPTY + i), 1 + (p.res_smem ? W_OUT_END - W_OUT0 : 0));
      mbar_init(BAR(BAR_RAW_FULL + i), 1);
    }
    for (int i = 0; i < MAX_SW; ++i) {
      mbar_init(BAR(BAR_W_FULL + i), 1 + cg2_lead);
      mbar_init(BAR(BAR_W_EMPTY + i), CG2 ? 1 : 1 + p.pair);   // multicast pair: released by both CTAs' MMAs
    }
    for (int i = 0; i < MAX_SD; ++i) {
      mbar_init(BAR(BAR_D_FULL + i), 1);
      // one arrival per WARP (lane 0 after __syncwarp): 32 per-thread arrivals on one mbarrier
      // serialise in the shared-memory atomic path -- at 768 arrivals per tile they, not the
      // math, set the tile rate of the narrow fused units
      mbar_init(BAR(BAR_D_EMPTY + i), (FUSED ? (HT ? 1 : E1_WARPS) : N_OUT / 32) + cg2_lead);
      mbar_init(BAR(BAR_H_FULL + i), E1_WARPS);
      mbar_init(BAR(BAR_H_EMPTY + i), 1);
      mbar_init(BAR(BAR_D2_FULL + i), 1);
      mbar_init(BAR(BAR_D2_EMPTY + i), N_OUT / 32);
    }
    fence_mbar_init();
  }
  for (int i = tid; i < p.n_ntiles * p.n_kchunks; i += N_THREADS) mask_s[i] = p.masks[i];
  for (int i = tid; i < N_total; i += N_THREADS) bias_s[i] = p.bias ? p.bias[i] : 0.f;
  if (FUSED) {
    for (int i = tid; i < p.N_TILE; i += N_THREADS) {
      bias2_s[i] = p.bias2 ? p.bias2[i] : 0.f;
      alpha2_s[i] = p.alpha2 ? p.alpha2[i] : 1.f;
      alpha2_s[i + p.N_TILE] = p.alpha2 ? p.alpha2[i + p.N_TILE] : 1.f;
    }
  }
  if (warp == W_MMA) {
    if (CG2) {
      tmem_alloc2(smem_u32(tmem_slot), p.tmem_cols);
      tmem_relinquish2();
    } else {
      tmem_alloc(smem_u32(tmem_slot), p.tmem_cols);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  if (p.pair) cluster_sync();                     // the peer's barriers exist before anything targets them
  // the prologue above (barriers, TMEM, bias / masks) and the weight loads of the W producer do not
  // depend on the previous kernel; activations do
  if (warp != W_WPROD) asm volatile("griddepcontrol.wait;" ::: "memory");

  const uint32_t lbo_a = (uint32_t)p.R_ALLOC * 16u;
  const uint32_t lbo_w = (uint32_t)(CG2 ? p.N_TILE / 2 : p.N_TILE) * 16u;
  const int ksteps = p.KC >> 4;
  const int C2 = p.N_TILE;                        // fused: channels of the unit
  const uint32_t d_stage_cols = FUSED ? 2u * C2 : (uint32_t)(p.MT * p.N_TILE);

  if (warp == W_APROD) {
    // ============================================================ A producer (one warp, TMA)
    {
      if (lane == 0) prefetch_tmap(&tmx);
      Ring ra;
      WorkIter wi;
      uint32_t ti = 0;
      for (wi.init(p); wi.left > 0; wi.next(p), ++ti) {
        const Work wk = wi.work(p);
        for (int kc = 0; kc < p.n_kchunks; ++kc) {
          if (mask_s[wk.nt * p.n_kchunks + kc] == 0) continue;
          const uint32_t sa = ra.idx;
          mbar_wait_t<CG2>(BAR(BAR_A_EMPTY + sa), ra.ph ^ 1, p.cg2_spin);
          TRACE(ti, 10);
          issue_a_tma(p, &tmx, wk.b, wk.t0 - p.off, kc, smem_u32(a_smem + (size_t)sa * p.a_stage_bytes),
                      BAR((PRE == RB_ACT_NONE ? BAR_A_FULL : BAR_RAW_FULL) + sa), lane);
          TRACE(ti, 11);
          ra.next(p.SA);
        }
      }
    }
  } else if (PRE != RB_ACT_NONE && warp >= W_LOAD0 && warp < W_MMA) {
    // =============================== pre-activation pass in smem (un-activated inputs only)
    {
      const int ltid = tid - W_LOAD0 * 32;
      Ring ra;
      WorkIter wi;
      for (wi.init(p); wi.left > 0; wi.next(p)) {
        for (int kc = 0; kc < p.n_kchunks; ++kc) {
          if (mask_s[wi.nt * p.n_kchunks + kc] == 0) continue;
          const uint32_t sa = ra.idx;
          mbar_wait(BAR(BAR_RAW_FULL + sa), ra.ph);
          finish_a_chunk<PRE>(p, kc, a_smem + (size_t)sa * p.a_stage_bytes, ltid);
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) mbar_arrive(BAR(BAR_A_FULL + sa));
          ra.next(p.SA);
        }
      }
    }
  } else if (warp == W_WPROD && !p.w_resident) {
    // ============================================== W producer (one thread, streamed weights)
    if (lane == 0) {
      {
        Ring rw;
        WorkIter wi;
        for (wi.init(p); wi.left > 0; wi.next(p)) {
          const Work wk = wi.work(p);
          int gblk = p.nt_base[wk.nt];
          for (int kc = 0; kc < p.n_kchunks; ++kc) {
            const uint64_t mask = mask_s[wk.nt * p.n_kchunks + kc];
            const int nb = __popcll(mask);
            for (int j = 0; j < nb; ++j) {
              const uint32_t sw = rw.idx;
              mbar_wait_t<CG2>(BAR(BAR_W_EMPTY + sw), rw.ph ^ 1, p.cg2_spin);
              mbar_expect_tx(BAR(BAR_W_FULL + sw), p.w_blk_bytes);
              if (CG2) {
                // this CTA's 128 rows of the block (tc_engine.cu packs the two halves one after the other)
                bulk_g2s(smem_u32(w_smem + (size_t)sw * p.w_blk_bytes),
                         reinterpret_cast<const uint8_t*>(p.w) +
                             ((size_t)gblk * 2 + ((uint32_t)blockIdx.x & 1u)) * p.w_blk_bytes,
                         p.w_blk_bytes, BAR(BAR_W_FULL + sw));
              } else if (p.pair) {
                // this CTA's half of the block, written into both CTAs' rings (the peer sends the other half)
                const uint32_t half = p.w_blk_bytes >> 1, o = ((uint32_t)blockIdx.x & 1u) * half;
                bulk_g2s_mc2(smem_u32(w_smem + (size_t)sw * p.w_blk_bytes) + o,
                             reinterpret_cast<const uint8_t*>(p.w) + (size_t)gblk * p.w_blk_bytes + o, half,
                             BAR(BAR_W_FULL + sw));
              } else {
                bulk_g2s(smem_u32(w_smem + (size_t)sw * p.w_blk_bytes),
                         reinterpret_cast<const uint8_t*>(p.w) + (size_t)gblk * p.w_blk_bytes, p.w_blk_bytes,
                         BAR(BAR_W_FULL + sw));
              }
              ++gblk;
              rw.next(p.SW);
            }
          }
        }
      }
    }
  } else if (warp == W_MMA || warp == W_MMA2 || warp == W_WPROD) {
    // ====================================== MMA issuer (warp-uniform control, elected lane issues)
    // Resident weights: the W producer warp loads them once; for narrow layers (dual_g1) it then
    // issues GEMM1 of the odd local tiles (W_MMA takes the even ones): there the serial issue
    // stream of one warp, sharing its scheduler with four output warps, bounds the tile rate.
    if (warp == W_WPROD && lane == 0) {
      const uint32_t total = (uint32_t)p.nt_base[p.n_ntiles] * p.w_blk_bytes +
                             (FUSED ? (uint32_t)(C2 / p.KC) * p.w_blk_bytes : 0u);
      mbar_expect_tx(BAR(BAR_W_FULL), total);
      for (uint32_t o = 0; o < total; o += 32768u) {
        const uint32_t n = total - o < 32768u ? total - o : 32768u;
        bulk_g2s(smem_u32(w_smem + o), reinterpret_cast<const uint8_t*>(p.w) + o, n, BAR(BAR_W_FULL));
      }
    }
    __syncwarp();
    {
      const bool leader = elect_one();
      const uint32_t idesc = make_idesc_bf16(128, p.N_TILE);
      const uint32_t a0 = smem_u32(a_smem), w0 = smem_u32(w_smem), h0 = smem_u32(h_smem);
      // descriptor halves: hi is constant (SBO = 128 B, version 1); lo = start>>4 | LBO>>4 << 16
      const uint32_t dhi = desc_hi(128u);
      const uint32_t a_lbo_bits = ((lbo_a >> 4) & 0x3FFFu) << 16, w_lbo_bits = ((lbo_w >> 4) & 0x3FFFu) << 16;
      const uint32_t a_kadd = (2u * lbo_a) >> 4, w_kadd = (2u * lbo_w) >> 4;   // one K=16 step
      const uint32_t tap_add = (uint32_t)p.dil;                                // rows -> 16 B units
      const uint32_t blk_units = p.w_blk_bytes >> 4;
      Ring ra, rw, rd1, rd2, rh;
      if (p.w_resident) {
        mbar_wait(BAR(BAR_W_FULL), 0);
        tc_fence_after();
      }
      // GEMM1 of local work item i (accumulator stage i % SD)
      uint32_t t1 = 0, t2 = 0;                  // trace: tile counters of GEMM1 / GEMM2
      auto gemm1 = [&](const Work& wk) {
        const uint32_t sd = rd1.idx;
        mbar_wait(BAR(BAR_D_EMPTY + sd), rd1.ph ^ 1);
        TRACE(t1, 0);
        rd1.next(p.SD);
        tc_fence_after();
        const uint32_t dcol = tmem + sd * d_stage_cols;
        uint32_t acc = 0;
        uint32_t wres = ((w0 >> 4) & 0x3FFFu) | w_lbo_bits;      // resident: next block's desc.lo
        const uint64_t* mrow = mask_s + wk.nt * p.n_kchunks;
        for (int kc = 0; kc < p.n_kchunks; ++kc) {
          uint64_t mask = mrow[kc];
          if (mask == 0) continue;
          const uint32_t sa = ra.idx;
          mbar_wait(BAR(BAR_A_FULL + sa), ra.ph);
          TRACE(t1, 1);
          ra.next(p.SA);
          tc_fence_after();
          const uint32_t alo0 = (((a0 + sa * p.a_stage_bytes) >> 4) & 0x3FFFu) | a_lbo_bits;
          const int j0 = __ffsll((long long)mask) - 1, n_act = __popcll(mask);
          const bool run = (mask >> j0) == (n_act >= 64 ? ~0ull : ((1ull << n_act) - 1ull));
          if (p.w_resident && run) {
            // fast path (one run of active taps, weights in smem, MT == 1): adds between MMAs only
            uint32_t al = alo0 + (uint32_t)j0 * tap_add, bl = wres;
            if (ksteps == 1) {
#pragma unroll 4
              for (int j = 0; j < n_act; ++j) {
                MMA(dcol, al, bl, dhi, idesc, acc | (uint32_t)j);
                al += tap_add;
                bl += blk_units;
              }
            } else if (ksteps == 4) {
#pragma unroll 2
              for (int j = 0; j < n_act; ++j) {
                MMA(dcol, al, bl, dhi, idesc, acc | (uint32_t)j);
                MMA(dcol, al + a_kadd, bl + w_kadd, dhi, idesc, 1u);
                MMA(dcol, al + 2u * a_kadd, bl + 2u * w_kadd, dhi, idesc, 1u);
                MMA(dcol, al + 3u * a_kadd, bl + 3u * w_kadd, dhi, idesc, 1u);
                al += tap_add;
                bl += blk_units;
              }
            } else {
              for (int j = 0; j < n_act; ++j) {
                uint32_t a2 = al, b2 = bl;
                MMA(dcol, a2, b2, dhi, idesc, acc | (uint32_t)j);
                for (int kk = 1; kk < ksteps; ++kk) {
                  a2 += a_kadd;
                  b2 += w_kadd;
                  MMA(dcol, a2, b2, dhi, idesc, 1u);
                }
                al += tap_add;
                bl += blk_units;
              }
            }
            wres = bl;
            acc = 1;
            mask = 0;
          }
          while (mask) {
            const int j = __ffsll((long long)mask) - 1;
            mask &= mask - 1;
            uint32_t blo, sw = 0;
            if (p.w_resident) {
              blo = wres;
              wres += blk_units;
            } else {
              sw = rw.idx;
              mbar_wait(BAR(BAR_W_FULL + sw), rw.ph);
              rw.next(p.SW);
              tc_fence_after();
              blo = (((w0 + sw * p.w_blk_bytes) >> 4) & 0x3FFFu) | w_lbo_bits;
            }
            uint32_t alo = alo0 + (uint32_t)j * tap_add;
            for (int mt = 0; mt < p.MT; ++mt) {
              uint32_t al = alo, bl = blo;
              const uint32_t dc = dcol + (uint32_t)(mt * p.N_TILE);
              MMA(dc, al, bl, dhi, idesc, acc);
#pragma unroll 4
              for (int kk = 1; kk < ksteps; ++kk) {
                al += a_kadd;
                bl += w_kadd;
                MMA(dc, al, bl, dhi, idesc, 1u);
              }
              alo += 128u;                                        // next 128-row sub-tile
            }
            acc = 1;
            if (!p.w_resident) {
              if (p.pair) { if (leader) mma_commit_mc2(BAR(BAR_W_EMPTY + sw)); }
              else COMMIT(BAR(BAR_W_EMPTY + sw));
            }
          }
          COMMIT(BAR(BAR_A_EMPTY + sa));
        }
        COMMIT(BAR(BAR_D_FULL + sd));
        TRACE(t1, 2);
        ++t1;
      };
      // GEMM2 (fused): D2 = h * W1^T  (h: [128][C2] core-matrix layout, LBO = 2048)
      const uint32_t h_lbo_bits = ((2048u >> 4) & 0x3FFFu) << 16, h_kadd = (2u * 2048u) >> 4;
      auto gemm2 = [&]() {
        const uint32_t sd = rd2.idx, sh = rh.idx;
        mbar_wait(BAR(BAR_H_FULL + sh), rh.ph);
        TRACE(t2, 3);
        mbar_wait(BAR(BAR_D2_EMPTY + sd), rd2.ph ^ 1);
        TRACE(t2, 4);
        rd2.next(p.SD);
        rh.next(p.SH);
        tc_fence_after();
        const uint32_t dcol = tmem + sd * d_stage_cols + (uint32_t)C2;
        const int nkc = C2 / p.KC;
        uint32_t accum = 0;
        if (HT) {
          // A operand from tensor memory: K step k of h sits at columns [16 k, 16 k + 8) of this stage's D1 region
          uint32_t at = tmem + sd * d_stage_cols;
          for (int kc = 0; kc < nkc; ++kc) {
            uint32_t bl = (((w0 + (uint32_t)(p.n_blk1 + kc) * p.w_blk_bytes) >> 4) & 0x3FFFu) | w_lbo_bits;
#pragma unroll 4
            for (int kk = 0; kk < ksteps; ++kk) {
              if (leader) mma_bf16_ts(dcol, at, bl, dhi, idesc, accum);
              accum = 1;
              at += 16u;
              bl += w_kadd;
            }
          }
          COMMIT(BAR(BAR_D_EMPTY + sd));
        } else {
          uint32_t al = (((h0 + sh * p.h_stage_bytes) >> 4) & 0x3FFFu) | h_lbo_bits;
          for (int kc = 0; kc < nkc; ++kc) {
            uint32_t bl = (((w0 + (uint32_t)(p.n_blk1 + kc) * p.w_blk_bytes) >> 4) & 0x3FFFu) | w_lbo_bits;
#pragma unroll 4
            for (int kk = 0; kk < ksteps; ++kk) {
              MMA(dcol, al, bl, dhi, idesc, accum);
              accum = 1;
              al += h_kadd;
              bl += w_kadd;
            }
          }
          COMMIT(BAR(BAR_H_EMPTY + sh));
        }
        COMMIT(BAR(BAR_D2_FULL + sd));
        TRACE(t2, 5);
        ++t2;
      };
      // GEMM1 and GEMM2 are issued by two different warps: each is a serial instruction stream
      // (waits, descriptor set-up, commits), and for narrow units that stream is the bottleneck.
      // the rings of a GEMM1 tile issued by the other warp: same advances, no instructions
      auto skip1 = [&](const Work& wk) {
        rd1.next(p.SD);
        const uint64_t* mrow = mask_s + wk.nt * p.n_kchunks;
        for (int kc = 0; kc < p.n_kchunks; ++kc)
          if (mrow[kc] != 0) ra.next(p.SA);
        ++t1;
      };
      // ---- cta_group::2 (CG2): leader issues for the pair, the peer relays
      // GEMM1 of one work item of the PAIR (streamed weights, MT = 1): the waits see the leader's own producers
      // and, through the relay, the peer's; every commit arrives in both CTAs.
      auto gemm1_cg2 = [&](const Work& wk) {
        const uint32_t idesc2 = make_idesc_bf16(256, p.N_TILE);
        const uint32_t sd = rd1.idx;
        const uint32_t phe = rd1.ph ^ 1;
        if (!p.split) mbar_wait_t<true>(BAR(BAR_D_EMPTY + sd), phe, p.cg2_spin & 2);
        rd1.next(p.SD);
        tc_fence_after();
        const uint32_t dcol = tmem + sd * d_stage_cols;
        uint32_t acc = 0;
        const uint64_t* mrow = mask_s + wk.nt * p.n_kchunks;
        int kf = -1, kl = -1;                                 // first / last chunk with work (split)
        if (p.split)
          for (int kc = 0; kc < p.n_kchunks; ++kc)
            if (mrow[kc] != 0) { if (kf < 0) kf = kc; kl = kc; }
        for (int kc = 0; kc < p.n_kchunks; ++kc) {
          uint64_t mask = mrow[kc];
          if (mask == 0) continue;
          const uint32_t sa = ra.idx;
          mbar_wait_t<true>(BAR(BAR_A_FULL + sa), ra.ph, p.cg2_spin & 2);
          ra.next(p.SA);
          tc_fence_after();
          const uint32_t alo0 = (((a0 + sa * p.a_stage_bytes) >> 4) & 0x3FFFu) | a_lbo_bits;
          if (p.split && (kc == kf || kc == kl)) {
            // Split accumulator (MT = 2 fills tensor memory: one stage, the drain would be exposed).  The two
            // sub-tiles have their own FULL / EMPTY barriers (indices 0, 1); the first and the last chunk of a
            // work item run sub-tile-major, so sub-tile 0 is published one chunk-pass before sub-tile 1 and taken
            // back one pass earlier: its drain overlaps the other's MMAs at both ends.  The weight stages of the
            // chunk stay until the second pass has used them (taps <= SW).
            Ring rw2 = rw;
            for (int mt = 0; mt < 2; ++mt) {
              if (kc == kf) {
                mbar_wait_t<true>(BAR(BAR_D_EMPTY + mt), phe, p.cg2_spin & 2);
                tc_fence_after();
              }
              rw2 = rw;
              uint32_t acc2 = kc == kf ? 0u : 1u;
              const uint32_t dc = dcol + (uint32_t)(mt * p.N_TILE);
              for (uint64_t m2 = mask; m2; m2 &= m2 - 1) {
                const int j = __ffsll((long long)m2) - 1;
                const uint32_t sw = rw2.idx;
                if (mt == 0) {
                  mbar_wait_t<true>(BAR(BAR_W_FULL + sw), rw2.ph, p.cg2_spin & 2);
                  tc_fence_after();
                }
                rw2.next(p.SW);
                uint32_t al = alo0 + (uint32_t)j * tap_add + (uint32_t)mt * 128u;
                uint32_t bl = (((w0 + sw * p.w_blk_bytes) >> 4) & 0x3FFFu) | w_lbo_bits;
                if (leader) mma2_bf16_lohi(dc, al, bl, dhi, idesc2, acc2);
#pragma unroll 4
                for (int kk = 1; kk < ksteps; ++kk) {
                  al += a_kadd;
                  bl += w_kadd;
                  if (leader) mma2_bf16_lohi(dc, al, bl, dhi, idesc2, 1u);
                }
                acc2 = 1;
                if (mt == 1 && leader) mma2_commit(BAR(BAR_W_EMPTY + sw));   // second pass: the stage is done
              }
              if (kc == kl && leader) mma2_commit(BAR(BAR_D_FULL + mt));
            }
            rw = rw2;
            acc = 1;
            if (leader) mma2_commit(BAR(BAR_A_EMPTY + sa));
            continue;
          }
          while (mask) {
            const int j = __ffsll((long long)mask) - 1;
            mask &= mask - 1;
            const uint32_t sw = rw.idx;
            mbar_wait_t<true>(BAR(BAR_W_FULL + sw), rw.ph, p.cg2_spin & 2);
            rw.next(p.SW);
            tc_fence_after();
            const uint32_t blo = (((w0 + sw * p.w_blk_bytes) >> 4) & 0x3FFFu) | w_lbo_bits;
            uint32_t alo = alo0 + (uint32_t)j * tap_add;
            for (int mt = 0; mt < p.MT; ++mt) {
              uint32_t al = alo, bl = blo;
              const uint32_t dc = dcol + (uint32_t)(mt * p.N_TILE);
              if (leader) mma2_bf16_lohi(dc, al, bl, dhi, idesc2, acc);
#pragma unroll 4
              for (int kk = 1; kk < ksteps; ++kk) {
                al += a_kadd;
                bl += w_kadd;
                if (leader) mma2_bf16_lohi(dc, al, bl, dhi, idesc2, 1u);
              }
              alo += 128u;                                        // next 128-row sub-tile
            }
            acc = 1;
            if (leader) mma2_commit(BAR(BAR_W_EMPTY + sw));
          }
          if (leader) mma2_commit(BAR(BAR_A_EMPTY + sa));
        }
        if (!p.split && leader) mma2_commit(BAR(BAR_D_FULL + sd));
      };
      // The peer's MMA warp: follows the same sequence of stages and, as each of ITS stages becomes ready
      // (A tile landed, weight half landed, accumulator drained), arrives on the leader's barrier of the same
      // name.  The first use of an accumulator stage needs no hand-over (the leader's wait passes on the fresh
      // barrier, and an arrival now would be counted for the NEXT drain).
      uint32_t n_item = 0;
      auto relay_cg2 = [&](const Work& wk) {
        const uint32_t sd = rd1.idx;
        const bool hand = n_item >= (uint32_t)p.SD;
        const uint32_t phe = rd1.ph ^ 1;
        if (hand) {
          mbar_wait_t<true>(BAR(BAR_D_EMPTY + sd), phe, p.cg2_spin);
          if (leader) mbar_arrive_cluster(mapa_shared(BAR(BAR_D_EMPTY + sd), 0));
        }
        ++n_item;
        rd1.next(p.SD);
        bool second = p.split && hand;                      // split: sub-tile 1 is handed over after the first chunk's stages
        const uint64_t* mrow = mask_s + wk.nt * p.n_kchunks;
        for (int kc = 0; kc < p.n_kchunks; ++kc) {
          uint64_t mask = mrow[kc];
          if (mask == 0) continue;
          const uint32_t sa = ra.idx;
          mbar_wait_t<true>(BAR(BAR_A_FULL + sa), ra.ph, p.cg2_spin);
          ra.next(p.SA);
          if (leader) mbar_arrive_cluster(mapa_shared(BAR(BAR_A_FULL + sa), 0));
          for (int nb = __popcll(mask); nb > 0; --nb) {
            const uint32_t sw = rw.idx;
            mbar_wait_t<true>(BAR(BAR_W_FULL + sw), rw.ph, p.cg2_spin);
            rw.next(p.SW);
            if (leader) mbar_arrive_cluster(mapa_shared(BAR(BAR_W_FULL + sw), 0));
          }
          if (second) {
            second = false;
            mbar_wait_t<true>(BAR(BAR_D_EMPTY + 1), phe, p.cg2_spin);
            if (leader) mbar_arrive_cluster(mapa_shared(BAR(BAR_D_EMPTY + 1), 0));
          }
        }
      };
      WorkIter wi;
      if (CG2) {
        if (warp == W_MMA) {
          if (cg2_lead) for (wi.init(p); wi.left > 0; wi.next(p)) gemm1_cg2(wi.work(p));
          else for (wi.init(p); wi.left > 0; wi.next(p)) relay_cg2(wi.work(p));
        }
      } else if (FUSED && warp == W_MMA2) {
        for (wi.init(p); wi.left > 0; wi.next(p)) gemm2();
      } else if (p.dual_g1 <= 1) {
        if (warp == W_MMA)
          for (wi.init(p); wi.left > 0; wi.next(p)) gemm1(wi.work(p));
      } else {
        // dual_g1 = 2 or 3 issuers: W_MMA, the weight-producer warp (idle once the resident weights are in) and,
        // in kernels without a second GEMM, W_MMA2 take the work items in turn
        const int me = warp == W_MMA ? 0 : warp == W_WPROD ? 1 : 2;
        if (me < p.dual_g1) {
          int k = 0;
          for (wi.init(p); wi.left > 0; wi.next(p)) {
            if (k == me) gemm1(wi.work(p)); else skip1(wi.work(p));
            if (++k == p.dual_g1) k = 0;
          }
        }
      }
    }
  } else {
    // ==================================================== epilogue (warps 0-7, or 0-11)
    const int rit = (warp & 3) * 32 + lane;                 // row within a 128-row sub-tile
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    Ring rd, rh;
    WorkIter wi;
    if (!FUSED) {
      const int part = warp >> 2;
      uint32_t ti = 0;
      for (wi.init(p); wi.left > 0; wi.next(p), rd.next(p.SD), ++ti) {
        const Work wk = wi.work(p);
        const uint32_t sd = rd.idx, phd = rd.ph;
        if (warp == 0) TRACE(ti, 8);
        const uint32_t trc = (warp == 0 ? ti : 0xffffffffu) | tmask;
        const uint32_t tb = tmem + lane_base + sd * d_stage_cols;
        if (DR == 3)
          drain_pair<CG2>(p, BAR(BAR_D_FULL + sd), phd, tb, bias_s, wk, rit, part, OUT_PARTS, p.MT, p.N_TILE);
        else if (DR == 0)
          drain(p, BAR(BAR_D_FULL + sd), phd, tb, bias_s, wk, rit, part, OUT_PARTS, wk.nt * p.N_TILE, p.MT,
                p.N_TILE, nullptr, trc);
        else if (DR == 4)
          drain_fast<0, CG2, true>(p, BAR(BAR_D_FULL + sd), phd, tb, bias_s, wk, rit, part, OUT_PARTS, wk.nt * p.N_TILE,
                                   p.MT, p.N_TILE, nullptr, trc, CG2 && p.split ? BAR(BAR_D_FULL + 1) : 0u,
                                   BAR(BAR_D_EMPTY));
        else if (p.res.p)
          drain_fast<1, CG2>(p, BAR(BAR_D_FULL + sd), phd, tb, bias_s, wk, rit, part, OUT_PARTS, wk.nt * p.N_TILE, p.MT,
                             p.N_TILE, nullptr, trc, CG2 && p.split ? BAR(BAR_D_FULL + 1) : 0u, BAR(BAR_D_EMPTY));
        else
          drain_fast<0, CG2>(p, BAR(BAR_D_FULL + sd), phd, tb, bias_s, wk, rit, part, OUT_PARTS, wk.nt * p.N_TILE, p.MT,
                             p.N_TILE, nullptr, trc, CG2 && p.split ? BAR(BAR_D_FULL + 1) : 0u, BAR(BAR_D_EMPTY));
        __syncwarp();
        if (lane == 0) mbar_arrive(BAR(BAR_D_EMPTY + (CG2 && p.split ? 1u : sd)));
        if (warp == 0) TRACE(ti, 9);
      }
    } else {
      // Fused unit.  E1(i): h = act2(D1 + bias) -> smem (A operand of the 1x1 conv);
      // E2(i): y = D2 + bias2 + x.  Without an activation pass the 16 output warps form two
      // sets that alternate: set s runs E1 on tiles = s (mod 2) and, after each, E2 on the tile
      // before it, so both sets carry the same load; otherwise warps 0-7 run E1 and 8-11 run E2.
      constexpr bool ALT = PRE == RB_ACT_NONE;
      const int part = ALT ? ((warp >> 2) & 1) : (warp < E1_WARPS ? (warp >> 2) : 0);
      const bool tr_warp = (warp & 7) == 0;
      // E1 of local tile `ti` (ring positions rd / rh of that tile)
      auto run_e1 = [&](const Ring& rd_, const Ring& rh_, uint32_t ti) {
        const uint32_t sd = rd_.idx, sh = rh_.idx;
        mbar_wait(BAR(BAR_D_FULL + sd), rd_.ph);
        if (tr_warp) TRACE(ti, 6);
        if (!HT) mbar_wait(BAR(BAR_H_EMPTY + sh), rh_.ph ^ 1);
        tc_fence_after();
        uint8_t* hb = h_smem + (size_t)sh * p.h_stage_bytes + (size_t)rit * 16;
        {
          // rolled 16-column pieces (see drain_smem16: instruction-cache footprint), both halves of this warp's
          // 32-column chunks
          const uint32_t tb1 = tmem + lane_base + sd * d_stage_cols;
#pragma unroll 1
          for (int c = part * 32; c < C2;) {
            h_chunk<ACT2, 16, HT>(tb1 + (uint32_t)c, bias_s + c, alpha2_s + c, C2, hb + (size_t)(c >> 3) * 2048);
            c = (c & 16) ? c - 16 + E1_PARTS * 32 : c + 16;
          }
        }
        if (HT) tmem_st_wait();
        else fence_proxy_async();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(BAR(BAR_H_FULL + sh));
          if (!HT) mbar_arrive(BAR(BAR_D_EMPTY + sd));    // h in TMEM: GEMM2 releases the stage (it reads h there)
        }
        if (tr_warp) TRACE(ti, 7);
      };
      // E2 of the tile at work position wk (ring positions rd_ / ra_ of that tile)
      auto run_e2 = [&](const Work& wk, const Ring& rd_, const Ring& ra_, uint32_t ti) {
        const uint32_t sd = rd_.idx, sa = ra_.idx;
        if (tr_warp) TRACE(ti, 8);
        const uint32_t trc = (tr_warp ? ti : 0xffffffffu) | tmask;
        const uint32_t tb = tmem + lane_base + sd * d_stage_cols + (uint32_t)C2;
        const uint8_t* rs = p.res_smem ? a_smem + (size_t)sa * p.a_stage_bytes : nullptr;
        if (DR == 0)
          drain(p, BAR(BAR_D2_FULL + sd), rd_.ph, tb, bias2_s, wk, rit, part, OUT_PARTS, 0, 1, C2, rs, trc);
        else if (DR == 2 && !TR)
          drain_smem16(p, BAR(BAR_D2_FULL + sd), rd_.ph, tb, bias2_s, wk, rit, part, OUT_PARTS, C2, rs);
        else if (DR == 2)
          drain_fast<2>(p, BAR(BAR_D2_FULL + sd), rd_.ph, tb, bias2_s, wk, rit, part, OUT_PARTS, 0, 1, C2, rs, trc);
        else      // DR == 1: residual from global memory
          drain_res16(p, BAR(BAR_D2_FULL + sd), rd_.ph, tb, bias2_s, wk, rit, part, OUT_PARTS, C2, trc);
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(BAR(BAR_D2_EMPTY + sd));
          if (p.res_smem) mbar_arrive(BAR(BAR_A_EMPTY + sa));   // this warp no longer needs the A stage
        }
        if (tr_warp) TRACE(ti, 9);
      };
      Ring ra;                                              // A stage of a tile (one chunk per tile when res_smem)
      wi.init(p);
      if (ALT) {
        // Set s (8 warps) runs E1 on tiles = s (mod NS) and, after each, E2 on the tile before it
        // (whose E1 belongs to the previous set): every set carries the same load and NS tiles are
        // in the output stage at once.  Two iterators: (wi, rd, rh) at the E1 tile, (w2, rd2, ra2)
        // at the E2 tile; both advance NS tiles at a time.
        const int set = warp >> 3;
        const int n = wi.left;
        WorkIter w2 = wi;
        Ring rd2;
        int t1 = set, t2 = set - 1;
        bool skip2 = false;
        if (t2 < 0) { t2 += NS; skip2 = true; }
        for (int i = 0; i < t1 && i < n; ++i) { wi.next(p); rd.next(p.SD); rh.next(p.SH); }
        for (int i = 0; i < t2 && i < n; ++i) { w2.next(p); rd2.next(p.SD); ra.next(p.SA); }
        for (;;) {
          bool did = false;
          if (t1 < n) {
            run_e1(rd, rh, (uint32_t)t1);
            t1 += NS;
            rd.advance(NS, p.SD);
            rh.advance(NS, p.SH);
            did = true;
          }
          if (!skip2 && t2 < n) {
            run_e2(w2.work(p), rd2, ra, (uint32_t)t2);
            t2 += NS;
            if (NS == 2) w2.next2(p);
            else
#pragma unroll
              for (int i = 0; i < NS; ++i) w2.next(p);
            rd2.advance(NS, p.SD);
            ra.advance(NS, p.SA);
            did = true;
          }
          skip2 = false;
          if (!did) break;
        }
      } else if (warp < E1_WARPS) {
        for (uint32_t ti = 0; wi.left > 0; wi.next(p), rd.next(p.SD), rh.next(p.SH), ++ti) run_e1(rd, rh, ti);
      } else {
        for (uint32_t ti = 0; wi.left > 0; wi.next(p), rd.next(p.SD), ra.next(p.SA), ++ti)
          run_e2(wi.work(p), rd, ra, ti);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (p.pair) cluster_sync();                     // no CTA leaves while its peer may still signal it
  if (warp == W_MMA) {
    tc_fence_after();
    if (CG2) tmem_dealloc2(tmem, p.tmem_cols);
    else tmem_dealloc(tmem, p.tmem_cols);
  }
}

// ---------------------------------------------------------------------------------------------------------
// Residual(DilatedUnit) (blocks.py:84-113 + :32-46) at C = 64 / 128 / 256 in ONE launch by a cluster pair: both
// GEMMs are cta_group::2 MMAs (M = 256: each CTA's own 128-row tile; each CTA holds N / 2 rows of every weight
// block), h never leaves the SM:
//     GEMM1  D1 = conv3_d(z)            A from shared memory (taps = start-row shifts)   -> TMEM stage s, [0, N)
//     E1     h  = leaky(D1 + b3)        packed bf16 written back into D1's own columns (tcgen05.st)
//     GEMM2  D2 = h W1^T                A from TENSOR MEMORY                              -> TMEM stage s, [N, 2N)
//     E2     y  = leaky(D2 + b1 + x)    residual from global memory (the unit's input, L2-hot), 16-byte stores
// At C = 256 the unfused 1x1 conv + residual launch sat at its HBM roofline (h written and read back: 0.54 GB per
// unit); at C <= 128 the one-CTA fused kernel above is bound by its output warps' latency chain (8 warps per phase,
// E1 -> GEMM2 -> E2 in series).  Here all 16 output warps work on one phase at a time and, for C <= 128, TMEM
// holds TWO tiles (2 x 2N <= 512 columns): the leader issues G1(t+1) before G2(t), so E1(t) runs under G1(t+1) and
// E2(t) under the GEMMs after it.  One warp issues every MMA: tcgen05.mma executes in issue order, so G1(t+NS)
// may overwrite D1 / h of tile t without a barrier; only D2 is handed back by the epilogue (D2_EMPTY).  The peer's MMA
// warp relays its CTA's barriers to the leader like the CG2 kernel above.  Weights: conv1's half blocks resident;
// conv3's resident too for C <= 128 (64 KB per CTA at C = 128), streamed through a ring at C = 256.
// Warps: 0-15 output, 16 MMA / relay, 17 W producer, 18 A producer.
template <int NS>
__global__ void __launch_bounds__(TC_THREADS, 1)
    tc_du_pair_kernel(const __grid_constant__ TcParams p, const __grid_constant__ TcTmap tmx) {
  using WorkIter = WorkIterT<false>;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + BAR_COUNT * 8);
  float* bias_s = reinterpret_cast<float*>(smem + p.off_bias);
  float* bias2_s = bias_s + 256;
  uint8_t* a_smem = smem + p.off_a;
  uint8_t* w_smem = smem + p.off_w;                       // conv3 half blocks: ring (streamed) or all of them (resident)
  uint8_t* w1_smem = smem + p.off_h;                      // conv1 half blocks, resident
  const uint32_t bar0 = smem_u32(bars);
  auto BAR = [&](int i) { return bar0 + 8u * (uint32_t)i; };
  int tid;
  asm volatile("mov.u32 %0, %%tid.x;" : "=r"(tid));
  const int warp = tid >> 5, lane = tid & 31;
  asm volatile("griddepcontrol.launch_dependents;");
  const uint32_t rank = cluster_ctarank();
  const uint32_t lead = rank == 0 ? 1u : 0u;
  constexpr int W_MMA = 16, W_WPROD = 17, W_APROD = 18;
  const int N = p.N_TILE;                                 // channels of the unit
  const int nkc = p.n_kchunks;                            // K chunks (KC = 64) of either GEMM
  const int nblk1 = nkc * p.n_taps;                       // conv3 weight blocks per tile
  const bool resident = p.w_resident != 0;
  const uint32_t stage_cols = 2u * (uint32_t)N;
  if (tid == 0) {
    for (int i = 0; i < MAX_SA; ++i) {
      mbar_init(BAR(BAR_A_FULL + i), 1 + lead);
      mbar_init(BAR(BAR_A_EMPTY + i), 1);
    }
    for (int i = 0; i < MAX_SW; ++i) {
      mbar_init(BAR(BAR_W_FULL + i), 1 + lead);
      mbar_init(BAR(BAR_W_EMPTY + i), 1);
    }
    mbar_init(BAR(BAR_RAW_FULL), 1 + lead);               // resident weights landed (once)
    for (int st = 0; st < NS; ++st) {
      mbar_init(BAR(BAR_D_FULL + st), 1);
      mbar_init(BAR(BAR_H_FULL + 2 * st), 16 + lead);      // lower K half of h
      mbar_init(BAR(BAR_H_FULL + 2 * st + 1), 16 + lead);  // upper K half
      mbar_init(BAR(BAR_D2_FULL + st), 1);
      mbar_init(BAR(BAR_D2_EMPTY + st), 16 + lead);
    }
    fence_mbar_init();
  }
  for (int i = tid; i < N; i += TC_THREADS) {
    bias_s[i] = p.bias ? p.bias[i] : 0.f;
    bias2_s[i] = p.bias2 ? p.bias2[i] : 0.f;
  }
  if (warp == W_MMA) {
    tmem_alloc2(smem_u32(tmem_slot), 512);
    tmem_relinquish2();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  cluster_sync();
  if (warp != W_WPROD) asm volatile("griddepcontrol.wait;" ::: "memory");

  if (warp == W_APROD) {
    if (lane == 0) prefetch_tmap(&tmx);
    Ring ra;
    WorkIter wi;
    for (wi.init(p); wi.left > 0; wi.next(p)) {
      const Work wk = wi.work(p);
      for (int kc = 0; kc < nkc; ++kc) {
        const uint32_t sa = ra.idx;
        mbar_wait(BAR(BAR_A_EMPTY + sa), ra.ph ^ 1);
        issue_a_tma(p, &tmx, wk.b, wk.t0 - p.off, kc, smem_u32(a_smem + (size_t)sa * p.a_stage_bytes),
                    BAR(BAR_A_FULL + sa), lane);
        ra.next(p.SA);
      }
    }
  } else if (warp == W_WPROD) {
    if (lane == 0) {
      // resident: conv1's blocks (behind conv3's in the packed order) and, for C <= 128, conv3's as well, once
      const int r0 = resident ? 0 : nblk1, rn = nblk1 + nkc - r0;
      mbar_expect_tx(BAR(BAR_RAW_FULL), (uint32_t)rn * p.w_blk_bytes);
      for (int i = 0; i < rn; ++i) {
        const int blk = r0 + i;
        uint8_t* dst = blk < nblk1 ? w_smem + (size_t)blk * p.w_blk_bytes : w1_smem + (size_t)(blk - nblk1) * p.w_blk_bytes;
        bulk_g2s(smem_u32(dst), reinterpret_cast<const uint8_t*>(p.w) + ((size_t)blk * 2 + rank) * p.w_blk_bytes,
                 p.w_blk_bytes, BAR(BAR_RAW_FULL));
      }
      if (!resident) {
        // conv3's blocks of every tile through the ring: while E1 / GEMM2 of a tile run it fills with the first
        // blocks of the next tile
        Ring rw;
        WorkIter wi;
        for (wi.init(p); wi.left > 0; wi.next(p)) {
          for (int blk = 0; blk < nblk1; ++blk) {          // (K chunk, tap) order
            const uint32_t sw = rw.idx;
            mbar_wait(BAR(BAR_W_EMPTY + sw), rw.ph ^ 1);
            mbar_expect_tx(BAR(BAR_W_FULL + sw), p.w_blk_bytes);
            bulk_g2s(smem_u32(w_smem + (size_t)sw * p.w_blk_bytes),
                     reinterpret_cast<const uint8_t*>(p.w) + ((size_t)blk * 2 + rank) * p.w_blk_bytes, p.w_blk_bytes,
                     BAR(BAR_W_FULL + sw));
            rw.next(p.SW);
          }
        }
      }
    }
  } else if (warp == W_MMA) {
    const bool leader = elect_one();
    Ring ra, rw;
    WorkIter wi;
    wi.init(p);
    const int n_items = wi.left;
    if (lead) {
      const uint32_t idesc2 = make_idesc_bf16(256, N);
      const uint32_t a0 = smem_u32(a_smem), w0 = smem_u32(w_smem), w10 = smem_u32(w1_smem);
      const uint32_t dhi = desc_hi(128u);
      const uint32_t lbo_a = (uint32_t)p.R_ALLOC * 16u, lbo_w = (uint32_t)(N / 2) * 16u;
      const uint32_t a_lbo_bits = ((lbo_a >> 4) & 0x3FFFu) << 16, w_lbo_bits = ((lbo_w >> 4) & 0x3FFFu) << 16;
      const uint32_t a_kadd = (2u * lbo_a) >> 4, w_kadd = (2u * lbo_w) >> 4;
      mbar_wait(BAR(BAR_RAW_FULL), 0);
      tc_fence_after();
      // GEMM1 of tile t -> D1 of stage t % NS
      auto g1 = [&](int t) {
        const uint32_t dcol = tmem + (uint32_t)(t % NS) * stage_cols;
        uint32_t acc = 0;
        for (int kc = 0; kc < nkc; ++kc) {
          const uint32_t sa = ra.idx;
          mbar_wait(BAR(BAR_A_FULL + sa), ra.ph);
          ra.next(p.SA);
          tc_fence_after();
          const uint32_t alo0 = (((a0 + sa * p.a_stage_bytes) >> 4) & 0x3FFFu) | a_lbo_bits;
          for (int j = 0; j < p.n_taps; ++j) {
            uint32_t bl, sw = 0;
            if (resident) {
              bl = (((w0 + (uint32_t)(kc * p.n_taps + j) * p.w_blk_bytes) >> 4) & 0x3FFFu) | w_lbo_bits;
            } else {
              sw = rw.idx;
              mbar_wait(BAR(BAR_W_FULL + sw), rw.ph);
              rw.next(p.SW);
              tc_fence_after();
              bl = (((w0 + sw * p.w_blk_bytes) >> 4) & 0x3FFFu) | w_lbo_bits;
            }
            uint32_t al = alo0 + (uint32_t)(j * p.dil);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              if (leader) mma2_bf16_lohi(dcol, al, bl, dhi, idesc2, acc);
              acc = 1;
              al += a_kadd;
              bl += w_kadd;
            }
            if (!resident && leader) mma2_commit(BAR(BAR_W_EMPTY + sw));
          }
          if (leader) mma2_commit(BAR(BAR_A_EMPTY + sa));
        }
        if (leader) mma2_commit(BAR(BAR_D_FULL + t % NS));
      };
      // GEMM2 of tile t: A = h in tensor memory (K step k at columns [16 k, 16 k + 8) of D1), D2 behind D1; the
      // lower K half as soon as E1 has written it
      auto g2 = [&](int t) {
        const int st = t % NS;
        const uint32_t par = (uint32_t)(t / NS) & 1u;
        const uint32_t dcol = tmem + (uint32_t)st * stage_cols + (uint32_t)N;
        mbar_wait(BAR(BAR_H_FULL + 2 * st), par);
        mbar_wait(BAR(BAR_D2_EMPTY + st), par ^ 1);
        tc_fence_after();
        uint32_t at = tmem + (uint32_t)st * stage_cols, acc2 = 0;
        const int ks = nkc * 4;
        for (int k = 0; k < ks; ++k) {
          if (k == ks / 2) {
            mbar_wait(BAR(BAR_H_FULL + 2 * st + 1), par);
            tc_fence_after();
          }
          const uint32_t bl = ((((w10 + (uint32_t)(k >> 2) * p.w_blk_bytes) >> 4) & 0x3FFFu) | w_lbo_bits) +
                              (uint32_t)(k & 3) * w_kadd;
          if (leader) mma2_bf16_ts(dcol, at, bl, dhi, idesc2, acc2);
          acc2 = 1;
          at += 16u;
        }
        if (leader) mma2_commit(BAR(BAR_D2_FULL + st));
      };
      if (NS == 1) {
        for (int t = 0; t < n_items; ++t) { g1(t); g2(t); }
      } else {
        if (n_items > 0) g1(0);
        for (int t = 0; t < n_items; ++t) {
          if (t + 1 < n_items) g1(t + 1);
          g2(t);
        }
      }
    } else {
      // relay: this CTA's stages, in the order the leader waits for them
      auto fwd = [&](int b) { if (leader) mbar_arrive_cluster(mapa_shared(BAR(b), 0)); };
      mbar_wait(BAR(BAR_RAW_FULL), 0);
      fwd(BAR_RAW_FULL);
      auto r1 = [&](int) {
        for (int kc = 0; kc < nkc; ++kc) {
          const uint32_t sa = ra.idx;
          mbar_wait(BAR(BAR_A_FULL + sa), ra.ph);
          ra.next(p.SA);
          fwd(BAR_A_FULL + sa);
          if (!resident)
            for (int j = 0; j < p.n_taps; ++j) {
              const uint32_t sw = rw.idx;
              mbar_wait(BAR(BAR_W_FULL + sw), rw.ph);
              rw.next(p.SW);
              fwd(BAR_W_FULL + sw);
            }
        }
      };
      auto r2 = [&](int t) {
        const int st = t % NS;
        const uint32_t par = (uint32_t)(t / NS) & 1u;
        mbar_wait(BAR(BAR_H_FULL + 2 * st), par);
        fwd(BAR_H_FULL + 2 * st);
        if (t >= NS) {                                     // the first use of a stage needs no drained D2 (see relay_cg2)
          mbar_wait(BAR(BAR_D2_EMPTY + st), par ^ 1);
          fwd(BAR_D2_EMPTY + st);
        }
        mbar_wait(BAR(BAR_H_FULL + 2 * st + 1), par);
        fwd(BAR_H_FULL + 2 * st + 1);
      };
      if (NS == 1) {
        for (int t = 0; t < n_items; ++t) { r1(t); r2(t); }
      } else {
        if (n_items > 0) r1(0);
        for (int t = 0; t < n_items; ++t) {
          if (t + 1 < n_items) r1(t + 1);
          r2(t);
        }
      }
    }
  } else if (warp < 16) {
    const int rit = (warp & 3) * 32 + lane;
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    const int part = warp >> 2;                            // 4 column parts
    const int ppw = N >> 6;                                // 16-column pieces per warp (1, 2 or 4)
    WorkIter wi;
    wi.init(p);
    const int n_items = wi.left;
    // E1 of tile t: pieces 16 part + 64 i -- every warp finishes the lower K half of h before the upper one
    auto e1 = [&](int t) {
      const int st = t % NS;
      const uint32_t d1 = tmem + lane_base + (uint32_t)st * stage_cols;
      mbar_wait(BAR(BAR_D_FULL + st), (uint32_t)(t / NS) & 1u);
      tc_fence_after();
      int i = 0;
      const bool on = !(p.dbg & 1);
#pragma unroll 1
      for (int half = 0; half < 2; ++half) {
        const int iend = half ? ppw : (ppw + 1) / 2;
#pragma unroll 1
        for (; i < iend && on; ++i) {
          const int c = 16 * part + 64 * i;
          h_chunk<RB_ACT_LEAKY, 16, true>(d1 + (uint32_t)c, bias_s + c, nullptr, 0, nullptr);
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(BAR(BAR_H_FULL + 2 * st + half));
      }
    };
    auto e2 = [&](int t, const Work& wk) {
      const int st = t % NS;
      const uint32_t par = (uint32_t)(t / NS) & 1u;
      if (p.dbg & 2) mbar_wait(BAR(BAR_D2_FULL + st), par);   // (timing experiments: no drain)
      else
        drain_res16(p, BAR(BAR_D2_FULL + st), par, tmem + lane_base + (uint32_t)st * stage_cols + (uint32_t)N, bias2_s,
                    wk, rit, part, 4, N, 0xffffffffu, N < 128);
      __syncwarp();
      if (lane == 0) mbar_arrive(BAR(BAR_D2_EMPTY + st));
    };
    if (NS == 1) {
      for (int t = 0; t < n_items; ++t, wi.next(p)) { e1(t); e2(t, wi.work(p)); }
    } else {
      if (n_items > 0) e1(0);
      for (int t = 0; t < n_items; ++t, wi.next(p)) {
        if (t + 1 < n_items) e1(t + 1);
        e2(t, wi.work(p));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync();
  if (warp == W_MMA) {
    tc_fence_after();
    tmem_dealloc2(tmem, 512);
  }
}

template <int PRE, int ACT2, bool FUSED, int DR, bool TR = false, int NS = 2, bool CG2 = false, bool HT = false>
static int launch_variant(const TcParams& p, const TcTmap& tm, int grid, cudaStream_t s) {
  static bool attr_set = false;
  if (!attr_set) {
    RB_CUDA(cudaFuncSetAttribute(tc_conv_kernel<PRE, ACT2, FUSED, DR, TR, NS, CG2, HT>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(NS == 3 ? 896 : TC_THREADS);
  cfg.dynamicSmemBytes = p.smem_bytes;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  attr[1].id = cudaLaunchAttributeClusterDimension;
  attr[1].val.clusterDim.x = 1u << p.pair;
  attr[1].val.clusterDim.y = 1;
  attr[1].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = p.pair ? 2 : 1;
  RB_CUDA(cudaLaunchKernelEx(&cfg, tc_conv_kernel<PRE, ACT2, FUSED, DR, TR, NS, CG2, HT>, p, tm));
  return RB_OK;
}

// 5-D tensor map over the input view (16-byte unit u(b,t,g) of tc_conv.cuh):
//   dim0 8 bf16 | dim1 phase g/G1 (stride sB) | dim2 row t - lo (stride sT) | dim3 g%G1 (stride sA)
//   | dim4 stream (stride sb);   box = {8, 1, a_box_rows, a_gb, 1}, zero fill out of range.
static int make_tmap(const TcParams& p, TcTmap* tm) {
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                               const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                               CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeFn encode = nullptr;
  if (!encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qr;
    RB_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr));
    RB_CHECK(fn && qr == cudaDriverEntryPointSuccess, RB_ERR_CUDA, "cuTensorMapEncodeTiled is not available");
    encode = (EncodeFn)fn;
  }
  static_assert(sizeof(TcTmap) == sizeof(CUtensorMap), "tensor map size");
  const TcView& x = p.x;
  const int groups = x.C / 8;
  const int phases = (groups + x.G1 - 1) / x.G1;
  RB_CHECK(groups % x.G1 == 0 || phases == 1, RB_ERR_INVALID, "view groups %d not a multiple of G1 %d", groups, x.G1);
  const int g1 = phases == 1 ? groups : x.G1;
  const long long base_units = (long long)(x.hist + x.lo) * x.sT;
  void* base = (void*)(reinterpret_cast<const uint4*>(x.p) + base_units);
  auto stride = [](long long units) { return (cuuint64_t)(units > 0 ? units : 1) * 16ull; };
  const cuuint64_t dims[5] = {8, (cuuint64_t)phases, (cuuint64_t)(x.n - x.lo), (cuuint64_t)g1, (cuuint64_t)p.B};
  const cuuint64_t strides[4] = {stride(x.sB), stride(x.sT), stride(x.sA), stride(x.sb)};
  const cuuint32_t box[5] = {8, 1, (cuuint32_t)p.a_box_rows, (cuuint32_t)p.a_gb, 1};
  const cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  const CUresult r = encode(reinterpret_cast<CUtensorMap*>(tm), CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, base, dims,
                            strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                            CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  RB_CHECK(r == CUDA_SUCCESS, RB_ERR_CUDA,
           "cuTensorMapEncodeTiled failed (%d): dims {8,%d,%d,%d,%d} strides {%lld,%lld,%lld,%lld} box {%d,%d}", (int)r,
           phases, x.n - x.lo, g1, p.B, (long long)strides[0], (long long)strides[1], (long long)strides[2],
           (long long)strides[3], p.a_box_rows, p.a_gb);
  return RB_OK;
}

// Launches tc_du_pair_kernel: p as for a plain cta_group::2 conv3 launch (x / y / res views, w = N-split weight
// halves of conv3 followed by the 1x1 conv's, bias / bias2), geometry filled by the engine.
int launch_tc_du_pair(TcParams& p, cudaStream_t s) {
  if (p.n_work <= 0) return RB_OK;
  static int n_sm = 0;
  if (!n_sm) {
    int dev = 0;
    RB_CUDA(cudaGetDevice(&dev));
    RB_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
  }
  RB_CHECK(p.smem_bytes <= 227u * 1024u && (p.N_TILE == 256 || p.N_TILE == 128 || p.N_TILE == 64) && p.KC == 64 &&
               p.n_kchunks * 64 == p.N_TILE && p.n_ntiles == 1 && p.pair == 1 && p.MT == 1 && p.res.p &&
               p.pre_act == RB_ACT_NONE && p.SW <= MAX_SW && p.SA <= MAX_SA && (p.w_resident || p.N_TILE == 256),
           RB_ERR_INVALID, "pair DilatedUnit launch outside its envelope");
  const bool two = p.N_TILE <= 128;                        // two tiles in TMEM
  const int slots = n_sm / 2;
  const int stride = p.n_work < slots ? p.n_work : slots;
  {
    const int rt = stride / p.n_ntiles;
    p.step_nt = stride - rt * p.n_ntiles;
    p.step_b = rt / p.tiles_per_stream;
    p.step_tile = rt - p.step_b * p.tiles_per_stream;
    const int rt2 = 2 * stride / p.n_ntiles;
    p.step2_nt = 2 * stride - rt2 * p.n_ntiles;
    p.step2_b = rt2 / p.tiles_per_stream;
    p.step2_tile = rt2 - p.step2_b * p.tiles_per_stream;
  }
  TcTmap tm;
  int rc = make_tmap(p, &tm);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    RB_CUDA(cudaFuncSetAttribute(tc_du_pair_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    RB_CUDA(cudaFuncSetAttribute(tc_du_pair_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * stride);
  cfg.blockDim = dim3(TC_THREADS);
  cfg.dynamicSmemBytes = p.smem_bytes;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  attr[1].id = cudaLaunchAttributeClusterDimension;
  attr[1].val.clusterDim.x = 2;
  attr[1].val.clusterDim.y = 1;
  attr[1].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  if (two) RB_CUDA(cudaLaunchKernelEx(&cfg, tc_du_pair_kernel<2>, p, tm));
  else RB_CUDA(cudaLaunchKernelEx(&cfg, tc_du_pair_kernel<1>, p, tm));
  ++g_launches;
  RB_CUDA(cudaGetLastError());
  return RB_OK;
}

int launch_tc_conv(TcParams& p, cudaStream_t s) {
  if (p.n_work <= 0) return RB_OK;
  static int n_sm = 0;
  if (!n_sm) {
    int dev = 0;
    RB_CUDA(cudaGetDevice(&dev));
    RB_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
  }
  RB_CHECK(p.smem_bytes <= 227u * 1024u, RB_ERR_UNSUPPORTED, "tc conv needs %u B of shared memory",
           p.smem_bytes);
  RB_CHECK(p.out_act == RB_ACT_NONE || p.out_act == RB_ACT_LEAKY, RB_ERR_UNSUPPORTED,
           "only LeakyReLU can be stored activated");
  RB_CHECK(!p.pair || (!p.fused && !p.w_resident && p.pre_act == RB_ACT_NONE), RB_ERR_INVALID,
           "pair mode is for plain streamed layers");
  RB_CHECK(p.dual_g1 <= 1 || (p.w_resident && p.SA >= p.dual_g1 * p.n_kchunks && p.SD >= p.dual_g1), RB_ERR_INVALID,
           "%d GEMM1 issuers need resident weights, %d A stages and as many accumulator stages", p.dual_g1,
           p.dual_g1 * p.n_kchunks);
  RB_CHECK((p.cg2 != 0) == (p.flat_units != 0), RB_ERR_INVALID, "flat work numbering is the cta_group::2 variants'");
  // pair mode: one cluster of two CTAs per work slot
  const int slots = n_sm >> p.pair;
  const int stride = p.n_work < slots ? p.n_work : slots;
  const int grid = stride << p.pair;
  {
    const int rt = stride / p.n_ntiles;                    // grid stride, decomposed (WorkIter::next)
    p.step_nt = stride - rt * p.n_ntiles;
    p.step_b = rt / p.tiles_per_stream;
    p.step_tile = rt - p.step_b * p.tiles_per_stream;
    const int rt2 = 2 * stride / p.n_ntiles;
    p.step2_nt = 2 * stride - rt2 * p.n_ntiles;
    p.step2_b = rt2 / p.tiles_per_stream;
    p.step2_tile = rt2 - p.step2_b * p.tiles_per_stream;
  }
  TcTmap tm;
  int rc = make_tmap(p, &tm);
  if (rc) return rc;
  rc = RB_ERR_UNSUPPORTED;
  // drain_fast: whole 32-column chunks whose four channel groups sit sA apart, plain stores
  const bool fast = !p.head && !p.out_mask && (p.N_TILE & 31) == 0 && (p.y.G1 & 3) == 0 &&
                    (p.res.p == nullptr || p.res_smem || (p.res.G1 & 3) == 0);
  // h in tensor memory is compiled into one variant (and its traced twin): the fast fused unit of a LeakyReLU
  // net with its residual in global memory -- tc_engine.cu sets h_tmem under the same conditions
  RB_CHECK(!p.sbias || (fast && !p.fused && !p.res.p && !p.pair_mode && p.pre_act == RB_ACT_NONE && !p.trace &&
                        (p.sbias_n & 3) == 0), RB_ERR_INVALID, "per-stream bias on a layer without that kernel variant");
  RB_CHECK(!p.h_tmem || (p.fused && fast && p.res.p && !p.res_smem && p.pre_act == RB_ACT_NONE && p.act2 == RB_ACT_LEAKY),
           RB_ERR_INVALID, "h_tmem set for a layer without that kernel variant");
  if (p.trace && p.pre_act == RB_ACT_NONE && fast && (!p.fused || (p.act2 == RB_ACT_LEAKY && p.res.p))) {
    // traced variants (debug hook): the headline kernels only
    if (!p.fused) rc = launch_variant<RB_ACT_NONE, RB_ACT_NONE, false, 1, true>(p, tm, grid, s);
    else if (p.res_smem) rc = launch_variant<RB_ACT_NONE, RB_ACT_LEAKY, true, 2, true>(p, tm, grid, s);
    else if (p.h_tmem) rc = launch_variant<RB_ACT_NONE, RB_ACT_LEAKY, true, 1, true, 2, false, true>(p, tm, grid, s);
    else rc = launch_variant<RB_ACT_NONE, RB_ACT_LEAKY, true, 1, true>(p, tm, grid, s);
  } else if (!p.fused) {
    if (p.cg2) {
      RB_CHECK(p.pair && p.pre_act == RB_ACT_NONE && !p.w_resident && p.N_TILE == 256 &&
                   (fast || p.pair_mode), RB_ERR_INVALID, "cta_group::2 layer outside its envelope");
      rc = p.pair_mode ? launch_variant<RB_ACT_NONE, RB_ACT_NONE, false, 3, false, 2, true>(p, tm, grid, s)
           : p.sbias   ? launch_variant<RB_ACT_NONE, RB_ACT_NONE, false, 4, false, 2, true>(p, tm, grid, s)
                       : launch_variant<RB_ACT_NONE, RB_ACT_NONE, false, 1, false, 2, true>(p, tm, grid, s);
    } else if (p.pre_act == RB_ACT_NONE && p.sbias) {
      rc = launch_variant<RB_ACT_NONE, RB_ACT_NONE, false, 4>(p, tm, grid, s);
    } else if (p.pre_act == RB_ACT_NONE && p.pair_mode) {
      rc = launch_variant<RB_ACT_NONE, RB_ACT_NONE, false, 3>(p, tm, grid, s);
    } else if (p.pre_act == RB_ACT_NONE) {
      rc = fast ? launch_variant<RB_ACT_NONE, RB_ACT_NONE, false, 1>(p, tm, grid, s)
                : launch_variant<RB_ACT_NONE, RB_ACT_NONE, false, 0>(p, tm, grid, s);
    } else if (p.pre_act == RB_ACT_LEAKY) {
      rc = launch_variant<RB_ACT_LEAKY, RB_ACT_NONE, false, 0>(p, tm, grid, s);
    } else if (p.pre_act == RB_ACT_SNAKE) {
      rc = launch_variant<RB_ACT_SNAKE, RB_ACT_NONE, false, 0>(p, tm, grid, s);
    }
  } else {
    if (p.pre_act == RB_ACT_LEAKY && p.act2 == RB_ACT_LEAKY) {
      rc = launch_variant<RB_ACT_LEAKY, RB_ACT_LEAKY, true, 0>(p, tm, grid, s);
    } else if (p.pre_act == RB_ACT_NONE && p.act2 == RB_ACT_LEAKY) {
      // three output-warp sets measured equal to two (0.57 vs 0.56 ms on the C = 64 units): the tile
      // rate is set upstream of the output warps; kept selectable for experiments
      static const bool three = getenv("RB_TC_THREESETS") != nullptr;
      if (fast && p.res.p && p.res_smem && three)
        rc = launch_variant<RB_ACT_NONE, RB_ACT_LEAKY, true, 2, false, 3>(p, tm, grid, s);
      else if (fast && p.res.p && p.res_smem) rc = launch_variant<RB_ACT_NONE, RB_ACT_LEAKY, true, 2>(p, tm, grid, s);
      else if (fast && p.res.p && p.h_tmem)
        rc = launch_variant<RB_ACT_NONE, RB_ACT_LEAKY, true, 1, false, 2, false, true>(p, tm, grid, s);
      else if (fast && p.res.p) rc = launch_variant<RB_ACT_NONE, RB_ACT_LEAKY, true, 1>(p, tm, grid, s);
      else rc = launch_variant<RB_ACT_NONE, RB_ACT_LEAKY, true, 0>(p, tm, grid, s);
    } else if (p.pre_act == RB_ACT_SNAKE && p.act2 == RB_ACT_SNAKE) {
      rc = launch_variant<RB_ACT_SNAKE, RB_ACT_SNAKE, true, 0>(p, tm, grid, s);
    } else if (p.pre_act == RB_ACT_NONE && p.act2 == RB_ACT_NONE) {
      rc = launch_variant<RB_ACT_NONE, RB_ACT_NONE, true, 0>(p, tm, grid, s);
    }
  }
  if (rc == RB_ERR_UNSUPPORTED) {
    set_error("tc conv: activation combination (%d, %d, fused=%d) has no kernel variant", p.pre_act, p.act2, p.fused);
    return rc;
  }
  if (rc) return rc;
  ++g_launches;
  RB_CUDA(cudaGetLastError());
  return RB_OK;
}

}  // namespace tc
}  // namespace rb
