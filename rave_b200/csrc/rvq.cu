// rvq.cu -- residual vector quantisation (discrete config), bit-exact code indices.
//
// Reference: ResidualVectorQuantization.encode/decode (rave/quantization.py:302-318) ->
// VectorQuantization.encode/decode (:239-249) -> EuclideanCodebook.quantize (:131-136):
//     dist = -(x.pow(2).sum(1) - 2 * x @ E^T + E.pow(2).sum(1));  ind = dist.max(-1).indices
// Index parity needs the same fp32 roundings as the reference's CPU evaluation, so the
// kernel reproduces its arithmetic ORDER (established by probing torch 2.11 CPU, see
// DESIGN.md "RVQ bit-exactness"):
//   * `(2*x) @ E^T` (python precedence) is one sequential FMA chain over k = 0..D-1 per
//     (row, code), starting from 0  -> `dot_chain`;
//   * `.pow(2).sum()` over a contiguous row multiplies first (rounded), then adds 8-lane
//     vectors v_i = a[8i..8i+7] into four interleaved accumulators p_k = sum_i v[4i+k]
//     (in i order), folds ((p0+p1)+p2)+p3 and finally adds the 8 lanes left to right
//     -> `torch_row_sumsq` (valid for D < 512, D % 32 == 0);
//   * ties resolve to the lowest index (torch.max returns the first maximal element).
// All adds / muls use the _rn intrinsics so that nvcc cannot contract them into FMAs.
#include "common.cuh"

namespace rb {

constexpr int RVQ_ROWS = 128;   // rows (latent vectors) per CTA
constexpr int RVQ_CODES = 64;   // codes per smem chunk
constexpr int RVQ_MAXD = 256;

// sum of squares of v[0..D) (stride `st`) in torch's CPU summation order
__device__ __forceinline__ float torch_row_sumsq(const float* v, int D, int st) {
  float s[8];
  const int nv = D >> 3;        // number of 8-lane vectors, multiple of 4
#pragma unroll
  for (int l = 0; l < 8; ++l) {
    float p[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float acc = 0.f;
      for (int i = 0; i < nv / 4; ++i) {
        const float a = v[((4 * i + k) * 8 + l) * st];
        acc = __fadd_rn(acc, __fmul_rn(a, a));
      }
      p[k] = acc;
    }
    s[l] = __fadd_rn(__fadd_rn(__fadd_rn(p[0], p[1]), p[2]), p[3]);
  }
  float t = s[0];
#pragma unroll
  for (int l = 1; l < 8; ++l) t = __fadd_rn(t, s[l]);
  return t;
}

__global__ void rvq_norms_kernel(const float* __restrict__ cb, int rows, int D, float* __restrict__ e2) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < rows) e2[r] = torch_row_sumsq(cb + (size_t)r * D, D, 1);
}

int launch_rvq_norms(const float* cb, int rows, int D, float* e2, cudaStream_t s) {
  RB_CHECK(D % 32 == 0 && D < 512 && D <= RVQ_MAXD, RB_ERR_UNSUPPORTED,
           "RVQ dim %d: bit-exact summation order is implemented for D %% 32 == 0, D <= %d", D, RVQ_MAXD);
  rvq_norms_kernel<<<(rows + 127) / 128, 128, 0, s>>>(cb, rows, D, e2);
  ++g_launches;
  RB_CUDA(cudaGetLastError());
  return RB_OK;
}

// One CTA = 128 latent vectors through all Q stages.  Residuals stay in shared memory across
// the stages ([k][row]: a thread's 8 rows are two float4); codebook chunks of 64 codes are staged
// transposed ([k][code], the 4-code groups of row k XOR-swizzled by (k >> 2) & 7 so that the
// coalesced global read -- a warp takes 128 contiguous bytes of four codes -- stores without
// bank conflicts).  Thread (ty, tx) owns rows 8ty..8ty+7 x codes 4tx..4tx+3 of each chunk:
// 32 independent FMA chains, 3 LDS.128 per 32 FMA.
constexpr int RVQ_RPT = RVQ_ROWS / 16;      // rows per thread
static_assert(RVQ_ROWS == 128 && RVQ_CODES == 64 && RVQ_RPT == 8, "index math below assumes 128 rows x 64 codes");

__global__ void __launch_bounds__(256)
rvq_encode_kernel(const float* __restrict__ z, int B, int D, int F, const float* __restrict__ cb,
                  const float* __restrict__ e2, int Q, int K, long long* __restrict__ idx) {
  extern __shared__ float smem[];
  float* rs = smem;                         // [D][128] residuals, row-minor
  float* es = rs + (size_t)D * RVQ_ROWS;    // [D][64] codebook chunk, code-minor (swizzled)
  float* t1 = es + (size_t)D * RVQ_CODES;   // [128]
  float* bestv = t1 + RVQ_ROWS;             // [128][16]
  int* besti = (int*)(bestv + RVQ_ROWS * 16);
  int* win = besti + RVQ_ROWS * 16;         // [128]
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const long long N = (long long)B * F;
  const long long row0 = (long long)blockIdx.x * RVQ_ROWS;
  // load residual = z[b, :, f] for row n = b*F + f   (permute(0,2,1).reshape(-1, D))
  for (int i = tid; i < D * RVQ_ROWS; i += 256) {
    const int k = i >> 7, r = i & (RVQ_ROWS - 1);
    const long long n = row0 + r;
    float v = 0.f;
    if (n < N) {
      const long long b = n / F, f = n - b * F;
      v = z[(b * D + k) * F + f];
    }
    rs[i] = v;
  }
  __syncthreads();
  for (int q = 0; q < Q; ++q) {
    if (tid < RVQ_ROWS) t1[tid] = torch_row_sumsq(rs + tid, D, RVQ_ROWS);
    float bv[RVQ_RPT];
    int bi[RVQ_RPT];
#pragma unroll
    for (int r = 0; r < RVQ_RPT; ++r) { bv[r] = INFINITY; bi[r] = 0x7fffffff; }
    const float* cq = cb + (size_t)q * K * D;
    for (int c0 = 0; c0 < K; c0 += RVQ_CODES) {
      __syncthreads();                      // previous chunk consumed, t1 visible
      // stage the chunk: lane = (code-in-group, k4 within a run of 8) reads 16 bytes; the four words go
      // to rows 4k4..4k4+3 at group (g ^ (k4 & 7)): 8 runs x 4 codes cover the 32 banks exactly once
      for (int i = tid; i < (D >> 2) * RVQ_CODES; i += 256) {
        const int cc = i & 3, kk = (i >> 2) & 7, w = i >> 5;
        const int g = w & 15, k4 = ((w >> 4) << 3) + kk;
        const int c = 4 * g + cc;
        float4 e = make_float4(0.f, 0.f, 0.f, 0.f);
        if (k4 < (D >> 2) && c0 + c < K) e = __ldg(reinterpret_cast<const float4*>(cq + (size_t)(c0 + c) * D) + k4);
        if (k4 < (D >> 2)) {
          float* d = es + (size_t)(4 * k4) * RVQ_CODES + 4 * (g ^ (k4 & 7)) + cc;
          d[0] = e.x; d[RVQ_CODES] = e.y; d[2 * RVQ_CODES] = e.z; d[3 * RVQ_CODES] = e.w;
        }
      }
      __syncthreads();
      // The reference's chain is acc <- fma(2 x_k, e_k, acc).  Doubling commutes with every rounding
      // (no overflow / subnormals at these magnitudes), so the chain runs on x_k and the result is
      // doubled once: bit-identical, and nothing but FMAs and three LDS.128 in the loop.
      float acc[RVQ_RPT][4];
#pragma unroll
      for (int r = 0; r < RVQ_RPT; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[r][c] = 0.f;
      for (int k0 = 0; k0 < D; k0 += 4) {
        const float* ep = es + (size_t)k0 * RVQ_CODES + 4 * (tx ^ ((k0 >> 2) & 7));
        const float* rp = rs + (size_t)k0 * RVQ_ROWS + ty * RVQ_RPT;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const float4 ra = *reinterpret_cast<const float4*>(rp + kk * RVQ_ROWS);
          const float4 rb = *reinterpret_cast<const float4*>(rp + kk * RVQ_ROWS + 4);
          const float4 ev = *reinterpret_cast<const float4*>(ep + kk * RVQ_CODES);
          const float rr[8] = {ra.x, ra.y, ra.z, ra.w, rb.x, rb.y, rb.z, rb.w};
          const float e[4] = {ev.x, ev.y, ev.z, ev.w};
#pragma unroll
          for (int r = 0; r < RVQ_RPT; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c) acc[r][c] = __fmaf_rn(rr[r], e[c], acc[r][c]);
        }
      }
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int code = c0 + tx * 4 + c;
        if (code < K) {
          const float t3 = __ldg(e2 + (size_t)q * K + code);
#pragma unroll
          for (int r = 0; r < RVQ_RPT; ++r) {
            // w = (|x|^2 - 2x.e) + |e|^2 ; dist = -w ; argmax dist = first argmin w
            const float w = __fadd_rn(__fsub_rn(t1[ty * RVQ_RPT + r], __fmul_rn(2.f, acc[r][c])), t3);
            if (w < bv[r]) { bv[r] = w; bi[r] = code; }   // codes visited in increasing order
          }
        }
      }
    }
#pragma unroll
    for (int r = 0; r < RVQ_RPT; ++r) {
      bestv[(ty * RVQ_RPT + r) * 16 + tx] = bv[r];
      besti[(ty * RVQ_RPT + r) * 16 + tx] = bi[r];
    }
    __syncthreads();
    if (tid < RVQ_ROWS) {
      float v = bestv[tid * 16];
      int ix = besti[tid * 16];
      for (int j = 1; j < 16; ++j) {
        const float vj = bestv[tid * 16 + j];
        const int ij = besti[tid * 16 + j];
        if (vj < v || (vj == v && ij < ix)) { v = vj; ix = ij; }
      }
      win[tid] = ix;
      const long long n = row0 + tid;
      if (n < N) {
        const long long b = n / F, f = n - b * F;
        idx[(b * Q + q) * F + f] = ix;
      }
    }
    __syncthreads();
    // residual -= E[q][win]   (quantization.py:307-308); lane -> row: conflict-free smem, 16-byte reads
    for (int i = tid; i < (D >> 2) * RVQ_ROWS; i += 256) {
      const int r = i & (RVQ_ROWS - 1), k4 = i >> 7;
      const float4 e = __ldg(reinterpret_cast<const float4*>(cq + (size_t)win[r] * D) + k4);
      float* d = rs + (size_t)(4 * k4) * RVQ_ROWS + r;
      d[0] = __fsub_rn(d[0], e.x);
      d[RVQ_ROWS] = __fsub_rn(d[RVQ_ROWS], e.y);
      d[2 * RVQ_ROWS] = __fsub_rn(d[2 * RVQ_ROWS], e.z);
      d[3 * RVQ_ROWS] = __fsub_rn(d[3 * RVQ_ROWS], e.w);
    }
    __syncthreads();
  }
}

int launch_rvq_encode(const float* z, int B, int D, int F, const float* cb, const float* e2, int Q, int K,
                      int64_t* idx, float* /*ws*/, cudaStream_t s) {
  RB_CHECK(D % 32 == 0 && D <= RVQ_MAXD, RB_ERR_UNSUPPORTED, "RVQ dim %d not supported", D);
  const long long N = (long long)B * F;
  const size_t smem = ((size_t)D * (RVQ_ROWS + RVQ_CODES) + RVQ_ROWS + RVQ_ROWS * 16) * sizeof(float) +
                      (RVQ_ROWS * 16 + RVQ_ROWS) * sizeof(int);
  static bool attr_set = false;
  if (!attr_set) {
    RB_CUDA(cudaFuncSetAttribute(rvq_encode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    attr_set = true;
  }
  const int grid = (int)((N + RVQ_ROWS - 1) / RVQ_ROWS);
  rvq_encode_kernel<<<grid, 256, smem, s>>>(z, B, D, F, cb, e2, Q, K, (long long*)idx);
  ++g_launches;
  RB_CUDA(cudaGetLastError());
  return RB_OK;
}

// ResidualVectorQuantization.decode (quantization.py:313-318): zq[b,:,f] = sum_q E_q[idx[b,q,f]]
// accumulated in stage order (0.0 + e_0 + e_1 + ...).  One CTA = 32 frames x D dims, staged in
// smem so that codebook reads are coalesced along d and the store along f.
__global__ void __launch_bounds__(256)
rvq_decode_kernel(const long long* __restrict__ idx, int D, int F, const float* __restrict__ cb, int Q,
                  int K, float* __restrict__ zq) {
  extern __shared__ float tile[];             // [D][33]
  __shared__ int ids[32];
  const int b = blockIdx.y, f0 = blockIdx.x * 32, tid = threadIdx.x;
  for (int i = tid; i < D * 33; i += 256) tile[i] = 0.f;
  for (int q = 0; q < Q; ++q) {
    __syncthreads();
    if (tid < 32) {
      long long v = f0 + tid < F ? idx[((long long)b * Q + q) * F + f0 + tid] : 0;
      v = v < 0 ? 0 : (v >= K ? K - 1 : v);   // export.py:511-513 clamps before decode
      ids[tid] = (int)v;
    }
    __syncthreads();
    for (int i = tid; i < D * 32; i += 256) {
      const int f = i / D, d = i - f * D;
      tile[d * 33 + f] = __fadd_rn(tile[d * 33 + f], __ldg(cb + ((size_t)q * K + ids[f]) * D + d));
    }
  }
  __syncthreads();
  for (int i = tid; i < D * 32; i += 256) {
    const int d = i >> 5, f = i & 31;
    if (f0 + f < F) zq[((size_t)b * D + d) * F + f0 + f] = tile[d * 33 + f];
  }
}

int launch_rvq_decode(const int64_t* idx, int B, int D, int F, const float* cb, int Q, int K, float* zq,
                      cudaStream_t s) {
  dim3 grid((F + 31) / 32, B);
  const size_t smem = (size_t)D * 33 * sizeof(float);
  rvq_decode_kernel<<<grid, 256, smem, s>>>((const long long*)idx, D, F, cb, Q, K, zq);
  ++g_launches;
  RB_CUDA(cudaGetLastError());
  return RB_OK;
}

}  // namespace rb
