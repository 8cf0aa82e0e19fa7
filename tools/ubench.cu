// ubench.cu -- per-SM micro-benchmarks behind the epilogue / MMA budget of tc_conv.cu:
//   ldtm   tcgen05.ld 32x32b.x32 rate with 4 / 8 / 16 warps reading at once
//   mma    back-to-back tcgen05.mma (M=128, K=16) rate for N = 64 / 128 / 256, operands in smem
//   alu    issue rate of the epilogue's instruction mix (FADD2, FMUL2, FMNMX, F2FP, HMUL2, LOP3)
// build:  nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I rave_b200/csrc tools/ubench.cu -o tools/ubench -lcuda
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "tc_ptx.cuh"

using namespace rb::tc;

__global__ void k_ldtm(long long* out, int iters) {
  __shared__ uint32_t slot;
  __shared__ long long t[2];
  const int warp = threadIdx.x >> 5;
  if (warp == 0) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot + ((uint32_t)((warp & 3) * 32) << 16);
  uint32_t acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    uint32_t r[32];
    tmem_ld32(tm + (uint32_t)(((i + warp) * 32) & 511), r);
    tmem_ld_wait();
#pragma unroll
    for (int e = 0; e < 32; ++e) acc ^= r[e];
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) { out[0] = t1 - t0; }
  if (acc == 0x12345) out[1] = acc;
  __syncthreads();
  if (warp == 0) tmem_dealloc(slot, 512);
  (void)t;
}

// x2 in flight per warp (two loads before the wait)
__global__ void k_ldtm2(long long* out, int iters) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot + ((uint32_t)((warp & 3) * 32) << 16);
  uint32_t acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int i = 0; i < iters; i += 2) {
    uint32_t r[32], q[32];
    tmem_ld32(tm + (uint32_t)(((i + warp) * 32) & 511), r);
    tmem_ld32(tm + (uint32_t)(((i + 1 + warp) * 32) & 511), q);
    tmem_ld_wait();
#pragma unroll
    for (int e = 0; e < 32; ++e) acc ^= r[e] + q[e];
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) { out[0] = t1 - t0; }
  if (acc == 0x12345) out[1] = acc;
  __syncthreads();
  if (warp == 0) tmem_dealloc(slot, 512);
}

__global__ void k_mma(long long* out, int N, int n_mma, int rows_alloc) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint32_t slot;
  __shared__ uint64_t bar;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (warp == 0) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 0) {
    const bool leader = elect_one();
    const uint32_t idesc = make_idesc_bf16(128, N);
    const uint32_t a0 = smem_u32(smem), w0 = a0 + 96 * 1024;
    const uint32_t lbo_a = (uint32_t)rows_alloc * 16u, lbo_w = (uint32_t)N * 16u;
    const uint32_t dhi = desc_hi(128u);
    const uint32_t al = desc_lo(a0, lbo_a), bl = desc_lo(w0, lbo_w);
    const long long t0 = clock64();
    for (int i = 0; i < n_mma; ++i) {
      if (leader) mma_bf16_lohi(slot, al + (uint32_t)(i & 3) * ((2u * lbo_a) >> 4), bl + (uint32_t)(i & 3) * ((2u * lbo_w) >> 4), dhi, idesc, i ? 1u : 0u);
    }
    if (leader) mma_commit(smem_u32(&bar));
    const long long t1 = clock64();
    mbar_wait(smem_u32(&bar), 0);
    const long long t2 = clock64();
    if (leader) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc(slot, 512);
}


// groups of `g` MMAs each followed by `ncommit` tcgen05.commit (barriers nobody waits on until the end)
__global__ void k_mma_commit(long long* out, int N, int groups, int g, int ncommit) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint32_t slot;
  __shared__ uint64_t bar[4];
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (warp == 0) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
  if (threadIdx.x == 0) { for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&bar[i]), 1 << 20); mbar_init(smem_u32(&bar[3]), 1); fence_mbar_init(); }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 0) {
    const bool leader = elect_one();
    const uint32_t idesc = make_idesc_bf16(128, N);
    const uint32_t a0 = smem_u32(smem), w0 = a0 + 96 * 1024;
    const uint32_t lbo_a = 130u * 16u, lbo_w = (uint32_t)N * 16u;
    const uint32_t dhi = desc_hi(128u);
    const uint32_t al = desc_lo(a0, lbo_a), bl = desc_lo(w0, lbo_w);
    const long long t0 = clock64();
    for (int q = 0; q < groups; ++q) {
      for (int i = 0; i < g; ++i)
        if (leader) mma_bf16_lohi(slot, al + (uint32_t)(i & 3) * ((2u * lbo_a) >> 4), bl + (uint32_t)(i & 3) * ((2u * lbo_w) >> 4), dhi, idesc, 1u);
      for (int c = 0; c < ncommit; ++c)
        if (leader) mma_commit(smem_u32(&bar[c & 1]));
    }
    if (leader) mma_commit(smem_u32(&bar[3]));
    mbar_wait(smem_u32(&bar[3]), 0);
    const long long t2 = clock64();
    if (leader) { out[0] = t2 - t0; }
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc(slot, 512);
}

// `nw` warps issue MMAs at the same time (different accumulator columns), `n_mma` each, in groups of g + commit
__global__ void k_mma_multi(long long* out, int N, int n_mma, int g, int nw) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint32_t slot;
  __shared__ uint64_t bar[8];
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (warp == 0) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
  if (threadIdx.x == 0) { for (int i = 0; i < 8; ++i) mbar_init(smem_u32(&bar[i]), i < 4 ? 1 : (1 << 20)); fence_mbar_init(); }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const long long t0 = clock64();
  if (warp < nw) {
    const bool leader = elect_one();
    const uint32_t idesc = make_idesc_bf16(128, N);
    const uint32_t a0 = smem_u32(smem) + warp * 20480, w0 = smem_u32(smem) + 96 * 1024 + warp * 8192;
    const uint32_t lbo_a = 130u * 16u, lbo_w = (uint32_t)N * 16u;
    const uint32_t dhi = desc_hi(128u);
    const uint32_t al = desc_lo(a0, lbo_a), bl = desc_lo(w0, lbo_w);
    for (int i = 0; i < n_mma; ++i) {
      if (leader) mma_bf16_lohi(slot + warp * 128, al + (uint32_t)(i & 3) * ((2u * lbo_a) >> 4), bl + (uint32_t)(i & 3) * ((2u * lbo_w) >> 4), dhi, idesc, 1u);
      if ((i % g) == g - 1 && leader) mma_commit(smem_u32(&bar[4 + warp]));
    }
    if (leader) mma_commit(smem_u32(&bar[warp]));
    mbar_wait(smem_u32(&bar[warp]), 0);
  }
  __syncthreads();
  const long long t2 = clock64();
  if (threadIdx.x == 0) out[0] = t2 - t0;
  if (warp == 0) tmem_dealloc(slot, 512);
}

// cp.async.bulk global -> smem: one thread issues `ncopy` copies of `bytes` per round onto one mbarrier
// (4 rounds in flight), `rounds` rounds; every CTA streams its own region of a buffer larger than L2.
__global__ void k_bulk(long long* out, const uint8_t* src, size_t cta_stride, int bytes, int ncopy, int rounds) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar[4];
  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&bar[i]), 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint8_t* g = src + (size_t)blockIdx.x * cta_stride;
    const uint32_t round_bytes = (uint32_t)bytes * ncopy;
    long long t_issue = 0;
    const long long t0 = clock64();
    for (int r = 0; r < rounds; ++r) {
      const int st = r & 3;
      if (r >= 4) mbar_wait(smem_u32(&bar[st]), ((r >> 2) - 1) & 1);
      const long long a = clock64();
      mbar_expect_tx(smem_u32(&bar[st]), round_bytes);
      for (int c = 0; c < ncopy; ++c)
        bulk_g2s(smem_u32(smem) + (uint32_t)st * 40960u + (uint32_t)c * bytes, g + (size_t)r * round_bytes + (size_t)c * bytes,
                 bytes, smem_u32(&bar[st]));
      t_issue += clock64() - a;
    }
    for (int r = rounds - 4; r < rounds; ++r) mbar_wait(smem_u32(&bar[r & 3]), (r >> 2) & 1);
    const long long t1 = clock64();
    if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t_issue; }
  }
}

// HBM streams: mode 0 write-only (STG.128), 1 read-only (LDG.128), 2 copy; every thread strides the buffer
__global__ void k_stream(uint4* dst, const uint4* src, size_t n, int mode, uint4* sink) {
  uint4 acc = make_uint4(0, 0, 0, 0);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    if (mode == 0) dst[i] = make_uint4((uint32_t)i, 1u, 2u, 3u);
    else if (mode == 1) { const uint4 v = src[i]; acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w; }
    else dst[i] = src[i];
  }
  if (mode == 1 && acc.x == 0x12345678u) *sink = acc;
}

// instruction-mix issue rates: every warp runs `iters` rounds of 32 independent ops of one kind
template <int KIND>
__global__ void k_alu(long long* out, float* sink, int iters) {
  float v[32];
#pragma unroll
  for (int e = 0; e < 32; ++e) v[e] = (float)(threadIdx.x + e) * 1e-3f;
  uint32_t u[16];
#pragma unroll
  for (int e = 0; e < 16; ++e) u[e] = threadIdx.x * 2654435761u + e;
  __syncthreads();
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    if (KIND == 0) {            // FADD2 x16
#pragma unroll
      for (int e = 0; e < 32; e += 2) add2(v[e], v[e + 1], 1.0001f, 0.9999f);
    } else if (KIND == 1) {     // FMNMX x32
#pragma unroll
      for (int e = 0; e < 32; ++e) v[e] = fmaxf(v[e], v[(e + 1) & 31] * 0.5f);
    } else if (KIND == 2) {     // F2FP x16 (+ feed back)
#pragma unroll
      for (int e = 0; e < 32; e += 2) {
        __nv_bfloat162 h = __floats2bfloat162_rn(v[e], v[e + 1]);
        u[e >> 1] ^= *reinterpret_cast<uint32_t*>(&h);
      }
    } else if (KIND == 3) {     // HMUL2.BF16 + HMNMX2.BF16 x16
#pragma unroll
      for (int e = 0; e < 16; ++e) {
        __nv_bfloat162 x = *reinterpret_cast<__nv_bfloat162*>(&u[e]);
        x = __hmax2(x, __hmul2(x, __floats2bfloat162_rn(0.2f, 0.2f)));
        u[e] = *reinterpret_cast<uint32_t*>(&x);
      }
    } else if (KIND == 4) {     // LOP3 / shift unpack x32
#pragma unroll
      for (int e = 0; e < 16; ++e) {
        v[2 * e] += __uint_as_float(u[e] << 16);
        v[2 * e + 1] += __uint_as_float(u[e] & 0xffff0000u);
      }
    } else if (KIND == 5) {     // FFMA x32
#pragma unroll
      for (int e = 0; e < 32; ++e) v[e] = fmaf(v[e], 1.0001f, 0.5f);
    }
  }
  const long long t1 = clock64();
  __syncthreads();
  float s = 0;
#pragma unroll
  for (int e = 0; e < 32; ++e) s += v[e];
#pragma unroll
  for (int e = 0; e < 16; ++e) s += (float)u[e];
  sink[threadIdx.x] = s;
  if (threadIdx.x == 0) out[0] = t1 - t0;
}

int main() {
  long long* d;
  float* sink;
  cudaMalloc(&d, 64);
  cudaMalloc(&sink, 4096 * 4);
  long long h[2];
  const int iters = 512;
  for (int w : {4, 8, 16}) {
    k_ldtm<<<1, w * 32>>>(d, iters);
    cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    printf("ldtm  x32, %2d warps, 1 in flight: %.1f cyc / warp-load  -> %.1f B/clk/SM\n", w, (double)h[0] / iters,
           (double)w * iters * 4096 / h[0]);
    k_ldtm2<<<1, w * 32>>>(d, iters);
    cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    printf("ldtm  x32, %2d warps, 2 in flight: %.1f cyc / warp-load  -> %.1f B/clk/SM\n", w, (double)h[0] / iters,
           (double)w * iters * 4096 / h[0]);
  }
  cudaFuncSetAttribute(k_mma, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int N : {32, 64, 128, 256}) {
    const int n = 1024;
    k_mma<<<1, 128, 200 * 1024>>>(d, N, n, 130);
    cudaError_t e = cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    printf("mma   M=128 N=%3d K=16: issue %.1f cyc / mma, complete %.1f cyc / mma (%s)\n", N, (double)h[0] / n,
           (double)h[1] / n, cudaGetErrorString(e));
  }



  {
    const size_t bytes = (size_t)2 << 30, n = bytes / 16;
    uint4 *a, *b;
    cudaMalloc(&a, bytes); cudaMalloc(&b, bytes);
    cudaMemset(a, 1, bytes); cudaMemset(b, 2, bytes);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const char* mn[3] = {"write-only", "read-only", "copy (r+w)"};
    for (int mode = 0; mode < 3; ++mode) {
      float best = 1e9f;
      for (int it = 0; it < 5; ++it) {
        cudaEventRecord(e0);
        k_stream<<<148 * 8, 512>>>(a, b, n, mode, (uint4*)sink);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
      }
      printf("hbm   %-10s 2 GiB: %.3f ms -> %.0f GB/s\n", mn[mode], best, (mode == 2 ? 2.0 : 1.0) * bytes / best / 1e6);
    }
    cudaFree(a); cudaFree(b);
  }
  cudaFuncSetAttribute(k_mma_commit, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int nc : {0, 1, 2}) {
    for (int g : {4, 12}) {
      k_mma_commit<<<1, 128, 200 * 1024>>>(d, 64, 256, g, nc);
      cudaError_t e = cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
      printf("mma+commit N=64: %2d MMAs + %d commits per group: %.0f cyc / group (%s)\n", g, nc, (double)h[0] / 256, cudaGetErrorString(e));
    }
  }

  cudaFuncSetAttribute(k_mma_multi, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int nw : {1, 2, 3}) {
    for (int g : {4, 1024}) {
      const int n = 768 / nw;
      k_mma_multi<<<1, 128, 200 * 1024>>>(d, 64, n, g, nw);
      cudaError_t e = cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
      printf("mma multi N=64: %d warps x %3d MMAs, commit every %4d: %.1f cyc / mma overall (%s)\n", nw, n, g, (double)h[0] / 768, cudaGetErrorString(e));
    }
  }
  {
    uint8_t* big;
    const size_t per_cta = 16u << 20;
    cudaMalloc(&big, per_cta * 148);
    cudaMemset(big, 1, per_cta * 148);
    cudaFuncSetAttribute(k_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (int grid : {1, 148}) {
      for (int bytes : {2080, 4160, 8320, 16640, 33280}) {
        const int ncopy = 33280 / bytes, rounds = 256;
        k_bulk<<<grid, 32, 200 * 1024>>>(d, big, per_cta, bytes, ncopy, rounds);
        cudaError_t e = cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        printf("bulk  grid %3d, %5d B x %2d per round: %.0f cyc / round (%.1f B/clk/SM), issue %.0f cyc / copy (%s)\n", grid,
               bytes, ncopy, (double)h[0] / rounds, (double)bytes * ncopy * rounds / h[0], (double)h[1] / rounds / ncopy,
               cudaGetErrorString(e));
      }
    }
    cudaFree(big);
  }
  const char* names[6] = {"FADD2 x16", "FMUL+FMNMX x32", "F2FP.BF16 x16", "HMUL2+HMNMX2 bf16 x16", "unpack+FADD x32", "FFMA x32"};
  for (int w : {4, 16}) {
    for (int k = 0; k < 6; ++k) {
      const int it = 256;
      switch (k) {
        case 0: k_alu<0><<<1, w * 32>>>(d, sink, it); break;
        case 1: k_alu<1><<<1, w * 32>>>(d, sink, it); break;
        case 2: k_alu<2><<<1, w * 32>>>(d, sink, it); break;
        case 3: k_alu<3><<<1, w * 32>>>(d, sink, it); break;
        case 4: k_alu<4><<<1, w * 32>>>(d, sink, it); break;
        default: k_alu<5><<<1, w * 32>>>(d, sink, it); break;
      }
      cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
      printf("alu   %-24s %2d warps: %.1f cyc / round / warp-per-scheduler\n", names[k], w,
             (double)h[0] / it / (w / 4));
    }
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
