// tc_conv.cuh -- parameters of the tcgen05 implicit-GEMM convolution kernel (tc_conv.cu).
//
// Every convolution of the path (cc.Conv1d, strided cc.Conv1d, cc.ConvTranspose1d as r
// phase-convs, both PQMF banks) is expressed in the bf16 engine as a ROW-SHIFT GEMM over
// row views of the activations:
//
//     D[t, n] = sum_taps j  sum_c  A[t + j*dil - off, c] * W_j[n, c]
//
// where a "row" is one time step of the (possibly reshaped) tensor: a stride-s conv reads
// its input as [T/s][s*C] and a transposed conv writes its output as [T][r*C'] -- the same
// bytes in memory, so striding and phase interleave cost nothing.  M = 128 rows (time) per
// MMA, N = output channels, K = input channels of one tap.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace rb {
namespace tc {

constexpr int TC_MAX_NT = 8;        // output-channel tiles per layer
// (n-tile, K-chunk) tap-mask entries the kernel keeps in smem: the widest layer of the path is the discrete
// config's 768 -> 1536 strided conv (CAPACITY 96): 6 n-tiles x 24 chunks of 64 channels = 144
constexpr int TC_MAX_MASKS = 256;
constexpr int TC_THREADS = 640;     // warps 0-15 output (12-15: activation pass if needed), 16 MMA, 17 W, 18 A producer, 19 MMA (fused GEMM2)

// A row view of an activation tensor, addressed in 16-byte units (8 bf16 = one channel group):
//   unit(b, t, g) = b*sb + (hist + t)*sT + (g % G1)*sA + (g / G1)*sB        t in [lo, n)
// g = view channel / 8.  The engine's internal layout is group-major [stream][C/8][row][8]
// (sA = rows per stream, sB = 1, sT = view factor): a 128-row tile of one channel group is one
// contiguous run, so loads are bulk copies and epilogue stores are fully coalesced.  Audio
// ([stream][T], read / written as rows of 16 samples) uses sA = 1, sT = 2.
struct TcView {
  const __nv_bfloat16* p;
  long long sb;
  int sA, sB, sT, G1;
  int g1_shift;             // log2(G1) when G1 is a power of two, else -1
  int C;                    // view channels (8 * number of groups)
  int hist, n, lo;
};

struct TcParams {
  TcView x;                 // input (row view)
  TcView y;                 // output (row view); written rows [0, y.n)
  TcView res;               // residual added to the output (res.p == nullptr: none)
  int res_delay;            // residual row = t - res_delay (AlignBranches delay, streaming)
  const __nv_bfloat16* w;   // packed weight blocks, consumption order (see tc_engine.cu)
  const uint64_t* masks;    // [n_ntiles][n_kchunks] active-tap bit masks
  const float* bias;        // [N_total] or nullptr
  const float* alpha;       // Snake alpha of the pre-activation per view channel, then 1 / (alpha + 1e-9) (alpha_n
                            // entries each), or nullptr; alpha2 likewise with N_TILE entries each
  int alpha_n;
  const float* aff_a;       // AdaIN scale  [rows][c_in] applied before the pre-activation, or nullptr
  const float* aff_b;       // AdaIN shift
  int aff_C;                // channels of the affine tables (un-reshaped tensor)
  int pre_act;              // rb_act applied to the input tile in smem (NONE: TMA feeds the MMA directly)
  int out_act;              // rb_act applied to the stored output (consumers then skip their pre_act)
  int res_act;              // the residual tensor is stored LeakyReLU-activated: invert it on read
  int res_smem;             // fused unit: the residual is the A tile itself, read from smem (no reload)
  int B;
  int n_taps, dil, off;     // row taps: shift_j = j*dil - off
  int KC, n_kchunks;        // input channels per K chunk (multiple of 16), chunks per row
  int N_TILE, n_ntiles;     // output channels per tile (multiple of 16, <= 256)
  int nt_base[TC_MAX_NT + 1];   // first weight block of each n-tile
  int MT;                   // 128-row sub-tiles per work item (1 or 2)
  int R, R_ALLOC;           // A rows per stage (MT*128 + max shift), allocated (odd)
  int w_resident;           // all blocks live in smem for the whole kernel
  int pair;                 // 1: clusters of two CTAs share each streamed weight block (half each, multicast)
  int cg2;                  // pair + 1: the pair runs tcgen05.mma.cta_group::2 (M = 256; w holds the N-split block layout,
                            // w_blk_bytes = one CTA's half block)
  int split;                // cta_group::2, MT = 2 (one accumulator stage fills tensor memory): the two sub-tiles are
                            // published and taken back separately so their drains overlap MMAs (gemm1_cg2)
  int reverse;              // walk the row tiles from the end of the tensor (alternate launches: L2 reuse, see WorkIter)
  int flat_units;           // > 0 (cta_group::2 clusters): 128*MT-row units per stream; work slots are 2^pair consecutive
                            // (stream, unit) pairs, stream-major -- see WorkIter
  int dbg;                  // timing experiments only (RB_TC_PD_DBG): 1 skip E1 math, 2 skip E2 drain of the pair unit
  int cg2_spin;             // cta_group::2 waits: bit 0 producers / epilogue / relay spin, bit 1 the leader's MMA warp spins
  int dual_g1;              // resident + short tiles: 2 or 3 warps issue GEMM1 of the work items in turn (0 / 1: one issuer)
  // A tile TMA boxes (tensor map over the input view, see launch_tc_conv): a chunk of KC/8
  // channel groups x R_ALLOC rows is fetched as boxes of a_gb groups x a_box_rows rows
  int a_gb, a_box_rows, a_rsplit;
  int a_flat;               // unit-stride view: interior tiles are fetched as linear bulk copies per group
  uint32_t a_tx_bytes;      // bytes one chunk adds to its mbarrier
  int step_nt, step_b, step_tile;   // work-item increment of a persistent CTA (grid stride, decomposed)
  int step2_nt, step2_b, step2_tile;   // twice the grid stride, decomposed (the output-warp sets advance two items at a time)
  int SA, SW, SD;           // A stages, W stages (streamed), accumulator stages
  int out_mask;             // 1: negate odd channels at even rows (pqmf.py:13-17 reverse_half)
  // > 0: every 32-column chunk is [16 wave | 16 amplitude] of one real row (head = chunks per view
  // row); the epilogue stores tanh(wave * sigmoid(amp)), reverse_half applied, as 16 channels
  int head;
  // Sector-complete stores (drain_pair): 0 off; 1 the output view has an even number of phases one unit apart
  // (row batching, transposed-conv phases): the pieces of phase 2i and 2i+1 (pair_dcol columns apart in the
  // accumulator) are neighbours in memory and leave as one 32-byte store; 2 consecutive channel groups are
  // neighbours (audio rows); 3 waveform head with two real rows per view row
  // Per-stream bias (the speaker channels of the decoder's first conv, folded: tc_engine.cu fold): [B][3][sbias_n]
  // fp32 -- row classes 0 interior, 1 first row of the stream, 2 last row (zero padding drops one tap there).
  // Kernel variants DR = 4 only.
  const float* sbias;
  int sbias_n;
  int pair_mode, pair_dcol;
  int pair_ph_tile;         // pair_mode 1 with several n-tiles: phases per n-tile (0: one n-tile)
  int pair_units;           // pair_mode 1: 32-byte units of a row tile, (N_TILE / (2 pair_dcol)) * (pair_dcol / 16)
  int tiles_per_stream, n_work;
  // fused DilatedUnit (blocks.py:84-113 + Residual :32-46): second GEMM k=1 on act(D1 + bias)
  int fused;
  int act2;                 // activation between the two convs
  const float* bias2;
  const float* alpha2;
  int n_blk1;               // weight blocks of GEMM1 (GEMM2 blocks follow)
  int SH;                   // h stages (h_tmem: = SD, the h ring follows the accumulator ring)
  int h_tmem;               // fused: act(conv3 + b) is written back into the accumulator's own TMEM columns as packed
                            // bf16 and the 1x1 conv reads its A operand from there (no h tile in shared memory)
  // shared memory carve-up (bytes from the 1024-aligned base)
  uint32_t off_bias, off_a, off_w, off_h, smem_bytes;
  uint32_t a_stage_bytes, w_blk_bytes, h_stage_bytes;
  uint32_t tmem_cols;       // power of two >= 32
  unsigned long long* trace;  // debug: CTA 0 writes clock stamps [tile < 64][event < 16] (nullptr: off)
};

// Opaque CUtensorMap (128 bytes, 64-byte aligned) passed to the kernel by value.
struct alignas(64) TcTmap {
  unsigned long long opaque[16];
};

// Fills the step_* fields and the tensor map from p.x, then launches.
int launch_tc_conv(TcParams& p, cudaStream_t s);
// Residual(DilatedUnit) at C = 256 as one cluster-pair launch (tc_conv.cu: tc_du_pair_kernel).
int launch_tc_du_pair(TcParams& p, cudaStream_t s);

}  // namespace tc
}  // namespace rb
