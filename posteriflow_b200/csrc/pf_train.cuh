// Fused training step of the tensor-core tiers: shared definitions of the forward-with-saves kernel
// (pf_tc_dense.cu, SAVE instantiation), the backward chain kernel and the weight-gradient kernel (pf_tc_train.cu),
// and the host orchestration (pf_fp32.cu: pf_flow_forward(saved) / pf_flow_backward).
//
// Replaces, for one NLL training step of the flow head (experiments/train_lean_npe.py:357-369, flows.py:727-779):
// 70 forward + 200 backward kernel launches of the GEMM-per-launch form by
//   2  context projection                  fp16 hi / lo images of the context rows, then ctx . Wctx^T on tcgen05 for all L*3
//                                          context Linear layers (two accumulators: epilogue under the next block's MMAs)
//   1  forward kernel                      all L layers per 128-row tile on chip, activations saved as bf16 images
//   1  backward chain kernel               all L layers per tile: spline VJP -> 7 data-gradient MMA stages per layer
//   2  weight-gradient kernel              bf16 context images, then one CTA per (layer, matrix, 128-row block of the
//                                          output) -- the 3 context Linear layers included: contraction over ALL flow rows
//                                          by MN-major tcgen05.mma on the saved images, bias gradients as column sums of
//                                          the same tiles, plain read-modify-write stores (no atomics)
//   1  d ctx kernel                        gate / context gradients (bf16 hi / lo images) . Wctx over all L*3 layers
#pragma once
#include "pf_common.cuh"

namespace pf {

constexpr int TT_ROWS = 128;
constexpr int TT_IMG = 65536;      // one [128 x 256] bf16 tile image (8-row x 16-byte core matrices, 2048 B per 8 columns)
constexpr int TT_XT = 4096;        // one [128 x 16] bf16 x-tile image
constexpr int TT_MAX_L = 32;

// forward images (written by tc_dense_forward_kernel<3, true>)
enum { TT_A_H0 = 0, TT_A_U1 = 1, TT_A_H1 = 2, TT_A_U2 = 3, TT_A_H2 = 4, TT_A_V1 = 5, TT_A_V1L = 6, TT_A_V2 = 7, TT_A_V2L = 8,
       TT_NA = 9 };
// gradient images (written by tt_backward_kernel): d params of dims 1..ga | ga+1.. , then the five hidden gradients
// ... and the three context-layer gradients (pre-activation of relu(context_layer), gate 1, gate 2), hi and lo halves
enum { TT_G_PA = 0, TT_G_PB = 1, TT_G_V2 = 2, TT_G_U2 = 3, TT_G_V1 = 4, TT_G_U1 = 5, TT_G_H0 = 6,
       TT_G_C0 = 7, TT_G_G1 = 8, TT_G_G2 = 9, TT_G_C0L = 10, TT_G_G1L = 11, TT_G_G2L = 12, TT_NG = 13 };
constexpr int TT_CIMG_COLS = 288;                 // context image: [128 x C] bf16 in the tile-image layout (C <= 288)
constexpr int TT_CIMG = TT_CIMG_COLS / 8 * 2048;  // 73 728 bytes

__host__ __device__ inline size_t tt_img_off(int l, int which, int nwhich, int tile, int tiles) {
  return (((size_t)l * nwhich + which) * (size_t)tiles + tile) * TT_IMG;
}
// Per-row fp32 arrays are kept TILE-TRANSPOSED: [tile][column][128 rows].  A chain-kernel thread owns one row of its tile
// (tcgen05.ld offers no other mapping), so with the row index minor every load / store of a warp is one 128-byte line;
// row-major [R][columns] would cost one line per LANE (32 LSU wavefronts per instruction, partial-sector writes).
__host__ __device__ inline size_t tt_t_off(int tile, int ncols, int col, int row_in_tile) {
  return ((size_t)tile * ncols + col) * TT_ROWS + row_in_tile;
}

static inline size_t tt_align(size_t b) { return (b + 255) / 256 * 256; }

// carve of the `saved` buffer handed from pf_flow_forward to pf_flow_backward (fused form)
struct TtSaved {
  float* proj;    // [tiles][L*3*H][128]  relu(c0) | gate 1 | gate 2 per layer (sorted hidden order), tile-transposed
  float* XS;      // [L][tiles][D][128]   input of layer l (before its ReversePermutation), tile-transposed
  float* P;       // [L][tiles][PLD][128] spline parameters, tile-transposed
  uint8_t* img;   // [L][TT_NA][tiles][TT_IMG]
  uint8_t* xt;    // [L][tiles][TT_XT]
  uint8_t* c16;   // [tiles][2][TT_CIMG]  fp16 hi | lo images of the context rows (operand of the projection kernel)
  size_t bytes;
  TtSaved(const pf_flow* f, void* base, int64_t R) {
    const size_t tiles = (size_t)((R + TT_ROWS - 1) / TT_ROWS);
    char* p = reinterpret_cast<char*>(base);
    size_t o = 0;
    auto take = [&](size_t b) { char* r = p ? p + o : nullptr; o += tt_align(b); return r; };
    proj = reinterpret_cast<float*>(take(tiles * TT_ROWS * (size_t)f->L * 3 * f->H * 4));
    XS = reinterpret_cast<float*>(take((size_t)f->L * tiles * TT_ROWS * f->D * 4));
    P = reinterpret_cast<float*>(take((size_t)f->L * tiles * TT_ROWS * f->PLD * 4));
    img = reinterpret_cast<uint8_t*>(take((size_t)f->L * TT_NA * tiles * TT_IMG));
    xt = reinterpret_cast<uint8_t*>(take((size_t)f->L * tiles * TT_XT));
    c16 = reinterpret_cast<uint8_t*>(take(tiles * 2 * (size_t)TT_CIMG));
    bytes = o;
  }
};

// carve of the backward workspace (fused form)
struct TtWork {
  uint8_t* gimg;   // [L][TT_NG][tiles][TT_IMG]  bf16 gradient images
  uint8_t* cimg;   // [tiles][TT_CIMG]  bf16 image of the context rows (weight-gradient operand of the context layers)
  size_t bytes;
  TtWork(const pf_flow* f, void* base, int64_t R) {
    const size_t tiles = (size_t)((R + TT_ROWS - 1) / TT_ROWS);
    char* p = reinterpret_cast<char*>(base);
    size_t o = 0;
    auto take = [&](size_t b) { char* r = p ? p + o : nullptr; o += tt_align(b); return r; };
    gimg = reinterpret_cast<uint8_t*>(take((size_t)f->L * TT_NG * tiles * TT_IMG));
    cimg = reinterpret_cast<uint8_t*>(take(tiles * (size_t)TT_CIMG));
    bytes = o;
  }
};

}  // namespace pf

// pf_tc_train.cu
bool pf_tt_supported(const pf_flow* f);
pf_status pf_tt_backward(pf_flow* f, cudaStream_t st, const pf::TtSaved& sv, const pf::TtWork& ws, int64_t R, const float* gz,
                         const float* glogdet, float* gx, const pf_layer_grads* grads);
pf_status pf_tt_wgrad(pf_flow* f, cudaStream_t st, const pf::TtSaved& sv, const pf::TtWork& ws, int64_t R,
                      const float* ctx, const pf_layer_grads* grads);
pf_status pf_tt_ctxproj(pf_flow* f, cudaStream_t st, const float* ctx, int64_t R, float* out, uint8_t* scratch);
size_t pf_tt_ctxproj_scratch(int64_t R);
pf_status pf_tt_dctx(pf_flow* f, cudaStream_t st, const pf::TtWork& ws, int64_t R, float* gctx);
