// Layer table of the UmeTrack network and the conv launchers shared by the encoder and tail.
#pragma once
#include "at_common.cuh"

namespace abt {

struct ConvLayer {
  const char* name;
  int cin, cout, ksize, stride;
  int64_t weight_offset;  // float index into the fp32 blob: [k*k][cin][coutP]
  int64_t bias_offset;    // [coutP]
  int64_t bf16_offset;    // byte offset into the bf16 operand blob (tcgen05 path), -1 if unused
};

// Layer ids (order of the blob).  Backbone = resnet_layers_2352-f32 (model_opts.py:12).
enum LayerId {
  L_STEM = 0,
  L_LAYER1 = 1,    // 2 blocks x (conv1, conv2)                         -> 4 layers
  L_LAYER2 = 5,    // block0: conv1, conv2, down; blocks 1-2: conv1, conv2 -> 7 layers
  L_LAYER3 = 12,   // 5 blocks                                           -> 11 layers
  L_LAYER4 = 23,   // 2 blocks                                           -> 5 layers
  L_PROJ = 28,
  L_FUSION = 29,   // 3 layers
  L_TEMPORAL = 32, // 3 layers
  L_SKEL = 35,
  L_HEAD_K = 36,   // b0.conv1 b0.conv2 b1.conv1 b1.conv2 out
  L_HEAD_U = 41,
  L_COUNT = 46
};

const ConvLayer* layer_table();  // L_COUNT entries
int64_t blob_floats();

// out[pixel, out_coff + n] = act(conv(in)[pixel, n] + bias[n] + res[pixel, n])
int conv_nhwc_f32(const ConvLayer& L, const float* blob, const float* in, int in_ld, int in_coff, const float* res,
                  int res_ld, float* out, int out_ld, int out_coff, int N, int H, int W, bool relu,
                  cudaStream_t stream);
int conv_nhwc_tf32x3(const ConvLayer& L, const float* blob, const float* in, int in_ld, int in_coff, const float* res,
                     int res_ld, float* out, int out_ld, int out_coff, int N, int H, int W, bool relu,
                     cudaStream_t stream);
// fp16 x 3 (error-compensated, fp32-grade) variant: weights pre-split by conv_f16x3_split_weights into
// conv_f16x3_weight_halves(L) halves at `wsplit`.
int64_t conv_f16x3_weight_halves(const ConvLayer& L);
int conv_f16x3_split_weights(const ConvLayer& L, const float* blob, void* wsplit, cudaStream_t stream);
int conv_nhwc_f16x3(const ConvLayer& L, const float* blob, const void* wsplit, const float* in, int in_ld, int in_coff,
                    const float* res, int res_ld, float* out, int out_ld, int out_coff, int N, int H, int W, bool relu,
                    cudaStream_t stream);
// Pose-head BasicBlocks (4 3x3 convs) as error-compensated fp16 x 3 shift-GEMMs on tcgen05 (conv_tc.cu).
int64_t head_tc_weight_bytes();            // per conv
int64_t head_tc_workspace_bytes(int n);
int head_tc_prepack(const ConvLayer& L, const float* blob, void* dst, cudaStream_t stream);
int head_tc_forward(const ConvLayer* L4, const float* blob, const void* const* wq, float* x, float* h2, float* y, int hc,
                    int n, void* workspace, cudaStream_t stream);
int stem_f32(const ConvLayer& L, const float* blob, const void* crops, int crop_dtype, float* out, int c,
             cudaStream_t stream);

}  // namespace abt
