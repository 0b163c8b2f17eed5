// Layer table + per-view encoder orchestration (C-ABI: at_model_*, at_encoder_*).
//
// Encoder = model_utils.create_backbone (model_utils.py:107-138): stem, ResNet [2,3,5,2]
// BasicBlocks 32/64/128/256 with strides 1/2/2/2 (backbone_resnet.py:14-158), 1x1 proj to 72.
#include <mutex>
#include <string>
#include <string.h>
#include <vector>

#include "conv.cuh"

namespace abt {

namespace {

std::vector<ConvLayer> g_layers;
std::vector<std::string>* g_names = nullptr;
int64_t g_blob_floats = 0;
std::once_flag g_once;

void add_layer(const std::string& name, int cin, int cout, int k, int stride) {
  g_names->push_back(name);
  ConvLayer L;
  L.name = nullptr;
  L.cin = cin; L.cout = cout; L.ksize = k; L.stride = stride;
  const int64_t coutp = round_up(cout, 4);
  L.weight_offset = g_blob_floats;
  g_blob_floats += (int64_t)k * k * cin * coutp;
  L.bias_offset = g_blob_floats;
  g_blob_floats += coutp;
  L.bf16_offset = -1;
  g_layers.push_back(L);
}

void add_block(const std::string& prefix, int cin, int cout, int stride) {
  add_layer(prefix + ".conv1", cin, cout, 3, stride);
  add_layer(prefix + ".conv2", cout, cout, 3, 1);
  if (stride != 1 || cin != cout) add_layer(prefix + ".down", cin, cout, 1, stride);
}

void build_table() {
  g_names = new std::vector<std::string>();
  add_layer("backbone.stem", 1, 32, 3, 1);
  const int nblocks[4] = {2, 3, 5, 2};
  const int cin[4] = {32, 32, 64, 128}, cout[4] = {32, 64, 128, 256}, stride[4] = {1, 2, 2, 2};
  for (int l = 0; l < 4; ++l)
    for (int b = 0; b < nblocks[l]; ++b)
      add_block("backbone.layer" + std::to_string(l + 1) + "." + std::to_string(b), b == 0 ? cin[l] : cout[l], cout[l],
                b == 0 ? stride[l] : 1);
  add_layer("backbone.proj", 256, 72, 1, 1);
  add_layer("fusion.0", 144, 108, 1, 1);
  add_layer("fusion.1", 108, 72, 1, 1);
  add_layer("fusion.2", 72, 72, 1, 1);
  for (int i = 0; i < 3; ++i) add_layer("temporal." + std::to_string(i), 90, 90, 1, 1);
  add_layer("skeleton", 132, 144, 1, 1);
  for (int h = 0; h < 2; ++h) {
    const std::string p = h == 0 ? "head_k" : "head_u";
    const int c = h == 0 ? 76 : 72;
    add_layer(p + ".0.conv1", c, c, 3, 1);
    add_layer(p + ".0.conv2", c, c, 3, 1);
    add_layer(p + ".1.conv1", c, c, 3, 1);
    add_layer(p + ".1.conv2", c, c, 3, 1);
    add_layer(p + ".out", c, h == 0 ? 62 : 63, 1, 1);
  }
  for (size_t i = 0; i < g_layers.size(); ++i) g_layers[i].name = (*g_names)[i].c_str();
}

}  // namespace

const ConvLayer* layer_table() {
  std::call_once(g_once, build_table);
  return g_layers.data();
}
int64_t blob_floats() {
  layer_table();
  return g_blob_floats;
}

namespace {

// BasicBlock (backbone_resnet.py:56-72): out = relu(conv2(relu(conv1(x))) + shortcut(x))
int basic_block(const ConvLayer* T, int id, bool has_down, const float* blob, const float* x, float* tmp, float* down,
                float* y, int N, int H, int W, cudaStream_t s) {
  const ConvLayer& c1 = T[id];
  const ConvLayer& c2 = T[id + 1];
  const int Ho = (H - 1) / c1.stride + 1, Wo = (W - 1) / c1.stride + 1;
  int rc = conv_nhwc_f32(c1, blob, x, c1.cin, 0, nullptr, 0, tmp, c1.cout, 0, N, H, W, true, s);
  if (rc) return rc;
  const float* res = x;
  if (has_down) {
    const ConvLayer& d = T[id + 2];
    rc = conv_nhwc_f32(d, blob, x, d.cin, 0, nullptr, 0, down, d.cout, 0, N, H, W, false, s);
    if (rc) return rc;
    res = down;
  }
  return conv_nhwc_f32(c2, blob, tmp, c2.cin, 0, res, c2.cout, y, c2.cout, 0, N, Ho, Wo, true, s);
}

constexpr int64_t kActFloatsPerCrop = 48 * 48 * 32;  // largest activation (after the stem)

}  // namespace

int encoder_forward_f32(const float* blob, const void* crops, int crop_dtype, int c, float* features, void* workspace,
                        cudaStream_t s) {
  const ConvLayer* T = layer_table();
  float* X = (float*)workspace;
  float* Tm = X + kActFloatsPerCrop * c;
  float* Y = Tm + kActFloatsPerCrop * c;
  float* D = Y + kActFloatsPerCrop * c;  // downsample shortcut, at most half an activation
  int rc = stem_f32(T[L_STEM], blob, crops, crop_dtype, X, c, s);
  if (rc) return rc;
  const int nblocks[4] = {2, 3, 5, 2};
  int id = L_LAYER1;
  int H = 48, W = 48;
  for (int l = 0; l < 4; ++l) {
    for (int b = 0; b < nblocks[l]; ++b) {
      const bool has_down = (l > 0 && b == 0);
      rc = basic_block(T, id, has_down, blob, X, Tm, D, Y, c, H, W, s);
      if (rc) return rc;
      if (has_down) {
        H = (H - 1) / 2 + 1;
        W = (W - 1) / 2 + 1;
      }
      id += has_down ? 3 : 2;
      float* t = X; X = Y; Y = t;
    }
  }
  // proj: 1x1 256->72 with bias, no activation (model_utils.py:134)
  return conv_nhwc_f32(T[L_PROJ], blob, X, 256, 0, nullptr, 0, features, 72, 0, c, H, W, false, s);
}

}  // namespace abt

namespace abt { int64_t encoder_workspace_bytes_bf16(int c); }

extern "C" int at_model_layer_table(AtLayerInfo* table, int capacity) {
  const abt::ConvLayer* T = abt::layer_table();
  if (table != nullptr) {
    for (int i = 0; i < abt::L_COUNT && i < capacity; ++i) {
      memset(&table[i], 0, sizeof(AtLayerInfo));
      strncpy(table[i].name, T[i].name, sizeof(table[i].name) - 1);
      table[i].cin = T[i].cin; table[i].cout = T[i].cout; table[i].ksize = T[i].ksize; table[i].stride = T[i].stride;
      table[i].weight_offset = T[i].weight_offset;
      table[i].bias_offset = T[i].bias_offset;
    }
  }
  return abt::L_COUNT;
}

extern "C" int64_t at_model_blob_floats(void) { return abt::blob_floats(); }

extern "C" int64_t at_encoder_workspace_bytes(int c, int precision) {
  if (c < 0) return AT_ERR_INVALID_ARGUMENT;
  if (precision == AT_PREC_BF16 || precision == AT_PREC_FP16) return abt::encoder_workspace_bytes_bf16(c);
  // X, T, Y (full) + D (half)
  return (int64_t)sizeof(float) * (abt::kActFloatsPerCrop * 3 + abt::kActFloatsPerCrop / 2) * (int64_t)c + 256;
}

namespace abt {
// tcgen05 path (conv_tc.cu)
int encoder_forward_bf16(const float* blob, const void* tc_blob, const void* crops, int crop_dtype, int c,
                         float* features, void* workspace, int64_t workspace_bytes, int fmt, int prepared_c,
                         int32_t* status, cudaStream_t front_stream, cudaStream_t release_stream, cudaStream_t s);
int64_t encoder_workspace_bytes_bf16(int c);
int encoder_chunk_crops();
}  // namespace abt

extern "C" int at_encoder_chunk_crops(void) { return abt::encoder_chunk_crops(); }

extern "C" int at_encoder_forward(const float* blob, const void* bf16_blob, const void* crops, int crop_dtype, int c,
                                  float* features, void* workspace, int64_t workspace_bytes, int precision,
                                  void* stream) {
  return at_encoder_forward_prepared(blob, bf16_blob, crops, crop_dtype, c, features, workspace, workspace_bytes,
                                     precision, 0, stream);
}

extern "C" int at_encoder_forward_prepared(const float* blob, const void* bf16_blob, const void* crops, int crop_dtype,
                                           int c, float* features, void* workspace, int64_t workspace_bytes,
                                           int precision, int prepared_chunk, void* stream) {
  return at_encoder_forward_status(blob, bf16_blob, crops, crop_dtype, c, features, workspace, workspace_bytes, precision,
                                   prepared_chunk, nullptr, stream);
}

extern "C" int at_encoder_forward_status(const float* blob, const void* bf16_blob, const void* crops, int crop_dtype,
                                         int c, float* features, void* workspace, int64_t workspace_bytes,
                                         int precision, int prepared_chunk, int32_t* status, void* stream) {
  return at_encoder_forward_release(blob, bf16_blob, crops, crop_dtype, c, features, workspace, workspace_bytes, precision,
                                    prepared_chunk, status, nullptr, stream);
}

extern "C" int at_encoder_forward_release(const float* blob, const void* bf16_blob, const void* crops, int crop_dtype,
                                          int c, float* features, void* workspace, int64_t workspace_bytes,
                                          int precision, int prepared_chunk, int32_t* status, void* release_stream,
                                          void* stream) {
  return at_encoder_forward_front(blob, bf16_blob, crops, crop_dtype, c, features, workspace, workspace_bytes, precision,
                                  prepared_chunk, status, nullptr, release_stream, stream);
}

extern "C" int at_encoder_forward_front(const float* blob, const void* bf16_blob, const void* crops, int crop_dtype,
                                        int c, float* features, void* workspace, int64_t workspace_bytes,
                                        int precision, int prepared_chunk, int32_t* status, void* front_stream,
                                        void* release_stream, void* stream) {
  AT_REQUIRE(c >= 0, "at_encoder_forward: c < 0");
  if (c == 0) return AT_OK;
  if (front_stream == stream) front_stream = nullptr;
  AT_REQUIRE(blob && crops && features && workspace, "at_encoder_forward: NULL pointer");
  AT_REQUIRE(crop_dtype == AT_U8 || crop_dtype == AT_F32, "at_encoder_forward: crops must be AT_U8 or AT_F32");
  AT_REQUIRE(c <= 200000, "at_encoder_forward: at most 200000 crops per call");
  if (workspace_bytes < at_encoder_workspace_bytes(c, precision)) {
    abt::set_error("at_encoder_forward: workspace too small (%lld < %lld)", (long long)workspace_bytes,
                  (long long)at_encoder_workspace_bytes(c, precision));
    return AT_ERR_WORKSPACE_TOO_SMALL;
  }
  if (precision == AT_PREC_FP32) {
    if (front_stream != nullptr) {  // no split on this path: the encoder stream waits for the crops' producer
      cudaEvent_t e;
      if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess || cudaEventRecord(e, (cudaStream_t)front_stream) != cudaSuccess ||
          cudaStreamWaitEvent((cudaStream_t)stream, e, 0) != cudaSuccess || cudaEventDestroy(e) != cudaSuccess) {
        abt::set_error("at_encoder_forward_front: %s", cudaGetErrorString(cudaGetLastError()));
        return AT_ERR_CUDA;
      }
    }
    const int rc = abt::encoder_forward_f32(blob, crops, crop_dtype, c, features, workspace, (cudaStream_t)stream);
    if (rc == AT_OK && release_stream != nullptr) {  // no mid-encoder point on this path: release behind the whole encoder
      cudaEvent_t e;
      if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess || cudaEventRecord(e, (cudaStream_t)stream) != cudaSuccess ||
          cudaStreamWaitEvent((cudaStream_t)release_stream, e, 0) != cudaSuccess || cudaEventDestroy(e) != cudaSuccess) {
        abt::set_error("at_encoder_forward_release: %s", cudaGetErrorString(cudaGetLastError()));
        return AT_ERR_CUDA;
      }
    }
    return rc;
  }
  if (precision == AT_PREC_BF16 || precision == AT_PREC_FP16) {
    AT_REQUIRE(bf16_blob != nullptr, "at_encoder_forward: the prepacked 16-bit blob is required for AT_PREC_BF16/FP16");
    return abt::encoder_forward_bf16(blob, bf16_blob, crops, crop_dtype, c, features, workspace, workspace_bytes,
                                    precision == AT_PREC_BF16 ? 1 : 0, prepared_chunk, status, (cudaStream_t)front_stream,
                                    (cudaStream_t)release_stream, (cudaStream_t)stream);
  }
  abt::set_error("at_encoder_forward: bad precision %d", precision);
  return AT_ERR_INVALID_ARGUMENT;
}
