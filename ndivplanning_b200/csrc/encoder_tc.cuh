// interface of the tensor-core convolution layers (k5_encoder_tc.cu) used by the encoder handle (k5_encoder.cu)
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>
#include "common.cuh"

namespace ndp {
namespace enc {

struct ConvTcP {
  const __half* in_hi;      // NHWC fp16 planes of the scaled input activation
  const __half* in_lo;
  int B, H, W, Cin, OH, OW, Cout, KW, stride, pad;
  int M, K;                 // M = B*OH*OW, K = KH*KW*Cin
  const uint8_t* wpack;     // conv_tc_pack
  const float* wsinv;       // [Cout] 1 / weight scale
  const float* ainv;        // [1]    1 / activation scale (conv_tc_split)
  const float* bias;        // [Cout]
  int relu;
  float* out;               // fp32 NHWC [M][Cout]
  float* part;              // split-K partial slots [splits][M][Cout] (splits > 1)
  float* amax_out;          // nullable [1]: running maximum of the (post-ReLU) outputs, for the next layer's split
  int n_chunks, chunks_per_split, splits;
};

size_t conv_tc_pack_bytes(int Cout, int K);
int conv_tc_pack(const float* W, int Cout, int K, uint8_t* wpack, float* wsinv, float* ws_scratch, cudaStream_t s);
int conv_tc_split(const float* act, size_t n, const float* amax, float* ainv, __half* hi, __half* lo, cudaStream_t s);
// fills n_chunks / chunks_per_split / splits from the geometry and launches the layer (+ finalize when split)
int conv_tc_forward(ConvTcP p, size_t part_floats, cudaStream_t s);

}  // namespace enc
}  // namespace ndp
