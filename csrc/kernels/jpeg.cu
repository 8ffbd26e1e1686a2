// b200ps — GPU back end of the JPEG decoder (the host half is csrc/io/jpeg_parse.cpp): dequantisation + 8x8
// inverse DCT into component planes, then chroma upsampling, YCbCr -> RGB and an optional bilinear resize.
// Replaces tf.image.decode_jpeg(channels=3) of input_pipeline.py:68-78.
//
// IDCT: libjpeg's default integer transform ("islow": LL&M factorisation, 13-bit constants) restated exactly, so
// the planes — and with the integer upsampling and colour conversion below, the RGB output — equal what libjpeg /
// PIL produce for the same file.  8 blocks per CTA, one thread per column (pass 1) and per row (pass 2).
// Chroma: libjpeg's "fancy" upsampling, reproduced in integer arithmetic —
//   2x1: out[2i] = (3 s[i] + s[i-1] + 1) >> 2, out[2i+1] = (3 s[i] + s[i+1] + 2) >> 2
//   2x2: column sums t = 3 near_row + far_row, then out[2i] = (3 t[i] + t[i-1] + 8) >> 4, out[2i+1] = (3 t[i] + t[i+1] + 7) >> 4
// with edge samples replicated; other ratios replicate samples.
#include "../common.h"

namespace b200ps {

struct JpegPlan {
  const int16_t* coef;
  uint8_t* planes;
  b200ps_jpeg_desc d;
};

// libjpeg's "islow" 8-point inverse DCT (Loeffler-Ligtenberg-Moschytz, 13-bit constants), bit for bit: pass 1 keeps
// 2 extra fraction bits, pass 2 descales by 13 + 2 + 3.  in[] / out[] are strided so the same routine serves the
// column pass and the row pass.
template <int kShift>
__device__ __forceinline__ void idct_islow_1d(const int (&in)[8], int (&out)[8]) {
  constexpr int C = 13;
  int z2 = in[2], z3 = in[6];
  int z1 = (z2 + z3) * 4433;
  int tmp2 = z1 + z3 * (-15137), tmp3 = z1 + z2 * 6270;
  z2 = in[0]; z3 = in[4];
  int tmp0 = (z2 + z3) * (1 << C), tmp1 = (z2 - z3) * (1 << C);
  const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  tmp0 = in[7]; tmp1 = in[5]; tmp2 = in[3]; tmp3 = in[1];
  z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2;
  int z4 = tmp1 + tmp3;
  const int z5 = (z3 + z4) * 9633;
  tmp0 *= 2446; tmp1 *= 16819; tmp2 *= 25172; tmp3 *= 12299;
  z1 *= -7373; z2 *= -20995; z3 *= -16069; z4 *= -3196;
  z3 += z5; z4 += z5;
  tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
  constexpr int R = 1 << (kShift - 1);
  out[0] = (tmp10 + tmp3 + R) >> kShift; out[7] = (tmp10 - tmp3 + R) >> kShift;
  out[1] = (tmp11 + tmp2 + R) >> kShift; out[6] = (tmp11 - tmp2 + R) >> kShift;
  out[2] = (tmp12 + tmp1 + R) >> kShift; out[5] = (tmp12 - tmp1 + R) >> kShift;
  out[3] = (tmp13 + tmp0 + R) >> kShift; out[4] = (tmp13 - tmp0 + R) >> kShift;
}

// 8 blocks per CTA: thread = (block in CTA, column) in pass 1 and (block in CTA, row) in pass 2
__global__ void __launch_bounds__(64) jpeg_idct_kernel(const JpegPlan p) {
  __shared__ int s_ws[8][64];
  const int bl = threadIdx.x >> 3, k = threadIdx.x & 7;
  const int blk = blockIdx.x * 8 + bl;
  const bool live = blk < p.d.total_blocks;
  int c = 0;
  if (live && p.d.ncomp > 1 && (int64_t)blk * 64 >= p.d.comp[1].coef_offset) c = 1;
  if (live && p.d.ncomp > 2 && (int64_t)blk * 64 >= p.d.comp[2].coef_offset) c = 2;
  const b200ps_jpeg_comp& cc = p.d.comp[c];
  if (live) {
    int in[8], out[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) in[r] = (int)p.coef[(int64_t)blk * 64 + r * 8 + k] * (int)cc.quant[r * 8 + k];
    idct_islow_1d<11>(in, out);                           // columns: CONST_BITS - PASS1_BITS
#pragma unroll
    for (int r = 0; r < 8; ++r) s_ws[bl][r * 8 + k] = out[r];
  }
  __syncthreads();
  if (live) {
    int in[8], out[8];
#pragma unroll
    for (int x = 0; x < 8; ++x) in[x] = s_ws[bl][k * 8 + x];
    idct_islow_1d<18>(in, out);                           // rows: CONST_BITS + PASS1_BITS + 3
    const int local = blk - (int)(cc.coef_offset / 64);
    const int by = local / cc.blocks_w, bx = local % cc.blocks_w;
    uint8_t* dst = p.planes + cc.plane_offset + (int64_t)(by * 8 + k) * (cc.blocks_w * 8) + bx * 8;
#pragma unroll
    for (int x = 0; x < 8; ++x) dst[x] = (uint8_t)min(255, max(0, out[x] + 128));
  }
}

__device__ __forceinline__ int jp_sample(const uint8_t* plane, int pitch, int y, int x) { return plane[(int64_t)y * pitch + x]; }

// component c at full-resolution pixel (y, x)
__device__ int jp_component(const JpegPlan& p, int c, int y, int x) {
  const b200ps_jpeg_comp& cc = p.d.comp[c];
  const uint8_t* plane = p.planes + cc.plane_offset;
  const int pitch = cc.blocks_w * 8;
  const int hr = p.d.hmax / cc.h, vr = p.d.vmax / cc.v;
  if (hr == 1 && vr == 1) return jp_sample(plane, pitch, y, x);
  const int dw = (p.d.width * cc.h + p.d.hmax - 1) / p.d.hmax, dh = (p.d.height * cc.v + p.d.vmax - 1) / p.d.vmax;
  const int sx = x / hr, sy = y / vr;
  if (hr == 2 && vr == 1) {
    const int s = jp_sample(plane, pitch, sy, sx);
    if (x & 1) return (3 * s + jp_sample(plane, pitch, sy, min(sx + 1, dw - 1)) + 2) >> 2;
    return (3 * s + jp_sample(plane, pitch, sy, max(sx - 1, 0)) + 1) >> 2;
  }
  if (hr == 2 && vr == 2) {
    const int fy = (y & 1) ? min(sy + 1, dh - 1) : max(sy - 1, 0);        // the farther of the two source rows
    const int xo = (x & 1) ? min(sx + 1, dw - 1) : max(sx - 1, 0);
    const int t_this = 3 * jp_sample(plane, pitch, sy, sx) + jp_sample(plane, pitch, fy, sx);
    const int t_other = 3 * jp_sample(plane, pitch, sy, xo) + jp_sample(plane, pitch, fy, xo);
    return (3 * t_this + t_other + ((x & 1) ? 7 : 8)) >> 4;
  }
  return jp_sample(plane, pitch, min(sy, dh - 1), min(sx, dw - 1));        // other ratios: replicate
}

__device__ __forceinline__ void jp_rgb(const JpegPlan& p, int y, int x, float (&rgb)[3]) {
  const int Y = jp_component(p, 0, y, x);
  if (p.d.ncomp == 1) { rgb[0] = rgb[1] = rgb[2] = (float)Y; return; }
  // libjpeg's 16-bit fixed-point YCbCr -> RGB (jdcolor.c), bit for bit
  const int cb = jp_component(p, 1, y, x) - 128, cr = jp_component(p, 2, y, x) - 128;
  const int r = Y + ((91881 * cr + 32768) >> 16);
  const int g = Y + ((-22554 * cb + 32768 - 46802 * cr) >> 16);
  const int b = Y + ((116130 * cb + 32768) >> 16);
  rgb[0] = (float)min(255, max(0, r));
  rgb[1] = (float)min(255, max(0, g));
  rgb[2] = (float)min(255, max(0, b));
}

__global__ void __launch_bounds__(256) jpeg_color_kernel(const JpegPlan p, uint8_t* out, int out_h, int out_w) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= out_h * out_w) return;
  const int oy = idx / out_w, ox = idx % out_w;
  float rgb[3];
  if (out_h == p.d.height && out_w == p.d.width) {
    jp_rgb(p, oy, ox, rgb);
  } else {                                   // bilinear, half-pixel centres (tf.image.resize_bilinear's later default)
    const float fy = fminf(fmaxf((oy + 0.5f) * p.d.height / out_h - 0.5f, 0.f), (float)(p.d.height - 1));
    const float fx = fminf(fmaxf((ox + 0.5f) * p.d.width / out_w - 0.5f, 0.f), (float)(p.d.width - 1));
    const int y0 = (int)fy, x0 = (int)fx, y1 = min(y0 + 1, p.d.height - 1), x1 = min(x0 + 1, p.d.width - 1);
    const float wy = fy - y0, wx = fx - x0;
    float a[3], b[3], c[3], d[3];
    jp_rgb(p, y0, x0, a); jp_rgb(p, y0, x1, b); jp_rgb(p, y1, x0, c); jp_rgb(p, y1, x1, d);
#pragma unroll
    for (int k = 0; k < 3; ++k)
      rgb[k] = rintf((a[k] * (1.f - wx) + b[k] * wx) * (1.f - wy) + (c[k] * (1.f - wx) + d[k] * wx) * wy);
  }
  out[(int64_t)idx * 3 + 0] = (uint8_t)rgb[0];
  out[(int64_t)idx * 3 + 1] = (uint8_t)rgb[1];
  out[(int64_t)idx * 3 + 2] = (uint8_t)rgb[2];
}

}  // namespace b200ps

using namespace b200ps;

extern "C" int b200ps_jpeg_idct_color(const int16_t* coef_dev, const b200ps_jpeg_desc* desc, uint8_t* planes_dev,
                                      uint8_t* out_dev, int32_t out_h, int32_t out_w, b200ps_stream_t stream) {
  B200PS_REQUIRE(coef_dev && desc && planes_dev && out_dev && out_h > 0 && out_w > 0, "jpeg_idct_color: bad arguments");
  B200PS_REQUIRE(desc->total_blocks > 0 && (desc->ncomp == 1 || desc->ncomp == 3), "jpeg_idct_color: empty descriptor");
  JpegPlan p;
  p.coef = coef_dev; p.planes = planes_dev; p.d = *desc;
  cudaStream_t s = as_stream(stream);
  jpeg_idct_kernel<<<(desc->total_blocks + 7) / 8, 64, 0, s>>>(p);
  B200PS_LAUNCH_CHECK();
  jpeg_color_kernel<<<(out_h * out_w + 255) / 256, 256, 0, s>>>(p, out_dev, out_h, out_w);
  B200PS_LAUNCH_CHECK();
  return 0;
}
