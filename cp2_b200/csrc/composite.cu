// K1 — copy-paste compositing (reference builder.py:1146-1159).
// HBM-bound elementwise: 2 reads + 1 write of [B,3,H,W] f32 per view
// (algorithmic bytes 3*B*3*H*W*4 + B*hs*ws*4).  One thread handles 4 consecutive
// pixels of all three channels with 128-bit loads/stores; the background's channel 0
// is read once and yields the mask for all channels.
#include "common.cuh"

namespace {

struct CompositeView {
  const float *img, *bg;
  float *out, *mask_full, *mask_ds;
};
struct CompositeViews { CompositeView v[2]; };     // blockIdx.y selects the view (1 or 2 views per launch)

template <bool kVec>
__global__ void __launch_bounds__(256)
composite_kernel(const CompositeViews views, int B, int H, int W, int os, int hs, int ws) {
  constexpr int V = kVec ? 4 : 1;
  const CompositeView &vw = views.v[blockIdx.y];
  const float *__restrict__ img = vw.img;
  const float *__restrict__ bg = vw.bg;
  float *__restrict__ out = vw.out;
  float *__restrict__ mask_full = vw.mask_full;
  float *__restrict__ mask_ds = vw.mask_ds;
  const long long plane = (long long)H * W;
  const long long groups_per_img = plane / V;
  const long long total = (long long)B * groups_per_img;
  const int o = os / 2;
  for (long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x; g < total;
       g += (long long)gridDim.x * blockDim.x) {
    const long long b = g / groups_per_img;
    const long long pix = (g - b * groups_per_img) * V;   // first pixel of the group
    const long long base = b * 3 * plane + pix;
    float m[V], bgv[3][V], im[3][V];
    if constexpr (kVec) {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float4 t = __ldcs(reinterpret_cast<const float4 *>(bg + base + c * plane));
        bgv[c][0] = t.x; bgv[c][1] = t.y; bgv[c][2] = t.z; bgv[c][3] = t.w;
        const float4 u = __ldcs(reinterpret_cast<const float4 *>(img + base + c * plane));
        im[c][0] = u.x; im[c][1] = u.y; im[c][2] = u.z; im[c][3] = u.w;
      }
    } else {
#pragma unroll
      for (int c = 0; c < 3; ++c) { bgv[c][0] = bg[base + c * plane]; im[c][0] = img[base + c * plane]; }
    }
#pragma unroll
    for (int v = 0; v < V; ++v) m[v] = (bgv[0][v] == 0.0f) ? 1.0f : 0.0f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float r[V];
#pragma unroll
      for (int v = 0; v < V; ++v) r[v] = im[c][v] * m[v] + bgv[c][v];   // product is exact (m in {0,1})
      if constexpr (kVec)
        __stcs(reinterpret_cast<float4 *>(out + base + c * plane), make_float4(r[0], r[1], r[2], r[3]));
      else
        out[base + c * plane] = r[0];
    }
    if (mask_full != nullptr) {
      if constexpr (kVec)
        __stcs(reinterpret_cast<float4 *>(mask_full + b * plane + pix), make_float4(m[0], m[1], m[2], m[3]));
      else
        mask_full[b * plane + pix] = m[0];
    }
    const int h = (int)(pix / W);
    if (h >= o && (h - o) % os == 0) {
      const int w0 = (int)(pix - (long long)h * W);
#pragma unroll
      for (int v = 0; v < V; ++v) {
        const int w = w0 + v;
        if (w >= o && (w - o) % os == 0)
          mask_ds[(b * hs + (h - o) / os) * ws + (w - o) / os] = m[v];
      }
    }
  }
}

}  // namespace

namespace {

int launch_composite(const CompositeViews &views, int nviews, int B, int H, int W, int os, cudaStream_t st) {
  const int o = os / 2;
  const int hs = H > o ? (H - o + os - 1) / os : 0;
  const int ws = W > o ? (W - o + os - 1) / os : 0;
  auto aligned = [](const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
  bool vec = (W % 4 == 0);
  for (int i = 0; i < nviews; ++i) {
    const CompositeView &v = views.v[i];
    if (!(v.img && v.bg && v.out)) { cp2_set_error("cp2_composite: null pointer"); return CP2_E_INVALID; }
    if (!(v.mask_ds || hs * ws == 0)) { cp2_set_error("cp2_composite: null mask_ds"); return CP2_E_INVALID; }
    vec = vec && aligned(v.img) && aligned(v.bg) && aligned(v.out) && (v.mask_full == nullptr || aligned(v.mask_full));
  }
  const long long groups = (long long)B * H * W / (vec ? 4 : 1);
  const int threads = 256;
  long long blocks = (groups + threads - 1) / threads;
  const long long cap = (long long)cp2_num_sms() * 16 / nviews;     // grid-stride; whole waves of 148 SMs
  if (blocks > cap) blocks = cap;
  dim3 grid((unsigned)blocks, nviews);
  if (vec)
    composite_kernel<true><<<grid, threads, 0, st>>>(views, B, H, W, os, hs, ws);
  else
    composite_kernel<false><<<grid, threads, 0, st>>>(views, B, H, W, os, hs, ws);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

}  // namespace

extern "C" int cp2_composite(const float *img, const float *bg, float *out, float *mask_full,
                             float *mask_ds, int B, int H, int W, int os, void *stream) {
  CP2_REQUIRE(B >= 0 && H > 0 && W > 0 && os > 0, "cp2_composite: bad shape B=%d H=%d W=%d os=%d", B, H, W, os);
  if (B == 0) return CP2_OK;                 // empty batch: pointers may be null
  CompositeViews views;
  views.v[0] = CompositeView{img, bg, out, mask_full, mask_ds};
  views.v[1] = views.v[0];
  return launch_composite(views, 1, B, H, W, os, static_cast<cudaStream_t>(stream));
}

extern "C" int cp2_composite2(const float *img_a, const float *bg_a, float *out_a, float *mask_full_a,
                              float *mask_ds_a, const float *img_b, const float *bg_b, float *out_b,
                              float *mask_full_b, float *mask_ds_b, int B, int H, int W, int os, void *stream) {
  CP2_REQUIRE(B >= 0 && H > 0 && W > 0 && os > 0, "cp2_composite2: bad shape B=%d H=%d W=%d os=%d", B, H, W, os);
  if (B == 0) return CP2_OK;
  CompositeViews views;
  views.v[0] = CompositeView{img_a, bg_a, out_a, mask_full_a, mask_ds_a};
  views.v[1] = CompositeView{img_b, bg_b, out_b, mask_full_b, mask_ds_b};
  return launch_composite(views, 2, B, H, W, os, static_cast<cudaStream_t>(stream));
}
