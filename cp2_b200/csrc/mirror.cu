// a13 / f5 — MirrorModule loss (reference networks/mirror_network.py:40-63, with the
// bilinear `resize` of networks/segment_network.py:219-231 optionally fused in).
//
//   class   = CE(cat(s,t), cat(masks,masks))                  mean over valid pixels
//   compare = CE(input = softmax(s/T), target = softmax(t/T)) soft-target CE whose *input*
//             is already a probability map (the reference's double softmax), mean over pixels
//   loss    = class + lmbd * compare
//
// Two kernels, one per-pixel device function:
//   mirror_full_kernel : logits already at label resolution.  HBM-bound elementwise:
//                        reads s,t (2*Cls*4 B/px) + labels (8 B/px), writes ds,dt (2*Cls*4 B/px).
//   mirror_up_kernel   : logits at the decoder's resolution [B,Cls,h,w]; the bilinear
//                        upsample (align_corners=False) is evaluated on the fly from a
//                        shared-memory patch and its backward is accumulated on chip, so
//                        the [B,Cls,H,W] logits and their gradients never exist in HBM:
//                        algorithmic traffic = the labels (8 B/px).
// Losses are reduced per CTA and summed in fixed order by mirror_finalize_kernel.
#include "common.cuh"

namespace {

constexpr int kMaxCls = 8;
constexpr long long kIgnoreIndex = -100;     // nn.CrossEntropyLoss() default (mirror_network.py:30,32)

struct MirrorScale {
  float inv_temp;   // 1/T
  float w_cls;      // 1 / (number of valid class-CE terms)
  float w_cmp;      // lmbd / (number of pixels)
};

__device__ __forceinline__ float ex2_fast(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float lg2_fast(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_fast(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// One pixel.  s,t: logits; label in [0,CLS) or -1 (ignored).  Adds the un-normalised loss terms to
// cls_sum / cmp_sum and returns d loss / d s, d loss / d t (already normalised by w_cls / w_cmp).
// All exponentials / logarithms are base-2 MUFU approximations (rel. error ~2^-22) with the
// log2(e) factors folded into the FFMA that applies the max shift.
// Two classes (the polyp masks of mirror_pretrain.py): every soft-max over two logits is a sigmoid of their
// difference, so the pixel needs 13 MUFU operations instead of 18 (class CE: ex2 + lg2 + rcp per network; the
// two temperature soft-maxes: ex2 + rcp each; the soft-max over u: ex2 + lg2 + rcp).  Same mathematics.
template <bool HAS_T, bool GRAD>
__device__ __forceinline__ void mirror_pixel2(const float (&s)[2], const float (&t)[2], int label,
                                              const MirrorScale &sc, float &cls_sum, float &cmp_sum,
                                              float (&gs)[2], float (&gt)[2]) {
  constexpr float kL2E = 1.4426950408889634f, kLn2 = 0.6931471805599453f;
  const float wv = label >= 0 ? sc.w_cls : 0.f;
  const int lab = label > 0 ? 1 : 0;                       // ignored pixels: any label, weight 0
  // class CE: softplus(d), d = logit(other) - logit(label); d/d logit(other) = sigmoid(d) = -d/d logit(label)
  auto class_ce = [&](const float (&x)[2], float (&g)[2]) -> float {
    const float d = lab ? x[0] - x[1] : x[1] - x[0];
    const float e = ex2_fast(-fabsf(d) * kL2E);
    const float term = fmaxf(d, 0.f) + kLn2 * lg2_fast(1.f + e);
    if constexpr (GRAD) {
      const float r = rcp_fast(1.f + e);
      const float sig = (d >= 0.f ? r : e * r) * wv;
      g[0] = lab ? sig : -sig;
      g[1] = lab ? -sig : sig;
    }
    return term;
  };
  float term = class_ce(s, gs);
  if constexpr (HAS_T) {
    term += class_ce(t, gt);
    const float k2 = kL2E * sc.inv_temp;
    // softmax(x / T) of two logits: (larger, smaller) = (r, e r), e = exp(-|x1 - x0| / T), r = 1 / (1 + e)
    auto soft2 = [&](const float (&x)[2], float (&p)[2]) {
      const float dx = x[1] - x[0];
      const float e = ex2_fast(-fabsf(dx) * k2), r = rcp_fast(1.f + e);
      p[1] = dx >= 0.f ? r : e * r;
      p[0] = dx >= 0.f ? e * r : r;
    };
    float u[2], tau[2];
    soft2(s, u);
    soft2(t, tau);
    // CE with u as the LOGITS and tau as the target: lse(u) - sum tau_c u_c
    const float du = u[1] - u[0];
    const float e2 = ex2_fast(-fabsf(du) * kL2E);
    const float lse2 = fmaxf(u[0], u[1]) + kLn2 * lg2_fast(1.f + e2);
    const float tu = fmaf(tau[0], u[0], tau[1] * u[1]);
    cmp_sum += lse2 - tu;
    if constexpr (GRAD) {
      const float r2 = rcp_fast(1.f + e2);
      float q[2];
      q[1] = du >= 0.f ? r2 : e2 * r2;
      q[0] = du >= 0.f ? e2 * r2 : r2;
      const float gu0 = q[0] - tau[0], gu1 = q[1] - tau[1];
      const float dud = fmaf(gu0, u[0], gu1 * u[1]);
      const float k = sc.w_cmp * sc.inv_temp;
      gs[0] = fmaf(k * u[0], gu0 - dud, gs[0]);
      gs[1] = fmaf(k * u[1], gu1 - dud, gs[1]);
      gt[0] = fmaf(k * tau[0], tu - u[0], gt[0]);
      gt[1] = fmaf(k * tau[1], tu - u[1], gt[1]);
    }
  }
  cls_sum += label >= 0 ? term : 0.f;
}

template <int CLS, bool HAS_T, bool GRAD>
__device__ __forceinline__ void mirror_pixel(const float (&s)[CLS], const float (&t)[CLS], int label,
                                             const MirrorScale &sc, float &cls_sum, float &cmp_sum,
                                             float (&gs)[CLS], float (&gt)[CLS]) {
  if constexpr (CLS == 2) {
    mirror_pixel2<HAS_T, GRAD>(s, t, label, sc, cls_sum, cmp_sum, gs, gt);
    return;
  }
  constexpr float kL2E = 1.4426950408889634f, kLn2 = 0.6931471805599453f;
  const float wv = label >= 0 ? sc.w_cls : 0.f;          // class weight of this pixel (0 if ignored)
  const float k2 = kL2E * sc.inv_temp;
  // ---- class CE on s (and t): lse(x) - x[label]
  float ms = s[0];
#pragma unroll
  for (int c = 1; c < CLS; ++c) ms = fmaxf(ms, s[c]);
  float es[CLS], zs = 0.f, s_lab = ms;
  const float nms = -ms * kL2E;
#pragma unroll
  for (int c = 0; c < CLS; ++c) {
    es[c] = ex2_fast(fmaf(s[c], kL2E, nms));
    zs += es[c];
    s_lab = (c == label) ? s[c] : s_lab;
  }
  float term = fmaf(kLn2, lg2_fast(zs), ms - s_lab);
  if constexpr (GRAD) {
    const float rzs = rcp_fast(zs) * wv;
#pragma unroll
    for (int c = 0; c < CLS; ++c) gs[c] = fmaf(es[c], rzs, c == label ? -wv : 0.f);
  }
  if constexpr (HAS_T) {
    float mt = t[0];
#pragma unroll
    for (int c = 1; c < CLS; ++c) mt = fmaxf(mt, t[c]);
    float et[CLS], zt = 0.f, t_lab = mt;
    const float nmt = -mt * kL2E;
#pragma unroll
    for (int c = 0; c < CLS; ++c) {
      et[c] = ex2_fast(fmaf(t[c], kL2E, nmt));
      zt += et[c];
      t_lab = (c == label) ? t[c] : t_lab;
    }
    term += fmaf(kLn2, lg2_fast(zt), mt - t_lab);
    if constexpr (GRAD) {
      const float rzt = rcp_fast(zt) * wv;
#pragma unroll
      for (int c = 0; c < CLS; ++c) gt[c] = fmaf(et[c], rzt, c == label ? -wv : 0.f);
    }
    // ---- compare: u = softmax(s/T) used as CE *logits*, target tau = softmax(t/T)
    float u[CLS], tau[CLS], zu = 0.f, ztau = 0.f;
    const float nms2 = -ms * k2, nmt2 = -mt * k2;
#pragma unroll
    for (int c = 0; c < CLS; ++c) {
      u[c] = ex2_fast(fmaf(s[c], k2, nms2));
      tau[c] = ex2_fast(fmaf(t[c], k2, nmt2));
      zu += u[c];
      ztau += tau[c];
    }
    const float rzu = rcp_fast(zu), rztau = rcp_fast(ztau);
    float e2[CLS], z2 = 0.f;
#pragma unroll
    for (int c = 0; c < CLS; ++c) {
      u[c] *= rzu;
      tau[c] *= rztau;
      e2[c] = ex2_fast(u[c] * kL2E);   // u in [0,1]: no shift needed
      z2 += e2[c];
    }
    const float lse2 = kLn2 * lg2_fast(z2);
    // cmp = -sum tau_c (u_c - lse2) = lse2 - sum tau_c u_c   (sum tau = 1)
    float tu = 0.f;
#pragma unroll
    for (int c = 0; c < CLS; ++c) tu = fmaf(tau[c], u[c], tu);
    cmp_sum += lse2 - tu;
    if constexpr (GRAD) {
      // d cmp/d u_c = softmax(u)_c - tau_c ;  d cmp/d tau_c = lse2 - u_c; both through a softmax/T
      const float rz2 = rcp_fast(z2);
      float gu[CLS], du = 0.f;
#pragma unroll
      for (int c = 0; c < CLS; ++c) {
        gu[c] = fmaf(e2[c], rz2, -tau[c]);
        du = fmaf(gu[c], u[c], du);
      }
      // sum_c (lse2 - u_c) tau_c = lse2 - tu, so gtau_c - dtau = tu - u_c
      const float k = sc.w_cmp * sc.inv_temp;
#pragma unroll
      for (int c = 0; c < CLS; ++c) {
        gs[c] = fmaf(k * u[c], gu[c] - du, gs[c]);
        gt[c] = fmaf(k * tau[c], tu - u[c], gt[c]);
      }
    }
  }
  cls_sum += label >= 0 ? term : 0.f;
}

__device__ __forceinline__ int mirror_label(long long l, int cls) {
  return (l >= 0 && l < cls) ? (int)l : -1;       // ignore_index (-100) and out-of-range -> ignored
}

// Block-level sum of three per-thread values -> partial[blockIdx][0..2] (fixed order: deterministic).
template <int THREADS>
__device__ __forceinline__ void block_partials(float a, float b, float c, float *partial_row) {
  __shared__ float red[3][THREADS / 32];
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    a += __shfl_down_sync(0xffffffffu, a, off);
    b += __shfl_down_sync(0xffffffffu, b, off);
    c += __shfl_down_sync(0xffffffffu, c, off);
  }
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) { red[0][wid] = a; red[1][wid] = b; red[2][wid] = c; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float x = 0.f, y = 0.f, z = 0.f;
    for (int i = 0; i < THREADS / 32; ++i) { x += red[0][i]; y += red[1][i]; z += red[2][i]; }
    partial_row[0] = x; partial_row[1] = y; partial_row[2] = z;
  }
}

// --------------------------------------------------------------------- full resolution
// s,t,ds,dt: [B,CLS,HW]; labels [B,HW].  One thread = V consecutive pixels, grid-stride.
template <int CLS, bool HAS_T, bool GRAD, int V>
__global__ void __launch_bounds__(256)
mirror_full_kernel(const float *__restrict__ s, const float *__restrict__ t,
                   const long long *__restrict__ labels, float *__restrict__ partial,
                   float *__restrict__ ds, float *__restrict__ dt, const float *__restrict__ gscale,
                   MirrorScale sc, int B, long long HW) {
  if constexpr (GRAD) {
    if (gscale != nullptr) { const float g = __ldg(gscale); sc.w_cls *= g; sc.w_cmp *= g; }
  }
  const long long groups_per_img = HW / V;
  const long long total = (long long)B * groups_per_img;
  float cls_sum = 0.f, cmp_sum = 0.f, n_valid = 0.f;
  for (long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x; g < total;
       g += (long long)gridDim.x * blockDim.x) {
    const long long b = g / groups_per_img;
    const long long pix = (g - b * groups_per_img) * V;
    const long long base = b * CLS * HW + pix;
    float sv[CLS][V], tv[CLS][V];
    long long lab[V];
    if constexpr (V == 4) {
      const longlong2 l0 = __ldcs(reinterpret_cast<const longlong2 *>(labels + b * HW + pix));
      const longlong2 l1 = __ldcs(reinterpret_cast<const longlong2 *>(labels + b * HW + pix + 2));
      lab[0] = l0.x; lab[1] = l0.y; lab[2] = l1.x; lab[3] = l1.y;
#pragma unroll
      for (int c = 0; c < CLS; ++c) {
        const float4 a = __ldcs(reinterpret_cast<const float4 *>(s + base + c * HW));
        sv[c][0] = a.x; sv[c][1] = a.y; sv[c][2] = a.z; sv[c][3] = a.w;
        if constexpr (HAS_T) {
          const float4 q = __ldcs(reinterpret_cast<const float4 *>(t + base + c * HW));
          tv[c][0] = q.x; tv[c][1] = q.y; tv[c][2] = q.z; tv[c][3] = q.w;
        }
      }
    } else {
      lab[0] = labels[b * HW + pix];
#pragma unroll
      for (int c = 0; c < CLS; ++c) {
        sv[c][0] = s[base + c * HW];
        if constexpr (HAS_T) tv[c][0] = t[base + c * HW];
      }
    }
    float gsv[CLS][V], gtv[CLS][V];
#pragma unroll
    for (int v = 0; v < V; ++v) {
      float sp[CLS], tp[CLS], gs[CLS], gt[CLS];
#pragma unroll
      for (int c = 0; c < CLS; ++c) { sp[c] = sv[c][v]; tp[c] = HAS_T ? tv[c][v] : 0.f; }
      const int label = mirror_label(lab[v], CLS);
      n_valid += label >= 0 ? 1.f : 0.f;
      mirror_pixel<CLS, HAS_T, GRAD>(sp, tp, label, sc, cls_sum, cmp_sum, gs, gt);
      if constexpr (GRAD) {
#pragma unroll
        for (int c = 0; c < CLS; ++c) { gsv[c][v] = gs[c]; gtv[c][v] = HAS_T ? gt[c] : 0.f; }
      }
    }
    if constexpr (GRAD) {
#pragma unroll
      for (int c = 0; c < CLS; ++c) {
        if constexpr (V == 4) {
          __stcs(reinterpret_cast<float4 *>(ds + base + c * HW),
                 make_float4(gsv[c][0], gsv[c][1], gsv[c][2], gsv[c][3]));
          if constexpr (HAS_T)
            __stcs(reinterpret_cast<float4 *>(dt + base + c * HW),
                   make_float4(gtv[c][0], gtv[c][1], gtv[c][2], gtv[c][3]));
        } else {
          ds[base + c * HW] = gsv[c][0];
          if constexpr (HAS_T) dt[base + c * HW] = gtv[c][0];
        }
      }
    }
  }
  block_partials<256>(cls_sum, cmp_sum, n_valid, partial + 3ll * blockIdx.x);
}

// out[0] = loss, out[1] = class, out[2] = compare (un-weighted), out[3] = valid labelled pixels counted
__global__ void __launch_bounds__(256)
mirror_finalize_kernel(const float *__restrict__ partial, int rows, float w_cls, float inv_npix,
                       float lmbd, float *__restrict__ out) {
  __shared__ double red[3][256];
  double a = 0, b = 0, c = 0;
  for (int i = threadIdx.x; i < rows; i += 256) { a += partial[3 * i]; b += partial[3 * i + 1]; c += partial[3 * i + 2]; }
  red[0][threadIdx.x] = a; red[1][threadIdx.x] = b; red[2][threadIdx.x] = c;
  __syncthreads();
  for (int off = 128; off > 0; off >>= 1) {
    if (threadIdx.x < off) {
      red[0][threadIdx.x] += red[0][threadIdx.x + off];
      red[1][threadIdx.x] += red[1][threadIdx.x + off];
      red[2][threadIdx.x] += red[2][threadIdx.x + off];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const double cls = red[0][0] * (double)w_cls, cmp = red[1][0] * (double)inv_npix;
    out[0] = (float)(cls + (double)lmbd * cmp);
    out[1] = (float)cls;
    out[2] = (float)cmp;
    out[3] = (float)red[2][0];
  }
}

// ------------------------------------------------------------- fused bilinear upsample
// ATen's upsample_bilinear2d (align_corners=False): src = r*(dst+0.5)-0.5 clamped at 0,
// i0 = (int)src, i1 = i0 + (i0 < in-1), l1 = src - i0, l0 = 1 - l1, with r = (float)in/out.
__host__ __device__ __forceinline__ void bilinear_src(int dst, float r, int in, int &i0, int &i1,
                                                      float &l0, float &l1) {
  float src = r * ((float)dst + 0.5f) - 0.5f;
  if (src < 0.f) src = 0.f;
  i0 = (int)src;
  if (i0 > in - 1) i0 = in - 1;
  i1 = i0 + (i0 < in - 1 ? 1 : 0);
  l1 = src - (float)i0;
  l0 = 1.f - l1;
}

constexpr int kUpTW = 128;   // output columns per CTA = threads per CTA
constexpr int kUpTH = 16;    // output rows per CTA

// s_lo,t_lo,ds_lo,dt_lo: [B,CLS,h,w]; labels [B,H,W].  CTA = kUpTH x kUpTW output pixels of one
// image, thread = one output column.  Shared memory: the low-resolution patch the tile reads
// ([2][CLS][PH][PW] values) and, with GRAD, a same-shaped gradient accumulator.
template <int CLS, bool HAS_T, bool GRAD>
__global__ void __launch_bounds__(kUpTW)
mirror_up_kernel(const float *__restrict__ s_lo, const float *__restrict__ t_lo,
                 const long long *__restrict__ labels, float *__restrict__ partial,
                 float *__restrict__ ds_lo, float *__restrict__ dt_lo, MirrorScale sc, int h, int w,
                 int H, int W, float ry, float rx, int PH, int PW) {
  extern __shared__ float smem[];
  constexpr int NS = HAS_T ? 2 : 1;
  const int cell = PH * PW;
  float *val = smem;                       // [NS][CLS][PH][PW]
  float *acc = smem + NS * CLS * cell;     // [NS][CLS][PH][PW]  (GRAD only)
  const int b = blockIdx.z, tid = threadIdx.x, lane = tid & 31;
  const int x_tile = blockIdx.x * kUpTW, y_tile = blockIdx.y * kUpTH;
  const int rows = min(kUpTH, H - y_tile);

  int py0, px0, tmp;
  float f0, f1;
  bilinear_src(y_tile, ry, h, py0, tmp, f0, f1);
  bilinear_src(x_tile, rx, w, px0, tmp, f0, f1);
  for (int i = tid; i < NS * CLS * cell; i += kUpTW) {
    const int k = i / (CLS * cell), c = (i / cell) % CLS, pi = (i % cell) / PW, pj = i % PW;
    const int yy = py0 + pi, xx = px0 + pj;
    const float *src = (k == 0) ? s_lo : t_lo;
    val[i] = (yy < h && xx < w) ? __ldg(src + (((long long)b * CLS + c) * h + yy) * w + xx) : 0.f;
    if constexpr (GRAD) acc[i] = 0.f;
  }

  const int x = x_tile + tid;
  const bool live = x < W;
  const int xc = live ? x : W - 1;          // dead lanes keep a sorted key and contribute zeros
  int ix0, ix1;
  float l0x, l1x;
  bilinear_src(xc, rx, w, ix0, ix1, l0x, l1x);
  ix0 -= px0; ix1 -= px0;
  // labels of this thread's column: all loads issued before the first use
  // (then packed to 4-bit codes, label+1, so that the row loop needs no unrolling)
  long long lab[kUpTH];
#pragma unroll
  for (int r = 0; r < kUpTH; ++r)
    lab[r] = (live && r < rows) ? __ldcs(labels + ((long long)b * H + y_tile + r) * W + x) : kIgnoreIndex;
  unsigned long long codes = 0;
#pragma unroll
  for (int r = 0; r < kUpTH; ++r)
    codes |= (unsigned long long)(mirror_label(lab[r], CLS) + 1) << (4 * r);
  // segmented-reduction plan over the (sorted) key ix0 within the warp
  unsigned same = 0;
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const int other = __shfl_down_sync(0xffffffffu, ix0, 1 << i);
    if (lane + (1 << i) < 32 && other == ix0) same |= 1u << i;
  }
  const int prev = __shfl_up_sync(0xffffffffu, ix0, 1);
  const bool head = lane == 0 || prev != ix0;
  __syncthreads();

  float top[NS][CLS], bot[NS][CLS], a_top[NS][CLS], a_bot[NS][CLS];
  int cur0 = -1, cur1 = -1;
  float cls_sum = 0.f, cmp_sum = 0.f, n_valid = 0.f;

  auto flush = [&]() {
    if constexpr (GRAD) {
#pragma unroll
      for (int k = 0; k < NS; ++k)
#pragma unroll
        for (int c = 0; c < CLS; ++c)
#pragma unroll
          for (int tb = 0; tb < 2; ++tb) {
            const float a = tb == 0 ? a_top[k][c] : a_bot[k][c];
            float v0 = l0x * a, v1 = l1x * a;
#pragma unroll
            for (int i = 0; i < 5; ++i) {
              const float o0 = __shfl_down_sync(0xffffffffu, v0, 1 << i);
              const float o1 = __shfl_down_sync(0xffffffffu, v1, 1 << i);
              if (same & (1u << i)) { v0 += o0; v1 += o1; }
            }
            if (head) {
              float *row = acc + (k * CLS + c) * cell + (tb == 0 ? cur0 : cur1) * PW;
              atomicAdd(row + ix0, v0);
              atomicAdd(row + ix1, v1);
            }
          }
    }
  };

#pragma unroll 1
  for (int r = 0; r < rows; ++r) {            // CTA-uniform trip count and branches
    int iy0, iy1;
    float l0y, l1y;
    bilinear_src(y_tile + r, ry, h, iy0, iy1, l0y, l1y);
    iy0 -= py0; iy1 -= py0;
    if (iy0 != cur0) {
      if (cur0 >= 0) flush();
      cur0 = iy0; cur1 = iy1;
#pragma unroll
      for (int k = 0; k < NS; ++k)
#pragma unroll
        for (int c = 0; c < CLS; ++c) {
          const float *p = val + (k * CLS + c) * cell;
          top[k][c] = l0x * p[iy0 * PW + ix0] + l1x * p[iy0 * PW + ix1];
          bot[k][c] = l0x * p[iy1 * PW + ix0] + l1x * p[iy1 * PW + ix1];
          a_top[k][c] = 0.f; a_bot[k][c] = 0.f;
        }
    }
    float sp[CLS], tp[CLS], gs[CLS], gt[CLS];
#pragma unroll
    for (int c = 0; c < CLS; ++c) {
      sp[c] = l0y * top[0][c] + l1y * bot[0][c];
      tp[c] = HAS_T ? l0y * top[NS - 1][c] + l1y * bot[NS - 1][c] : 0.f;
    }
    const int label = (int)((codes >> (4 * r)) & 15ull) - 1;   // -1: ignored / dead lane
    float cs = 0.f, cm = 0.f;
    mirror_pixel<CLS, HAS_T, GRAD>(sp, tp, label, sc, cs, cm, gs, gt);
    if (live) { cls_sum += cs; cmp_sum += cm; n_valid += label >= 0 ? 1.f : 0.f; }
    if constexpr (GRAD) {
      const float m = live ? 1.f : 0.f;
#pragma unroll
      for (int c = 0; c < CLS; ++c) {
        a_top[0][c] += m * l0y * gs[c]; a_bot[0][c] += m * l1y * gs[c];
        if constexpr (HAS_T) { a_top[1][c] += m * l0y * gt[c]; a_bot[1][c] += m * l1y * gt[c]; }
      }
    }
  }
  if (cur0 >= 0) flush();
  if constexpr (GRAD) {
    __syncthreads();
    for (int i = tid; i < NS * CLS * cell; i += kUpTW) {
      const int k = i / (CLS * cell), c = (i / cell) % CLS, pi = (i % cell) / PW, pj = i % PW;
      const int yy = py0 + pi, xx = px0 + pj;
      const float v = acc[i];
      if (yy < h && xx < w && v != 0.f)
        atomicAdd((k == 0 ? ds_lo : dt_lo) + (((long long)b * CLS + c) * h + yy) * w + xx, v);
    }
  }
  const long long cta = ((long long)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
  block_partials<kUpTW>(cls_sum, cmp_sum, n_valid, partial + 3 * cta);
}

// ------------------------------------------------- fused bilinear upsample, two classes (the polyp masks)
// With two classes every term of the loss depends on the logits only through their DIFFERENCE a = s1 - s0,
// b = t1 - t0, and bilinear interpolation is linear: the kernel interpolates the difference maps (half the patch,
// half the lerps), evaluates one pixel in ~70 arithmetic instructions (mirror_pixel2d) and accumulates ONE gradient
// per network (d/ds1 = -d/ds0 = d/da): half the segmented reductions and shared-memory atomics.  A CTA walks
// th = 16, 32 or 64 rows in batches of 16 labels, so the per-CTA set-up (patch, reduction plan, block sums, gradient
// write-back) is paid once per 64 rows on large batches; the per-row vertical weights come from a table in smem.
constexpr int kUpMaxTH = 64;
constexpr int kUpLB = 8;          // labels a thread has in flight (batch of rows)

// a, b: interpolated logit differences; label in {0, 1} or -1 (ignored).  lm = 1 for live lanes, 0 for lanes past W.
// cls_sum / cmp_sum gain the un-normalised loss terms; ga, gb = d loss / d a, d loss / d b (normalised).
template <bool HAS_T, bool GRAD>
__device__ __forceinline__ void mirror_pixel2d(float a, float b, int label, float w_cls, float k2, float kcmp,
                                               float lm, float &cls_sum, float &cmp_sum, float &n_valid,
                                               float &ga, float &gb) {
  constexpr float kL2E = 1.4426950408889634f, kLn2 = 0.6931471805599453f;
  const float lv = label >= 0 ? 1.f : 0.f;
  const float sgn = label > 0 ? -1.f : 1.f;                // d = logit(other) - logit(label) = sgn * (x1 - x0)
  const float gsc = sgn * (lv * w_cls);
  // class CE = softplus(d); d/dx = sgn * sigmoid(d)
  auto class_ce = [&](float x, float &g) -> float {
    const float d = sgn * x;
    const float e = ex2_fast(-fabsf(d) * kL2E);
    const float o = 1.f + e;
    const float term = fmaf(kLn2, lg2_fast(o), fmaxf(d, 0.f));
    if constexpr (GRAD) {
      const float r = rcp_fast(o);
      g = gsc * (d >= 0.f ? r : e * r);
    }
    return term;
  };
  float term = class_ce(a, ga);
  if constexpr (HAS_T) {
    term += class_ce(b, gb);
    // softmax(x / T) over two logits = (big, small) = (r, e r), e = exp(-|x| / T), the big one on the side of sign(x)
    const float es = ex2_fast(-fabsf(a) * k2), rs = rcp_fast(1.f + es), ss = es * rs;     // u   = softmax(s / T)
    const float et = ex2_fast(-fabsf(b) * k2), rt = rcp_fast(1.f + et), st = et * rt;     // tau = softmax(t / T)
    // CE with u as the LOGITS and tau as the target: lse(u) - sum_c tau_c u_c;  |u1 - u0| = rs - ss
    const float dus = rs - ss;
    const float e2 = ex2_fast(-dus * kL2E), o2 = 1.f + e2;
    const float lse2 = fmaf(kLn2, lg2_fast(o2), rs);
    const bool same = (__float_as_int(a) ^ __float_as_int(b)) >= 0;       // big entries in the same class
    const float x = same ? rt : st, y = same ? st : rt;
    const float tu = fmaf(rs, x, ss * y);
    cmp_sum = fmaf(lm, lse2 - tu, cmp_sum);
    if constexpr (GRAD) {
      // d/da = (u0 u1 / T) ((q1 - q0) - (tau1 - tau0)),  q = softmax(u);   d/db = -(tau0 tau1 / T) (u1 - u0)
      const float r2 = rcp_fast(o2), dq = fmaf(-e2, r2, r2);               // |q1 - q0|, sign(a)
      const float dts = rt - st;                                           // |tau1 - tau0|, sign(b)
      const float inner = dq - (same ? dts : -dts);                        // in units of sign(a)
      const unsigned sa = __float_as_uint(a) & 0x80000000u;
      const float ca = kcmp * (rs * ss) * inner, cb = kcmp * (rt * st) * dus;
      ga += __uint_as_float(__float_as_uint(ca) ^ sa);
      gb -= __uint_as_float(__float_as_uint(cb) ^ sa);
    }
  }
  cls_sum = fmaf(lv, term, cls_sum);
  n_valid += lv;
}

// s_lo,t_lo,ds_lo,dt_lo: [B,2,h,w]; labels [B,H,W].  CTA = th x kUpTW output pixels of one image, thread = one
// output column.  Shared memory: the low-resolution patch of logit DIFFERENCES the tile reads ([NS][PH][PW]) and,
// with GRAD, a same-shaped accumulator of d loss / d difference.
template <bool HAS_T, bool GRAD>
__global__ void __launch_bounds__(kUpTW)
mirror_up2_kernel(const float *__restrict__ s_lo, const float *__restrict__ t_lo,
                  const long long *__restrict__ labels, float *__restrict__ partial,
                  float *__restrict__ ds_lo, float *__restrict__ dt_lo, MirrorScale sc, int h, int w,
                  int H, int W, float ry, float rx, int PH, int PW, int th) {
  extern __shared__ float smem[];
  __shared__ int row_i[kUpMaxTH];          // (patch row of iy0) * 2 + (iy1 - iy0) of every output row of the tile
  __shared__ float row_l[kUpMaxTH];        // l1y
  constexpr int NS = HAS_T ? 2 : 1;
  const int cell = PH * PW;
  float *val = smem;                       // [NS][PH][PW]
  float *acc = smem + NS * cell;           // [NS][PH][PW]  (GRAD only)
  const int b = blockIdx.z, tid = threadIdx.x, lane = tid & 31;
  const int x_tile = blockIdx.x * kUpTW, y_tile = blockIdx.y * th;
  const int rows = min(th, H - y_tile);
  const long long plane = (long long)h * w;

  int py0, px0, tmp;
  float f0, f1;
  bilinear_src(y_tile, ry, h, py0, tmp, f0, f1);
  bilinear_src(x_tile, rx, w, px0, tmp, f0, f1);
  for (int i = tid; i < NS * cell; i += kUpTW) {
    const int k = i / cell, rem = i - k * cell, pi = rem / PW, pj = rem - pi * PW;
    const int yy = py0 + pi, xx = px0 + pj;
    const float *src = ((k == 0) ? s_lo : t_lo) + ((long long)b * 2 * h + yy) * w + xx;
    val[i] = (yy < h && xx < w) ? __ldg(src + plane) - __ldg(src) : 0.f;
    if constexpr (GRAD) acc[i] = 0.f;
  }
  if (tid < th) {
    int i0, i1;
    float l0, l1;
    bilinear_src(min(y_tile + tid, H - 1), ry, h, i0, i1, l0, l1);
    row_i[tid] = ((i0 - py0) << 1) | (i1 - i0);
    row_l[tid] = l1;
  }

  const int x = x_tile + tid;
  const bool live = x < W;
  const int xc = live ? x : W - 1;          // dead lanes keep a sorted key and contribute zeros
  int ix0, ix1;
  float l0x, l1x;
  bilinear_src(xc, rx, w, ix0, ix1, l0x, l1x);
  ix0 -= px0; ix1 -= px0;
  // segmented-reduction plan over the (sorted) key ix0 within the warp
  unsigned same = 0;
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const int other = __shfl_down_sync(0xffffffffu, ix0, 1 << i);
    if (lane + (1 << i) < 32 && other == ix0) same |= 1u << i;
  }
  const int prev = __shfl_up_sync(0xffffffffu, ix0, 1);
  const bool head = lane == 0 || prev != ix0;
  __syncthreads();

  constexpr float kL2E = 1.4426950408889634f;
  const float lm = live ? 1.f : 0.f;
  const float k2 = kL2E * sc.inv_temp, kcmp = live ? sc.w_cmp * sc.inv_temp : 0.f, w_cls = sc.w_cls;
  float top[NS], bot[NS], a_top[NS], a_bot[NS];
  int cur = -1, cur0 = 0, cur1 = 0;
  float cls_sum = 0.f, cmp_sum = 0.f, n_valid = 0.f;

  auto flush = [&]() {
    if constexpr (GRAD) {
#pragma unroll
      for (int k = 0; k < NS; ++k)
#pragma unroll
        for (int tb = 0; tb < 2; ++tb) {
          const float a = tb == 0 ? a_top[k] : a_bot[k];
          float v0 = l0x * a, v1 = l1x * a;
#pragma unroll
          for (int i = 0; i < 5; ++i) {
            const float o0 = __shfl_down_sync(0xffffffffu, v0, 1 << i);
            const float o1 = __shfl_down_sync(0xffffffffu, v1, 1 << i);
            if (same & (1u << i)) { v0 += o0; v1 += o1; }
          }
          if (head) {
            float *row = acc + k * cell + (tb == 0 ? cur0 : cur1) * PW;
            atomicAdd(row + ix0, v0);
            atomicAdd(row + ix1, v1);
          }
        }
    }
  };

  const long long *lab_col = labels + ((long long)b * H + y_tile) * W + x;
#pragma unroll 1
  for (int r0 = 0; r0 < rows; r0 += kUpLB) {          // CTA-uniform trip counts and branches throughout
    // labels of this thread's column for kUpLB rows: all loads issued before the first use, then packed to 2-bit
    // codes (label + 1; 0 = ignored) so that the row loop needs no unrolling
    const int rend = min(r0 + kUpLB, rows);
    const int nlab = live ? rend - r0 : 0;
    unsigned long long lab[kUpLB];
#pragma unroll
    for (int r = 0; r < kUpLB; ++r) {
      lab[r] = r < nlab ? (unsigned long long)__ldcs(lab_col) : ~0ull;
      lab_col += W;
    }
    unsigned codes = 0;
#pragma unroll
    for (int r = 0; r < kUpLB; ++r) codes |= (lab[r] < 2ull ? (unsigned)lab[r] + 1u : 0u) << (2 * r);
#pragma unroll 1
    for (int r = r0; r < rend; ++r) {
      const int ri = row_i[r];
      const float l1y = row_l[r], l0y = 1.f - l1y;
      if (ri != cur) {
        if (cur >= 0) flush();
        cur = ri; cur0 = ri >> 1; cur1 = cur0 + (ri & 1);
#pragma unroll
        for (int k = 0; k < NS; ++k) {
          const float *p = val + k * cell;
          top[k] = l0x * p[cur0 * PW + ix0] + l1x * p[cur0 * PW + ix1];
          bot[k] = l0x * p[cur1 * PW + ix0] + l1x * p[cur1 * PW + ix1];
          a_top[k] = 0.f; a_bot[k] = 0.f;
        }
      }
      const float a = l0y * top[0] + l1y * bot[0];
      const float bb = HAS_T ? l0y * top[NS - 1] + l1y * bot[NS - 1] : 0.f;
      const int label = (int)((codes >> (2 * (r - r0))) & 3u) - 1;       // -1: ignored / dead lane
      float ga = 0.f, gb = 0.f;
      mirror_pixel2d<HAS_T, GRAD>(a, bb, label, w_cls, k2, kcmp, lm, cls_sum, cmp_sum, n_valid, ga, gb);
      if constexpr (GRAD) {
        a_top[0] = fmaf(l0y, ga, a_top[0]); a_bot[0] = fmaf(l1y, ga, a_bot[0]);
        if constexpr (HAS_T) { a_top[1] = fmaf(l0y, gb, a_top[1]); a_bot[1] = fmaf(l1y, gb, a_bot[1]); }
      }
    }
  }
  if (cur >= 0) flush();
  if constexpr (GRAD) {
    __syncthreads();
    for (int i = tid; i < NS * cell; i += kUpTW) {
      const int k = i / cell, rem = i - k * cell, pi = rem / PW, pj = rem - pi * PW;
      const int yy = py0 + pi, xx = px0 + pj;
      const float v = acc[i];
      if (yy < h && xx < w && v != 0.f) {
        float *dst = (k == 0 ? ds_lo : dt_lo) + ((long long)b * 2 * h + yy) * w + xx;
        atomicAdd(dst + plane, v);           // d/d logit(1) = d/d difference
        atomicAdd(dst, -v);                  // d/d logit(0) = -d/d difference
      }
    }
  }
  const long long cta = ((long long)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
  block_partials<kUpTW>(cls_sum, cmp_sum, n_valid, partial + 3 * cta);
}

struct UpGeom { int gx, gy, PH, PW; };

UpGeom up_geometry(int h, int w, int H, int W, int th = kUpTH) {
  UpGeom g;
  g.gx = cp2_cdiv(W, kUpTW);
  g.gy = cp2_cdiv(H, th);
  const float ry = (float)h / (float)H, rx = (float)w / (float)W;
  g.PH = 1; g.PW = 1;
  int a0, a1, b0, b1; float f0, f1;
  for (int ty = 0; ty < g.gy; ++ty) {
    bilinear_src(ty * th, ry, h, a0, a1, f0, f1);
    const int last = (ty * th + th - 1 < H) ? ty * th + th - 1 : H - 1;
    bilinear_src(last, ry, h, b0, b1, f0, f1);
    if (b1 - a0 + 1 > g.PH) g.PH = b1 - a0 + 1;
  }
  for (int tx = 0; tx < g.gx; ++tx) {
    bilinear_src(tx * kUpTW, rx, w, a0, a1, f0, f1);
    const int last = (tx * kUpTW + kUpTW - 1 < W) ? tx * kUpTW + kUpTW - 1 : W - 1;
    bilinear_src(last, rx, w, b0, b1, f0, f1);
    if (b1 - a0 + 1 > g.PW) g.PW = b1 - a0 + 1;
  }
  return g;
}

}  // namespace

// ------------------------------------------------------------------------ entry points
namespace {

int full_blocks(int B, long long HW, bool vec) {
  const long long groups = (long long)B * HW / (vec ? 4 : 1);
  long long blocks = (groups + 255) / 256;
  const long long cap = (long long)cp2_num_sms() * 8;    // grid-stride; whole waves of 148 SMs
  if (blocks > cap) blocks = cap;
  return blocks < 1 ? 1 : (int)blocks;
}

bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

template <int CLS>
int launch_full(const float *s, const float *t, const long long *labels, float *partial, float *ds,
                float *dt, const float *gscale, MirrorScale sc, int B, long long HW, bool vec,
                int blocks, cudaStream_t st) {
  const bool has_t = t != nullptr, grad = ds != nullptr;
#define CP2_FULL(HT, GR, VV)                                                                     \
  mirror_full_kernel<CLS, HT, GR, VV><<<blocks, 256, 0, st>>>(s, t, labels, partial, ds, dt,     \
                                                               gscale, sc, B, HW)
  if (vec) {
    if (has_t) { if (grad) CP2_FULL(true, true, 4); else CP2_FULL(true, false, 4); }
    else       { if (grad) CP2_FULL(false, true, 4); else CP2_FULL(false, false, 4); }
  } else {
    if (has_t) { if (grad) CP2_FULL(true, true, 1); else CP2_FULL(true, false, 1); }
    else       { if (grad) CP2_FULL(false, true, 1); else CP2_FULL(false, false, 1); }
  }
#undef CP2_FULL
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

template <int CLS>
int launch_up(const float *s, const float *t, const long long *labels, float *partial, float *ds,
              float *dt, MirrorScale sc, int B, int h, int w, int H, int W, const UpGeom &g,
              cudaStream_t st) {
  const bool has_t = t != nullptr, grad = ds != nullptr;
  const size_t smem = sizeof(float) * (size_t)(has_t ? 2 : 1) * CLS * g.PH * g.PW * (grad ? 2 : 1);
  CP2_REQUIRE(smem <= 200 * 1024, "cp2_mirror_loss_upsampled: low-resolution patch of %zu bytes per tile "
              "does not fit shared memory (h=%d w=%d H=%d W=%d)", smem, h, w, H, W);
  const float ry = (float)h / (float)H, rx = (float)w / (float)W;
  const dim3 grid(g.gx, g.gy, B);
#define CP2_UP(HT, GR)                                                                            \
  do {                                                                                            \
    if (smem > 48 * 1024)                                                                         \
      CP2_CUDA(cudaFuncSetAttribute(mirror_up_kernel<CLS, HT, GR>,                                \
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));     \
    mirror_up_kernel<CLS, HT, GR><<<grid, kUpTW, smem, st>>>(s, t, labels, partial, ds, dt, sc, h, \
                                                             w, H, W, ry, rx, g.PH, g.PW);        \
  } while (0)
  if (has_t) { if (grad) CP2_UP(true, true); else CP2_UP(true, false); }
  else       { if (grad) CP2_UP(false, true); else CP2_UP(false, false); }
#undef CP2_UP
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

// Two classes: the difference kernel.  Rows per CTA: 16 while the grid is small (occupancy first), 32 / 64 once
// there are enough CTAs to keep every SM full (amortises the per-CTA set-up and gradient write-back).
int up2_rows_per_cta(int B, int H, int W) {
  const long long ctas16 = (long long)B * cp2_cdiv(H, kUpTH) * cp2_cdiv(W, kUpTW);
  const long long wave = 8LL * cp2_num_sms();
  return ctas16 >= 3 * wave ? 64 : ctas16 >= 3 * wave / 2 ? 32 : 16;
}

int launch_up2(const float *s, const float *t, const long long *labels, float *partial, float *ds,
               float *dt, MirrorScale sc, int B, int h, int w, int H, int W, const UpGeom &g, int th,
               cudaStream_t st) {
  const bool has_t = t != nullptr, grad = ds != nullptr;
  const size_t smem = sizeof(float) * (size_t)(has_t ? 2 : 1) * g.PH * g.PW * (grad ? 2 : 1);
  CP2_REQUIRE(smem <= 200 * 1024, "cp2_mirror_loss_upsampled: low-resolution patch of %zu bytes per tile "
              "does not fit shared memory (h=%d w=%d H=%d W=%d)", smem, h, w, H, W);
  const float ry = (float)h / (float)H, rx = (float)w / (float)W;
  const dim3 grid(g.gx, g.gy, B);
#define CP2_UP2(HT, GR)                                                                           \
  do {                                                                                            \
    if (smem > 48 * 1024)                                                                         \
      CP2_CUDA(cudaFuncSetAttribute(mirror_up2_kernel<HT, GR>,                                    \
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));     \
    mirror_up2_kernel<HT, GR><<<grid, kUpTW, smem, st>>>(s, t, labels, partial, ds, dt, sc, h, w, \
                                                         H, W, ry, rx, g.PH, g.PW, th);           \
  } while (0)
  if (has_t) { if (grad) CP2_UP2(true, true); else CP2_UP2(true, false); }
  else       { if (grad) CP2_UP2(false, true); else CP2_UP2(false, false); }
#undef CP2_UP2
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

int finalize(const float *partial, int rows, bool has_t, long long n_pix, long long n_valid,
             float lmbd, float *out, cudaStream_t st) {
  const float w_cls = 1.0f / (float)((has_t ? 2 : 1) * (double)n_valid);
  mirror_finalize_kernel<<<1, 256, 0, st>>>(partial, rows, w_cls, 1.0f / (float)n_pix, has_t ? lmbd : 0.f, out);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

MirrorScale make_scale(bool has_t, long long n_pix, long long n_valid, float lmbd, float inv_temp) {
  MirrorScale sc;
  sc.inv_temp = inv_temp;
  sc.w_cls = 1.0f / (float)((has_t ? 2 : 1) * (double)n_valid);
  sc.w_cmp = lmbd / (float)n_pix;
  return sc;
}

}  // namespace

extern "C" int cp2_mirror_partials(int B, int64_t HW) {
  if (B <= 0 || HW <= 0) return 1;
  return full_blocks(B, HW, HW % 4 == 0);
}

extern "C" int cp2_mirror_loss(const float *s_logits, const float *t_logits, const int64_t *labels,
                               float *partial, float *out, float *ds, float *dt, const float *gscale,
                               float lmbd_compare, float inv_temp, int64_t n_valid, int B, int Cls,
                               int64_t HW, void *stream) {
  CP2_REQUIRE(B > 0 && HW > 0, "cp2_mirror_loss: empty input (B=%d HW=%lld): the mean is undefined", B, (long long)HW);
  CP2_REQUIRE(Cls >= 2 && Cls <= kMaxCls, "cp2_mirror_loss: Cls=%d not in [2,%d]", Cls, kMaxCls);
  CP2_REQUIRE(s_logits && labels && partial && out, "cp2_mirror_loss: null pointer");
  CP2_REQUIRE((ds == nullptr) == (dt == nullptr) || t_logits == nullptr, "cp2_mirror_loss: ds and dt go together");
  CP2_REQUIRE(inv_temp > 0.f, "cp2_mirror_loss: inv_temp=%g", inv_temp);
  const bool has_t = t_logits != nullptr;
  const long long n_pix = (long long)B * HW;
  if (n_valid <= 0) n_valid = n_pix;
  const bool vec = HW % 4 == 0 && aligned16(s_logits) && (!has_t || aligned16(t_logits)) && aligned16(labels) &&
                   (ds == nullptr || (aligned16(ds) && (!has_t || aligned16(dt))));
  // the partial buffer is sized by cp2_mirror_partials (vector geometry): never launch more CTAs
  int blocks = full_blocks(B, HW, vec);
  const int cap = cp2_mirror_partials(B, HW);
  if (blocks > cap) blocks = cap;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const MirrorScale sc = make_scale(has_t, n_pix, n_valid, lmbd_compare, inv_temp);
  const long long *lab = reinterpret_cast<const long long *>(labels);
  int rc = CP2_OK;
#define CP2_CALL_FULL(N) launch_full<N>(s_logits, t_logits, lab, partial, ds, dt, gscale, sc, B, HW, vec, blocks, st)
  switch (Cls) {
    case 2: rc = CP2_CALL_FULL(2); break;
    case 3: rc = CP2_CALL_FULL(3); break;
    case 4: rc = CP2_CALL_FULL(4); break;
    case 5: rc = CP2_CALL_FULL(5); break;
    case 6: rc = CP2_CALL_FULL(6); break;
    case 7: rc = CP2_CALL_FULL(7); break;
    default: rc = CP2_CALL_FULL(8); break;
  }
#undef CP2_CALL_FULL
  if (rc != CP2_OK) return rc;
  return finalize(partial, blocks, has_t, n_pix, n_valid, lmbd_compare, out, st);
}

extern "C" int cp2_mirror_up_partials(int B, int H, int W) {
  if (B <= 0 || H <= 0 || W <= 0) return 1;
  return B * cp2_cdiv(H, kUpTH) * cp2_cdiv(W, kUpTW);
}

extern "C" int cp2_mirror_loss_upsampled(const float *s_lo, const float *t_lo, const int64_t *labels,
                                         float *partial, float *out, float *ds_lo, float *dt_lo,
                                         float lmbd_compare, float inv_temp, int64_t n_valid, int B,
                                         int Cls, int h, int w, int H, int W, void *stream) {
  CP2_REQUIRE(B > 0 && h > 0 && w > 0 && H > 0 && W > 0,
              "cp2_mirror_loss_upsampled: empty input (B=%d h=%d w=%d H=%d W=%d)", B, h, w, H, W);
  CP2_REQUIRE(Cls >= 2 && Cls <= kMaxCls, "cp2_mirror_loss_upsampled: Cls=%d not in [2,%d]", Cls, kMaxCls);
  CP2_REQUIRE(s_lo && labels && partial && out, "cp2_mirror_loss_upsampled: null pointer");
  CP2_REQUIRE((ds_lo == nullptr) == (dt_lo == nullptr) || t_lo == nullptr,
              "cp2_mirror_loss_upsampled: ds_lo and dt_lo go together");
  CP2_REQUIRE(inv_temp > 0.f, "cp2_mirror_loss_upsampled: inv_temp=%g", inv_temp);
  CP2_REQUIRE(B <= 65535, "cp2_mirror_loss_upsampled: B=%d > 65535", B);
  const bool has_t = t_lo != nullptr;
  const long long n_pix = (long long)B * H * W;
  if (n_valid <= 0) n_valid = n_pix;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t lo_bytes = sizeof(float) * (size_t)B * Cls * h * w;
  if (ds_lo && dt_lo && has_t && dt_lo == ds_lo + (size_t)B * Cls * h * w) {
    CP2_CUDA(cudaMemsetAsync(ds_lo, 0, 2 * lo_bytes, st));        // adjacent gradients: one memset node
  } else {
    if (ds_lo) CP2_CUDA(cudaMemsetAsync(ds_lo, 0, lo_bytes, st));
    if (dt_lo && has_t) CP2_CUDA(cudaMemsetAsync(dt_lo, 0, lo_bytes, st));
  }
  const int th = Cls == 2 ? up2_rows_per_cta(B, H, W) : kUpTH;
  const UpGeom g = up_geometry(h, w, H, W, th);
  const MirrorScale sc = make_scale(has_t, n_pix, n_valid, lmbd_compare, inv_temp);
  const long long *lab = reinterpret_cast<const long long *>(labels);
  int rc = CP2_OK;
#define CP2_CALL_UP(N) launch_up<N>(s_lo, t_lo, lab, partial, ds_lo, dt_lo, sc, B, h, w, H, W, g, st)
  switch (Cls) {
    case 2: rc = launch_up2(s_lo, t_lo, lab, partial, ds_lo, dt_lo, sc, B, h, w, H, W, g, th, st); break;
    case 3: rc = CP2_CALL_UP(3); break;
    case 4: rc = CP2_CALL_UP(4); break;
    case 5: rc = CP2_CALL_UP(5); break;
    case 6: rc = CP2_CALL_UP(6); break;
    case 7: rc = CP2_CALL_UP(7); break;
    default: rc = CP2_CALL_UP(8); break;
  }
#undef CP2_CALL_UP
  if (rc != CP2_OK) return rc;
  return finalize(partial, B * g.gx * g.gy, has_t, n_pix, n_valid, lmbd_compare, out, st);
}
