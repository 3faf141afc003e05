// Feature extraction on the device: reference src/v2/frame.py:5-14
// (`FeatureExtractor.compute_features`, ORB build variant of frame.py:8):
//   goodFeaturesToTrack(mean-of-channels gray, 3000, 0.01, 7) -> KeyPoint(size 20, angle -1) ->
//   ORB.compute -> KeyPoint_convert.
// Every stage reproduces OpenCV 4.13's arithmetic bit for bit (the operation orders were pinned against
// cv2 in oracle/features.py): fused multiply-adds exactly where OpenCV's AVX2 filter loops have them,
// explicit _rn intrinsics everywhere else so that nvcc cannot contract, a float64 RUNNING column sum in the
// 3x3 box filter (serial down each column, one thread per column), and the greedy minimum-distance
// selection of goodFeaturesToTrack as the lexicographically-first maximal independent set, resolved in
// parallel rounds by a thread-block cluster (8 CTAs per image, DSMEM termination vote).
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace vslam {

__device__ __forceinline__ int reflect101(int i, int n) {
  if (i < 0) i = -i;
  if (i >= n) i = 2 * n - 2 - i;
  return i;
}
// order-preserving map float -> int (for atomicMax over signed floats)
__device__ __forceinline__ int float_key(float v) {
  const int b = __float_as_int(v);
  return b ^ ((b >> 31) & 0x7fffffff);
}
__device__ __forceinline__ float key_float(int k) { return __int_as_float(k ^ ((k >> 31) & 0x7fffffff)); }

// ---- K1: the two grey images -------------------------------------------------------------------
// gm = floor((B+G+R)/3)  (frame.py:11, np.mean(...).astype(uint8));  gb = cvtColor(BGR2GRAY):
// (B*3735 + G*19235 + R*9798 + 2^14) >> 15.  Four pixels per thread (three 32-bit loads).
__device__ __forceinline__ void gray_px(uint32_t b, uint32_t g, uint32_t r, uint32_t& gm, uint32_t& gb) {
  gm = (b + g + r) / 3u;
  gb = (b * 3735u + g * 19235u + r * 9798u + 16384u) >> 15;
}

__global__ void __launch_bounds__(256)
feat_gray_kernel(const uint8_t* __restrict__ img, long long n_px, int px_per_img, uint8_t* __restrict__ gm,
                 uint8_t* __restrict__ gb, int* __restrict__ counters, int batch) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < batch) {  // per-image bookkeeping: [0] max key, [1] n_cand, [2] n_acc, [3] error
    counters[4 * t + 0] = INT_MIN;
    counters[4 * t + 1] = 0;
    counters[4 * t + 2] = 0;
    counters[4 * t + 3] = 0;
  }
  const long long p0 = 4 * t;
  if (p0 >= n_px) return;
  if (p0 + 3 < n_px) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(img + 3 * p0);
    const uint32_t w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2);
    uint32_t m0, m1, m2, m3, b0, b1, b2, b3;
    gray_px(w0 & 255u, (w0 >> 8) & 255u, (w0 >> 16) & 255u, m0, b0);
    gray_px(w0 >> 24, w1 & 255u, (w1 >> 8) & 255u, m1, b1);
    gray_px((w1 >> 16) & 255u, w1 >> 24, w2 & 255u, m2, b2);
    gray_px((w2 >> 8) & 255u, (w2 >> 16) & 255u, w2 >> 24, m3, b3);
    *reinterpret_cast<uint32_t*>(gm + p0) = m0 | (m1 << 8) | (m2 << 16) | (m3 << 24);
    *reinterpret_cast<uint32_t*>(gb + p0) = b0 | (b1 << 8) | (b2 << 16) | (b3 << 24);
  } else {
    for (long long p = p0; p < n_px; ++p) {
      uint32_t m, b;
      gray_px(img[3 * p], img[3 * p + 1], img[3 * p + 2], m, b);
      gm[p] = (uint8_t)m;
      gb[p] = (uint8_t)b;
    }
  }
}

// ---- K2: cornerMinEigenVal(blockSize 3, Sobel 3, REFLECT_101) --------------------------------------
// One thread per column (plus one halo column either side of the strip), walking down the image.
//   Sobel d/dx: rows (-1,0,1) exact; columns (s,2s,s): fma(up + down, s1, mid * s2)
//   Sobel d/dy: rows (s,2s,s): fma(s1, right, fma(s2, mid, s1 * left)); columns (-1,0,1) exact
//   cov = (dx*dx, dx*dy, dy*dy) float; box rows: (l + m) + r in double; box columns: running double sum
//   eig = (a + c) - sqrt((a - c)^2 + b^2), a = A/2, c = C/2, float, not contracted.
constexpr int kEigT = 64;

__global__ void __launch_bounds__(kEigT)
feat_eig_kernel(const uint8_t* __restrict__ gray, int H, int W, float* __restrict__ eig, int* __restrict__ counters) {
  __shared__ float sm[2][3][kEigT];
  const int b = blockIdx.y, tid = threadIdx.x;
  const int c = blockIdx.x * (kEigT - 2) + tid - 1;
  const bool owner = tid >= 1 && tid <= kEigT - 2 && c < W;
  const int cc = reflect101(min(c, W), W);
  const int xm = reflect101(cc - 1, W), xp = reflect101(cc + 1, W);
  const uint8_t* g = gray + (size_t)b * H * W;
  float* out = eig + (size_t)b * H * W;
  const float s1 = __int_as_float(0x39ab5601), s2 = __int_as_float(0x3a2b5601);
  const bool vec_col = cc < (W / 32) * 32;

  auto row_terms = [&](int q, float& rx, float& R) {
    const uint8_t* gr = g + (size_t)q * W;
    const float l = (float)__ldg(gr + xm), m = (float)__ldg(gr + cc), r = (float)__ldg(gr + xp);
    rx = __fsub_rn(r, l);
    // OpenCV's uint8 -> float32 row filter fuses in its 32-pixel vector body and rounds every product in the
    // scalar tail (oracle/features.py: row_vector_width)
    R = vec_col ? __fmaf_rn(s1, r, __fmaf_rn(s2, m, __fmul_rn(s1, l)))
                : __fadd_rn(__fadd_rn(__fmul_rn(s1, l), __fmul_rn(s2, m)), __fmul_rn(s1, r));
  };
  float rx0, R0, rx1, R1, rx2, R2;
  row_terms(1, rx0, R0);  // row -1 reflects to row 1
  row_terms(0, rx1, R1);
  rx2 = rx0; R2 = R0;
  double rsa[3] = {0, 0, 0}, rsb[3] = {0, 0, 0}, sum[3] = {0, 0, 0};
  float vmax = -INFINITY;

  auto emit = [&](int y, const double* s0) {
    const float A = (float)s0[0], B = (float)s0[1], C = (float)s0[2];
    const float a = __fmul_rn(A, 0.5f), cq = __fmul_rn(C, 0.5f);
    float t = __fsub_rn(a, cq);
    t = __fadd_rn(__fmul_rn(t, t), __fmul_rn(B, B));
    const float e = __fsub_rn(__fadd_rn(a, cq), __fsqrt_rn(t));
    out[(size_t)y * W + c] = e;
    vmax = fmaxf(vmax, e);
  };

  for (int r = 0; r < H; ++r) {
    float rxn, Rn;
    row_terms(reflect101(r + 2, H), rxn, Rn);  // prefetch the row after next
    const float dx = __fmaf_rn(__fadd_rn(rx0, rx2), s1, __fmul_rn(rx1, s2));
    const float dy = __fsub_rn(R2, R0);
    const int buf = r & 1;
    sm[buf][0][tid] = __fmul_rn(dx, dx);
    sm[buf][1][tid] = __fmul_rn(dx, dy);
    sm[buf][2][tid] = __fmul_rn(dy, dy);
    __syncthreads();
    if (owner) {
      double rsc[3];
#pragma unroll
      for (int k = 0; k < 3; ++k)
        rsc[k] = ((double)sm[buf][k][tid - 1] + (double)sm[buf][k][tid]) + (double)sm[buf][k][tid + 1];
      if (r == 1) {
#pragma unroll
        for (int k = 0; k < 3; ++k) sum[k] = (0.0 + rsc[k]) + rsb[k];  // padded rows -1 (= row 1) and 0
      }
      if (r >= 1) {
        double s0[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          s0[k] = sum[k] + rsc[k];
          sum[k] = s0[k] - (r == 1 ? rsc[k] : rsa[k]);  // minus row y-1 (row -1 reflects to row 1)
        }
        emit(r - 1, s0);
      }
#pragma unroll
      for (int k = 0; k < 3; ++k) { rsa[k] = rsb[k]; rsb[k] = rsc[k]; }
    }
    rx0 = rx1; R0 = R1; rx1 = rx2; R1 = R2; rx2 = rxn; R2 = Rn;
  }
  if (owner) {  // last row: row H reflects to row H-2, which is rsa after the final shift
    double s0[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) s0[k] = sum[k] + rsa[k];
    emit(H - 1, s0);
  }
  int key = owner ? float_key(vmax) : INT_MIN;
  key = __reduce_max_sync(0xffffffffu, key);
  if ((tid & 31) == 0) atomicMax(&counters[4 * b + 0], key);
}

// ---- K3: threshold + strict 3x3 local maxima -> candidate list and status image ------------------
__global__ void __launch_bounds__(256)
feat_candidates_kernel(const float* __restrict__ eig, int H, int W, double quality, uint8_t* __restrict__ stat,
                       int* __restrict__ cand, int* __restrict__ counters) {
  const int b = blockIdx.y;
  const int pix = blockIdx.x * blockDim.x + threadIdx.x;
  const float* e = eig + (size_t)b * H * W;
  bool is_cand = false;
  if (pix < H * W) {
    const int y = pix / W, x = pix - y * W;
    if (y >= 1 && y < H - 1 && x >= 1 && x < W - 1) {
      const float thr = (float)((double)key_float(counters[4 * b + 0]) * quality);
      const float v = e[pix];
      const float ev = v > thr ? v : 0.0f;  // threshold(THRESH_TOZERO)
      if (ev != 0.0f) {
        float mx = ev;                      // dilate 3x3 of the thresholded image
#pragma unroll
        for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
          for (int dx = -1; dx <= 1; ++dx) {
            const float n = e[pix + dy * W + dx];
            mx = fmaxf(mx, n > thr ? n : 0.0f);
          }
        is_cand = ev == mx && ev > 0.0f;    // (a negative response cannot be a corner)
      }
    }
    stat[(size_t)b * H * W + pix] = is_cand ? 1 : 0;
  }
  const unsigned m = __ballot_sync(0xffffffffu, is_cand);
  if (m) {
    const int lane = threadIdx.x & 31;
    int base = 0;
    if (lane == __ffs(m) - 1) base = atomicAdd(&counters[4 * b + 1], __popc(m));
    base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
    if (is_cand) cand[(size_t)b * H * W + base + __popc(m & ((1u << lane) - 1))] = pix;
  }
}

// ---- K4: greedy minimum-distance selection as parallel rounds over a cluster ---------------------
// cv2 walks the candidates in order (value desc, address desc) and keeps one iff no kept candidate of
// higher order lies at squared distance < d2.  Equivalent fixed point: a candidate is KEPT once every
// higher-order candidate within range is DROPPED, and DROPPED once one of them is KEPT.  Decisions are
// final, so rounds may read each other's writes.  status: 0 none, 1 undecided, 2 kept, 3 dropped.
constexpr int kSupCtas = 8;
constexpr int kSupThreads = 512;
constexpr int kRankTile = 1024;

__global__ void __cluster_dims__(kSupCtas, 1, 1) __launch_bounds__(kSupThreads)
feat_suppress_kernel(const float* __restrict__ eig, uint8_t* stat, const int* __restrict__ cand, int H, int W,
                     int radius, int d2, unsigned long long* acc_key, int acc_cap, int max_corners,
                     float2* __restrict__ corners, int* __restrict__ n_corners, int* counters) {
  cg::cluster_group cluster = cg::this_cluster();
  __shared__ int slot[2];
  __shared__ unsigned long long tile[kRankTile];
  const int b = blockIdx.y;
  const int rank = (int)cluster.block_rank();
  const int gtid = rank * kSupThreads + threadIdx.x, gsize = kSupCtas * kSupThreads;
  const size_t img_off = (size_t)b * H * W;
  const unsigned* ebits = reinterpret_cast<const unsigned*>(eig) + img_off;
  uint8_t* st = stat + img_off;
  const int* cl = cand + img_off;
  unsigned long long* acc = acc_key + (size_t)b * acc_cap;
  const int n = counters[4 * b + 1];

  for (int round = 0;; ++round) {
    int pending = 0;
    for (int i = gtid; i < n; i += gsize) {
      const int pix = cl[i];
      if (__ldcg(st + pix) != 1) continue;
      const int y = pix / W, x = pix - y * W;
      const unsigned v = ebits[pix];
      bool drop = false, wait = false;
      const int y0 = max(0, y - radius), y1 = min(H - 1, y + radius);
      const int x0 = max(0, x - radius), x1 = min(W - 1, x + radius);
      for (int yy = y0; yy <= y1 && !drop; ++yy) {
        const int dy2 = (yy - y) * (yy - y);
        for (int xx = x0; xx <= x1; ++xx) {
          if (dy2 + (xx - x) * (xx - x) >= d2) continue;
          const int q = yy * W + xx;
          const int sq = __ldcg(st + q);
          if (sq == 0 || sq == 3 || q == pix) continue;
          const unsigned vq = ebits[q];
          if (vq > v || (vq == v && q > pix)) {
            if (sq == 2) { drop = true; break; }
            wait = true;
          }
        }
      }
      if (drop) {
        __stcg(st + pix, (uint8_t)3);
      } else if (!wait) {
        __stcg(st + pix, (uint8_t)2);
        const int k = atomicAdd(&counters[4 * b + 2], 1);
        if (k < acc_cap) acc[k] = ((unsigned long long)v << 32) | (unsigned)pix;
      } else {
        ++pending;
      }
    }
    const int any = __syncthreads_or(pending > 0);
    if (threadIdx.x == 0) slot[round & 1] = any;
    cluster.sync();
    int total = 0;
#pragma unroll
    for (int k = 0; k < kSupCtas; ++k) total |= *cluster.map_shared_rank(&slot[round & 1], k);
    if (!total) break;
  }
  cluster.sync();
  // order of the kept corners = number of kept corners of higher order (counting rank, O(A^2))
  const int n_acc_raw = __ldcg(&counters[4 * b + 2]);
  const int A = min(n_acc_raw, acc_cap);
  if (gtid == 0) {
    if (n_acc_raw > acc_cap) counters[4 * b + 3] = 1;
    n_corners[b] = min(A, max_corners);
  }
  const int passes = (A + gsize - 1) / gsize;
  for (int p = 0; p < passes; ++p) {
    const int i = p * gsize + gtid;
    const unsigned long long key = i < A ? __ldcg(acc + i) : 0ull;
    int higher = 0;
    for (int t0 = 0; t0 < A; t0 += kRankTile) {
      const int m = min(kRankTile, A - t0);
      __syncthreads();
      for (int j = threadIdx.x; j < m; j += kSupThreads) tile[j] = __ldcg(acc + t0 + j);
      __syncthreads();
      if (i < A)
        for (int j = 0; j < m; ++j) higher += tile[j] > key;
    }
    if (i < A && higher < max_corners) {
      const int pix = (int)(key & 0xffffffffu);
      const int y = pix / W;
      corners[(size_t)b * max_corners + higher] = make_float2((float)(pix - y * W), (float)y);
    }
  }
  cluster.sync();  // no CTA leaves while its shared memory may still be read
}

// ---- K5: the blur ORB applies to its level image ---------------------------------------------------
// float32 separable 7x7, sigma 2 (taps = cv2.getGaussianKernel(7, 2, CV_32F)), REFLECT_101;
// rows: s = k0*p0, s = fma(k_i, p_i, s); columns: s = k3*c, s = fma(k_{3+j}, up_j + down_j, s);
// saturate_cast<uchar> = round-half-even.
constexpr int kBlurW = 64, kBlurH = 32;

__global__ void __launch_bounds__(256)
feat_blur_kernel(const uint8_t* __restrict__ gray, int H, int W, uint8_t* __restrict__ blur) {
  __shared__ float in[kBlurH + 6][kBlurW + 6 + 2];
  __shared__ float hz[kBlurH + 6][kBlurW];
  const float k0 = __int_as_float(0x3d8fafb1), k1 = __int_as_float(0x3e06387e), k2 = __int_as_float(0x3e434a39),
              k3 = __int_as_float(0x3e5d4ae0);
  const float kk[7] = {k0, k1, k2, k3, k2, k1, k0};
  const int b = blockIdx.z, x0 = blockIdx.x * kBlurW, y0 = blockIdx.y * kBlurH;
  const uint8_t* g = gray + (size_t)b * H * W;
  for (int i = threadIdx.x; i < (kBlurH + 6) * (kBlurW + 6); i += 256) {
    const int r = i / (kBlurW + 6), c = i - r * (kBlurW + 6);
    const int y = reflect101(min(y0 + r - 3, H + 2), H), x = reflect101(min(x0 + c - 3, W + 2), W);
    in[r][c] = (float)__ldg(g + (size_t)y * W + x);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < (kBlurH + 6) * kBlurW; i += 256) {
    const int r = i / kBlurW, c = i - r * kBlurW;
    float s = __fmul_rn(kk[0], in[r][c]);
    if (x0 + c < (W / 32) * 32) {  // vector body of OpenCV's row filter: fused; scalar tail: products rounded
#pragma unroll
      for (int t = 1; t < 7; ++t) s = __fmaf_rn(kk[t], in[r][c + t], s);
    } else {
#pragma unroll
      for (int t = 1; t < 7; ++t) s = __fadd_rn(s, __fmul_rn(kk[t], in[r][c + t]));
    }
    hz[r][c] = s;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kBlurH * kBlurW; i += 256) {
    const int r = i / kBlurW, c = i - r * kBlurW;
    const int y = y0 + r, x = x0 + c;
    if (y >= H || x >= W) continue;
    float s = __fmul_rn(kk[3], hz[r + 3][c]);
#pragma unroll
    for (int j = 1; j <= 3; ++j) s = __fmaf_rn(kk[3 + j], __fadd_rn(hz[r + 3 + j][c], hz[r + 3 - j][c]), s);
    const int v = __float2int_rn(s);
    blur[(size_t)b * H * W + (size_t)y * W + x] = (uint8_t)min(255, max(0, v));
  }
}

// ---- K6a: KeyPointsFilter::runByImageBorder(31), order kept ---------------------------------------
__global__ void __launch_bounds__(1024)
feat_border_compact_kernel(const float2* __restrict__ corners, const int* __restrict__ n_corners, int max_corners,
                           int H, int W, float2* __restrict__ kp, int* __restrict__ n_kp) {
  __shared__ int warp_tot[32];
  __shared__ int base_s;
  const int b = blockIdx.x, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int n = n_corners[b];
  if (threadIdx.x == 0) base_s = 0;
  __syncthreads();
  for (int i0 = 0; i0 < n; i0 += 1024) {
    const int i = i0 + threadIdx.x;
    float2 p = make_float2(0.f, 0.f);
    bool keep = false;
    if (i < n) {
      p = corners[(size_t)b * max_corners + i];
      keep = p.x >= 31.f && p.x < (float)(W - 31) && p.y >= 31.f && p.y < (float)(H - 31);
    }
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) warp_tot[wid] = __popc(m);
    __syncthreads();
    int before = 0, total = 0;
    for (int w = 0; w < 32; ++w) {
      const int c = warp_tot[w];
      before += w < wid ? c : 0;
      total += c;
    }
    const int base = base_s;
    if (keep) kp[(size_t)b * max_corners + base + before + __popc(m & ((1u << lane) - 1))] = p;
    __syncthreads();
    if (threadIdx.x == 0) base_s = base + total;
    __syncthreads();
  }
  if (threadIdx.x == 0) n_kp[b] = base_s;
}

// ---- K6b: rBRIEF, WTA_K = 2 -----------------------------------------------------------------------
// One warp per key point, one descriptor byte per lane (16 pattern points = 8 tests).  Offsets:
// x = px*a - py*b, y = px*b + py*a in float32 without contraction, round-half-even (cvRound).
constexpr int kOrbWarps = 8;

__global__ void __launch_bounds__(kOrbWarps * 32)
feat_orb_kernel(const uint8_t* __restrict__ blur, int H, int W, const float2* __restrict__ kp,
                const int* __restrict__ n_kp, int max_corners, const int8_t* __restrict__ pattern, float ca, float sa,
                uint8_t* __restrict__ desc) {
  __shared__ int off[512];  // rotated offsets (the angle is shared by all key points of a call)
  const int b = blockIdx.y;
  for (int i = threadIdx.x; i < 512; i += blockDim.x) {
    const float px = (float)pattern[2 * i], py = (float)pattern[2 * i + 1];
    const float x = __fsub_rn(__fmul_rn(px, ca), __fmul_rn(py, sa));
    const float y = __fadd_rn(__fmul_rn(px, sa), __fmul_rn(py, ca));
    off[i] = __float2int_rn(y) * W + __float2int_rn(x);
  }
  __syncthreads();
  const int n = n_kp[b];
  const int lane = threadIdx.x & 31;
  const uint8_t* img = blur + (size_t)b * H * W;
  for (int k = blockIdx.x * kOrbWarps + (threadIdx.x >> 5); k < n; k += gridDim.x * kOrbWarps) {
    const float2 p = kp[(size_t)b * max_corners + k];
    const uint8_t* c = img + (size_t)__float2int_rn(p.y) * W + __float2int_rn(p.x);
    unsigned byte = 0;
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      const int t0 = c[off[lane * 16 + 2 * t]], t1 = c[off[lane * 16 + 2 * t + 1]];
      byte |= (unsigned)(t0 < t1) << t;
    }
    desc[((size_t)b * max_corners + k) * 32 + lane] = (uint8_t)byte;
  }
}

static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

struct FeatLayout {
  size_t gm, gb, stat, cand, acc, counters, total;
  int acc_cap;
};

static FeatLayout feat_layout(int batch, int H, int W, double min_distance) {
  FeatLayout L;
  const size_t px = (size_t)batch * H * W;
  long long cap = (long long)H * W;
  if (min_distance >= 1.0) {
    // circles of radius d/2 around kept corners are disjoint and lie inside the image grown by d/2
    const double d = min_distance;
    const double bound = ((W + d) * (H + d)) / (0.8660254 * d * d) + 64.0;
    if (bound < (double)cap) cap = (long long)bound;
  }
  L.acc_cap = (int)cap;
  size_t o = 0;
  L.gm = o; o = align256(o + px);
  L.gb = o; o = align256(o + px);
  L.stat = o; o = align256(o + px);
  L.cand = o; o = align256(o + px * sizeof(int));
  L.acc = o; o = align256(o + (size_t)batch * L.acc_cap * sizeof(unsigned long long));
  L.counters = o; o = align256(o + (size_t)batch * 4 * sizeof(int));
  L.total = o;
  return L;
}

}  // namespace vslam

extern "C" long long vslam_features_workspace_bytes(int batch, int H, int W, double min_distance) {
  if (batch < 1 || H < 16 || W < 16 || (long long)H * W > (1ll << 30)) return -1;
  return (long long)vslam::feat_layout(batch, H, W, min_distance).total;
}

extern "C" int vslam_extract_features(const uint8_t* img, int batch, int H, int W, int max_corners, double quality,
                                      double min_distance, float cos_a, float sin_a, const int8_t* pattern,
                                      void* workspace, long long workspace_bytes, float* eig, uint8_t* blur,
                                      float* corners, int32_t* n_corners, float* kp, uint8_t* desc,
                                      int32_t* n_kp, int32_t* status, void* stream) {
  using namespace vslam;
  if (batch < 1 || H < 16 || W < 16 || max_corners < 1 || (long long)H * W > (1ll << 30)) return VSLAM_ERR_BAD_ARG;
  if (!img || !pattern || !workspace || !eig || !blur || !corners || !n_corners || !kp || !desc || !n_kp)
    return VSLAM_ERR_BAD_ARG;
  if (!(quality > 0.0) || min_distance < 0.0) return VSLAM_ERR_BAD_ARG;
  // cv2 looks for conflicts only in the 3x3 grid cells of size cvRound(d) around a corner, which equals the
  // plain distance test exactly when d is an integer (or < 1: no test at all)
  if (min_distance >= 1.0 && min_distance != (double)(int)min_distance) return VSLAM_ERR_BAD_ARG;
  const FeatLayout L = feat_layout(batch, H, W, min_distance);
  if (workspace_bytes < (long long)L.total) return VSLAM_ERR_BAD_ARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  uint8_t* ws = static_cast<uint8_t*>(workspace);
  uint8_t* gm = ws + L.gm;
  uint8_t* gb = ws + L.gb;
  uint8_t* stat = ws + L.stat;
  int* cand = reinterpret_cast<int*>(ws + L.cand);
  unsigned long long* acc = reinterpret_cast<unsigned long long*>(ws + L.acc);
  int* counters = reinterpret_cast<int*>(ws + L.counters);
  const long long n_px = (long long)batch * H * W;
  const int px = H * W;

  const long long gthreads = (n_px + 3) / 4 > batch ? (n_px + 3) / 4 : batch;
  feat_gray_kernel<<<(unsigned)((gthreads + 255) / 256), 256, 0, st>>>(img, n_px, px, gm, gb, counters, batch);
  feat_eig_kernel<<<dim3((W + kEigT - 3) / (kEigT - 2), batch), kEigT, 0, st>>>(gm, H, W, eig, counters);
  feat_blur_kernel<<<dim3((W + kBlurW - 1) / kBlurW, (H + kBlurH - 1) / kBlurH, batch), 256, 0, st>>>(gb, H, W, blur);
  feat_candidates_kernel<<<dim3((px + 255) / 256, batch), 256, 0, st>>>(eig, H, W, quality, stat, cand, counters);
  int radius = 0, d2 = 0;
  if (min_distance >= 1.0) {
    d2 = (int)(min_distance * min_distance);
    while ((radius + 1) * (radius + 1) < d2) ++radius;
  }
  feat_suppress_kernel<<<dim3(kSupCtas, batch), kSupThreads, 0, st>>>(
      eig, stat, cand, H, W, radius, d2, acc, L.acc_cap, max_corners, reinterpret_cast<float2*>(corners),
      n_corners, counters);
  feat_border_compact_kernel<<<batch, 1024, 0, st>>>(reinterpret_cast<const float2*>(corners), n_corners,
                                                     max_corners, H, W, reinterpret_cast<float2*>(kp), n_kp);
  feat_orb_kernel<<<dim3((max_corners + kOrbWarps * 4 - 1) / (kOrbWarps * 4), batch), kOrbWarps * 32, 0, st>>>(
      blur, H, W, reinterpret_cast<const float2*>(kp), n_kp, max_corners, pattern, cos_a, sin_a, desc);
  if (status)
    VSLAM_CUDA_TRY(cudaMemcpy2DAsync(status, sizeof(int32_t), counters + 3, 4 * sizeof(int), sizeof(int32_t), batch,
                                     cudaMemcpyDeviceToDevice, st));
  return check_launch("vslam_extract_features");
}
