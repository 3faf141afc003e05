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
#include "tuning.h"

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
//   Sobel d/dx: rows (-1,0,1) exact; columns (s,2s,s): fma(up + down, s1, mid * s2)
//   Sobel d/dy: rows (s,2s,s): fma(s1, right, fma(s2, mid, s1 * left)); columns (-1,0,1) exact
//   cov = (dx*dx, dx*dy, dy*dy) float; box rows: (l + m) + r in double; box columns: running double sum
//   eig = (a + c) - sqrt((a - c)^2 + b^2), a = A/2, c = C/2, float, not contracted.
// The box filter's column sum is a RUNNING float64 sum (sum += new row; out = float(sum); sum -= old row), so the
// rows of one column cannot be split -- but everything up to the row sums can.  Warp-specialised: a CTA owns a
// strip of 62 columns; fourteen PRODUCER warps take the rows round-robin and turn them into float64 row sums in a
// 14-slot shared-memory ring (full/empty mbarriers per slot); two CONSUMER warps (one thread per column) walk
// down the frame doing ONLY the running sum (six dependent float64 adds per row) and leave the column sums in the
// slot; the slot's producer turns them into eigenvalues before it reuses the slot.
constexpr int kEigT = 64;       // strip width including one halo column either side
constexpr int kEigProd = 14;    // producer warps
constexpr int kEigRing = 14;    // ring slots (a multiple of kEigProd)
constexpr int kEigThreads = (2 + kEigProd) * 32;

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// eig = (a + c) - sqrt((a - c)^2 + b^2) with a = A/2, c = C/2 from the three float64 column sums of one pixel
__device__ __forceinline__ float min_eig_from_sums(double sA, double sB, double sC) {
  const float A = (float)sA, B = (float)sB, C = (float)sC;
  const float a = __fmul_rn(A, 0.5f), cq = __fmul_rn(C, 0.5f);
  float t = __fsub_rn(a, cq);
  t = __fadd_rn(__fmul_rn(t, t), __fmul_rn(B, B));
  return __fsub_rn(__fadd_rn(a, cq), __fsqrt_rn(t));
}

__global__ void __launch_bounds__(kEigThreads)
feat_eig_kernel(const uint8_t* __restrict__ gray, int H, int W, float* __restrict__ eig, int* __restrict__ counters) {
  __shared__ double ring[kEigRing][3][kEigT];
  __shared__ double prod[kEigProd][3][kEigT];  // float products, widened once
  __shared__ uint64_t full[kEigRing], empty[kEigRing];
  const int b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int c0 = blockIdx.x * (kEigT - 2) - 1;  // image column of strip index 0
  const uint8_t* g = gray + (size_t)b * H * W;
  if (tid == 0) {
    for (int s = 0; s < kEigRing; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 2); }
    mbar_fence_init();
  }
  __syncthreads();

  if (wid >= 2) {
    // ------------------------------------------------------------------ producers: row -> row sums
    const int w = wid - 2;
    const float s1 = __int_as_float(0x39ab5601), s2 = __int_as_float(0x3a2b5601);  // float(1/3060), float(2/3060)
    int cc[2], xm[2], xp[2];
    bool vec_col[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      cc[h] = reflect101(min(c0 + h * 32 + lane, W), W);
      xm[h] = reflect101(cc[h] - 1, W);
      xp[h] = reflect101(cc[h] + 1, W);
      // OpenCV's uint8 -> float32 row filter fuses in its 32-pixel vector body and rounds every product in the
      // scalar tail (oracle/features.py: row_vector_width)
      vec_col[h] = cc[h] < (W / 32) * 32;
    }
    float* out = eig + (size_t)b * H * W;
    float vmax = -INFINITY;
    // the consumers leave the float64 column sums of a row in its ring slot; the slot's producer turns them into
    // eigenvalues (output row = ring row - 1) before it reuses the slot, so the serial consumer chain stays short
    auto finish = [&](int slot, int row) {
      if (row < 1) return;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int i = h * 32 + lane, c = c0 + i;
        if (i >= 1 && i <= kEigT - 2 && c < W) {
          const float e = min_eig_from_sums(ring[slot][0][i], ring[slot][1][i], ring[slot][2][i]);
          out[(size_t)(row - 1) * W + c] = e;
          vmax = fmaxf(vmax, e);
        }
      }
    };
    int last_row = -1;
    for (int r = w; r < H; r += kEigProd) {
      const int slot = r % kEigRing, use = r / kEigRing;
      const uint8_t* gm_ = g + (size_t)reflect101(r - 1, H) * W;
      const uint8_t* gc_ = g + (size_t)r * W;
      const uint8_t* gp_ = g + (size_t)reflect101(r + 1, H) * W;
      float p[2][3];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const float ul = (float)__ldg(gm_ + xm[h]), um = (float)__ldg(gm_ + cc[h]), ur = (float)__ldg(gm_ + xp[h]);
        const float ml = (float)__ldg(gc_ + xm[h]), mr = (float)__ldg(gc_ + xp[h]);
        const float dl = (float)__ldg(gp_ + xm[h]), dm = (float)__ldg(gp_ + cc[h]), dr = (float)__ldg(gp_ + xp[h]);
        const float rx_u = __fsub_rn(ur, ul), rx_m = __fsub_rn(mr, ml), rx_d = __fsub_rn(dr, dl);
        float R_u, R_d;
        if (vec_col[h]) {
          R_u = __fmaf_rn(s1, ur, __fmaf_rn(s2, um, __fmul_rn(s1, ul)));
          R_d = __fmaf_rn(s1, dr, __fmaf_rn(s2, dm, __fmul_rn(s1, dl)));
        } else {
          R_u = __fadd_rn(__fadd_rn(__fmul_rn(s1, ul), __fmul_rn(s2, um)), __fmul_rn(s1, ur));
          R_d = __fadd_rn(__fadd_rn(__fmul_rn(s1, dl), __fmul_rn(s2, dm)), __fmul_rn(s1, dr));
        }
        const float dx = __fmaf_rn(__fadd_rn(rx_u, rx_d), s1, __fmul_rn(rx_m, s2));
        const float dy = __fsub_rn(R_d, R_u);
        p[h][0] = __fmul_rn(dx, dx); p[h][1] = __fmul_rn(dx, dy); p[h][2] = __fmul_rn(dy, dy);
      }
      mbar_wait(&empty[slot], (uint32_t)((use & 1) ^ 1));  // the consumers are done with the slot's previous row
      if (use > 0) finish(slot, r - kEigRing);
      __syncwarp();
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int k = 0; k < 3; ++k) prod[w][k][h * 32 + lane] = (double)p[h][k];
      __syncwarp();
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int i = h * 32 + lane;
        if (i >= 1 && i <= kEigT - 2) {
#pragma unroll
          for (int k = 0; k < 3; ++k)
            ring[slot][k][i] = (prod[w][k][i - 1] + prod[w][k][i]) + prod[w][k][i + 1];
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&full[slot]);
      last_row = r;
    }
    if (last_row >= 0) {  // the last row of this warp's slot
      mbar_wait(&empty[last_row % kEigRing], (uint32_t)((last_row / kEigRing) & 1));
      finish(last_row % kEigRing, last_row);
    }
    int pkey = float_key(vmax);
    pkey = __reduce_max_sync(0xffffffffu, pkey);
    if (lane == 0) atomicMax(&counters[4 * b + 0], pkey);
    return;
  }

  // -------------------------------------------------------------------- consumers: the running column sum
  const int c = c0 + tid;
  const bool owner = tid >= 1 && tid <= kEigT - 2 && c < W;
  float* out = eig + (size_t)b * H * W;
  double rsa[3] = {0, 0, 0}, rsb[3] = {0, 0, 0}, sum[3] = {0, 0, 0};
  float vmax = -INFINITY;
  for (int r = 0; r < H; ++r) {
    const int slot = r % kEigRing, use = r / kEigRing;
    mbar_wait(&full[slot], (uint32_t)(use & 1));
    if (owner) {
      double rsc[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) rsc[k] = ring[slot][k][tid];
      if (r == 1) {
#pragma unroll
        for (int k = 0; k < 3; ++k) sum[k] = (0.0 + rsc[k]) + rsb[k];  // padded rows -1 (= row 1) and 0
      }
      if (r >= 1) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const double s0 = sum[k] + rsc[k];
          sum[k] = s0 - (r == 1 ? rsc[k] : rsa[k]);  // minus row y-1 (row -1 reflects to row 1)
          ring[slot][k][tid] = s0;                    // column sums of output row r-1, finished by the producer
        }
      }
#pragma unroll
      for (int k = 0; k < 3; ++k) { rsa[k] = rsb[k]; rsb[k] = rsc[k]; }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[slot]);
  }
  if (owner) {  // last row: row H reflects to row H-2, which is rsa after the final shift
    double s0[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) s0[k] = sum[k] + rsa[k];
    const float e = min_eig_from_sums(s0[0], s0[1], s0[2]);
    out[(size_t)(H - 1) * W + c] = e;
    vmax = e;
  }
  int key = owner ? float_key(vmax) : INT_MIN;
  key = __reduce_max_sync(0xffffffffu, key);
  if (lane == 0) atomicMax(&counters[4 * b + 0], key);
}

// Throughput variant of K2 for large batches (enough strips to fill the GPU on their own): one thread per column
// does everything for its column -- Sobel in a three-row register window, the three products exchanged through
// double-buffered shared memory (one barrier per row), input rows fetched eight rows ahead.  Fewer instructions
// per row than the warp-specialised kernel (no ring, no mbarrier waits), but a ~6x longer serial chain.
constexpr int kEigChunk = 8;

__global__ void __launch_bounds__(kEigT)
feat_eig_walk_kernel(const uint8_t* __restrict__ gray, int H, int W, float* __restrict__ eig, int* __restrict__ counters) {
  __shared__ double sm[2][3][kEigT];  // the float products, already widened (one conversion per value, not three)
  const int b = blockIdx.y, tid = threadIdx.x;
  const int c = blockIdx.x * (kEigT - 2) + tid - 1;
  const bool owner = tid >= 1 && tid <= kEigT - 2 && c < W;
  const int cc = reflect101(min(c, W), W);
  const int xm = reflect101(cc - 1, W), xp = reflect101(cc + 1, W);
  const uint8_t* g = gray + (size_t)b * H * W;
  float* out = eig + (size_t)b * H * W;
  const float s1 = __int_as_float(0x39ab5601), s2 = __int_as_float(0x3a2b5601);
  const bool vec_col = cc < (W / 32) * 32;

  auto terms = [&](unsigned lb, unsigned mb, unsigned rb, float& rx, float& R) {
    const float l = (float)lb, m = (float)mb, r = (float)rb;
    rx = __fsub_rn(r, l);
    // OpenCV's uint8 -> float32 row filter fuses in its 32-pixel vector body and rounds every product in the
    // scalar tail (oracle/features.py: row_vector_width)
    R = vec_col ? __fmaf_rn(s1, r, __fmaf_rn(s2, m, __fmul_rn(s1, l)))
                : __fadd_rn(__fadd_rn(__fmul_rn(s1, l), __fmul_rn(s2, m)), __fmul_rn(s1, r));
  };
  auto row_terms = [&](int q, float& rx, float& R) {
    const uint8_t* gr = g + (size_t)q * W;
    terms(__ldg(gr + xm), __ldg(gr + cc), __ldg(gr + xp), rx, R);
  };
  // the rows of the NEXT chunk of kEigChunk iterations are fetched while the current chunk is worked on
  unsigned cur[kEigChunk][3], nxt[kEigChunk][3];
  auto load_chunk = [&](int r0, unsigned (&buf)[kEigChunk][3]) {
#pragma unroll
    for (int j = 0; j < kEigChunk; ++j) {
      const uint8_t* gr = g + (size_t)reflect101(min(r0 + j + 2, H + 1), H) * W;
      buf[j][0] = __ldg(gr + xm); buf[j][1] = __ldg(gr + cc); buf[j][2] = __ldg(gr + xp);
    }
  };
  float rx0, R0, rx1, R1, rx2, R2;
  row_terms(1, rx0, R0);  // row -1 reflects to row 1
  row_terms(0, rx1, R1);
  rx2 = rx0; R2 = R0;
  load_chunk(0, cur);
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

  for (int r0 = 0; r0 < H; r0 += kEigChunk) {
   if (r0 + kEigChunk < H) load_chunk(r0 + kEigChunk, nxt);
#pragma unroll
   for (int j = 0; j < kEigChunk; ++j) {
    const int r = r0 + j;
    if (r >= H) break;
    float rxn, Rn;
    terms(cur[j][0], cur[j][1], cur[j][2], rxn, Rn);  // row r + 2
    const float dx = __fmaf_rn(__fadd_rn(rx0, rx2), s1, __fmul_rn(rx1, s2));
    const float dy = __fsub_rn(R2, R0);
    const int buf = r & 1;
    const double pxx = (double)__fmul_rn(dx, dx), pxy = (double)__fmul_rn(dx, dy), pyy = (double)__fmul_rn(dy, dy);
    sm[buf][0][tid] = pxx;
    sm[buf][1][tid] = pxy;
    sm[buf][2][tid] = pyy;
    __syncthreads();
    if (owner) {
      double rsc[3];
#pragma unroll
      for (int k = 0; k < 3; ++k)
        rsc[k] = (sm[buf][k][tid - 1] + (k == 0 ? pxx : k == 1 ? pxy : pyy)) + sm[buf][k][tid + 1];
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
#pragma unroll
   for (int j = 0; j < kEigChunk; ++j) { cur[j][0] = nxt[j][0]; cur[j][1] = nxt[j][1]; cur[j][2] = nxt[j][2]; }
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

// ---- K3: goodFeaturesToTrack's selection, one thread-block cluster per image ----------------------
// cv2: threshold the response at float(max * quality), keep strict-or-equal 3x3 local maxima of interior
// pixels (dilate), walk them in order (value desc, address desc) and keep one iff no kept corner of higher
// order lies at squared distance < d2.  Equivalent fixed point: a candidate is KEPT once every higher-order
// candidate within range is DROPPED, and DROPPED once one of them is KEPT; decisions are final, so the rounds
// may read each other's writes.  A kept neighbour within range is necessarily of higher order (it could not
// have been kept while this one was undecided otherwise), so only undecided neighbours need a key compare.
// The status image lives in the SHARED MEMORY OF THE CLUSTER: CTA k owns a band of H/8 rows (one byte per
// pixel: 0 none/dropped, 1 undecided, 2 kept); windows that cross a band edge read the neighbour CTA's band
// through distributed shared memory; the termination vote is a DSMEM read of eight flags.
constexpr int kSupCtas = 8;
constexpr int kSupThreads = 512;        // large batches: four CTAs per SM, every cluster of the batch resident
constexpr int kSupThreadsSmall = 1024;  // small batches: the SMs are idle anyway, halve the passes per round

template <int THREADS>
__global__ void __cluster_dims__(kSupCtas, 1, 1) __launch_bounds__(THREADS, 2048 / THREADS)
feat_select_kernel(const float* __restrict__ eig, int H, int W, int Hb, int Wp, double quality, int radius, int d2,
                   int* __restrict__ lists, unsigned long long* acc_key, int acc_cap, int* counters) {
  cg::cluster_group cluster = cg::this_cluster();
  extern __shared__ __align__(16) uint8_t band[];  // [Hb][Wp]
  __shared__ int slot[2];
  __shared__ int n_list[2];
  const int b = blockIdx.y;
  const int rank = (int)cluster.block_rank();
  const float* e = eig + (size_t)b * H * W;
  const unsigned* ebits = reinterpret_cast<const unsigned*>(e);
  unsigned long long* acc = acc_key + (size_t)b * acc_cap;
  const int y_lo = rank * Hb, y_hi = min(H, y_lo + Hb);
  const float thr = (float)((double)key_float(counters[4 * b + 0]) * quality);
  // two work lists of this CTA (undecided candidates of its band, as band offsets r * Wp + x), ping-pong
  int* list[2];
  list[0] = lists + ((size_t)(b * kSupCtas + rank) * 2) * Hb * Wp;
  list[1] = list[0] + (size_t)Hb * Wp;
  if (threadIdx.x < 2) n_list[threadIdx.x] = 0;
  __syncthreads();

  // candidates of this band: ev = v > thr ? v : 0 (THRESH_TOZERO); candidate iff ev != 0 and ev == dilate3x3(ev')
  // <=> v > thr, v > 0 and v >= every neighbour (a neighbour <= thr cannot exceed v).  Four pixels per thread.
  unsigned* bandw = reinterpret_cast<unsigned*>(band);
  const bool vec = (W & 3) == 0;
  const int Wq = Wp / 4;
  for (int r = threadIdx.x >> 5; r < Hb; r += THREADS / 32)
   for (int xq = threadIdx.x & 31; xq < Wq; xq += 32) {
    const int x = xq * 4, y = y_lo + r;
    unsigned word = 0;
    if (y < y_hi && y >= 1 && y < H - 1 && x < W) {
      float v4[4];
      if (vec) {
        const float4 t = *reinterpret_cast<const float4*>(e + (size_t)y * W + x);
        v4[0] = t.x; v4[1] = t.y; v4[2] = t.z; v4[3] = t.w;
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) v4[k] = x + k < W ? e[(size_t)y * W + x + k] : 0.0f;
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int xx = x + k;
        const float v = v4[k];
        bool maybe = v > thr && v > 0.0f && xx >= 1 && xx < W - 1;
        if (k > 0) maybe = maybe && v4[k - 1] <= v;       // horizontal neighbours already in registers
        if (k < 3) maybe = maybe && v4[k + 1] <= v;
        if (maybe) {
          float mx = v;
#pragma unroll
          for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
            for (int dx = -1; dx <= 1; ++dx) mx = fmaxf(mx, e[(size_t)(y + dy) * W + xx + dx]);
          if (mx == v) word |= 1u << (8 * k);
        }
      }
    }
    bandw[r * Wq + xq] = word;
    if (word) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if ((word >> (8 * k)) & 1u) list[0][atomicAdd(&n_list[0], 1)] = (r << 16) | (x + k);
    }
   }
  cluster.sync();

  for (int round = 0;; ++round) {
    const int src = round & 1, dst = src ^ 1;
    const int n = n_list[src];
    const int* in = list[src];
    int* outl = list[dst];
    for (int i0 = 0; i0 < n; i0 += THREADS) {   // uniform trip count: the ballots below need whole warps
      const int i = i0 + threadIdx.x;
      bool wait = false;
      int idx = 0;
      if (i < n) {
        idx = in[i];
        const int r = idx >> 16, x = idx & 0xffff, y = y_lo + r, pix = y * W + x;
        const unsigned v = ebits[pix];
        bool drop = false;
        const int yy0 = max(0, y - radius), yy1 = min(H - 1, y + radius);
        const int xa = max(0, x - radius) & ~15, xe = min(Wp, x + radius + 1);
        int owner = rank, rr = r - (y - yy0);   // band and row inside it of the first window row, no division
        while (rr < 0) { --owner; rr += Hb; }
        for (int yy = yy0; yy <= yy1 && !drop; ++yy, ++rr) {
          if (rr == Hb) { ++owner; rr = 0; }
          const uint8_t* row = (owner == rank ? band : cluster.map_shared_rank(band, owner)) + rr * Wp;
          const int dy2 = (yy - y) * (yy - y);
          for (int xw = xa; xw < xe && !drop; xw += 16) {
            const uint4 w16 = *reinterpret_cast<const uint4*>(row + xw);
            if (!(w16.x | w16.y | w16.z | w16.w)) continue;
            const unsigned wv[4] = {w16.x, w16.y, w16.z, w16.w};
#pragma unroll
            for (int jw = 0; jw < 4; ++jw) {
              unsigned m = wv[jw] & 0x03030303u;   // visit only the non-zero status bytes
              while (m) {
                const int bit = __ffs(m) - 1, j = bit >> 3;
                const unsigned st = (m >> (8 * j)) & 3u;
                m &= ~(0xffu << (8 * j));
                const int xx = xw + 4 * jw + j;
                if (dy2 + (xx - x) * (xx - x) >= d2 || (yy == y && xx == x)) continue;
                if (st == 2u) { drop = true; break; }
                const int q = yy * W + xx;
                const unsigned vq = ebits[q];
                if (vq > v || (vq == v && q > pix)) wait = true;
              }
              if (drop) break;
            }
          }
        }
        if (drop) {
          band[r * Wp + x] = 0;
          wait = false;
        } else if (!wait) {
          band[r * Wp + x] = 2;
          const int pos = atomicAdd(&counters[4 * b + 2], 1);
          if (pos < acc_cap) acc[pos] = ((unsigned long long)v << 32) | (unsigned)pix;
        }
      }
      const unsigned m = __ballot_sync(0xffffffffu, wait);
      if (m) {
        const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
        int base = 0;
        if (lane == leader) base = atomicAdd(&n_list[dst], __popc(m));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (wait) outl[base + __popc(m & ((1u << lane) - 1))] = idx;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) { slot[round & 1] = n_list[dst] > 0; n_list[src] = 0; }
    cluster.sync();
    int total = 0;
#pragma unroll
    for (int k = 0; k < kSupCtas; ++k) total |= *cluster.map_shared_rank(&slot[round & 1], k);
    if (!total) break;
  }
  cluster.sync();  // no CTA leaves while its shared memory may still be read
}

// Order of the kept corners: descending (response, address) = descending 64-bit key.  One CTA per image, bitonic
// sort in shared memory (the list holds at most ~7.3k corners at VGA); first max_corners are the result.
constexpr int kSortMax = 8192;
constexpr int kSortThreads = 1024;

__global__ void __launch_bounds__(kSortThreads)
feat_order_kernel(const unsigned long long* __restrict__ acc_key, int acc_cap, int W, int max_corners,
                  float2* __restrict__ corners, int* __restrict__ n_corners, int* __restrict__ counters) {
  extern __shared__ __align__(16) unsigned long long keys[];
  const int b = blockIdx.x;
  const unsigned long long* acc = acc_key + (size_t)b * acc_cap;
  const int n_raw = counters[4 * b + 2];
  const int A = min(n_raw, acc_cap);
  if (threadIdx.x == 0) {
    if (n_raw > acc_cap) counters[4 * b + 3] = 1;
    n_corners[b] = min(A, max_corners);
  }
  float2* out = corners + (size_t)b * max_corners;
  if (A > kSortMax) {  // (only reachable without a minimum distance on large frames) counting rank
    for (int i = threadIdx.x; i < A; i += kSortThreads) {
      const unsigned long long key = acc[i];
      int higher = 0;
      for (int j = 0; j < A; ++j) higher += acc[j] > key;
      if (higher < max_corners) {
        const int pix = (int)(key & 0xffffffffu), y = pix / W;
        out[higher] = make_float2((float)(pix - y * W), (float)y);
      }
    }
    return;
  }
  int N = 32;
  while (N < A) N <<= 1;
  for (int i = threadIdx.x; i < N; i += kSortThreads) keys[i] = i < A ? acc[i] : 0ull;
  __syncthreads();
  // Bitonic network, descending.  Every warp owns a contiguous chunk; the stages whose partner distance j stays
  // inside a chunk need only __syncwarp, so a block barrier is paid for 15 of the 78 stages at N = 4096.
  const int chunk = max(N / (kSortThreads / 32), 64);
  const int lane = threadIdx.x & 31, base = (threadIdx.x >> 5) * chunk, cw = min(chunk, N);
  auto cmpex = [&](int lo, int hi, int k) {
    const unsigned long long a = keys[lo], c = keys[hi];
    if ((a < c) == ((lo & k) == 0)) { keys[lo] = c; keys[hi] = a; }   // descending blocks first
  };
  for (int k = 2; k <= N; k <<= 1) {
    int j = k >> 1;
    for (; j >= chunk; j >>= 1) {          // partners in different chunks: whole block
      for (int t = threadIdx.x; t < N / 2; t += kSortThreads)
        cmpex(((t & ~(j - 1)) << 1) | (t & (j - 1)), (((t & ~(j - 1)) << 1) | (t & (j - 1))) | j, k);
      __syncthreads();
    }
    if (base < N) {
      for (; j > 0; j >>= 1) {             // partners inside this warp's chunk
        for (int t = lane; t < cw / 2; t += 32) {
          const int lo = base + (((t & ~(j - 1)) << 1) | (t & (j - 1)));
          cmpex(lo, lo | j, k);
        }
        __syncwarp();
      }
    }
    if ((k << 1) > chunk) __syncthreads();   // the next k starts with (or is followed by) a cross-chunk stage
  }
  const int n_out = min(A, max_corners);
  for (int i = threadIdx.x; i < n_out; i += kSortThreads) {
    const int pix = (int)(keys[i] & 0xffffffffu), y = pix / W;
    out[i] = make_float2((float)(pix - y * W), (float)y);
  }
}

// Latency variant of the ordering for small batches (the SMs are idle anyway): rank by counting, one key per thread,
// 64 threads per CTA so that one frame's ~3.5k kept corners spread over ~56 SMs; all keys stream through 8 KB
// tiles of shared memory (warp-wide broadcast reads).
constexpr int kCountThreads = 64;
constexpr int kCountTile = 1024;

__global__ void __launch_bounds__(kCountThreads)
feat_order_count_kernel(const unsigned long long* __restrict__ acc_key, int acc_cap, int W, int max_corners,
                        float2* __restrict__ corners, int* __restrict__ n_corners, int* __restrict__ counters) {
  __shared__ unsigned long long tile[kCountTile];
  const int b = blockIdx.y;
  const unsigned long long* acc = acc_key + (size_t)b * acc_cap;
  const int n_raw = counters[4 * b + 2];
  const int A = min(n_raw, acc_cap);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    if (n_raw > acc_cap) counters[4 * b + 3] = 1;
    n_corners[b] = min(A, max_corners);
  }
  if ((int)blockIdx.x * kCountThreads >= A) return;
  const int i = blockIdx.x * kCountThreads + threadIdx.x;
  const unsigned long long key = i < A ? acc[i] : 0ull;
  int h0 = 0, h1 = 0, h2 = 0, h3 = 0;
  for (int t0 = 0; t0 < A; t0 += kCountTile) {
    const int m = min(kCountTile, A - t0);
    __syncthreads();
    for (int j = threadIdx.x; j < kCountTile; j += kCountThreads) tile[j] = j < m ? acc[t0 + j] : 0ull;
    __syncthreads();
#pragma unroll 4
    for (int j = 0; j < kCountTile; j += 4) {   // padding keys are 0: never greater than a real key
      h0 += tile[j] > key; h1 += tile[j + 1] > key; h2 += tile[j + 2] > key; h3 += tile[j + 3] > key;
    }
  }
  const int higher = (h0 + h1) + (h2 + h3);
  if (i < A && higher < max_corners) {
    const int pix = (int)(key & 0xffffffffu), y = pix / W;
    corners[(size_t)b * max_corners + higher] = make_float2((float)(pix - y * W), (float)y);
  }
}

// ---- K4: the blur ORB applies to its level image ---------------------------------------------------
// float32 separable 7x7, sigma 2 (taps = cv2.getGaussianKernel(7, 2, CV_32F)), REFLECT_101;
// rows: s = k0*p0, s = fma(k_i, p_i, s); columns: s = k3*c, s = fma(k_{3+j}, up_j + down_j, s);
// saturate_cast<uchar> = round-half-even.
constexpr int kBlurW = 64, kBlurH = 32;

__global__ void __launch_bounds__(256)
feat_blur_kernel(const uint8_t* __restrict__ gray, int H, int W, uint8_t* __restrict__ blur) {
  __shared__ __align__(16) float in[kBlurH + 6][kBlurW + 8];
  __shared__ __align__(16) float hz[kBlurH + 6][kBlurW];
  const float k0 = __int_as_float(0x3d8fafb1), k1 = __int_as_float(0x3e06387e), k2 = __int_as_float(0x3e434a39),
              k3 = __int_as_float(0x3e5d4ae0);
  const float kk[7] = {k0, k1, k2, k3, k2, k1, k0};
  const int b = blockIdx.z, x0 = blockIdx.x * kBlurW, y0 = blockIdx.y * kBlurH;
  const uint8_t* g = gray + (size_t)b * H * W;
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
  {  // tile + 3-pixel apron, uint8 -> float
    const int xa = reflect101(min(x0 + tx - 3, W + 2), W), xb = reflect101(min(x0 + tx + 61, W + 2), W);
    for (int r = ty; r < kBlurH + 6; r += 4) {
      const uint8_t* gr = g + (size_t)reflect101(min(y0 + r - 3, H + 2), H) * W;
      in[r][tx] = (float)__ldg(gr + xa);
      if (tx < 6) in[r][64 + tx] = (float)__ldg(gr + xb);
    }
  }
  __syncthreads();
  {  // rows: four adjacent outputs per thread from twelve inputs (three 128-bit shared loads)
    const int q = threadIdx.x & 15;
    const bool fused = x0 + 4 * q < (W / 32) * 32;  // vector body of OpenCV's row filter; its scalar tail rounds products
    for (int r = threadIdx.x >> 4; r < kBlurH + 6; r += 16) {
      const float4 a = *reinterpret_cast<const float4*>(&in[r][4 * q]);
      const float4 c = *reinterpret_cast<const float4*>(&in[r][4 * q + 4]);
      const float4 d = *reinterpret_cast<const float4*>(&in[r][4 * q + 8]);
      const float v[12] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w, d.x, d.y, d.z, d.w};
      float o[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float s = __fmul_rn(kk[0], v[e]);
        if (fused) {
#pragma unroll
          for (int t = 1; t < 7; ++t) s = __fmaf_rn(kk[t], v[e + t], s);
        } else {
#pragma unroll
          for (int t = 1; t < 7; ++t) s = __fadd_rn(s, __fmul_rn(kk[t], v[e + t]));
        }
        o[e] = s;
      }
      *reinterpret_cast<float4*>(&hz[r][4 * q]) = make_float4(o[0], o[1], o[2], o[3]);
    }
  }
  __syncthreads();
  {  // columns: eight vertically adjacent outputs per thread from fourteen inputs
    float v[14];
#pragma unroll
    for (int j = 0; j < 14; ++j) v[j] = hz[ty * 8 + j][tx];
    const int x = x0 + tx;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int y = y0 + ty * 8 + e;
      float s = __fmul_rn(kk[3], v[e + 3]);
#pragma unroll
      for (int j = 1; j <= 3; ++j) s = __fmaf_rn(kk[3 + j], __fadd_rn(v[e + 3 + j], v[e + 3 - j]), s);
      const int o = __float2int_rn(s);
      if (y < H && x < W) blur[(size_t)b * H * W + (size_t)y * W + x] = (uint8_t)min(255, max(0, o));
    }
  }
}

// ---- K5a: KeyPointsFilter::runByImageBorder(31), order kept ---------------------------------------
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

// ---- K5b: rBRIEF, WTA_K = 2 -----------------------------------------------------------------------
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

// the status band of one CTA must fit its shared memory (8 CTAs per image: frames up to ~1.6 Mpixel)
constexpr size_t kMaxBandBytes = 200 * 1024;

static int g_eig_variant = 0;  // 0 auto, 1 warp-specialised, 2 column walk (vslam_features_set_variant; tests, tuning)

static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

struct FeatLayout {
  size_t gm, gb, lists, acc, counters, total;
  int acc_cap, Hb, Wp;
  size_t band_bytes;
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
  L.Hb = (H + kSupCtas - 1) / kSupCtas;
  L.Wp = (W + 15) & ~15;
  L.band_bytes = (size_t)L.Hb * L.Wp;
  L.lists = o; o = align256(o + (size_t)batch * kSupCtas * 2 * L.band_bytes * sizeof(int));
  L.acc = o; o = align256(o + (size_t)batch * L.acc_cap * sizeof(unsigned long long));
  L.counters = o; o = align256(o + (size_t)batch * 4 * sizeof(int));
  L.total = o;
  return L;
}

}  // namespace vslam

extern "C" int vslam_features_set_variant(int variant) {
  if (variant < 0 || variant > 2) return VSLAM_ERR_BAD_ARG;
  vslam::g_eig_variant = variant;
  return VSLAM_OK;
}

extern "C" long long vslam_features_workspace_bytes(int batch, int H, int W, double min_distance) {
  if (batch < 1 || H < 16 || W < 16 || (long long)H * W > (1ll << 30)) return -1;
  if (vslam::feat_layout(batch, H, W, min_distance).band_bytes > vslam::kMaxBandBytes) return -1;
  return (long long)vslam::feat_layout(batch, H, W, min_distance).total;
}

extern "C" int vslam_extract_features(const uint8_t* img, int batch, int H, int W, int max_corners, double quality,
                                      double min_distance, float cos_a, float sin_a, const int8_t* pattern,
                                      void* workspace, long long workspace_bytes, float* eig, uint8_t* blur,
                                      float* corners, int32_t* n_corners, float* kp, uint8_t* desc,
                                      int32_t* n_kp, int32_t* status, void* stream) {
  using namespace vslam;
  if (batch < 1 || H < 16 || W < 16 || W > 32768 || max_corners < 1 || (long long)H * W > (1ll << 30)) return VSLAM_ERR_BAD_ARG;
  if (!img || !pattern || !workspace || !eig || !blur || !corners || !n_corners || !kp || !desc || !n_kp)
    return VSLAM_ERR_BAD_ARG;
  if (!(quality > 0.0) || min_distance < 0.0) return VSLAM_ERR_BAD_ARG;
  // cv2 looks for conflicts only in the 3x3 grid cells of size cvRound(d) around a corner, which equals the
  // plain distance test exactly when d is an integer (or < 1: no test at all)
  if (min_distance >= 1.0 && min_distance != (double)(int)min_distance) return VSLAM_ERR_BAD_ARG;
  const FeatLayout L = feat_layout(batch, H, W, min_distance);
  if (workspace_bytes < (long long)L.total || L.band_bytes > kMaxBandBytes) return VSLAM_ERR_BAD_ARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  uint8_t* ws = static_cast<uint8_t*>(workspace);
  uint8_t* gm = ws + L.gm;
  uint8_t* gb = ws + L.gb;
  unsigned long long* acc = reinterpret_cast<unsigned long long*>(ws + L.acc);
  int* counters = reinterpret_cast<int*>(ws + L.counters);
  const long long n_px = (long long)batch * H * W;
  const int px = H * W;

  const long long gthreads = (n_px + 3) / 4 > batch ? (n_px + 3) / 4 : batch;
  feat_gray_kernel<<<(unsigned)((gthreads + 255) / 256), 256, 0, st>>>(img, n_px, px, gm, gb, counters, batch);
  const int strips = (W + kEigT - 3) / (kEigT - 2);
  const bool ws_variant = g_eig_variant == 1 || (g_eig_variant == 0 && (long long)strips * batch < 2ll * sm_count());
  if (ws_variant)                                     // latency variant: rows in parallel, short serial chain
    feat_eig_kernel<<<dim3(strips, batch), kEigThreads, 0, st>>>(gm, H, W, eig, counters);
  else                                                // throughput variant: one thread per column
    feat_eig_walk_kernel<<<dim3(strips, batch), kEigT, 0, st>>>(gm, H, W, eig, counters);
  feat_blur_kernel<<<dim3((W + kBlurW - 1) / kBlurW, (H + kBlurH - 1) / kBlurH, batch), 256, 0, st>>>(gb, H, W, blur);
  int radius = 0, d2 = 0;
  if (min_distance >= 1.0) {
    d2 = (int)(min_distance * min_distance);
    while ((radius + 1) * (radius + 1) < d2) ++radius;
  }
  // (set on every call: the attribute is per device and a process may drive several)
  const bool small_batch = (long long)batch * kSupCtas <= sm_count();
  auto select_small = feat_select_kernel<kSupThreadsSmall>;
  auto select_large = feat_select_kernel<kSupThreads>;
  if (L.band_bytes > 48 * 1024)
    VSLAM_CUDA_TRY(cudaFuncSetAttribute(small_batch ? select_small : select_large,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxBandBytes));
  if (small_batch)
    select_small<<<dim3(kSupCtas, batch), kSupThreadsSmall, L.band_bytes, st>>>(
        eig, H, W, L.Hb, L.Wp, quality, radius, d2, reinterpret_cast<int*>(ws + L.lists), acc, L.acc_cap, counters);
  else
    select_large<<<dim3(kSupCtas, batch), kSupThreads, L.band_bytes, st>>>(
        eig, H, W, L.Hb, L.Wp, quality, radius, d2, reinterpret_cast<int*>(ws + L.lists), acc, L.acc_cap, counters);
  if (batch <= 4) {   // measured: 27 us against 35 us for one frame, break-even at ~6 frames
    feat_order_count_kernel<<<dim3((L.acc_cap + kCountThreads - 1) / kCountThreads, batch), kCountThreads, 0, st>>>(
        acc, L.acc_cap, W, max_corners, reinterpret_cast<float2*>(corners), n_corners, counters);
  } else {
    VSLAM_CUDA_TRY(cudaFuncSetAttribute(feat_order_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)(kSortMax * sizeof(unsigned long long))));
    feat_order_kernel<<<batch, kSortThreads, kSortMax * sizeof(unsigned long long), st>>>(
        acc, L.acc_cap, W, max_corners, reinterpret_cast<float2*>(corners), n_corners, counters);
  }
  feat_border_compact_kernel<<<batch, 1024, 0, st>>>(reinterpret_cast<const float2*>(corners), n_corners,
                                                     max_corners, H, W, reinterpret_cast<float2*>(kp), n_kp);
  feat_orb_kernel<<<dim3((max_corners + kOrbWarps * 4 - 1) / (kOrbWarps * 4), batch), kOrbWarps * 32, 0, st>>>(
      blur, H, W, reinterpret_cast<const float2*>(kp), n_kp, max_corners, pattern, cos_a, sin_a, desc);
  if (status)
    VSLAM_CUDA_TRY(cudaMemcpy2DAsync(status, sizeof(int32_t), counters + 3, 4 * sizeof(int), sizeof(int32_t), batch,
                                     cudaMemcpyDeviceToDevice, st));
  return check_launch("vslam_extract_features");
}
