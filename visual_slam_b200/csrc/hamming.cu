// 256-bit Hamming brute-force kNN (k = 2) with the cross-check direction fused in.
//
// Replaces cv2.BFMatcher(NORM_HAMMING).knnMatch(k=2) (reference src/v2/frame.py:23)
// and the mutual-nearest test of src/slam.py:40-47.
//
// Layout / mapping (B200):
//   grid  = (query tiles, frame pairs); block = THREADS; each thread keeps QPT query
//           descriptors (8 x u32 each) and their running top-2 in registers.
//   train descriptors stream through shared memory in tiles of TT rows, moved by
//           1-D TMA bulk copies (cp.async.bulk -> UBLKCP) into a 2-stage mbarrier ring.
//   every lane of a warp reads the SAME train row (two LDS.128 broadcasts) and
//           compares it with its own QPT queries: no shuffles on the forward path.
//   (distance, trainIdx) is packed as distance<<22 | trainIdx, so the cv2 order --
//           ascending distance, lower trainIdx first on ties -- is a plain u32 min.
//   cross-check: per train row the warp reduces min(distance<<22 | queryIdx) with
//           REDUX.MIN and lane 0 folds it into a shared-memory slot; slots are merged
//           across query tiles with a global atomicMin.
//
// The popcount of the 8 XOR words uses a carry-save adder tree (Harley-Seal), which
// needs 4 POPC instead of 8: POPC is a quarter-rate pipe on sm_100, LOP3 is not.
#include "common.cuh"
#include "tuning.h"

namespace vslam {

constexpr int kKeyShift = 22;
constexpr uint32_t kIdxMask = (1u << kKeyShift) - 1u;
constexpr uint32_t kInvalidQ = 0x80000000u;  // bit 31: padded query lane, never wins

// The XOR and the two CSA LOP3s are written as PTX so that neither NVVM nor ptxas
// re-fuses the XORs into the adders (that duplicates them: 25 LOP3 instead of 16).
__device__ __forceinline__ uint32_t xor2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, 0, 0x3c;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ void csa(uint32_t& hi, uint32_t& lo, uint32_t a, uint32_t b,
                                    uint32_t c) {
  asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(lo) : "r"(a), "r"(b), "r"(c));  // a^b^c
  asm("lop3.b32 %0, %1, %2, %3, 0xe8;" : "=r"(hi) : "r"(a), "r"(b), "r"(c));  // majority
}

template <bool CSA>
__device__ __forceinline__ uint32_t hamming256(const uint32_t (&q)[8], const uint4& t0,
                                               const uint4& t1) {
  uint32_t x0 = xor2(q[0], t0.x), x1 = xor2(q[1], t0.y), x2 = xor2(q[2], t0.z),
           x3 = xor2(q[3], t0.w), x4 = xor2(q[4], t1.x), x5 = xor2(q[5], t1.y),
           x6 = xor2(q[6], t1.z), x7 = xor2(q[7], t1.w);
  if (CSA) {
    uint32_t c0, s0, c1, s1, c2, s2, c3, s3;
    csa(c0, s0, x0, x1, x2);
    csa(c1, s1, x3, x4, x5);
    csa(c2, s2, s0, s1, x6);
    csa(c3, s3, c0, c1, c2);
    return __popc(s2) + __popc(x7) + 2u * __popc(s3) + 4u * __popc(c3);
  } else {
    return __popc(x0) + __popc(x1) + __popc(x2) + __popc(x3) + __popc(x4) + __popc(x5) +
           __popc(x6) + __popc(x7);
  }
}

template <int THREADS, int QPT, int TT, bool CSA, bool REV>
__global__ void __launch_bounds__(THREADS)
hamming_knn2_kernel(const uint8_t* __restrict__ q, int nq_stride, const uint8_t* __restrict__ t, int nt_stride,
                    int32_t* __restrict__ idx, int32_t* __restrict__ dist,
                    uint32_t* __restrict__ rev_key, const int32_t* __restrict__ nq_arr,
                    const int32_t* __restrict__ nt_arr) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  auto tile = [&](int b) { return reinterpret_cast<uint4*>(smem_raw + (size_t)b * TT * 32); };
  uint32_t* rev_s = reinterpret_cast<uint32_t*>(smem_raw + (size_t)2 * TT * 32);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + (size_t)2 * TT * 32 + (REV ? TT * 4 : 0));

  const int tid = threadIdx.x;
  const int lane = tid & 31;
  const int pair = blockIdx.y;
  const int q_base = blockIdx.x * (THREADS * QPT);
  const uint8_t* qp = q + (size_t)pair * nq_stride * 32;
  const uint8_t* tp = t + (size_t)pair * nt_stride * 32;
  // ragged batches: rows [n, stride) of a pair are padding (per-pair counts from the feature extractor)
  const int nq = nq_arr ? min(max(nq_arr[pair], 0), nq_stride) : nq_stride;
  const int nt = nt_arr ? min(max(nt_arr[pair], 0), nt_stride) : nt_stride;

  if (q_base >= nq) {  // a CTA of padding queries only (ragged batch): nothing to match
#pragma unroll
    for (int j = 0; j < QPT; ++j) {
      const int qi = q_base + j * THREADS + tid;
      if (qi < nq_stride) {
        reinterpret_cast<int2*>(idx)[(size_t)pair * nq_stride + qi] = make_int2(-1, -1);
        reinterpret_cast<int2*>(dist)[(size_t)pair * nq_stride + qi] = make_int2(-1, -1);
      }
    }
    return;
  }
  const int n_tiles = (nt + TT - 1) / TT;
  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_fence_init();
  }
  if (REV) {
    for (int i = tid; i < TT; i += THREADS) rev_s[i] = 0xFFFFFFFFu;
  }
  __syncthreads();
  if (tid == 0 && n_tiles > 0) {
    uint32_t bytes = (uint32_t)min(TT, nt) * 32u;
    mbar_expect_tx(&bars[0], bytes);
    tma_bulk_g2s(tile(0), tp, bytes, &bars[0]);
  }

  // This thread's queries: rows q_base + j*THREADS + tid (coalesced across the warp).
  uint32_t qr[QPT][8];
  uint32_t qid[QPT];
  uint32_t b1[QPT], b2[QPT];
#pragma unroll
  for (int j = 0; j < QPT; ++j) {
    int qi = q_base + j * THREADS + tid;
    b1[j] = 0xFFFFFFFFu;
    b2[j] = 0xFFFFFFFFu;
    if (qi < nq) {
      const uint4* src = reinterpret_cast<const uint4*>(qp + (size_t)qi * 32);
      uint4 a = __ldg(src), b = __ldg(src + 1);
      qr[j][0] = a.x; qr[j][1] = a.y; qr[j][2] = a.z; qr[j][3] = a.w;
      qr[j][4] = b.x; qr[j][5] = b.y; qr[j][6] = b.z; qr[j][7] = b.w;
      qid[j] = (uint32_t)qi;
    } else {
#pragma unroll
      for (int w = 0; w < 8; ++w) qr[j][w] = 0u;
      qid[j] = kInvalidQ;
    }
  }

  for (int it = 0; it < n_tiles; ++it) {
    const int buf = it & 1;
    const int t0 = it * TT;
    const int rows = min(TT, nt - t0);
    if (tid == 0 && it + 1 < n_tiles) {
      // buffer buf^1 was released by the __syncthreads() closing iteration it-1
      uint32_t bytes = (uint32_t)min(TT, nt - (t0 + TT)) * 32u;
      mbar_expect_tx(&bars[buf ^ 1], bytes);
      tma_bulk_g2s(tile(buf ^ 1), tp + (size_t)(t0 + TT) * 32, bytes, &bars[buf ^ 1]);
    }
    mbar_wait(&bars[buf], (uint32_t)((it >> 1) & 1));

    const uint4* tl = tile(buf);
#pragma unroll 2
    for (int r = 0; r < rows; ++r) {
      const uint4 ta = tl[2 * r], tb = tl[2 * r + 1];  // warp-wide broadcast LDS.128 x2
      const uint32_t tkey = (uint32_t)(t0 + r);
      uint32_t rmin = 0xFFFFFFFFu;
#pragma unroll
      for (int j = 0; j < QPT; ++j) {
        const uint32_t d = hamming256<CSA>(qr[j], ta, tb);
        const uint32_t key = (d << kKeyShift) + tkey;
        const uint32_t lo = min(key, b1[j]);
        const uint32_t hi = max(key, b1[j]);
        b1[j] = lo;
        b2[j] = min(b2[j], hi);
        if (REV) rmin = min(rmin, (d << kKeyShift) + qid[j]);
      }
      if (REV) {
        rmin = __reduce_min_sync(0xffffffffu, rmin);
        if (lane == 0) atomicMin(&rev_s[r], rmin);
      }
    }
    __syncthreads();  // tile consumed by every warp; rev_s complete for this tile
    if (REV) {
      for (int i = tid; i < rows; i += THREADS) {
        uint32_t k = rev_s[i];
        rev_s[i] = 0xFFFFFFFFu;
        if (k < kInvalidQ) atomicMin(&rev_key[(size_t)pair * nt_stride + t0 + i], k);
      }
      __syncthreads();
    }
  }

#pragma unroll
  for (int j = 0; j < QPT; ++j) {
    int qi = q_base + j * THREADS + tid;
    if (qi < nq_stride) {
      if (qi >= nq) b1[j] = b2[j] = 0xFFFFFFFFu;   // padding query: no match
      int2 oi, od;
      oi.x = (b1[j] == 0xFFFFFFFFu) ? -1 : (int)(b1[j] & kIdxMask);
      od.x = (b1[j] == 0xFFFFFFFFu) ? -1 : (int)(b1[j] >> kKeyShift);
      oi.y = (b2[j] == 0xFFFFFFFFu) ? -1 : (int)(b2[j] & kIdxMask);
      od.y = (b2[j] == 0xFFFFFFFFu) ? -1 : (int)(b2[j] >> kKeyShift);
      size_t o = (size_t)pair * nq_stride + qi;
      reinterpret_cast<int2*>(idx)[o] = oi;
      reinterpret_cast<int2*>(dist)[o] = od;
    }
  }
}

// Ratio test + cross-check + order-preserving compaction; one CTA per frame pair.
__global__ void __launch_bounds__(256)
match_filter_kernel(const int32_t* __restrict__ idx, const int32_t* __restrict__ dist,
                    const uint32_t* __restrict__ rev_key, int nq, int nt, double ratio,
                    int cross_check, int32_t* __restrict__ keep_q, int32_t* __restrict__ keep_t,
                    int32_t* __restrict__ n_keep) {
  __shared__ int warp_cnt[8];
  __shared__ int base_s;
  const int pair = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int2* pi = reinterpret_cast<const int2*>(idx) + (size_t)pair * nq;
  const int2* pd = reinterpret_cast<const int2*>(dist) + (size_t)pair * nq;
  const uint32_t* rk = cross_check ? rev_key + (size_t)pair * nt : nullptr;
  int32_t* oq = keep_q + (size_t)pair * nq;
  int32_t* ot = keep_t + (size_t)pair * nq;
  if (tid == 0) base_s = 0;
  __syncthreads();
  for (int q0 = 0; q0 < nq; q0 += 256) {
    const int qi = q0 + tid;
    bool keep = false;
    int t0 = -1;
    if (qi < nq) {
      int2 i2 = pi[qi], d2 = pd[qi];
      t0 = i2.x;
      // frame.py:33  m.distance < ratio * n.distance  (float32 values compared in float64)
      keep = (i2.y >= 0) && ((double)d2.x < ratio * (double)d2.y);
      if (keep && cross_check) keep = ((rk[t0] & kIdxMask) == (uint32_t)qi);
    }
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) warp_cnt[wid] = __popc(bal);
    __syncthreads();
    int off = base_s;
    for (int w = 0; w < wid; ++w) off += warp_cnt[w];
    if (keep) {
      int o = off + __popc(bal & ((1u << lane) - 1u));
      oq[o] = qi;
      ot[o] = t0;
    }
    __syncthreads();
    if (tid == 0) {
      int s = 0;
      for (int w = 0; w < 8; ++w) s += warp_cnt[w];
      base_s += s;
    }
    __syncthreads();
  }
  const int total = base_s;
  for (int i = total + tid; i < nq; i += 256) { oq[i] = -1; ot[i] = -1; }
  if (tid == 0) n_keep[pair] = total;
}

__global__ void fill_u32_kernel(uint32_t* p, size_t n, uint32_t v) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) p[i] = v;
}

// popc throughput probe (roofline denominator for matching): 8 independent chains per thread,
// each step = one POPC (its input depends on the running sum, so ptxas cannot CSE it) + one IADD3.
__global__ void __launch_bounds__(256) popc_probe_kernel(int iters, uint32_t seed, uint32_t* out) {
  uint32_t acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = seed + threadIdx.x * 2654435761u + i * 40503u;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = acc[i] + __popc(acc[i]) + 0x9e3779b9u;
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s ^= acc[i];
  if (s == 0xdeadbeefu) out[0] = s;
}

// ALU-pipe probe: 8 chains per thread, each step = ONE lop3.b32 (majority of three chain values: no constant
// operand to fold, every input depends on the previous step).
__global__ void __launch_bounds__(256) lop3_probe_kernel(int iters, uint32_t seed, uint32_t* out) {
  uint32_t acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = seed * 747796405u + threadIdx.x * 2654435761u + i * 40503u;
  for (int it = 0; it < iters; ++it) {
    uint32_t nxt[8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
      asm volatile("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(nxt[i]) : "r"(acc[i]), "r"(acc[(i + 1) & 7]), "r"(acc[(i + 3) & 7]));
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = nxt[i];
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += acc[i];
  if (s == 0xdeadbeefu) out[0] = s;
}

template <int THREADS, int QPT, int TT, bool CSA>
static int launch_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx,
                       int32_t* dist, uint32_t* rev_key, int batch, cudaStream_t st,
                       const int32_t* nq_arr = nullptr, const int32_t* nt_arr = nullptr) {
  dim3 grid((nq + THREADS * QPT - 1) / (THREADS * QPT), batch);
  if (rev_key) {
    size_t smem = (size_t)2 * TT * 32 + TT * 4 + 16;
    auto k = hamming_knn2_kernel<THREADS, QPT, TT, CSA, true>;
    VSLAM_CUDA_TRY(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<grid, THREADS, smem, st>>>(q, nq, t, nt, idx, dist, rev_key, nq_arr, nt_arr);
  } else {
    size_t smem = (size_t)2 * TT * 32 + 16;
    auto k = hamming_knn2_kernel<THREADS, QPT, TT, CSA, false>;
    VSLAM_CUDA_TRY(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<grid, THREADS, smem, st>>>(q, nq, t, nt, idx, dist, rev_key, nq_arr, nt_arr);
  }
  return check_launch("hamming_knn2_kernel");
}

}  // namespace vslam

using namespace vslam;

// Tuning variant, selected by vslam_hamming_set_variant() (bench sweeps only).
static int g_variant = 0;
extern "C" int vslam_hamming_set_variant(int v) { g_variant = v; return VSLAM_OK; }

static int hamming_knn2_impl(const uint8_t* q, int nq, const int32_t* nq_arr, const uint8_t* t, int nt,
                            const int32_t* nt_arr, int32_t* idx, int32_t* dist, uint32_t* rev_key, int batch,
                            void* stream) {
  if (nq < 0 || nt < 0 || batch < 0) return VSLAM_ERR_BAD_ARG;
  if (nq > VSLAM_HAMMING_MAX_N || nt > VSLAM_HAMMING_MAX_N) return VSLAM_ERR_BAD_ARG;
  if (batch == 0 || nq == 0) return VSLAM_OK;
  if (!q || !idx || !dist || (nt > 0 && !t)) return VSLAM_ERR_BAD_ARG;
  if ((reinterpret_cast<uintptr_t>(q) & 15) || (reinterpret_cast<uintptr_t>(t) & 15))
    return VSLAM_ERR_ALIGNMENT;
  if ((reinterpret_cast<uintptr_t>(idx) & 7) || (reinterpret_cast<uintptr_t>(dist) & 7))
    return VSLAM_ERR_ALIGNMENT;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (rev_key && nt > 0) {
    size_t n = (size_t)batch * nt;
    int blocks = (int)((n + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    fill_u32_kernel<<<blocks, 256, 0, st>>>(rev_key, n, 0xFFFFFFFFu);
    int rc = check_launch("fill_u32_kernel");
    if (rc) return rc;
  }
  switch (g_variant) {
    case 1: return launch_knn2<128, 4, 512, false>(q, nq, t, nt, idx, dist, rev_key, batch, st, nq_arr, nt_arr);
    case 2: return launch_knn2<256, 4, 512, true>(q, nq, t, nt, idx, dist, rev_key, batch, st, nq_arr, nt_arr);
    case 3: return launch_knn2<128, 8, 512, true>(q, nq, t, nt, idx, dist, rev_key, batch, st, nq_arr, nt_arr);
    case 4: return launch_knn2<128, 2, 512, true>(q, nq, t, nt, idx, dist, rev_key, batch, st, nq_arr, nt_arr);
    case 5: return launch_knn2<256, 2, 512, true>(q, nq, t, nt, idx, dist, rev_key, batch, st, nq_arr, nt_arr);
    case 6: return launch_knn2<128, 4, 1024, true>(q, nq, t, nt, idx, dist, rev_key, batch, st, nq_arr, nt_arr);
    default: return launch_knn2<128, 4, 512, true>(q, nq, t, nt, idx, dist, rev_key, batch, st, nq_arr, nt_arr);
  }
}

extern "C" int vslam_hamming_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx,
                                  int32_t* dist, uint32_t* rev_key, int batch, void* stream) {
  return hamming_knn2_impl(q, nq, nullptr, t, nt, nullptr, idx, dist, rev_key, batch, stream);
}

extern "C" int vslam_hamming_knn2_ragged(const uint8_t* q, int nq_stride, const int32_t* nq_arr, const uint8_t* t,
                                         int nt_stride, const int32_t* nt_arr, int32_t* idx, int32_t* dist,
                                         uint32_t* rev_key, int batch, void* stream) {
  if (!nq_arr || !nt_arr) return VSLAM_ERR_BAD_ARG;
  return hamming_knn2_impl(q, nq_stride, nq_arr, t, nt_stride, nt_arr, idx, dist, rev_key, batch, stream);
}

extern "C" int vslam_match_filter(const int32_t* idx, const int32_t* dist, const uint32_t* rev_key,
                                  int nq, int nt, double ratio, int cross_check, int32_t* keep_q,
                                  int32_t* keep_t, int32_t* n_keep, int batch, void* stream) {
  if (nq < 0 || nt < 0 || batch < 0) return VSLAM_ERR_BAD_ARG;
  if (batch == 0) return VSLAM_OK;
  if (!n_keep || (nq > 0 && (!idx || !dist || !keep_q || !keep_t))) return VSLAM_ERR_BAD_ARG;
  if (cross_check && !rev_key && nt > 0 && nq > 0) return VSLAM_ERR_BAD_ARG;
  match_filter_kernel<<<batch, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      idx, dist, rev_key, nq, nt, ratio, cross_check, keep_q, keep_t, n_keep);
  return check_launch("match_filter_kernel");
}

template <typename Kernel>
static int run_probe(Kernel kernel, const char* name, int iters, double* per_s, void* stream);

extern "C" int vslam_microbench_lop3(int iters, double* lop3_per_s, void* stream) {
  if (!lop3_per_s || iters <= 0) return VSLAM_ERR_BAD_ARG;
  return run_probe(lop3_probe_kernel, "lop3_probe_kernel", iters, lop3_per_s, stream);
}

extern "C" int vslam_microbench_popc(int iters, double* popc_per_s, void* stream) {
  if (!popc_per_s || iters <= 0) return VSLAM_ERR_BAD_ARG;
  return run_probe(popc_probe_kernel, "popc_probe_kernel", iters, popc_per_s, stream);
}

template <typename Kernel>
static int run_probe(Kernel kernel, const char* name, int iters, double* popc_per_s, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int sms = sm_count();
  if (sms <= 0) return VSLAM_ERR_NO_DEVICE;
  uint32_t* out = nullptr;
  VSLAM_CUDA_TRY(cudaMalloc(&out, 4));
  cudaEvent_t e0, e1;
  VSLAM_CUDA_TRY(cudaEventCreate(&e0));
  VSLAM_CUDA_TRY(cudaEventCreate(&e1));
  const int blocks = sms * 8;
  kernel<<<blocks, 256, 0, st>>>(iters, 1u, out);  // warm-up
  float best = 1e30f;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0, st);
    kernel<<<blocks, 256, 0, st>>>(iters, 2u + rep, out);
    cudaEventRecord(e1, st);
    VSLAM_CUDA_TRY(cudaEventSynchronize(e1));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(out);
  *popc_per_s = (double)blocks * 256.0 * (double)iters * 8.0 / ((double)best * 1e-3);
  return check_launch(name);
}
