// (b) log_softmax over depth + expectation / variance / entropy / x0.25 carry for full-resolution volumes, with the
// volume moved by TMA (cp.async.bulk.tensor) instead of per-lane loads and stores.
//   models/models.py:351,435 (F.log_softmax(dim=1) of the decoder logits) ; utils/img_utils.py:197-215 ;
//   trainer/default_trainer.py:284,289,446-450.
//
// Why: in lsm_stats_kernel (dpv_softmax.cu) a pixel's depth column is split over 4 adjacent lanes, so one warp load
// touches 4 planes x 8 pixels = four separate 32-byte sectors; ncu showed the L1/TEX pipe 83 % busy at 47 % issue
// (profiles/ncu_r01_lsm_full_summary.txt): bound by L1 wavefronts, not by HBM.  Here a CTA owns 32 image columns and
// walks `rows` image rows.  Per row the TMA engine brings the [D planes x 32 pixels] tile (D rows of 128 bytes, plane
// stride H*W) into shared memory in the 128-byte-swizzle layout, two rows ahead of the arithmetic (two mbarrier
// stages); the results go back through a swizzled staging tile and one TMA store.  The register file never sees a
// global address of the volume.
//
// Lane <-> bin assignment.  The tile row r (= depth bin) holds pixel p at 16-byte chunk (p / 4) XOR (r mod 8) of its
// 128 bytes (CU_TENSOR_MAP_SWIZZLE_128B).  Lane (p, t) -- 4 lanes t per pixel p, as before -- owns the bins
// {8 j + 2 t + s : s = 0,1; j = 0 .. D/8-1} (log-softmax needs no neighbours along depth, so the bins of a lane need
// not be adjacent).  Its rows have r mod 8 = 2 t + s, so for fixed s the four lanes of a pixel read chunks
// (p / 4) XOR {0, 2, 4, 6} [+ s]: the warp's 8 pixels x 4 lanes hit 8 different chunks x 4 words = 32 different banks.
// Every shared-memory access of the kernel is conflict-free, for loads and for the staged stores alike.
#include <cuda.h>

#include <cstdlib>

#include "dpv_pixel.cuh"

namespace dpv {

constexpr int kTmaThreads = 128;     // 32 pixels x 4 lanes
constexpr int kTmaCols = 32;         // pixels per tile row = 128 bytes

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
      "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* tm, uint32_t src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(tm), "r"(src),
               "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}

// D = 64 or 128 bins; DPT = D / 4 per lane
template <int D>
__global__ void __launch_bounds__(kTmaThreads, D == 64 ? 6 : 3)
lsm_stats_tma_kernel(const __grid_constant__ CUtensorMap tm_in, const __grid_constant__ CUtensorMap tm_out, int H, int W,
                     const float* __restrict__ d_candi, float* __restrict__ mean_o, float* __restrict__ var_o,
                     float* __restrict__ ent_o, float* __restrict__ carry, int rows_per_cta) {
  constexpr int DPT = D / 4, NJ = D / 8;
  constexpr uint32_t kTile = D * kTmaCols * 4;                     // bytes of one [D x 32] tile
  extern __shared__ __align__(1024) unsigned char smraw[];        // the swizzle pattern is tied to 1024-byte phases
  unsigned char* sm = reinterpret_cast<unsigned char*>(((uintptr_t)smraw + 1023) & ~(uintptr_t)1023);
  __shared__ float dc[D];
  __shared__ __align__(8) unsigned long long bars[2];
  const uint32_t in0 = smem_u32(sm), outb = in0 + 2 * kTile;
  const uint32_t bar0 = smem_u32(&bars[0]);
  const int tid = threadIdx.x, p = tid >> 2, t = tid & 3;
  const int b = blockIdx.z, x0 = blockIdx.x * kTmaCols;
  const int v0 = blockIdx.y * rows_per_cta, v1 = min(H, v0 + rows_per_cta);
  for (int i = tid; i < D; i += kTmaThreads) dc[i] = d_candi[i];
  if (tid == 0) {
    mbar_init(bar0, 1);
    mbar_init(bar0 + 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (tid == 0) {                                                  // two rows in flight before any arithmetic
#pragma unroll
    for (int s = 0; s < 2; s++)
      if (v0 + s < v1) {
        mbar_expect_tx(bar0 + 8 * s, kTile);
        tma_load_3d(in0 + s * kTile, &tm_in, bar0 + 8 * s, x0, v0 + s, b * D);
      }
  }
  // byte offsets of this lane's two row families (s = 0, 1) inside a tile; bin j of family s is at off[s] + j * 1024
  uint32_t off[2];
#pragma unroll
  for (int s = 0; s < 2; s++) {
    const int r7 = 2 * t + s;
    off[s] = (uint32_t)(r7 * 128 + ((((p >> 2) ^ r7) & 7) << 4) + (p & 3) * 4);
  }
  const int x = x0 + p;
  const bool live = x < W;
  const unsigned HW = (unsigned)H * (unsigned)W;

#pragma unroll 1
  for (int v = v0; v < v1; v++) {
    const int st = (v - v0) & 1;
    const uint32_t par = ((v - v0) >> 1) & 1;
    while (!mbar_try_wait(bar0 + 8 * st, par)) {}
    float tt[DPT];                                                 // tt[2 j + s] = bin 8 j + 2 t + s
    {
      const uint32_t src = in0 + st * kTile;
#pragma unroll
      for (int j = 0; j < NJ; j++) {
#pragma unroll
        for (int s = 0; s < 2; s++)
          asm volatile("ld.shared.f32 %0, [%1];" : "=f"(tt[2 * j + s]) : "r"(src + off[s] + 1024u * j));
      }
    }
    __syncthreads();                                               // the stage has been read by everyone: refill it
    if (tid == 0 && v + 2 < v1) {
      mbar_expect_tx(bar0 + 8 * st, kTile);
      tma_load_3d(in0 + st * kTile, &tm_in, bar0 + 8 * st, x0, v + 2, b * D);
    }
    // ---- log-softmax + statistics of the pixel (same formulas as lsm_stats_kernel)
    float m0 = tt[0], m1 = tt[1];
#pragma unroll
    for (int d = 2; d < DPT; d += 2) { m0 = fmaxf(m0, tt[d]); m1 = fmaxf(m1, tt[d + 1]); }
    float m = fmaxf(m0, m1);
    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
    float e[DPT];
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int d = 0; d < DPT; d += 2) {
      tt[d] -= m;
      tt[d + 1] -= m;
      e[d] = exp2f_approx(tt[d] * 1.4426950408889634f);
      e[d + 1] = exp2f_approx(tt[d + 1] * 1.4426950408889634f);
      s0 += e[d];
      s1 += e[d + 1];
    }
    float ssum = s0 + s1;
    ssum += __shfl_xor_sync(0xffffffffu, ssum, 1);
    ssum += __shfl_xor_sync(0xffffffffu, ssum, 2);
    const float lse = log2f_approx(ssum) * 0.6931471805599453f;    // ssum in [1, D]
    const float inv = 1.f / ssum;
    float mean0 = 0.f, mean1 = 0.f, ent0 = 0.f, ent1 = 0.f;
#pragma unroll
    for (int j = 0; j < NJ; j++) {
      const float2 dq = *reinterpret_cast<const float2*>(dc + 8 * j + 2 * t);
      const float lp0 = tt[2 * j] - lse, lp1 = tt[2 * j + 1] - lse;
      e[2 * j] *= inv;
      e[2 * j + 1] *= inv;
      mean0 = fmaf(dq.x, e[2 * j], mean0);
      mean1 = fmaf(dq.y, e[2 * j + 1], mean1);
      ent0 = fmaf(-e[2 * j], lp0, ent0);
      ent1 = fmaf(-e[2 * j + 1], lp1, ent1);
      tt[2 * j] = lp0;
      tt[2 * j + 1] = lp1;
    }
    float mean = mean0 + mean1, ent = ent0 + ent1;
    mean += __shfl_xor_sync(0xffffffffu, mean, 1);
    mean += __shfl_xor_sync(0xffffffffu, mean, 2);
    ent += __shfl_xor_sync(0xffffffffu, ent, 1);
    ent += __shfl_xor_sync(0xffffffffu, ent, 2);
    const size_t pix = (size_t)b * HW + (size_t)v * W + x;
    const bool writer = live && t == 0;
    if (mean_o && writer) mean_o[pix] = mean;
    if (ent_o && writer) ent_o[pix] = ent;
    if (var_o) {
      float q0 = 0.f, q1 = 0.f;
#pragma unroll
      for (int j = 0; j < NJ; j++) {
        const float2 dq = *reinterpret_cast<const float2*>(dc + 8 * j + 2 * t);
        const float a0 = dq.x - mean, a1 = dq.y - mean;
        q0 = fmaf(a0 * a0, e[2 * j], q0);
        q1 = fmaf(a1 * a1, e[2 * j + 1], q1);
      }
      float var = q0 + q1;
      var += __shfl_xor_sync(0xffffffffu, var, 1);
      var += __shfl_xor_sync(0xffffffffu, var, 2);
      if (writer) var_o[pix] = var;
    }
    if (carry && live && (v & 3) == 0 && (x & 3) == 0) {           // F.interpolate(0.25, 'nearest'): dst (j,i) <- src (4j,4i)
      const int h4 = H / 4, w4 = W / 4;
      if ((v >> 2) < h4 && (x >> 2) < w4) {
        float* c = carry + (size_t)b * D * h4 * w4 + (size_t)(v >> 2) * w4 + (x >> 2);
#pragma unroll
        for (int j = 0; j < NJ; j++) {
          c[(size_t)(8 * j + 2 * t) * h4 * w4] = tt[2 * j];
          c[(size_t)(8 * j + 2 * t + 1) * h4 * w4] = tt[2 * j + 1];
        }
      }
    }
    // ---- log-probabilities back through the staging tile (same swizzle), one TMA store per row
    if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");    // the previous row's store has read it
    __syncthreads();
#pragma unroll
    for (int j = 0; j < NJ; j++) {
#pragma unroll
      for (int s = 0; s < 2; s++)
        asm volatile("st.shared.f32 [%0], %1;" ::"r"(outb + off[s] + 1024u * j), "f"(tt[2 * j + s]) : "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the TMA engine
    __syncthreads();
    if (tid == 0) {
      tma_store_3d(&tm_out, outb, x0, v, b * D);
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
  }
  if (tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");            // stores complete before the CTA exits
}

// ---- host side ----------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess) {
      (void)cudaGetLastError();
      f = nullptr;
    }
    return reinterpret_cast<EncodeTiledFn>(f);
  }();
  return fn;
}

// [B*D, H, W] fp32 volume as a 3-D tensor (W fastest), box = (32, 1, D) in the 128-byte swizzle
static bool make_volume_map(CUtensorMap* tm, const float* base, int B, int D, int H, int W) {
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) return false;
  const cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B * D};
  const cuuint64_t strides[2] = {(cuuint64_t)W * 4, (cuuint64_t)H * W * 4};      // bytes, dims 1 and 2
  const cuuint32_t box[3] = {(cuuint32_t)kTmaCols, 1u, (cuuint32_t)D};
  const cuuint32_t estr[3] = {1u, 1u, 1u};
  return enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims, strides, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// Returns DPV_OK when the TMA kernel was launched, -1 when this shape is not its case (the caller then runs
// lsm_stats_kernel), another DPV_E* code on a CUDA error.
int logsoftmax_stats_tma(const float* x, int B, int D, int H, int W, const float* d_candi, float* logp, float* mean,
                         float* var, float* entropy, float* carry, cudaStream_t st) {
  static const bool enabled = !(getenv("DPV_LSM_TMA") && atoi(getenv("DPV_LSM_TMA")) == 0);
  if (!enabled || (D != 64 && D != 128) || W < kTmaCols || (W & 3) != 0 || B > 65535) return -1;
  if (((uintptr_t)x & 15) != 0 || ((uintptr_t)logp & 15) != 0) return -1;           // TMA: 16-byte aligned base and strides
  CUtensorMap tin, tout;
  if (!make_volume_map(&tin, x, B, D, H, W) || !make_volume_map(&tout, logp, B, D, H, W)) return -1;
  static const int rows_env = getenv("DPV_LSM_ROWS") ? atoi(getenv("DPV_LSM_ROWS")) : 0;
  int rows = rows_env > 0 ? rows_env : 8;
  const long long per_row = (long long)((W + kTmaCols - 1) / kTmaCols) * B;
  if (rows_env <= 0)
    while (rows > 1 && per_row * ((H + rows - 1) / rows) < 8ll * 148 * 6) rows >>= 1;
  const dim3 grid((unsigned)((W + kTmaCols - 1) / kTmaCols), (unsigned)((H + rows - 1) / rows), (unsigned)B);
  const size_t smem = (size_t)3 * D * kTmaCols * 4 + 1024;                          // 2 stages + staging + alignment slack
  if (D == 64) {
    DPV_CUDA(cudaFuncSetAttribute(lsm_stats_tma_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    lsm_stats_tma_kernel<64><<<grid, kTmaThreads, smem, st>>>(tin, tout, H, W, d_candi, mean, var, entropy, carry, rows);
  } else {
    DPV_CUDA(cudaFuncSetAttribute(lsm_stats_tma_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    lsm_stats_tma_kernel<128><<<grid, kTmaThreads, smem, st>>>(tin, tout, H, W, d_candi, mean, var, entropy, carry, rows);
  }
  return check_launch("dpv_logsoftmax_stats (tma)");
}

}  // namespace dpv
