// (d3) Dynamic-programming curtain planner.
//   external/lcsim/lcsim/src/planning.cpp:14-73 (Planner ctor + constructGraph), :75-126
//   (optimizedDesignPts), :162-217 (Trajectory compare), :255-264 (PolarIdentityInterpolator);
//   python wrapper external/lcsim/python/planner.py:162-189 (PlannerRT).
//
// Graph: node (ray i, range k); edge (i,k)->(i+1,k') iff |theta_las(i+1,k') - theta_las(i,k)| <=
// laser_limit*laser_timestep.  theta_las is monotone along a ray, so each node's edge list is one
// contiguous window [lo,hi] of k' — two bytes per node instead of a vector (verified at creation).
// DP (right to left over rays) is sequential in i; one CTA solves one umap with LANES threads per
// node scanning the window, combined under the reference's total order
//   (unc larger) > (las smaller) > (first in edge order), "no edge" first of all,
// which is exactly what its strict-'>' sequential scan selects, so indices are bit-exact.
// Back-pointers for all W x R nodes stay in shared memory (W*R bytes).
#include <cmath>
#include <cstring>
#include <vector>

#include "lc_device.cuh"

#define LCP_PI 3.14159265 /* planning.h:11 (truncated literal) */

struct LcPlanner {
  int W = 0, R = 0;
  float max_d_las = 0.f;
  long long edges = 0;
  float* d_theta = nullptr;     // [W,R]
  float* d_xz = nullptr;        // [W,R,2]
  uint8_t* d_win = nullptr;     // [W,R,2] lo,hi (lo > hi = no edge); last ray has none
};

namespace dpv {

struct Cand {
  float u, l;
  int e;    // -1 = no edge (path ends here)
};

// reference order (planning.cpp:210-217 + first-in-edge-order on full ties)
__device__ __forceinline__ bool cand_better(const Cand& a, const Cand& b) {
  if (a.u > b.u) return true;
  if (a.u == b.u) {
    if (a.l < b.l) return true;
    if (a.l == b.l) return a.e < b.e;
  }
  return false;
}

template <int LANES>
__global__ void __launch_bounds__(1024)
planner_solve_kernel(const float* __restrict__ umaps, int W, int R, const float* __restrict__ theta,
                     const float* __restrict__ node_xz, const uint8_t* __restrict__ win, int32_t* __restrict__ ks,
                     float* __restrict__ xz, int32_t* __restrict__ counts) {
  extern __shared__ unsigned char smraw[];
  float* unc = reinterpret_cast<float*>(smraw);          // [2][R]
  float* las = unc + 2 * R;                              // [2][R]
  float* th = las + 2 * R;                               // [2][R]
  int16_t* path = reinterpret_cast<int16_t*>(th + 2 * R);   // [W]
  uint8_t* sub = reinterpret_cast<uint8_t*>(path + ((W + 1) & ~1));   // [W][R]
  __shared__ int s_cnt;

  const int k = threadIdx.x / LANES, s = threadIdx.x % LANES;
  const bool active = k < R;
  const float* umap = umaps + (size_t)blockIdx.x * R * W;
  const unsigned gmask = 0xffffffffu;

  // Rays are processed in groups of LANES: lane s of node k prefetches (u, theta, window) of ray
  // i0 - s one group ahead; values are handed round with shuffles.
  float pu = 0.f, pt = 0.f;
  int plo = 1, phi = 0;
  auto prefetch = [&](int i0) {
    int i = i0 - s;
    if (active && i >= 0) {
      pu = __ldg(umap + (size_t)k * W + i);
      pt = __ldg(theta + (size_t)i * R + k);
      if (i < W - 1) {
        plo = win[((size_t)i * R + k) * 2];
        phi = win[((size_t)i * R + k) * 2 + 1];
      } else {
        plo = 1; phi = 0;
      }
    }
  };
  prefetch(W - 1);
  int buf = 0;
  for (int i0 = W - 1; i0 >= 0; i0 -= LANES) {
    const float cu = pu, ct = pt;
    const int clo = plo, chi = phi;
    prefetch(i0 - LANES);
    for (int j = 0; j < LANES && i0 - j >= 0; j++) {
      const int i = i0 - j;
      const int src = (threadIdx.x & 31 & ~(LANES - 1)) + j;
      const float u = __shfl_sync(gmask, cu, src);
      const float tn = __shfl_sync(gmask, ct, src);
      const int lo = __shfl_sync(gmask, clo, src);
      const int hi = __shfl_sync(gmask, chi, src);
      const float* uncn = unc + (buf ^ 1) * R;
      const float* lasn = las + (buf ^ 1) * R;
      const float* thn = th + (buf ^ 1) * R;
      Cand best;
      best.u = u; best.l = 0.f; best.e = -1;                  // Trajectory(node): planning.cpp:169-184
      if (active) {
        for (int e = lo + s; e <= hi; e += LANES) {
          Cand c;
          c.u = __fadd_rn(u, uncn[e]);                        // :194
          float dth = __fsub_rn(thn[e], tn);
          c.l = __fadd_rn(__fmul_rn(dth, dth), lasn[e]);      // :197-198
          c.e = e;
          if (cand_better(c, best)) best = c;
        }
      }
#pragma unroll
      for (int off = LANES / 2; off > 0; off >>= 1) {
        Cand o;
        o.u = __shfl_xor_sync(gmask, best.u, off);
        o.l = __shfl_xor_sync(gmask, best.l, off);
        o.e = __shfl_xor_sync(gmask, best.e, off);
        if (cand_better(o, best)) best = o;
      }
      if (active && s == 0) {
        unc[buf * R + k] = best.u;
        las[buf * R + k] = best.l;
        th[buf * R + k] = tn;
        sub[(size_t)i * R + k] = best.e < 0 ? 255 : (uint8_t)best.e;
      }
      __syncthreads();
      buf ^= 1;
    }
  }
  // ray 0 results are in buffer buf^1.  Best start against a default (0,0,null) trajectory (:99-104)
  if (threadIdx.x == 0) {
    const float* u0 = unc + (buf ^ 1) * R;
    const float* l0 = las + (buf ^ 1) * R;
    float bu = 0.f, bl = 0.f;
    int bk = -1;
    for (int kk = 0; kk < R; kk++) {
      float tu = u0[kk], tl = l0[kk];
      bool gt = (tu > bu) ? true : ((tu == bu) ? (tl < bl) : false);
      if (gt) { bu = tu; bl = tl; bk = kk; }
    }
    int cnt = 0;
    if (bk >= 0) {                       // the reference dereferences a null node otherwise (:116)
      int i = 0, kk = bk;
      while (true) {
        path[cnt++] = (int16_t)kk;
        int nx = sub[(size_t)i * R + kk];
        if (nx == 255 || i == W - 1) break;
        kk = nx;
        i++;
      }
    }
    s_cnt = cnt;
    if (counts) counts[blockIdx.x] = cnt;
  }
  __syncthreads();
  const int cnt = s_cnt;
  for (int i = threadIdx.x; i < W; i += blockDim.x) {
    int kk = i < cnt ? path[i] : -1;
    if (ks) ks[(size_t)blockIdx.x * W + i] = kk;
    if (xz) {
      float x = 0.f, z = 0.f;
      if (kk >= 0) {
        x = node_xz[((size_t)i * R + kk) * 2];
        z = node_xz[((size_t)i * R + kk) * 2 + 1];
      }
      xz[((size_t)blockIdx.x * W + i) * 2] = x;
      xz[((size_t)blockIdx.x * W + i) * 2 + 1] = z;
    }
  }
}

}  // namespace dpv

using namespace dpv;

static float deg2rad(float deg) { return (float)(deg * LCP_PI / 180); }   // planning.cpp:6-11
static float rad2deg(float rad) { return (float)(rad * 180 / LCP_PI); }

extern "C" {

int lc_planner_create(const LcDevice* dev, const float* ranges, int R, LcPlanner** out) {
  DPV_REQUIRE(dev && ranges && out, "lc_planner_create: null pointer");
  DPV_REQUIRE(R > 0 && R <= 255, "lc_planner_create: R=%d outside [1,255]", R);
  const int W = dev->W;
  LcPlanner* p = new LcPlanner();
  p->W = W; p->R = R;
  p->max_d_las = dev->laser_limit * dev->laser_timestep;       // planning.cpp:24
  std::vector<float> th((size_t)W * R), xz((size_t)W * R * 2);
  for (int i = 0; i < W; i++)
    for (int k = 0; k < R; k++) {                                // constructGraph :35-58
      float r = ranges[k], a = dev->angles[i];
      float x = r * std::sin(deg2rad(a));
      float z = r * std::cos(deg2rad(a));
      float c4[4] = {x, 0.0f, z, 1.0f}, l4[4];
      for (int q = 0; q < 4; q++) {
        float s = dev->laser_to_cam.m[q][0] * c4[0];
        for (int b = 1; b < 4; b++) s = s + dev->laser_to_cam.m[q][b] * c4[b];
        l4[q] = s;
      }
      xz[((size_t)i * R + k) * 2] = x;
      xz[((size_t)i * R + k) * 2 + 1] = z;
      th[(size_t)i * R + k] = rad2deg(std::atan2(l4[0], l4[2]));
    }
  std::vector<uint8_t> win((size_t)W * R * 2);
  long long edges = 0;
  bool contiguous = true;
  for (int i = 0; i < W; i++)
    for (int a = 0; a < R; a++) {
      int lo = 1, hi = 0, n = 0;
      if (i + 1 < W)
        for (int b = 0; b < R; b++) {                            // :60-71
          float d = th[(size_t)(i + 1) * R + b] - th[(size_t)i * R + a];
          if ((-p->max_d_las <= d) && (d <= p->max_d_las)) {
            if (n == 0) lo = b;
            hi = b;
            n++;
          }
        }
      if (n > 0 && hi - lo + 1 != n) contiguous = false;
      edges += n;
      win[((size_t)i * R + a) * 2] = (uint8_t)lo;
      win[((size_t)i * R + a) * 2 + 1] = (uint8_t)hi;
    }
  if (!contiguous) {
    delete p;
    dpv::set_error("lc_planner_create: edge lists are not contiguous windows for this geometry");
    return DPV_EUNSUPPORTED;
  }
  p->edges = edges;
  size_t smem = (size_t)6 * R * 4 + (size_t)((W + 1) & ~1) * 2 + (size_t)W * R;
  if (smem > 227 * 1024 - 64) {
    delete p;
    dpv::set_error("lc_planner_create: W*R = %d*%d back-pointers exceed shared memory", W, R);
    return DPV_EUNSUPPORTED;
  }
  cudaError_t e;
  if ((e = cudaMalloc(&p->d_theta, th.size() * 4)) != cudaSuccess ||
      (e = cudaMalloc(&p->d_xz, xz.size() * 4)) != cudaSuccess ||
      (e = cudaMalloc(&p->d_win, win.size())) != cudaSuccess ||
      (e = cudaMemcpy(p->d_theta, th.data(), th.size() * 4, cudaMemcpyHostToDevice)) != cudaSuccess ||
      (e = cudaMemcpy(p->d_xz, xz.data(), xz.size() * 4, cudaMemcpyHostToDevice)) != cudaSuccess ||
      (e = cudaMemcpy(p->d_win, win.data(), win.size(), cudaMemcpyHostToDevice)) != cudaSuccess) {
    dpv::set_error("lc_planner_create: %s", cudaGetErrorString(e));
    lc_planner_destroy(p);
    return DPV_ECUDA;
  }
  *out = p;
  return DPV_OK;
}

void lc_planner_destroy(LcPlanner* p) {
  if (!p) return;
  cudaFree(p->d_theta);
  cudaFree(p->d_xz);
  cudaFree(p->d_win);
  delete p;
}

long long lc_planner_edge_count(const LcPlanner* p) { return p ? p->edges : 0; }

int lc_planner_solve(LcPlanner* p, const float* umaps, int n, int32_t* ks, float* xz, int32_t* counts,
                     void* stream) {
  DPV_REQUIRE(p && umaps, "lc_planner_solve: null pointer");
  DPV_REQUIRE(n > 0, "lc_planner_solve: n=%d", n);
  const int W = p->W, R = p->R;
  int lanes = 8;
  while (lanes > 1 && R * lanes > 1024) lanes >>= 1;
  int threads = ((R * lanes + 31) / 32) * 32;
  size_t smem = (size_t)6 * R * 4 + (size_t)((W + 1) & ~1) * 2 + (size_t)W * R;
  cudaStream_t st = (cudaStream_t)stream;
#define LAUNCH(L)                                                                                              \
  do {                                                                                                         \
    DPV_CUDA(cudaFuncSetAttribute(planner_solve_kernel<L>, cudaFuncAttributeMaxDynamicSharedMemorySize,        \
                                  (int)smem));                                                                 \
    planner_solve_kernel<L><<<n, threads, smem, st>>>(umaps, W, R, p->d_theta, p->d_xz, p->d_win, ks, xz,      \
                                                      counts);                                                 \
  } while (0)
  if (lanes == 8) LAUNCH(8);
  else if (lanes == 4) LAUNCH(4);
  else if (lanes == 2) LAUNCH(2);
  else LAUNCH(1);
#undef LAUNCH
  return check_launch("lc_planner_solve");
}

}  // extern "C"
