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
  if (smem > 227 * 1024 - 128)   /* the solve kernel also has a few static shared words */ {
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
  for (auto& kv : p->logs) {
    cudaFree(kv.second.d_log);
    cudaFree(kv.second.d_epoch);
  }
  delete p;
}

long long lc_planner_edge_count(const LcPlanner* p) { return p ? p->edges : 0; }

}  // extern "C"
