// TEST INFRASTRUCTURE -- C entry points around the reference's own lc::Planner (planning.cpp compiled verbatim from
// /root/reference; python/planner.py:162-189 PlannerRT builds it the same way: PolarIdentityInterpolator(num rays,
// num ranges), umap handed over as a [ranges x rays] matrix).  Used by tests/golden/make_planner_ref.py to produce
// tests/golden/planner_ref.npz; never linked into or called by the product.
#include <planning.h>

#include <cstring>

extern "C" {

// angles [W] degrees, ranges [R], laser_to_cam row-major 4x4 -> handle (NULL on failure)
void* ref_planner_create(const float* angles, int W, const float* ranges, int R, const float* laser_to_cam,
                         float laser_limit, float laser_timestep) {
  if (W > MAX_RAYS || R > MAX_NODES_PER_RAY) return nullptr;      // planning.h:12-13 (fixed-size arrays)
  try {
    auto dp = std::make_shared<lc::DatumProcessor>();
    auto datum = std::make_shared<lc::Datum>();
    datum->valid_angles.assign(angles, angles + W);
    lc::Laser laser;
    laser.laser_limit = laser_limit;
    laser.laser_timestep = laser_timestep;
    laser.laser_to_cam = Eigen::MatrixXf(4, 4);
    for (int i = 0; i < 4; i++)
      for (int j = 0; j < 4; j++) laser.laser_to_cam(i, j) = laser_to_cam[i * 4 + j];
    datum->laser_data["laser01"] = laser;
    dp->c_datums["camera01"] = datum;
    std::vector<float> rs(ranges, ranges + R);
    auto interp = std::make_shared<lc::PolarIdentityInterpolator>(W, R);
    return new lc::Planner(dp, rs, interp, false);
  } catch (...) {
    return nullptr;
  }
}

void ref_planner_destroy(void* h) { delete static_cast<lc::Planner*>(h); }

// umap row-major [R,W] -> xz [W,2]; returns the number of design points, -1 on an exception
int ref_planner_solve(void* h, const float* umap, int R, int W, float* xz) {
  try {
    Eigen::MatrixXf m(R, W);
    for (int i = 0; i < R; i++)
      for (int j = 0; j < W; j++) m(i, j) = umap[(size_t)i * W + j];
    auto pts = static_cast<lc::Planner*>(h)->optimizedDesignPts(m);
    for (size_t k = 0; k < pts.size() && k < (size_t)W; k++) {
      xz[2 * k] = pts[k].first;
      xz[2 * k + 1] = pts[k].second;
    }
    return (int)pts.size();
  } catch (...) {
    return -1;
  }
}

// edges leaving the nodes of one ray (for the edge-count check)
long long ref_planner_edge_count(void* h) {
  long long n = 0;
  for (auto& ray : static_cast<lc::Planner*>(h)->getVectorizedGraph())
    for (auto& node : ray) n += node.second;
  return n;
}

}  // extern "C"
