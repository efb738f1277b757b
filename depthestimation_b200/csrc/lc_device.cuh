// Light-curtain device model shared by the planner (lc_plan.cu) and the sensing kernels (lc_sense.cu).
// Host part mirrors DatumProcessor::setSensors (external/lcsim/lcsim/src/dprocessor.cpp:41-235);
// the per-geometry tables are computed once on the host and kept in HBM.
#pragma once
#include <map>
#include <mutex>
#include <vector>

#include "common.cuh"

struct LcMat4 {
  float m[4][4];
};

// Plain-old-data view passed to kernels by value.
struct LcDeviceView {
  int W, H;
  const float* nn_x;        // [W] un-normalised ray x per column ((u+1-cx)/fx, 1-based pixel grid)
  const float* nn_y;        // [H]
  const float* ztoramap;    // [H*W]
  const float* mid_x;       // [W] normalised mid-row ray
  const float* mid_z;       // [W]
  const float* angles;      // [W] camera angle per column, degrees
  const float* hit_lut;     // [hit_N]
  int hit_N, hit_mode;
  float limit, thickness, divergence;
  float cam_origin[2], p_left_cam[2], p_right_cam[2], laser_origin[2], p_left_laser[2], p_right_laser[2];
  LcMat4 cam_to_laser, laser_to_cam_inv;
};

// The four transforms of calculateSurface (dprocessor.cpp:841-857) that do not depend on the column: laser sheet
// offset by +-thickness/2 and rotated by -+divergence.  Built once on the host with the same routine the kernel used
// to run per column (30 double-precision sin / cos per column otherwise).
struct LcPlaneConsts {
  LcMat4 off_l, div_l, off_r, div_r;
};

struct LcDevice {
  int W = 0, H = 0;
  float K[9];
  LcMat4 cam_to_laser, laser_to_cam, laser_to_cam_inv;
  float fov = 80.f, laser_timestep = 1.5e-5f, laser_limit = 14000.f;
  float thickness = 0.00055f, divergence = 0.055f, limit = 1.0f;
  int hit_mode = 0, hit_N = 1000;
  float hit_std = 0.3f, hit_pow = 2.f;
  std::vector<float> nn_x, nn_y, ztoramap, mid_x, mid_z, angles, hit_lut;
  float t_max = 0.f;
  float* d_tables = nullptr;   // one device allocation holding all tables
  LcDeviceView view;
  LcPlaneConsts plane_consts;
};

// DP planner handle (lc_planner.cu builds it, lc_planner_solve.cu uses it)
struct LcPlanner {
  int W = 0, R = 0;
  float max_d_las = 0.f;
  long long edges = 0;
  float* d_theta = nullptr;     // [W,R]
  float* d_xz = nullptr;        // [W,R,2]
  uint8_t* d_win = nullptr;     // [W,R,2] lo,hi (lo > hi = no edge); last ray has none
  // lc_planner_solve's producer -> consumer log.  Solves issued on one stream are ordered and share a log; solves on
  // different streams may overlap on the GPU, so each stream gets its own (grown on demand, never while that stream
  // is being captured; freed with the handle).  The map is guarded by `mu` (the launch itself is outside the lock).
  struct Log {
    unsigned char* d_log = nullptr;
    size_t log_bytes = 0;
    // device-side solve counter {epoch, CTAs finished} of the DPV_PLANNER_DEVTAG=1 variant: a launch captured in a
    // CUDA graph takes the tag of its log entries from it (the default clears the log with a memset node instead)
    unsigned* d_epoch = nullptr;
    int host_epoch = 0;         // tags of directly launched solves (kernel parameter)
    int last_n = 0;
  };
  std::map<void*, Log> logs;    // keyed by cudaStream_t
  std::mutex mu;
  int pack = 1;                 // maps per CTA pair (lc_planner_set_pack): 1 = lowest latency, 2 = half the SMs per solve
};
