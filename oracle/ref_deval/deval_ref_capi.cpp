// TEST INFRASTRUCTURE -- C wrapper around the reference's own depthError / evaluateErrors
// (external/deval_lib/src/evaluate_depth.h:20-142), included VERBATIM from /root/reference by oracle/ref_deval.py.
// The pybind caster of the reference (python/pyevaluatedepth_lib.cpp:36-60) builds DepthImage(data, shape[1], shape[0])
// from a float32 [H,W] array; so does this.
#include "evaluate_depth.h"

extern "C" int ref_depth_error(const float* gt, const float* pred, int width, int height, float* out9) {
  DepthImage a(gt, width, height), b(pred, width, height);
  try {
    std::vector<float> e = depthError(a, b);
    for (int i = 0; i < 9; i++) out9[i] = e[i];
  } catch (int) {
    return 1;                                              // no valid ground-truth pixel (evaluate_depth.h:96-99)
  }
  return 0;
}

// errors [n,9] -> out [9,3] = (mean, min, max) per metric in the order of evaluate_depth.h:124-132
extern "C" void ref_evaluate_errors(const float* errors, int n, float* out27) {
  std::vector<std::vector<float>> all(n, std::vector<float>(9));
  for (int i = 0; i < n; i++)
    for (int k = 0; k < 9; k++) all[i][k] = errors[i * 9 + k];
  std::map<std::string, std::vector<float>> r = evaluateErrors(all);
  const char* metrics[] = {"mae", "rmse", "inverse mae", "inverse rmse", "log mae", "log rmse", "scale invariant log",
                           "abs relative", "squared relative"};
  for (int k = 0; k < 9; k++)
    for (int j = 0; j < 3; j++) out27[k * 3 + j] = r[metrics[k]][j];
}
