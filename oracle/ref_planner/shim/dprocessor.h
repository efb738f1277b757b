// TEST INFRASTRUCTURE -- stand-in for lcsim's dprocessor.h (see common.h in this directory): planning.cpp only calls
// DatumProcessor::getCDatum("camera01") (planning.cpp:18,21).
#ifndef DPV_REF_PLANNER_SHIM_DPROCESSOR_H
#define DPV_REF_PLANNER_SHIM_DPROCESSOR_H

#include <common.h>

namespace lc {

class DatumProcessor {
 public:
  std::map<std::string, std::shared_ptr<Datum>> c_datums;
  const std::shared_ptr<Datum> getCDatum(std::string camera_name) { return c_datums.at(camera_name); }
};

}  // namespace lc

#endif
