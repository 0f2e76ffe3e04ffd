// sirrt_sweep.cuh -- throughput kernel for capsule links vs box/sphere obstacles (placeholder:
// the generic k_collide_bits path serves every scene until the queue-compacted kernel lands).
#pragma once
#include <vector>

#include "sirrt_kernels.cuh"
#include "sirrt_tables.hpp"

struct SweepTables {
  int supported = 0;
};
static inline int sweep_build(const sirrt::HScene&, const sirrt::HostTables&, const DScene&, SweepTables* st, std::vector<void*>&) {
  st->supported = 0;
  return 0;
}
static inline int sweep_launch(const SweepTables&, const DRobot&, const DScene&, const QView&, const WorkList&, uint32_t*, int, float, int,
                               cudaStream_t) {
  return 0;
}
