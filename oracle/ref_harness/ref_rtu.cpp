// oracle/ref_harness/ref_rtu.cpp — TEST INFRASTRUCTURE ONLY.
//
// The simx RT-unit traversal of the REFERENCE (sim/simx/rt_traversal.cpp), compiled verbatim where it lies
// under /root/reference and linked behind a one-function shim: vortex::RTUnit::dcache_read, which the
// reference routes through the simulated core's data cache (sim/simx/rt_unit.cpp:42-44,294-296), is defined
// here as a memcpy out of a flat "device memory" image supplied by the caller.  Nothing is copied from the
// reference; this file only #includes its headers.
//
// rtu_trace_rays() drives BVHTraverser::traverse exactly as RTUnit does for opaque geometry
// (rt_unit.cpp:98-116 traverse, :199-202 commit(ACCEPT) issued by shaders/anyhit.cpp:28-35):
//     while (!traverse(ray, hit, trail, stack)) hit.dist = hit.pending_dist;
// with a zero-initialised trail and an empty short stack per ray (fresh map entries in RTUnit).
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>
#include <atomic>

#include "rt_unit.h"
#include "rt_traversal.h"

namespace {
thread_local const uint8_t *g_mem = nullptr;   // flat image; Vortex address A lives at g_mem[A - g_base]
thread_local uint64_t g_base = 0, g_size = 0;
}

// the shim: the only project symbol rt_traversal.o leaves unresolved
void vortex::RTUnit::dcache_read(void *data, uint64_t addr, uint32_t size) {
  if (addr < g_base || addr + size > g_base + g_size) {
    memset(data, 0, size);
    return;
  }
  memcpy(data, g_mem + (addr - g_base), size);
}

extern "C" {

struct rtu_hit_t {   // == raycast ray_hit_t (24 B): dist, bcoords, blasIdx, triIdx
  float dist, bx, by, bz;
  uint32_t blasIdx, triIdx;
};

// mem/base/size: flat memory image; tlas/blas/qbvh/tri: the four RTU DCR values (0x006..0x009), 32-bit addresses.
// rays: 6 floats per ray; hits: n records.  work (optional, 3 x u64): traverse() calls, restarts (calls-1 summed), 0.
void rtu_trace_rays(const uint8_t *mem, uint64_t base, uint64_t size, uint32_t tlas, uint32_t blas, uint32_t qbvh, uint32_t tri,
                    const float *rays, uint64_t n, rtu_hit_t *hits, uint32_t threads, uint64_t *work) {
  if (threads == 0) threads = 1;
  std::atomic<uint64_t> next{0}, calls{0};
  auto worker = [&]() {
    g_mem = mem, g_base = base, g_size = size;
    vortex::DCRS dcrs;
    dcrs.base_dcrs.write(VX_DCR_BASE_RTX_TLAS_PTR, tlas);
    dcrs.base_dcrs.write(VX_DCR_BASE_RTX_BLAS_PTR, blas);
    dcrs.base_dcrs.write(VX_DCR_BASE_RTX_BVH_PTR, qbvh);
    dcrs.base_dcrs.write(VX_DCR_BASE_RTX_TRI_PTR, tri);
    vortex::BVHTraverser trav(nullptr, dcrs);
    uint64_t my_calls = 0;
    for (;;) {
      uint64_t b = next.fetch_add(1024);
      if (b >= n) break;
      uint64_t e = b + 1024 < n ? b + 1024 : n;
      for (uint64_t i = b; i < e; ++i) {
        vortex::Ray ray{rays[6 * i], rays[6 * i + 1], rays[6 * i + 2], rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]};
        vortex::Hit hit;
        hit.bx = hit.by = hit.bz = 0.0f, hit.blasIdx = hit.triIdx = 0, hit.pending_dist = 0.0f;
        vortex::TraversalTrail trail{};
        vortex::TraversalStack stack;
        vortex::per_thread_info info;
        for (;;) {
          ++my_calls;
          bool done = trav.traverse(ray, hit, trail, stack, info);
          info.RT_mem_accesses.clear();
          if (done) break;
          hit.dist = hit.pending_dist;   // commit(ACCEPT)
        }
        hits[i] = rtu_hit_t{hit.dist, hit.bx, hit.by, hit.bz, hit.blasIdx, hit.triIdx};
      }
    }
    calls += my_calls;
  };
  if (threads == 1) worker();
  else {
    std::vector<std::thread> pool;
    for (uint32_t t = 0; t < threads; ++t) pool.emplace_back(worker);
    for (auto &t : pool) t.join();
  }
  if (work) work[0] = calls.load(), work[1] = calls.load() - n, work[2] = 0;
}

}  // extern "C"
