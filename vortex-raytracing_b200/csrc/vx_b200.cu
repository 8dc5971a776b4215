// vx_b200.cu — the b200 runtime backend: class Device = everything that belongs to ONE B200 (arena, stream, caches, launches),
// DeviceGroup (vx_group.inc) = what a vx_device_h points at, one Device per GPU of the box it spans, and the C ABI of
// include/vortex_b200.h.
//
// Mirrors the role of runtime/simx/vortex.cpp (class vx_device, :49-636) + runtime/common/callbacks.inc
// (the handle/glue layer, :14-225) + runtime/stub/utils.cpp (upload helpers, :25-122) of the reference,
// re-designed for CUDA: one arena allocation backs the Vortex address space, uploads/downloads are stream
// ordered copies, vx_start decodes kernel_arg_t and launches the sm_100a ray-tracing kernels, vx_ready_wait
// is a stream wait.  There is no simulator and no CPU path behind it.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../../include/vortex_b200.h"
#include "address_space.h"
#include "rt_raycast.h"
#include "rt_rtu.h"
#include "rt_build.h"
#include "rt_probe.h"
#include "rt_xfer.h"
#include <algorithm>

#define VXB_EXPORT extern "C" __attribute__((visibility("default")))

namespace vxb {

static thread_local std::string g_last_error;
static int g_selected_device = -1;

static int fail(const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_last_error = buf;
  if (getenv("VORTEX_B200_VERBOSE")) fprintf(stderr, "[VXDRV-b200] %s\n", buf);
  return -1;
}

#define CUDA_TRY(expr)                                                                   \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess) return fail("%s failed: %s", #expr, cudaGetErrorString(_e)); \
  } while (0)

static uint64_t env_u64(const char *name, uint64_t dflt) {
  const char *s = getenv(name);
  if (!s || !*s) return dflt;
  return strtoull(s, nullptr, 0);
}

class Device {
public:
  Device() : space_(0) {}
  ~Device() { shutdown(); }

  int init(int dev) {
    cuda_dev_ = dev;
    CUDA_TRY(cudaSetDevice(cuda_dev_));
    CUDA_TRY(cudaGetDeviceProperties(&prop_, cuda_dev_));
    if (prop_.major < 10 && !getenv("VORTEX_B200_ALLOW_ANY_ARCH"))
      return fail("device %d is sm_%d%d; this backend is built for sm_100a only", cuda_dev_, prop_.major, prop_.minor);
    uint64_t cap = env_u64("VORTEX_B200_MEM_GB", 4) << 30;   // GLOBAL_MEM_SIZE of the XLEN=32 build (common.h:36)
    space_ = AddressSpace(cap);
    CUDA_TRY(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));
    CUDA_TRY(cudaMalloc(&arena_, cap));
    CUDA_TRY(cudaMemsetAsync(arena_, 0, cap, stream_));
    CUDA_TRY(cudaMalloc(&counters_, sizeof(WorkCounters)));
    CUDA_TRY(cudaMalloc(&tile_counter_, sizeof(uint32_t)));
    CUDA_TRY(cudaMallocHost(&h_counters_, sizeof(WorkCounters)));
    CUDA_TRY(cudaEventCreate(&ev_begin_));
    CUDA_TRY(cudaEventCreate(&ev_prep_));
    CUDA_TRY(cudaEventCreate(&ev_end_));
    part_.rank = (uint32_t)env_u64("VORTEX_B200_RANK", 0);
    part_.world = (uint32_t)env_u64("VORTEX_B200_WORLD", 1);
    part_.band_rows = 8;
    if (part_.world == 0) part_.world = 1;
    if (part_.rank >= part_.world) return fail("VORTEX_B200_RANK %u >= VORTEX_B200_WORLD %u", part_.rank, part_.world);
    // defaults of the b200-private DCRs for hosts that cannot write them (an unmodified Vortex application)
    if (getenv("VORTEX_B200_MODE")) dcrs_[VX_DCR_B200_RT_MODE] = (uint32_t)env_u64("VORTEX_B200_MODE", 0);
    if (getenv("VORTEX_B200_ORDER")) dcrs_[VX_DCR_B200_RT_ORDER] = (uint32_t)env_u64("VORTEX_B200_ORDER", 0);
    if (getenv("VORTEX_B200_KERNEL")) dcrs_[VX_DCR_B200_RT_KERNEL] = (uint32_t)env_u64("VORTEX_B200_KERNEL", 0);
    if (getenv("VORTEX_B200_CAMERA")) dcrs_[VX_DCR_B200_RT_CAMERA] = (uint32_t)env_u64("VORTEX_B200_CAMERA", 0);
    CUDA_TRY(cudaStreamSynchronize(stream_));
    return 0;
  }

  void shutdown() {
    if (cuda_dev_ < 0) return;
    cudaSetDevice(cuda_dev_);
    if (stream_) cudaStreamSynchronize(stream_);
    if (remote_base_) cudaIpcCloseMemHandle(remote_base_);
    remote_base_ = nullptr, remote_fb_ = nullptr;
    if (stage_) {
      cudaFreeHost(stage_);
      for (auto &e : stage_ev_) if (e) cudaEventDestroy(e), e = nullptr;
      stage_ = nullptr, stage_next_ = 0;
    }
    build_release_scratch();
    if (gtri_) cudaFree(gtri_);
    if (wide_) cudaFree(wide_);
    if (wide_bases_) cudaFree(wide_bases_);
    if (wide_bases_host_) cudaFreeHost(wide_bases_host_);
    wide_ = nullptr, wide_bases_ = nullptr, wide_bases_host_ = nullptr, wide_valid_ = false;
    if (rtu_nodes_) cudaFree(rtu_nodes_);
    if (rtu_tri_) cudaFree(rtu_tri_);
    if (rtu_bases_) cudaFree(rtu_bases_);
    if (rtu_bases_host_) cudaFreeHost(rtu_bases_host_);
    rtu_bases_ = rtu_bases_host_ = nullptr, rtu_bases_cap_ = 0;
    pipe_free_locked();
    alt_free_locked();
    strips_free_locked();
    ray_order_free_locked();
    sched_free_locked();
    if (counters_) cudaFree(counters_);
    if (tile_counter_) cudaFree(tile_counter_);
    if (h_counters_) cudaFreeHost(h_counters_);
    if (arena_) cudaFree(arena_);
    if (ev_begin_) cudaEventDestroy(ev_begin_);
    if (ev_prep_) cudaEventDestroy(ev_prep_);
    if (ev_end_) cudaEventDestroy(ev_end_);
    if (stream_) cudaStreamDestroy(stream_);
    gtri_ = nullptr, rtu_nodes_ = nullptr, rtu_tri_ = nullptr, counters_ = nullptr, tile_counter_ = nullptr, h_counters_ = nullptr, arena_ = nullptr;
    stream_ = nullptr;
    cuda_dev_ = -1;
  }

  // ---- capabilities (reference ids: vortex.h:30-44; simx answers: runtime/simx/vortex.cpp:76-133) ----
  int get_caps(uint32_t id, uint64_t *value) {
    uint64_t v;
    switch (id) {
      case VX_CAPS_VERSION: v = 0xB200; break;
      case VX_CAPS_NUM_THREADS: v = (uint64_t)prop_.warpSize; break;
      case VX_CAPS_NUM_WARPS: v = (uint64_t)prop_.maxThreadsPerMultiProcessor / prop_.warpSize; break;
      case VX_CAPS_NUM_CORES: v = (uint64_t)prop_.multiProcessorCount; break;
      case VX_CAPS_NUM_CLUSTERS: v = 1; break;
      case VX_CAPS_SOCKET_SIZE: v = 1; break;
      case VX_CAPS_ISSUE_WIDTH: v = 4; break;   // 4 SM sub-partitions issue one warp instruction each per clock
      case VX_CAPS_CACHE_LINE_SIZE: v = 128; break;
      case VX_CAPS_GLOBAL_MEM_SIZE: v = space_.capacity(); break;
      case VX_CAPS_LOCAL_MEM_SIZE: v = (uint64_t)prop_.sharedMemPerBlockOptin; break;
      case VX_CAPS_ISA_FLAGS: v = 0; break;     // not a RISC-V device: no ISA extension bits to advertise
      case VX_CAPS_NUM_MEM_BANKS: v = (uint64_t)(prop_.memoryBusWidth / 64 > 0 ? prop_.memoryBusWidth / 64 : 1); break;
      case VX_CAPS_MEM_BANK_SIZE: v = (uint64_t)prop_.totalGlobalMem / (uint64_t)(prop_.memoryBusWidth / 64 > 0 ? prop_.memoryBusWidth / 64 : 1); break;
      case VX_CAPS_CLOCK_RATE: {
        int khz = 0;
        cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, cuda_dev_);
        v = (uint64_t)khz * 1000ull;
        break;
      }
      case VX_CAPS_PEAK_MEM_BW: {
        int mem_khz = 0;
        cudaDeviceGetAttribute(&mem_khz, cudaDevAttrMemoryClockRate, cuda_dev_);
        v = (uint64_t)(2.0 * mem_khz * 1e3 * (prop_.memoryBusWidth / 8) / 1e6);   // MB/s, DDR
        break;
      }
      default:
        return fail("invalid caps id %u", id);   // simx aborts here (runtime/simx/vortex.cpp:126-130); we report
    }
    *value = v;
    return 0;
  }

  // ---- memory ----
  int mem_alloc(uint64_t size, int flags, uint64_t *addr) {
    std::lock_guard<std::mutex> g(mu_);
    if (space_.allocate(size, flags, addr) != 0) return fail("out of device address space (size=%llu)", (unsigned long long)size);
    return 0;
  }
  int mem_reserve(uint64_t addr, uint64_t size, int flags) {
    std::lock_guard<std::mutex> g(mu_);
    if (space_.reserve(addr, size, flags) != 0)
      return fail("cannot reserve [0x%llx, +%llu)", (unsigned long long)addr, (unsigned long long)size);
    return 0;
  }
  int mem_free(uint64_t addr) {
    std::lock_guard<std::mutex> g(mu_);
    small_shadow_.erase(addr);
    if (space_.release(addr) != 0) return fail("release address not found: 0x%llx", (unsigned long long)addr);
    ++generation_;   // any cached GPU layout keyed on this block is stale now
    return 0;
  }
  int mem_access(uint64_t addr, uint64_t size, int flags) {
    (void)flags;   // ACL bits are recorded by simx's RAM model; CUDA memory has no per-range ACL
    uint64_t asize = AddressSpace::align_up(size, AddressSpace::kBlockAlign);
    if (addr + asize > space_.capacity()) return -1;
    return 0;
  }
  int mem_info(uint64_t *mem_free, uint64_t *mem_used) {
    std::lock_guard<std::mutex> g(mu_);
    if (mem_free) *mem_free = space_.free_bytes();
    if (mem_used) *mem_used = space_.allocated();
    return 0;
  }

  static constexpr uint64_t kShadowMax = 4096;
  static constexpr uint64_t kStageMax = 64 << 10;
  static constexpr int kStageSlots = 32;
  // defer_sync: the caller (DeviceGroup) waits for the copy itself, after it has issued the other devices' copies
  int upload(uint64_t dst, const void *src, uint64_t size, bool defer_sync = false) {
    if (size == 0) return 0;
    std::lock_guard<std::mutex> g(mu_);
    if (dst + size > space_.capacity() || dst + size < dst) return fail("upload out of range");
    CUDA_TRY(cudaSetDevice(cuda_dev_));
    uint64_t start = 0;
    if (auto *b = space_.find(dst, size, &start)) {
      b->write_gen = ++generation_;
      // The host shadow exists only while it is known to hold EVERY byte of the block: it is created by an upload that
      // covers the whole block (blocks of at most kShadowMax bytes), patched by later partial uploads, and dropped by
      // anything else that writes the block (a partial upload without a complete shadow, a GPU build, mark_written).
      auto sh = small_shadow_.find(start);
      if (dst == start && size == b->size && size <= kShadowMax) {
        auto &v = small_shadow_[start];
        v.assign((const uint8_t *)src, (const uint8_t *)src + size);
      } else if (sh != small_shadow_.end()) {
        if (sh->second.size() == b->size && dst - start + size <= sh->second.size()) memcpy(sh->second.data() + (dst - start), src, size);
        else small_shadow_.erase(sh);
      }
    }
    if (size <= kStageMax) {
      // small uploads (argument blocks, TLAS / instance records) are copied into a page-locked ring first: the caller's
      // memory has been read when this returns, the device copy is stream-ordered, and nothing waits for the GPU
      if (!stage_) {
        CUDA_TRY(cudaMallocHost(&stage_, kStageMax * kStageSlots));
        for (auto &e : stage_ev_) CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      }
      const int slot = stage_next_++ % kStageSlots;
      if (stage_next_ > kStageSlots) CUDA_TRY(cudaEventSynchronize(stage_ev_[slot]));   // the copy that last used this slot
      memcpy(stage_ + (uint64_t)slot * kStageMax, src, size);
      CUDA_TRY(cudaMemcpyAsync(arena_ + dst, stage_ + (uint64_t)slot * kStageMax, size, cudaMemcpyHostToDevice, stream_));
      CUDA_TRY(cudaEventRecord(stage_ev_[slot], stream_));
      return 0;
    }
    CUDA_TRY(cudaMemcpyAsync(arena_ + dst, src, size, cudaMemcpyHostToDevice, stream_));
    if (!defer_sync) CUDA_TRY(cudaStreamSynchronize(stream_));
    return 0;
  }
  // block containing `addr` (DeviceGroup replicates whole blocks)
  bool block_of(uint64_t addr, uint64_t *start, uint64_t *size) {
    std::lock_guard<std::mutex> g(mu_);
    auto *b = space_.find(addr, 1, start);
    if (!b) return false;
    *size = b->size;
    return true;
  }
  int download(void *dst, uint64_t src, uint64_t size) {
    if (size == 0) return 0;
    std::lock_guard<std::mutex> g(mu_);
    if (src + size > space_.capacity() || src + size < src) return fail("download out of range");
    CUDA_TRY(cudaSetDevice(cuda_dev_));
    CUDA_TRY(cudaMemcpyAsync(dst, arena_ + src, size, cudaMemcpyDeviceToHost, stream_));
    CUDA_TRY(cudaStreamSynchronize(stream_));
    return finish_run_locked();
  }

  // ---- launch ----
  int start(uint64_t krnl_addr, uint64_t args_addr) {
    (void)krnl_addr;   // the RISC-V image is accepted and ignored
    std::lock_guard<std::mutex> g(mu_);
    CUDA_TRY(cudaSetDevice(cuda_dev_));
    std::vector<uint8_t> args;
    if (fetch_args_locked(args_addr, args) != 0) return -1;
    uint32_t layout = layout_of_locked(args_addr, args.size());
    if (layout == 1) return start_raycast_locked(args, nullptr, nullptr, 0, 0);
    if (layout == 2) return start_rtu_locked(args);
    return fail("vx_start: argument block of %zu bytes is neither a raycast (192 B) nor a raytracing (216 B) kernel_arg_t", args.size());
  }

  int trace(uint64_t args_addr, const uint64_t *rays_addr, uint64_t count, uint64_t hits_addr) {
    std::lock_guard<std::mutex> g(mu_);
    CUDA_TRY(cudaSetDevice(cuda_dev_));
    std::vector<uint8_t> args;
    if (fetch_args_locked(args_addr, args) != 0) return -1;
    const uint32_t layout = layout_of_locked(args_addr, args.size());
    if (layout == 2) return trace_rtu_locked(args, rays_addr, count, hits_addr);
    if (layout != 1) return fail("trace: argument block too small");
    return start_raycast_locked(args, &hits_addr, rays_addr, count, 1);
  }

  // Host-resident rays in, host-resident hit records out: the same result as vx_copy_to_dev(rays) + vx_b200_trace_rays +
  // vx_ready_wait + vx_copy_from_dev(hits), with the three legs of consecutive chunks overlapped (PCIe is full duplex):
  //   upload stream:   H2D(i+1)            compute streams:   trace kernels(i-3 .. i)    download stream:   D2H(i-1 ..)
  // kPipeSlots staging slots; a slot's rays may be overwritten once its kernel finished, its hits once they were downloaded.
  // A launch over incoherent rays ends with a ~0.15 ms tail (its slowest rays, profiles/r01_sweep_host_trace.txt), so the
  // chunks rotate over kRunCtxs run contexts (stream + counters + events each): one chunk's tail runs under the next ones.
  int trace_host(uint64_t args_addr, const void *host_rays, uint64_t count, void *host_hits) {
    std::lock_guard<std::mutex> g(mu_);
    CUDA_TRY(cudaSetDevice(cuda_dev_));
    std::vector<uint8_t> args;
    if (fetch_args_locked(args_addr, args) != 0) return -1;
    const uint32_t layout = layout_of_locked(args_addr, args.size());
    const bool layout_b = layout == 2;
    if (layout == 0) return fail("trace: argument block too small");
    uint64_t chunk = env_u64("VORTEX_B200_HOST_CHUNK", 1ull << 20);   // sweep: profiles/r01_sweep_host_trace.txt
    if (chunk < 4096 || chunk > (1ull << 24)) chunk = 1ull << 20;
    if (chunk > count) chunk = (count + 31) & ~31ull;
    if (run_pending_) {
      CUDA_TRY(cudaStreamSynchronize(stream_));
      if (finish_run_locked() != 0) return -1;
    }
    if (pipe_init_locked(chunk) != 0) return -1;
    const uint64_t n_chunks = (count + chunk - 1) / chunk;
    const uint8_t *src = (const uint8_t *)host_rays;
    uint8_t *dst = (uint8_t *)host_hits;
    auto upload = [&](uint64_t i) -> int {
      int slot = (int)(i % kPipeSlots);
      uint64_t off = i * chunk, n = std::min(chunk, count - off);
      CUDA_TRY(cudaStreamWaitEvent(pipe_.s_up, pipe_.done[slot], 0));   // the kernel that last read this slot's rays finished
      CUDA_TRY(cudaMemcpyAsync(pipe_.rays[slot], src + off * 24, n * 24, cudaMemcpyHostToDevice, pipe_.s_up));
      CUDA_TRY(cudaEventRecord(pipe_.up[slot], pipe_.s_up));
      return 0;
    };
    vx_b200_stats_t total{};
    auto settle = [&]() -> int {   // wait for the chunk in flight on the current run context and add its counters
      CUDA_TRY(cudaStreamSynchronize(stream_));
      if (finish_run_locked() != 0) return -1;
      total.rays += stats_.rays, total.box_tests += stats_.box_tests, total.tri_tests += stats_.tri_tests;
      total.blas_entered += stats_.blas_entered, total.shaded_hits += stats_.shaded_hits;
      total.kernel_launches += stats_.kernel_launches, total.prep_launches += stats_.prep_launches;
      total.kernel_ms += stats_.kernel_ms, total.prep_ms += stats_.prep_ms;
      total.lane_steps += stats_.lane_steps, total.warp_steps += stats_.warp_steps, total.warps += stats_.warps;
      for (int a = 0; a < 15; a++) (&total.latency_dist[0][0])[a] += (&stats_.latency_dist[0][0])[a];
      for (int a = 0; a < 16; a++) total.steps_hist[a] += stats_.steps_hist[a];
      return 0;
    };
    auto body = [&]() -> int {
      if (upload(0) != 0) return -1;
      for (uint64_t i = 0; i < n_chunks; i++) {
        int slot = (int)(i % kPipeSlots);
        uint64_t off = i * chunk, n = std::min(chunk, count - off);
        if (i + 1 < n_chunks && upload(i + 1) != 0) return -1;
        use_ctx_locked((int)(i % kRunCtxs));            // chunks rotate over the run contexts
        if (run_pending_ && settle() != 0) return -1;   // chunk i - kRunCtxs
        CUDA_TRY(cudaStreamWaitEvent(stream_, pipe_.up[slot], 0));
        CUDA_TRY(cudaStreamWaitEvent(stream_, pipe_.down[slot], 0));    // the hit records that last sat in this slot were downloaded
        pipe_rays_ = pipe_.rays[slot], pipe_hits_ = pipe_.hits[slot];
        uint64_t zero = 0;
        int rc = layout_b ? trace_rtu_locked(args, &zero, n, 0) : start_raycast_locked(args, &zero, &zero, n, 1);
        pipe_rays_ = nullptr, pipe_hits_ = nullptr;
        if (rc != 0) return -1;
        if (i == 0)   // scene records (re)built by chunk 0 are visible to the other contexts
          for (int c = 1; c < kRunCtxs; c++) CUDA_TRY(cudaStreamWaitEvent(ctx_[c].stream, ev_prep_, 0));
        CUDA_TRY(cudaEventRecord(pipe_.done[slot], stream_));
        CUDA_TRY(cudaStreamWaitEvent(pipe_.s_down, pipe_.done[slot], 0));
        CUDA_TRY(cudaMemcpyAsync(dst + off * sizeof(HitA), pipe_.hits[slot], n * sizeof(HitA), cudaMemcpyDeviceToHost, pipe_.s_down));
        CUDA_TRY(cudaEventRecord(pipe_.down[slot], pipe_.s_down));
      }
      return 0;
    };
    int rc = body();
    for (int c = kRunCtxs - 1; c >= 0; c--) {   // ends on the device's own context
      use_ctx_locked(c);
      if (run_pending_ && settle() != 0) rc = -1;
    }
    cudaStreamSynchronize(pipe_.s_up);
    CUDA_TRY(cudaStreamSynchronize(pipe_.s_down));
    if (rc != 0) return -1;
    stats_ = total;   // kernel_ms is the sum over chunks; consecutive chunks overlap, so it exceeds the wall time of the call
    return 0;
  }

  int ready_wait(uint64_t timeout_ms) {
    std::lock_guard<std::mutex> g(mu_);
    if (!run_pending_) return 0;
    CUDA_TRY(cudaSetDevice(cuda_dev_));
    auto t0 = std::chrono::steady_clock::now();
    for (;;) {
      cudaError_t q = cudaEventQuery(ev_end_);
      if (q == cudaSuccess) break;
      if (q != cudaErrorNotReady) return fail("kernel failed: %s", cudaGetErrorString(q));
      auto ms = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count();
      if ((uint64_t)ms >= timeout_ms) return -1;   // reference: -1 on timeout (runtime/simx/vortex.cpp:350-364)
      if (ms < 2) std::this_thread::yield();
      else std::this_thread::sleep_for(std::chrono::microseconds(50));
    }
    return finish_run_locked();
  }

  int dcr_write(uint32_t addr, uint32_t value) {
    std::lock_guard<std::mutex> g(mu_);
    // the reference serialises a DCR write behind a running kernel (runtime/simx/vortex.cpp:366-373);
    // DCRs are sampled at vx_start here, so a write never races a kernel in flight.
    dcrs_[addr] = value;
    return 0;
  }
  int dcr_read(uint32_t addr, uint32_t *value) {
    std::lock_guard<std::mutex> g(mu_);
    auto it = dcrs_.find(addr);
    if (it == dcrs_.end()) return fail("DCR 0x%x was never written", addr);
    *value = it->second;
    return 0;
  }
  int mpm_query(uint32_t addr, uint32_t core_id, uint64_t *value) {
    std::lock_guard<std::mutex> g(mu_);
    uint32_t off = addr - VX_CSR_MPM_BASE;
    if (off > 31) return -1;
    if (core_id >= (uint32_t)prop_.multiProcessorCount) return -1;
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, cuda_dev_);
    if (addr == VX_CSR_MCYCLE) *value = (uint64_t)((double)stats_.kernel_ms * khz);                      // elapsed SM clocks
    else if (addr == VX_CSR_MINSTRET) *value = stats_.rays / (uint64_t)prop_.multiProcessorCount;        // "retired" = rays per SM
    else *value = 0;
    return 0;
  }

  // ---- extensions ----
  int set_partition(uint32_t rank, uint32_t world, uint32_t band_rows) {
    std::lock_guard<std::mutex> g(mu_);
    if (world == 0 || rank >= world) return fail("bad partition rank=%u world=%u", rank, world);
    if (band_rows == 0 || band_rows % 4 != 0) return fail("band_rows must be a positive multiple of 4");
    part_.rank = rank, part_.world = world, part_.band_rows = band_rows;
    return 0;
  }
  int build_bvh(uint64_t tri_addr, uint64_t tri_count, uint64_t tri_first, uint64_t bvh_addr, uint64_t idx_addr, uint64_t *nodes_used) {
    std::lock_guard<std::mutex> g(mu_);
    CUDA_TRY(cudaSetDevice(cuda_dev_));
    if (tri_count == 0 || tri_count > 0x7FFFFFFFull) return fail("vx_b200_build_bvh: bad triangle count");
    if (run_pending_) {
      CUDA_TRY(cudaStreamSynchronize(stream_));
      if (finish_run_locked() != 0) return -1;
    }
    const TriA *tri = nullptr;
    BvhNodeA *nodes = nullptr;
    uint32_t *idx = nullptr;
    if (resolve(tri_addr, tri_count * sizeof(TriA), 4, "tri", &tri)) return -1;
    if (resolve(bvh_addr, (2 * tri_count - 1) * sizeof(BvhNodeA), 16, "bvh", &nodes)) return -1;
    if (resolve(idx_addr, tri_count * 4, 4, "triIdx", &idx)) return -1;
    int nl = 0;
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0));
    CUDA_TRY(cudaEventCreate(&e1));
    CUDA_TRY(cudaEventRecord(e0, stream_));
    const char *ct = getenv("VORTEX_B200_BUILD_CTRAV");
    const float c_trav = ct && *ct ? (float)atof(ct) : 1.0f;
    int rc = build_lbvh(tri, (uint32_t)tri_count, nodes, idx, (uint32_t)tri_first, stream_, &nl, c_trav,
                        (uint32_t)env_u64("VORTEX_B200_BUILD_MAX_LEAF", 4));
    cudaEventRecord(e1, stream_);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0), cudaEventDestroy(e1);
    if (rc != 0) return fail("vx_b200_build_bvh: %s", cudaGetErrorString((cudaError_t)(-rc)));
    // the buffers changed under the caches keyed on upload generations
    uint64_t start = 0;
    if (auto *b = space_.find(bvh_addr, 1, &start)) b->write_gen = ++generation_, small_shadow_.erase(start);
    if (auto *b = space_.find(idx_addr, 1, &start)) b->write_gen = ++generation_, small_shadow_.erase(start);
    stats_.kernel_launches = (uint64_t)nl, stats_.prep_launches = 0, stats_.kernel_ms = ms, stats_.prep_ms = 0;
    if (nodes_used) *nodes_used = 2 * tri_count - 1;
    return 0;
  }
  // vx_b200_build_bvh_static: build_bvh with the builder's scratch in a static device array — no allocation of any kind here
  // (no events either: the statistics of this call carry no time)
  int build_bvh_static(uint64_t tri_addr, uint64_t tri_count, uint64_t tri_first, uint64_t bvh_addr, uint64_t idx_addr, uint64_t *nodes_used) {
    std::lock_guard<std::mutex> g(mu_);
    CUDA_TRY(cudaSetDevice(cuda_dev_));
    if (tri_count == 0) return fail("vx_b200_build_bvh_static: bad triangle count");
    if (tri_count > kBuildStaticMaxTris)
      return fail("vx_b200_build_bvh_static: %llu triangles exceed the static scratch array's limit of %u (use vx_b200_build_bvh)",
                  (unsigned long long)tri_count, kBuildStaticMaxTris);
    if (run_pending_) {
      CUDA_TRY(cudaStreamSynchronize(stream_));
      if (finish_run_locked() != 0) return -1;
    }
    const TriA *tri = nullptr;
    BvhNodeA *nodes = nullptr;
    uint32_t *idx = nullptr;
    if (resolve(tri_addr, tri_count * sizeof(TriA), 4, "tri", &tri)) return -1;
    if (resolve(bvh_addr, (2 * tri_count - 1) * sizeof(BvhNodeA), 16, "bvh", &nodes)) return -1;
    if (resolve(idx_addr, tri_count * 4, 4, "triIdx", &idx)) return -1;
    int nl = 0;
    const char *ct = getenv("VORTEX_B200_BUILD_CTRAV");
    const float c_trav = ct && *ct ? (float)atof(ct) : 1.0f;
    int rc = build_lbvh_static(tri, (uint32_t)tri_count, nodes, idx, (uint32_t)tri_first, stream_, &nl, c_trav,
                               (uint32_t)env_u64("VORTEX_B200_BUILD_MAX_LEAF", 4));
    if (rc != 0) return fail("vx_b200_build_bvh_static: %s", cudaGetErrorString((cudaError_t)(-rc)));
    uint64_t start = 0;
    if (auto *b = space_.find(bvh_addr, 1, &start)) b->write_gen = ++generation_, small_shadow_.erase(start);
    if (auto *b = space_.find(idx_addr, 1, &start)) b->write_gen = ++generation_, small_shadow_.erase(start);
    stats_.kernel_launches = (uint64_t)nl, stats_.prep_launches = 0, stats_.kernel_ms = 0, stats_.prep_ms = 0;
    if (nodes_used) *nodes_used = 2 * tri_count - 1;
    return 0;
  }
  int build_qbvh(uint64_t tri_addr, uint64_t tri_count, uint64_t tri_first, uint64_t ex_addr, uint64_t qbvh_addr, uint64_t capacity,
                 uint64_t *nodes_used, float *root_box6) {
    std::lock_guard<std::mutex> g(mu_);
    CUDA_TRY(cudaSetDevice(cuda_dev_));
    if (tri_count == 0 || tri_count > 0x7FFFFFFFull) return fail("vx_b200_build_qbvh: bad triangle count");
    if (run_pending_) {
      CUDA_TRY(cudaStreamSynchronize(stream_));
      if (finish_run_locked() != 0) return -1;
    }
    TriA *tri = nullptr;
    TriExB *ex = nullptr;
    QNodeB *q = nullptr;
    if (resolve(tri_addr, tri_count * sizeof(TriA), 4, "tri", &tri)) return -1;
    if (ex_addr != ~0ull && resolve(ex_addr, tri_count * sizeof(TriExB), 4, "triEx", &ex)) return -1;
    if (resolve(qbvh_addr, capacity * sizeof(QNodeB), 4, "qbvh", &q)) return -1;
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0));
    CUDA_TRY(cudaEventCreate(&e1));
    cudaEventRecord(e0, stream_);
    int nl = 0;
    uint32_t used = 0;
    int rc = build_qbvh4(tri, ex, (uint32_t)tri_count, (uint32_t)tri_first, q, capacity, stream_, &used, root_box6, &nl);
    cudaEventRecord(e1, stream_);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0), cudaEventDestroy(e1);
    if (rc != 0) return fail("vx_b200_build_qbvh: %s (capacity %llu nodes, needed %u)", cudaGetErrorString((cudaError_t)(-rc)),
                             (unsigned long long)capacity, used);
    uint64_t start = 0;   // the buffers changed under the caches keyed on upload generations
    if (auto *b = space_.find(tri_addr, 1, &start)) b->write_gen = ++generation_, small_shadow_.erase(start);
    if (ex_addr != ~0ull)
      if (auto *b = space_.find(ex_addr, 1, &start)) b->write_gen = ++generation_, small_shadow_.erase(start);
    if (auto *b = space_.find(qbvh_addr, 1, &start)) b->write_gen = ++generation_, small_shadow_.erase(start);
    stats_.kernel_launches = (uint64_t)nl, stats_.prep_launches = 0, stats_.kernel_ms = ms, stats_.prep_ms = 0;
    if (nodes_used) *nodes_used = used;
    return 0;
  }
  // the buffer's device memory was written behind the runtime's back (NCCL receive, a peer's store): bump its upload
  // generation so the caches keyed on it rebuild, and drop the host shadow of small uploads
  // roofline denominator: random-record gather bandwidth out of L2 (or HBM, for a buffer beyond L2); diagnostics only
  int probe_gather(uint64_t buffer_bytes, uint32_t record_bytes, double *gbs) {
    std::lock_guard<std::mutex> g(mu_);
    CUDA_TRY(cudaSetDevice(cuda_dev_));
    if (finish_run_locked() != 0) return -1;
    cudaError_t e = vxb::probe_gather(buffer_bytes, record_bytes, prop_.multiProcessorCount, stream_, gbs);
    if (e != cudaSuccess) return fail("vx_b200_probe_gather: %s", cudaGetErrorString(e));
    return 0;
  }
  int mark_written(uint64_t addr) {
    std::lock_guard<std::mutex> g(mu_);
    uint64_t start = 0;
    auto *b = space_.find(addr, 1, &start);
    if (!b) return fail("vx_b200_mark_written: not a live buffer");
    b->write_gen = ++generation_;
    small_shadow_.erase(start);
    return 0;
  }
  int set_remote_framebuffer(const void *handle, uint64_t offset, uint64_t bytes) {
    std::lock_guard<std::mutex> g(mu_);
    CUDA_TRY(cudaSetDevice(cuda_dev_));
    if (run_pending_) {
      CUDA_TRY(cudaStreamSynchronize(stream_));
      if (finish_run_locked() != 0) return -1;
    }
    if (remote_base_) {
      cudaIpcCloseMemHandle(remote_base_);
      remote_base_ = nullptr, remote_fb_ = nullptr, remote_bytes_ = 0;
    }
    if (!handle) return 0;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void *base = nullptr;
    CUDA_TRY(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
    remote_base_ = base;
    remote_fb_ = reinterpret_cast<uint32_t *>(reinterpret_cast<uint8_t *>(base) + offset);
    remote_bytes_ = bytes;
    return 0;
  }
  // ---- hooks for DeviceGroup (one vx_device_h over several GPUs) ----
  struct Output {            // what the run in flight / the last run wrote
    uint64_t addr = 0, bytes = 0;
    uint32_t W = 0, H = 0;   // frames: image size (bands of part_); hit records: W = H = 0
    bool frame = false, valid = false;
  };
  Output last_output() {
    std::lock_guard<std::mutex> g(mu_);
    return last_out_;
  }
  Partition partition() {
    std::lock_guard<std::mutex> g(mu_);
    return part_;
  }
  // frames are stored into `owner_arena + dst_addr` (a peer device's arena, same process) instead of this device's arena
  void set_peer_framebuffer(uint8_t *owner_arena) {
    std::lock_guard<std::mutex> g(mu_);
    peer_fb_arena_ = owner_arena;
  }
  // Asynchronous readback (vx_b200_set_readback): frames whose framebuffer is `fb_addr` are followed by the copy of this
  // GPU's bands into page-locked host memory at `host` (the image's first byte).  When the frame runs in row-major tile order
  // with the ray state machine, it is cut into kStrips strips launched alternately on two streams, and the bands of a finished
  // strip are copied (copy stream) while the next strips are traced.
  void set_readback_target(uint64_t fb_addr, uint8_t *host, uint64_t bytes) {
    std::lock_guard<std::mutex> g(mu_);
    rb_addr_ = fb_addr, rb_host_ = host, rb_bytes_ = bytes;
  }
  bool readback_issued() {
    std::lock_guard<std::mutex> g(mu_);
    return rb_issued_;
  }
  // bands [b0, b1) of this GPU's own bands (indices into ITS band list) of a W x H image at arena offset `fb`: one strided copy
  int copy_bands(uint8_t *base, uint64_t fb, uint32_t W, uint32_t H, uint32_t b0, uint32_t b1, cudaStream_t st) {
    const Partition &p = part_;
    const uint64_t row = (uint64_t)W * 4, band = row * p.band_rows, pitch = band * p.world;
    const uint32_t bands_total = (H + p.band_rows - 1) / p.band_rows;
    const uint32_t mine = bands_total > p.rank ? (bands_total - p.rank + p.world - 1) / p.world : 0;
    if (b1 > mine) b1 = mine;
    if (b0 >= b1) return 0;
    const uint8_t *src = arena_ + fb;
    // the last band of the image may be short; it belongs to this GPU iff (bands_total - 1) % world == rank
    const bool short_last = (H % p.band_rows) != 0 && (bands_total - 1) % p.world == p.rank && b1 == mine;
    const uint32_t full = (b1 - b0) - (short_last ? 1u : 0u);
    const uint64_t o = band * ((uint64_t)p.rank + (uint64_t)b0 * p.world);
    if (full == 1 || (p.world == 1 && full)) {
      CUDA_TRY(cudaMemcpyAsync(base + o, src + o, band * (p.world == 1 ? full : 1), cudaMemcpyDefault, st));
    } else if (full) {
      CUDA_TRY(cudaMemcpy2DAsync(base + o, pitch, src + o, pitch, band, full, cudaMemcpyDefault, st));
    }
    if (short_last) {
      const uint64_t lo = band * (uint64_t)(bands_total - 1);
      CUDA_TRY(cudaMemcpyAsync(base + lo, src + lo, row * (H % p.band_rows), cudaMemcpyDefault, st));
    }
    return 0;
  }
  cudaEvent_t end_event() const { return ev_end_; }
  bool run_pending() {
    std::lock_guard<std::mutex> g(mu_);
    return run_pending_;
  }
  // a range of the arena was (or is being, stream-ordered) written by the group: same bookkeeping as an upload
  int note_written(uint64_t dst, uint64_t size) {
    std::lock_guard<std::mutex> g(mu_);
    uint64_t start = 0;
    if (auto *b = space_.find(dst, size ? size : 1, &start)) {
      b->write_gen = ++generation_;
      small_shadow_.erase(start);
    }
    return 0;
  }
  const cudaDeviceProp &prop() const { return prop_; }
  uint64_t capacity() const { return space_.capacity(); }
  uint8_t *arena() const { return arena_; }
  cudaStream_t stream() const { return stream_; }
  int cuda_device() const { return cuda_dev_; }
  void enable_stats(int on) { stats_on_ = on; }
  int get_stats(vx_b200_stats_t *s) {
    std::lock_guard<std::mutex> g(mu_);
    *s = stats_;
    return 0;
  }
  float last_ms() const { return stats_.kernel_ms; }

private:
  uint32_t dcr_get(uint32_t addr) const {
    auto it = dcrs_.find(addr);
    return it == dcrs_.end() ? 0u : it->second;
  }

  // Which kernel_arg_t the argument block holds: 1 = raycast (layout A, 192 B), 2 = raytracing / RT unit (layout B, 216 B),
  // 0 = neither.  The reference dispatches by the kernel image; here, in this order: the explicit VX_DCR_B200_RT_LAYOUT; a
  // host that programmed the RT unit's scene-pointer DCRs 0x006-0x009 (every raytracing host does: tracer.cpp:252-256) is
  // layout B; an argument buffer of exactly 216 bytes is layout B; anything else of at least 192 bytes is layout A (a raycast
  // host may allocate its argument buffer with headroom).  `avail` is only a lower-bound check.
  uint32_t layout_of_locked(uint64_t args_addr, size_t avail) {
    uint32_t layout = dcr_get(VX_DCR_B200_RT_LAYOUT);
    if (layout == 0) {
      const bool rtx_dcrs = dcrs_.count(VX_DCR_BASE_RTX_TLAS_PTR) && dcrs_.count(VX_DCR_BASE_RTX_BVH_PTR);
      uint64_t start = 0;
      auto *b = space_.find(args_addr, 1, &start);
      const bool exact_b = b && b->size - (args_addr - start) == sizeof(KernelArgB);
      layout = (rtx_dcrs || exact_b) ? 2 : 1;
    }
    if (layout == 2 && avail < sizeof(KernelArgB)) return 0;
    if (layout == 1 && avail < sizeof(KernelArgA)) return 0;
    return layout <= 2 ? layout : 0;
  }

  int fetch_args_locked(uint64_t args_addr, std::vector<uint8_t> &out) {
    uint64_t start = 0;
    auto *b = space_.find(args_addr, 1, &start);
    if (!b) return fail("arguments address 0x%llx is not inside a live buffer", (unsigned long long)args_addr);
    uint64_t avail = b->size - (args_addr - start);
    if (avail > 4096) avail = 4096;
    out.resize(avail);
    auto it = small_shadow_.find(start);
    if (it != small_shadow_.end() && it->second.size() >= (args_addr - start) + avail) {
      memcpy(out.data(), it->second.data() + (args_addr - start), avail);   // host copy kept at upload time: no device round trip
    } else {
      CUDA_TRY(cudaMemcpyAsync(out.data(), arena_ + args_addr, avail, cudaMemcpyDeviceToHost, stream_));
      CUDA_TRY(cudaStreamSynchronize(stream_));
    }
    return 0;
  }

  // resolve a Vortex address to a device pointer, requiring `need` bytes inside one live buffer
  template <class T>
  int resolve(uint64_t addr, uint64_t need, uint64_t align, const char *what, T **out, uint64_t *avail = nullptr, uint64_t *gen = nullptr) {
    uint64_t start = 0;
    auto *b = space_.find(addr, need ? need : 1, &start);
    if (!b) return fail("%s address 0x%llx (+%llu) is not inside a live buffer", what, (unsigned long long)addr, (unsigned long long)need);
    if (align && (addr % align) != 0) return fail("%s address 0x%llx is not %llu-byte aligned", what, (unsigned long long)addr, (unsigned long long)align);
    *out = reinterpret_cast<T *>(arena_ + addr);
    if (avail) *avail = b->size - (addr - start);
    if (gen) *gen = b->write_gen;
    return 0;
  }

  int begin_run_locked() {
    if (run_pending_) {   // "ensure prior run completed" (runtime/simx/vortex.cpp:331-333)
      CUDA_TRY(cudaStreamSynchronize(stream_));
      if (finish_run_locked() != 0) return -1;
    }
    CUDA_TRY(cudaMemsetAsync(counters_, 0, sizeof(WorkCounters), stream_));
    CUDA_TRY(cudaMemsetAsync(tile_counter_, 0, sizeof(uint32_t), stream_));
    launches_ = 0, prep_launches_ = 0;
    rb_issued_ = false, kend_valid_ = false, half_valid_ = false;
    CUDA_TRY(cudaEventRecord(ev_begin_, stream_));
    return 0;
  }
  int end_run_locked() {
    CUDA_TRY(cudaEventRecord(ev_end_, stream_));
    if (sched_active_ >= 0) {   // group costs of this frame, read by the next sched_prepare_locked
      SchedState &st = sched_[sched_active_];
      CUDA_TRY(cudaMemcpyAsync(st.h_cost, st.d_cost, st.n_groups * 4, cudaMemcpyDeviceToHost, stream_));
      st.pending = true;
      sched_active_ = -1;
    }
    CUDA_TRY(cudaMemcpyAsync(h_counters_, counters_, sizeof(WorkCounters), cudaMemcpyDeviceToHost, stream_));
    run_pending_ = true;
    return 0;
  }
  int finish_run_locked() {
    if (!run_pending_) return 0;
    CUDA_TRY(cudaStreamSynchronize(stream_));
    run_pending_ = false;
    float ms_total = 0, ms_prep = 0;
    CUDA_TRY(cudaEventElapsedTime(&ms_total, ev_begin_, kend_valid_ ? ev_kend_ : ev_end_));
    if (kend_valid_ && half_valid_) {   // strip-pipelined frame: share of the kernel time taken by the strips that went first
      float ms_half = 0, ms_all = 0;
      CUDA_TRY(cudaEventElapsedTime(&ms_half, ev_prep_, ev_half_));
      CUDA_TRY(cudaEventElapsedTime(&ms_all, ev_prep_, ev_kend_));
      if (ms_all > 0 && ms_half / ms_all > strip_half_share_ + 0.08f) strip_reverse_ = !strip_reverse_;
    }
    kend_valid_ = false, half_valid_ = false;
    CUDA_TRY(cudaEventElapsedTime(&ms_prep, ev_begin_, ev_prep_));
    stats_.rays = h_counters_->rays;
    stats_.box_tests = h_counters_->box_tests;
    stats_.tri_tests = h_counters_->tri_tests;
    stats_.blas_entered = h_counters_->blas_entered;
    stats_.shaded_hits = h_counters_->shaded_hits;
    stats_.lane_steps = h_counters_->lane_steps;
    stats_.warp_steps = h_counters_->warp_steps;
    stats_.warps = h_counters_->warps;
    for (int a = 0; a < 3; a++)
      for (int b = 0; b < 5; b++) stats_.latency_dist[a][b] = h_counters_->latency[a * 5 + b];
    for (int a = 0; a < 16; a++) stats_.steps_hist[a] = h_counters_->steps_hist[a];
    if (h_counters_->errors) return fail("run failed on the device (error bits 0x%llx: 1 traversal stack overflow, 2 traceRay nesting too deep, 4 COMMIT_CONT in strict order)", (unsigned long long)h_counters_->errors);
    stats_.kernel_launches = launches_;
    stats_.prep_launches = prep_launches_;
    stats_.prep_ms = ms_prep;
    stats_.kernel_ms = ms_total - ms_prep;
    if (sched_timed_ >= 0) {   // the two halves of the schedule trial
      (sched_timed_phase_ == 1 ? sched_[sched_timed_].ms_identity : sched_[sched_timed_].ms_ordered) = stats_.kernel_ms;
      sched_timed_ = -1;
    }
    return 0;
  }

  // layout A: decode kernel_arg_t, (re)build the GPU triangle records if the scene changed, launch
  int start_raycast_locked(const std::vector<uint8_t> &args, const uint64_t *hits_addr, const uint64_t *rays_addr, uint64_t count, int trace) {
    KernelArgA a;
    memcpy(&a, args.data(), sizeof(a));
    if (a.dst_width == 0 || a.dst_height == 0) return fail("empty framebuffer %ux%u", a.dst_width, a.dst_height);
    SceneA sc{};
    const TriA *tri = nullptr;
    const uint32_t *triIdx = nullptr;
    uint64_t tri_bytes = 0, idx_bytes = 0, tri_gen = 0, idx_gen = 0;
    if (resolve(a.tri_addr, sizeof(TriA), 4, "tri", &tri, &tri_bytes, &tri_gen)) return -1;
    if (resolve(a.triIdx_addr, 4, 4, "triIdx", &triIdx, &idx_bytes, &idx_gen)) return -1;
    if (resolve(a.triEx_addr, sizeof(TriExA), 4, "triEx", &sc.triEx)) return -1;
    if (resolve(a.bvh_addr, sizeof(BvhNodeA), 16, "bvh", &sc.nodes)) return -1;
    if (resolve(a.blas_addr, sizeof(BlasNodeA), 8, "blas", &sc.blas)) return -1;
    if (resolve(a.tlas_addr, sizeof(TlasNodeA), 16, "tlas", &sc.tlas)) return -1;
    if (resolve(a.tex_addr, 4, 4, "tex", &sc.tex)) return -1;
    sc.tlas_root = a.tlas_root;
    uint32_t *out = nullptr;
    HitA *hits = nullptr;
    const float *rays = nullptr;
    if (!trace) {
      if (resolve(a.dst_addr, (uint64_t)a.dst_width * a.dst_height * 4, 4, "framebuffer", &out)) return -1;
      if (remote_fb_) {   // store this rank's bands straight into the image owner's framebuffer (peer memory)
        if (remote_bytes_ < (uint64_t)a.dst_width * a.dst_height * 4) return fail("remote framebuffer smaller than the image");
        out = remote_fb_;
      } else if (peer_fb_arena_) {
        out = reinterpret_cast<uint32_t *>(peer_fb_arena_ + a.dst_addr);
      }
      last_out_ = Output{a.dst_addr, (uint64_t)a.dst_width * a.dst_height * 4, a.dst_width, a.dst_height, true, true};
    } else {
      uint64_t n = rays_addr ? count : (uint64_t)a.dst_width * a.dst_height;
      if (pipe_hits_) {   // chunk of vx_b200_trace_rays_host: staging buffers outside the arena
        hits = pipe_hits_, rays = pipe_rays_;
      } else {
        if (resolve(*hits_addr, n * sizeof(HitA), 4, "hits", &hits)) return -1;
        if (rays_addr && resolve(*rays_addr, n * 24, 4, "rays", &rays)) return -1;
        last_out_ = Output{*hits_addr, n * sizeof(HitA), 0, 0, false, true};
      }
      count = n;
    }
    if (begin_run_locked() != 0) return -1;
    // scene re-flatten, cached on (addresses, last-upload generations)
    uint64_t n_slots = idx_bytes / 4, n_tris = tri_bytes / sizeof(TriA);
    if (!gtri_ || gtri_key_[0] != a.tri_addr || gtri_key_[1] != a.triIdx_addr || gtri_key_[2] != tri_gen || gtri_key_[3] != idx_gen ||
        gtri_slots_ != n_slots) {
      if (gtri_cap_ < n_slots) {
        if (gtri_) cudaFree(gtri_);
        gtri_ = nullptr, gtri_cap_ = 0;
        CUDA_TRY(cudaMalloc(&gtri_, n_slots * sizeof(GpuTri)));
        gtri_cap_ = n_slots;
      }
      CUDA_TRY(launch_gather_tris(tri, triIdx, gtri_, n_slots, n_tris, stream_));
      gtri_key_[0] = a.tri_addr, gtri_key_[1] = a.triIdx_addr, gtri_key_[2] = tri_gen, gtri_key_[3] = idx_gen;
      gtri_slots_ = n_slots;
      ++launches_, ++prep_launches_;
    }
    sc.gtri = gtri_;
    sc.wide = nullptr;
    uint32_t n_blas_launch = 0;
    {   // 4-wide records for the state machine's fast order, cached on (bvh, blas) addresses + upload generations
      uint64_t bvh_bytes = 0, bvh_gen = 0, blas_bytes = 0, blas_gen = 0;
      const uint8_t *nodes = nullptr, *blas_raw = nullptr;
      resolve(a.bvh_addr, sizeof(BvhNodeA), 16, "bvh", &nodes, &bvh_bytes, &bvh_gen);
      resolve(a.blas_addr, sizeof(BlasNodeA), 8, "blas", &blas_raw, &blas_bytes, &blas_gen);
      const uint64_t n_nodes = bvh_bytes / sizeof(BvhNodeA), n_blas = blas_bytes / sizeof(BlasNodeA);
      n_blas_launch = (uint32_t)std::min<uint64_t>(n_blas, 0xFFFFFFFFu);
      const bool machine = dcr_get(VX_DCR_B200_RT_KERNEL) != 1 && n_nodes < kMachineMaxNodes && n_slots < kMachineMaxNodes;
      const bool want = machine && dcr_get(VX_DCR_B200_RT_ORDER) == 0 && n_nodes > 0 && n_blas > 0;
      if (want) {
        uint64_t key[5] = {a.bvh_addr, a.blas_addr, bvh_gen, blas_gen, n_nodes};
        if (!wide_valid_ || memcmp(key, wide_key_, sizeof(key)) != 0) {
          // instance bases: the bvh_offset field of every blas_node_t (child indices are instance-local)
          std::vector<uint8_t> hb(n_blas * sizeof(BlasNodeA));
          uint64_t bstart = 0;
          space_.find(a.blas_addr, 1, &bstart);
          auto sh = small_shadow_.find(bstart);
          if (sh != small_shadow_.end() && sh->second.size() >= (a.blas_addr - bstart) + hb.size()) {
            memcpy(hb.data(), sh->second.data() + (a.blas_addr - bstart), hb.size());   // host copy kept at upload time (<= 4 KB uploads)
          } else {
            CUDA_TRY(cudaMemcpyAsync(hb.data(), blas_raw, hb.size(), cudaMemcpyDeviceToHost, stream_));
            CUDA_TRY(cudaStreamSynchronize(stream_));
          }
          std::vector<uint32_t> bases;
          for (uint64_t i = 0; i < n_blas; ++i) {
            uint32_t off;
            memcpy(&off, hb.data() + i * sizeof(BlasNodeA) + offsetof(BlasNodeA, bvh_offset), 4);
            if (off < n_nodes) bases.push_back(off);
          }
          std::sort(bases.begin(), bases.end());
          bases.erase(std::unique(bases.begin(), bases.end()), bases.end());
          if (bases.empty() || bases[0] != 0) bases.insert(bases.begin(), 0u);
          if (wide_cap_ < n_nodes) {
            if (wide_) cudaFree(wide_);
            wide_ = nullptr, wide_cap_ = 0;
            CUDA_TRY(cudaMalloc(&wide_, n_nodes * sizeof(GpuNode4)));
            wide_cap_ = n_nodes;
          }
          if (wide_bases_cap_ < bases.size()) {
            if (wide_bases_) cudaFree(wide_bases_);
            if (wide_bases_host_) cudaFreeHost(wide_bases_host_);
            wide_bases_ = nullptr, wide_bases_host_ = nullptr, wide_bases_cap_ = 0;
            CUDA_TRY(cudaMalloc(&wide_bases_, bases.size() * 4));
            CUDA_TRY(cudaMallocHost(&wide_bases_host_, bases.size() * 4));
            wide_bases_cap_ = bases.size();
          }
          // page-locked staging, rewritten only after the previous run was finished: the copy needs no host wait
          memcpy(wide_bases_host_, bases.data(), bases.size() * 4);
          CUDA_TRY(cudaMemcpyAsync(wide_bases_, wide_bases_host_, bases.size() * 4, cudaMemcpyHostToDevice, stream_));
          CUDA_TRY(launch_build_wide(nodes, n_nodes, wide_bases_, (uint32_t)bases.size(), wide_, stream_));
          memcpy(wide_key_, key, sizeof(key));
          wide_valid_ = true;
          ++launches_, ++prep_launches_;
        }
        sc.wide = wide_;
      }
    }
    CUDA_TRY(cudaEventRecord(ev_prep_, stream_));
    int rc;
    if (!trace) {
      RenderA p{};
      p.scene = sc, p.arg = a, p.out = out, p.part = part_;
      p.mode = dcr_get(VX_DCR_B200_RT_MODE), p.strict_order = dcr_get(VX_DCR_B200_RT_ORDER);
      p.counters = counters_, p.tile_counter = tile_counter_, p.stats = stats_on_;
      p.n_blas = n_blas_launch;
      // Fast-order frames: one pixel per lane over the 4-wide records (k_render_direct) — measured against the warp-scheduled
      // state machine on one B200 (profiles/r02_sweep_direct_a.txt): cfg1 11 284 vs 7 264 Mrays/s, cfg3 12 786 vs 7 783, 4K
      // primary 11 759 vs 8 519, cfg2 (4 diffuse bounces) 4 905 vs 4 773.  VORTEX_B200_DIRECT=0 selects the state machine.
      p.direct = (uint32_t)env_u64("VORTEX_B200_DIRECT", 1);
      p.quorum_refill = (uint32_t)env_u64("VORTEX_B200_QUORUM_REFILL", 12);
      p.quorum_inner = (uint32_t)env_u64("VORTEX_B200_QUORUM_INNER", 8);
      p.chunk = (uint32_t)env_u64("VORTEX_B200_CHUNK", 32);
      if (p.chunk < 1 || p.chunk > 65536) p.chunk = 32;
      if (p.quorum_refill < 2 || p.quorum_refill > 32) p.quorum_refill = 12;
      if (p.quorum_inner < 1 || p.quorum_inner >= p.quorum_refill) p.quorum_inner = p.quorum_refill / 2;
      uint64_t bvh_bytes = 0;
      const uint8_t *dummy = nullptr;
      resolve(a.bvh_addr, sizeof(BvhNodeA), 16, "bvh", &dummy, &bvh_bytes);
      bool simple = dcr_get(VX_DCR_B200_RT_KERNEL) == 1 || bvh_bytes / sizeof(BvhNodeA) >= kMachineMaxNodes || n_slots >= kMachineMaxNodes ||
                    (dcr_get(VX_DCR_B200_RT_ORDER) == 0 && !sc.wide);
      // the prep-time event is already recorded: the schedule's small copies count as kernel time
      const bool rb = rb_host_ && rb_addr_ == a.dst_addr && rb_bytes_ >= (uint64_t)a.dst_width * a.dst_height * 4 && !remote_fb_ && !peer_fb_arena_;
      const uint32_t tiles_per_band = (part_.band_rows / 4) * ((a.dst_width + 7) / 8), ntiles = rank_tiles(a.dst_width, a.dst_height);
      const uint32_t my_bands = tiles_per_band ? ntiles / tiles_per_band : 0;
      // strip count and taper (strip k+1 = taper % of strip k; only the last strip's copy is exposed).  One GPU, 8K frame
      // (profiles/r02_sweep_readback_taper.txt): start + wait + download 4.46 ms in 8 equal strips, 4.41 in 16, 4.31-4.34 in 16
      // strips tapered by 90 %; 70 % makes the first strip too long (4.55).  The defaults of a group's GPUs stay 8 equal strips.
      const bool big_single = part_.world == 1 && my_bands >= 512;
      const uint32_t strips = (uint32_t)env_u64("VORTEX_B200_READBACK_STRIPS", big_single ? 16 : kStrips);
      const uint32_t taper = (uint32_t)std::min<uint64_t>(std::max<uint64_t>(env_u64("VORTEX_B200_READBACK_TAPER", big_single ? 90 : 100), 25), 100);
      // A frame that is read back in strips keeps the row-major tile order (a strip must be a set of whole bands): hiding the
      // framebuffer copy under the tracing is worth more than the ordered schedule's shorter tail.
      // Measured (profiles/r02_sweep_readback_strips.txt, one GPU): the 8K frame's end-to-end time is 11.36 ms as one launch,
      // 10.20 / 10.06 / 9.85 / 9.79 ms in 2 / 4 / 6 / 8 strips; a 1080p frame (0.37 ms of tracing) only loses to the extra
      // launch tails, so small frames stay one launch.  Eight GPUs (129 600 tiles each, profiles/r02_scale/strips_n8.txt): 3.92 ms as
      // one launch per GPU, 3.90 / 3.44 / 3.33 ms in 2 / 4 / 8 strips — the copies (95 GB/s for the whole box) take longer than the
      // tracing there, so hiding the tracing under them pays although the strips give up the ordered schedule.
      const bool strip_ok = rb && !simple && strips > 1 && my_bands >= 4 * strips && ntiles >= env_u64("VORTEX_B200_READBACK_MIN_TILES", 100000);
      if (sched_prepare_locked(0, a.dst_width, a.dst_height, !simple && p.chunk == 32 && !strip_ok, &p.sched) != 0) return -1;
      if (strip_ok && !p.sched.order) {
        if (strips_init_locked() != 0) return -1;
        CUDA_TRY(cudaStreamWaitEvent(aux_stream_, ev_prep_, 0));   // scene records (re)built above, counters zeroed
        CUDA_TRY(cudaMemsetAsync(strip_counters_, 0, sizeof(uint32_t) * kMaxStrips, stream_));
        CUDA_TRY(cudaEventRecord(ev_strip_[kMaxStrips], stream_));
        CUDA_TRY(cudaStreamWaitEvent(aux_stream_, ev_strip_[kMaxStrips], 0));
        rc = 0;
        const uint32_t S = std::min<uint32_t>(strips, kMaxStrips);
        uint32_t edge[kMaxStrips + 1];
        {
          const double q = taper / 100.0;
          double total = 0, w = 1, acc = 0;
          for (uint32_t k = 0; k < S; k++) total += w, w *= q;
          w = 1, edge[0] = 0;
          for (uint32_t k = 0; k < S; k++) {   // every strip keeps at least one band (my_bands >= 4 * S)
            acc += w, w *= q;
            const uint32_t want = taper == 100 ? (uint32_t)((uint64_t)my_bands * (k + 1) / S) : (uint32_t)(my_bands * (acc / total) + 0.5);
            edge[k + 1] = std::max<uint32_t>(edge[k] + 1, std::min<uint32_t>(my_bands - (S - 1 - k), want));
          }
          edge[S] = my_bands;
        }
        // Which end of the image goes first: a strip's copy can only start when its kernel has ended, so strips that trace
        // quickly (sky, a far ground plane) should go FIRST — at the end their copies would queue behind each other after the
        // last kernel (profiles/r02_sweep_readback_taper.txt, timeline).  The previous strip-pipelined frame tells: if the half
        // of the bands that went first took 8 points more than its share of the kernel time, the order is turned round (and
        // stays: the same half then reads 8 points less).
        const uint64_t order_env = env_u64("VORTEX_B200_READBACK_ORDER", 2);   // 0 top-down, 1 bottom-up, 2 by feedback
        const bool rev = order_env == 2 ? strip_reverse_ : order_env == 1;
        strip_reverse_ = rev;
        uint32_t k_half = 0;   // the strip with which half of the bands have been launched (strips may be tapered)
        while (k_half + 1 < S && 2 * edge[k_half + 1] < my_bands) ++k_half;
        strip_half_share_ = (float)edge[k_half + 1] / (float)my_bands;
        for (uint32_t k = 0; k < S && rc >= 0; k++) {
          const uint32_t b0 = rev ? my_bands - edge[k + 1] : edge[k], b1 = rev ? my_bands - edge[k] : edge[k + 1];
          cudaStream_t st = (k & 1) ? aux_stream_ : stream_;
          RenderA q = p;
          q.tile_first = b0 * tiles_per_band, q.tile_end = b1 * tiles_per_band, q.tile_counter = strip_counters_ + k;
          q.blocks_per_sm_cap = 0;
          int r = launch_render_machine(q, prop_.multiProcessorCount, st);
          if (r < 0) { rc = r; break; }
          rc += r;
          CUDA_TRY(cudaEventRecord(ev_strip_[k], st));
          if (k == k_half) {
            CUDA_TRY(cudaEventRecord(ev_half_, st));
            half_valid_ = true;
          }
          CUDA_TRY(cudaStreamWaitEvent(copy_stream_, ev_strip_[k], 0));
          if (copy_bands(rb_host_, a.dst_addr, a.dst_width, a.dst_height, b0, b1 == my_bands ? my_bands + 1 : b1, copy_stream_) != 0) return -1;
        }
        if (rc >= 0) {   // the main stream joins the other two: kernel time ends where the last strip ends, the run where the last copy ends
          for (uint32_t k = 0; k < S; k++) CUDA_TRY(cudaStreamWaitEvent(stream_, ev_strip_[k], 0));
          CUDA_TRY(cudaEventRecord(ev_kend_, stream_));
          kend_valid_ = true;
          CUDA_TRY(cudaEventRecord(ev_strip_[kMaxStrips], copy_stream_));
          CUDA_TRY(cudaStreamWaitEvent(stream_, ev_strip_[kMaxStrips], 0));
          rb_issued_ = true;
        }
      } else {
        rc = simple ? launch_render_a(p, prop_.multiProcessorCount, stream_) : launch_render_machine(p, prop_.multiProcessorCount, stream_);
        if (rc >= 0 && rb) {   // one launch: the bands follow it on the same stream
          if (strips_init_locked() != 0) return -1;
          CUDA_TRY(cudaEventRecord(ev_kend_, stream_));
          kend_valid_ = true;
          if (copy_bands(rb_host_, a.dst_addr, a.dst_width, a.dst_height, 0, my_bands + 1, stream_) != 0) return -1;
          rb_issued_ = true;
        }
      }
    } else {
      uint64_t bvh_bytes = 0;
      const uint8_t *dummy = nullptr;
      resolve(a.bvh_addr, sizeof(BvhNodeA), 16, "bvh", &dummy, &bvh_bytes);
      bool simple = dcr_get(VX_DCR_B200_RT_KERNEL) == 1 || bvh_bytes / sizeof(BvhNodeA) >= kMachineMaxNodes || n_slots >= kMachineMaxNodes ||
                    count >= 0xFFFFFFE0ull || (dcr_get(VX_DCR_B200_RT_ORDER) == 0 && !sc.wide);
      if (simple) {
        TraceA p{};
        p.scene = sc, p.arg = a, p.rays = rays, p.count = count, p.hits = hits;
        p.strict_order = dcr_get(VX_DCR_B200_RT_ORDER), p.counters = counters_, p.stats = stats_on_;
        rc = launch_trace_a(p, prop_.multiProcessorCount, stream_);
      } else {   // the warp-scheduled ray state machine with rays as work items
        RenderA p{};
        p.scene = sc, p.arg = a, p.out = nullptr, p.part = Partition{};
        p.mode = 0, p.strict_order = dcr_get(VX_DCR_B200_RT_ORDER);
        p.counters = counters_, p.tile_counter = tile_counter_, p.stats = stats_on_;
        p.n_blas = n_blas_launch;
        p.quorum_refill = (uint32_t)env_u64("VORTEX_B200_TRACE_QUORUM_REFILL", 16);
        p.quorum_inner = (uint32_t)env_u64("VORTEX_B200_TRACE_QUORUM_INNER", 12);
        p.chunk = (uint32_t)env_u64("VORTEX_B200_TRACE_CHUNK", 64);   // sweep: profiles/r01_sweep_trace_quorum_chunk.txt
        if (p.chunk < 1 || p.chunk > 65536) p.chunk = 64;
        if (p.quorum_refill < 2 || p.quorum_refill > 32) p.quorum_refill = 16;
        if (p.quorum_inner < 1 || p.quorum_inner >= p.quorum_refill) p.quorum_inner = p.quorum_refill / 2;
        p.rays = rays, p.hits = hits, p.count = count;
        // Coherence ordering of large caller-supplied batches (opt-in, VORTEX_B200_RAY_ORDER=1): a probe counts how often
        // neighbouring rays fall into different coarse cells; an incoherent batch is traced in (direction octant, origin Morton)
        // order through an index array (results land at the rays' own positions).  Measured on BASELINE configs[4]
        // (profiles/r01_sweep_ray_order.txt): rays that ARRIVE in that order trace 34 % faster (2.37 -> 1.77 ms), but ordering
        // them here costs keys 0.05 + sort 0.27 ms and the indexed kernel runs 2.12 ms, so the call gets slower (2.66 vs
        // 2.38 ms) — hence off by default; callers that can generate rays in a coherent order should.
        if (rays && !pipe_hits_ && count >= env_u64("VORTEX_B200_RAY_ORDER_MIN", 1ull << 18) && env_u64("VORTEX_B200_RAY_ORDER", 0) != 0) {
          uint32_t bits = (uint32_t)env_u64("VORTEX_B200_RAY_ORDER_BITS", 7);
          if (bits < 2 || bits > 9) bits = 7;
          if (ray_order_reserve_locked(count) != 0) return -1;
          CUDA_TRY(ray_order_keys(sc, rays, count, bits, order_, stream_));
          uint32_t jumps = 0;
          CUDA_TRY(cudaMemcpyAsync(&jumps, order_.jumps, 4, cudaMemcpyDeviceToHost, stream_));
          CUDA_TRY(cudaStreamSynchronize(stream_));
          launches_ += 1;
          if ((double)jumps * 8.0 > 0.25 * (double)count) {   // the probe samples one block in eight
            const uint32_t *sorted = nullptr;
            CUDA_TRY(ray_order_sort(count, bits, order_, &sorted, stream_));
            launches_ += 1;   // the CUB sort counted as one library step
            p.perm = sorted;
          }
        }
        // chunk of the host pipeline: kRunCtxs kernels share the SMs, so that one chunk's tail (its slowest rays) runs
        // under the other chunks' bodies instead of idling the device
        if (pipe_hits_) p.blocks_per_sm_cap = (uint32_t)env_u64("VORTEX_B200_HOST_BLOCKS_PER_SM", 2);
        rc = launch_trace_machine(p, prop_.multiProcessorCount, stream_);
      }
    }
    if (rc < 0) {
      sched_active_ = -1, sched_timed_ = -1;   // no frame ran: nothing to read costs or a trial time from
      return fail("kernel launch failed: %s", cudaGetErrorString((cudaError_t)(-rc)));
    }
    launches_ += rc;
    return end_run_locked();
  }

  // layout B: resolve the scene, (re)build the decoded-node / edge-form-triangle arrays if the uploads changed.
  // Must run between begin_run_locked() and the ev_prep_ record (prep launches are timed separately).
  int rtu_scene_locked(const KernelArgB &a, RtuScene *sc) {
    // the RT unit reads its four scene pointers from DCRs 0x006..0x009 (sim/simx/rt_traversal.cpp:35-38);
    // hosts that never wrote them fall back to the kernel_arg_t fields
    uint64_t tlas_addr = dcrs_.count(VX_DCR_BASE_RTX_TLAS_PTR) ? dcr_get(VX_DCR_BASE_RTX_TLAS_PTR) : a.tlas_addr;
    uint64_t blas_addr = dcrs_.count(VX_DCR_BASE_RTX_BLAS_PTR) ? dcr_get(VX_DCR_BASE_RTX_BLAS_PTR) : a.blas_addr;
    uint64_t qbvh_addr = dcrs_.count(VX_DCR_BASE_RTX_BVH_PTR) ? dcr_get(VX_DCR_BASE_RTX_BVH_PTR) : a.qBvh_addr;
    uint64_t tri_addr = dcrs_.count(VX_DCR_BASE_RTX_TRI_PTR) ? dcr_get(VX_DCR_BASE_RTX_TRI_PTR) : a.tri_addr;
    const QNodeB *tlas = nullptr, *qbvh = nullptr;
    const TriA *tri = nullptr;
    uint64_t tlas_bytes = 0, qbvh_bytes = 0, tri_bytes = 0, blas_bytes = 0, ex_bytes = 0, mat_bytes = 0, tex_bytes = 0;
    uint64_t gen[3] = {0, 0, 0};
    if (resolve(tlas_addr, sizeof(QNodeB), 4, "tlas(q)", &tlas, &tlas_bytes, &gen[0])) return -1;
    uint64_t blas_gen = 0;
    if (resolve(blas_addr, sizeof(BlasNodeB), 4, "blas", &sc->blas, &blas_bytes, &blas_gen)) return -1;
    if (resolve(qbvh_addr, sizeof(QNodeB), 4, "qBvh", &qbvh, &qbvh_bytes, &gen[1])) return -1;
    if (resolve(tri_addr, sizeof(TriA), 4, "tri", &tri, &tri_bytes, &gen[2])) return -1;
    if (resolve(a.triEx_addr, sizeof(TriExB), 4, "triEx", &sc->triEx, &ex_bytes)) return -1;
    sc->mat = nullptr, sc->tex = nullptr;
    if (a.mat_addr && space_.find(a.mat_addr, 1, nullptr)) { if (resolve(a.mat_addr, 4, 4, "mat", &sc->mat, &mat_bytes)) return -1; }
    if (a.tex_addr && space_.find(a.tex_addr, 1, nullptr)) { if (resolve(a.tex_addr, 4, 4, "tex", &sc->tex, &tex_bytes)) return -1; }
    uint64_t n_tlas = tlas_bytes / sizeof(QNodeB), n_bvh = qbvh_bytes / sizeof(QNodeB), n_tri = tri_bytes / sizeof(TriA);
    if (n_tlas >= (1ull << 30) || n_bvh >= (1ull << 30) || n_tri > 0xFFFFFFFFull) return fail("layout B scene too large");
    if (ex_bytes / sizeof(TriExB) < n_tri) n_tri = ex_bytes / sizeof(TriExB);   // shading indexes triEx by triIdx
    uint64_t key[11] = {tlas_addr, qbvh_addr, tri_addr, gen[0], gen[1], gen[2], n_tlas, n_bvh, n_tri, blas_addr, blas_gen};
    if (!rtu_valid_ || memcmp(key, rtu_key_, sizeof(key)) != 0) {
      uint64_t need_nodes = n_tlas + n_bvh;
      if (rtu_nodes_cap_ < need_nodes) {
        if (rtu_nodes_) cudaFree(rtu_nodes_);
        rtu_nodes_ = nullptr, rtu_nodes_cap_ = 0;
        CUDA_TRY(cudaMalloc(&rtu_nodes_, need_nodes * sizeof(GpuNodeB)));
        rtu_nodes_cap_ = need_nodes;
      }
      if (rtu_tri_cap_ < n_tri) {
        if (rtu_tri_) cudaFree(rtu_tri_);
        rtu_tri_ = nullptr, rtu_tri_cap_ = 0;
        CUDA_TRY(cudaMalloc(&rtu_tri_, n_tri * sizeof(GpuTri)));
        rtu_tri_cap_ = n_tri;
      }
      CUDA_TRY(launch_rtu_decode_nodes(tlas, rtu_nodes_, n_tlas, stream_));
      CUDA_TRY(launch_rtu_decode_nodes(qbvh, rtu_nodes_ + n_tlas, n_bvh, stream_));
      CUDA_TRY(launch_rtu_gather_tris(tri, rtu_tri_, n_tri, stream_));
      {   // child cursors for the ray state machine: a BLAS node's children sit at (its instance's bvh_offset) + leftFirst + k
        const uint64_t n_blas = blas_bytes / sizeof(BlasNodeB);
        std::vector<uint8_t> hb(n_blas * sizeof(BlasNodeB));
        uint64_t bstart = 0;
        space_.find(blas_addr, 1, &bstart);
        auto sh = small_shadow_.find(bstart);
        if (sh != small_shadow_.end() && sh->second.size() >= (blas_addr - bstart) + hb.size()) {
          memcpy(hb.data(), sh->second.data() + (blas_addr - bstart), hb.size());
        } else if (!hb.empty()) {
          CUDA_TRY(cudaMemcpyAsync(hb.data(), sc->blas, hb.size(), cudaMemcpyDeviceToHost, stream_));
          CUDA_TRY(cudaStreamSynchronize(stream_));
        }
        std::vector<uint32_t> bases;
        for (uint64_t i = 0; i < n_blas; ++i) {
          uint32_t off;
          memcpy(&off, hb.data() + i * sizeof(BlasNodeB) + offsetof(BlasNodeB, bvh_offset), 4);
          if (off < n_bvh) bases.push_back(off);
        }
        std::sort(bases.begin(), bases.end());
        bases.erase(std::unique(bases.begin(), bases.end()), bases.end());
        if (bases.empty() || bases[0] != 0) bases.insert(bases.begin(), 0u);
        if (rtu_bases_cap_ < bases.size()) {
          if (rtu_bases_) cudaFree(rtu_bases_);
          if (rtu_bases_host_) cudaFreeHost(rtu_bases_host_);
          rtu_bases_ = rtu_bases_host_ = nullptr, rtu_bases_cap_ = 0;
          CUDA_TRY(cudaMalloc(&rtu_bases_, bases.size() * 4));
          CUDA_TRY(cudaMallocHost(&rtu_bases_host_, bases.size() * 4));
          rtu_bases_cap_ = bases.size();
        }
        memcpy(rtu_bases_host_, bases.data(), bases.size() * 4);
        CUDA_TRY(cudaMemcpyAsync(rtu_bases_, rtu_bases_host_, bases.size() * 4, cudaMemcpyHostToDevice, stream_));
        CUDA_TRY(launch_rtu_link(rtu_nodes_, n_tlas, n_bvh, rtu_bases_, (uint32_t)bases.size(), stream_));
      }
      launches_ += 4, prep_launches_ += 4;
      memcpy(rtu_key_, key, sizeof(key));
      rtu_valid_ = true;
    }
    sc->gtlas = rtu_nodes_, sc->gbvh = rtu_nodes_ + n_tlas, sc->gtri = rtu_tri_;
    sc->n_tlas = (uint32_t)n_tlas, sc->n_bvh = (uint32_t)n_bvh, sc->n_tri = (uint32_t)n_tri;
    sc->n_blas = (uint32_t)(blas_bytes / sizeof(BlasNodeB));
    sc->n_mat = (uint32_t)(mat_bytes / sizeof(MaterialB)), sc->_pad = 0;
    sc->tex_bytes = tex_bytes;
    return 0;
  }

  int start_rtu_locked(const std::vector<uint8_t> &args) {
    if (args.size() < sizeof(KernelArgB)) return fail("layout B needs a 216-byte kernel_arg_t");
    KernelArgB a;
    memcpy(&a, args.data(), sizeof(a));
    if (a.dst_width == 0 || a.dst_height == 0) return fail("empty framebuffer");
    RtuRender p{};
    if (resolve(a.dst_addr, (uint64_t)a.dst_width * a.dst_height * 4, 4, "framebuffer", &p.out)) return -1;
    if (remote_fb_) {
      if (remote_bytes_ < (uint64_t)a.dst_width * a.dst_height * 4) return fail("remote framebuffer smaller than the image");
      p.out = remote_fb_;
    } else if (peer_fb_arena_) {
      p.out = reinterpret_cast<uint32_t *>(peer_fb_arena_ + a.dst_addr);
    }
    last_out_ = Output{a.dst_addr, (uint64_t)a.dst_width * a.dst_height * 4, a.dst_width, a.dst_height, true, true};
    if (begin_run_locked() != 0) return -1;
    if (rtu_scene_locked(a, &p.scene)) return -1;
    CUDA_TRY(cudaEventRecord(ev_prep_, stream_));
    p.arg = a, p.part = part_, p.counters = counters_, p.tile_counter = tile_counter_, p.stats = stats_on_;
    p.camera = dcr_get(VX_DCR_B200_RT_CAMERA), p.strict_order = dcr_get(VX_DCR_B200_RT_ORDER);
    p.mode = dcr_get(VX_DCR_B200_RT_MODE) & 3u;
    // VX_DCR_B200_RT_KERNEL: 0 = the fused frame kernel — fast order: the warp-scheduled ray state machine (rt_machine_b.cu),
    // strict order (or VX_DCR_B200_RT_KERNEL = 1, or more nested mirror rays than a slot holds): one ray per lane (rt_rtu.cu);
    // 2 + n = shader program n through the traceRay/getWork/getAttr/commit shim (vx_raytrace_b200.cuh; program 0 = the three
    // shipped shaders, rt_rtu_programs.cuh)
    uint32_t kernel = dcr_get(VX_DCR_B200_RT_KERNEL);
    // Measured (tools/quick_rtu_machine.py, profiles/r02_sweep_rtu_machine.txt): on coherent primary rays the state machine is
    // 7 % SLOWER than one ray per lane (8 522 vs 9 128 Mrays/s at 1080p / 100 K triangles, 7 876 vs 8 447 at 4K / 1 M: with
    // 4-wide nodes a primary ray makes ~10 visits and 21.5 of 32 lanes are active anyway), on incoherent rays 2.4x faster
    // (4 203 vs 1 777 Mrays/s).  Frames therefore keep the one-ray-per-lane kernel unless VORTEX_B200_RTU_MACHINE=2 asks for
    // the machine; hit records (vx_b200_trace_*) use the machine (VORTEX_B200_RTU_MACHINE=0 turns it off).
    if (p.mode && kernel >= 2) return fail("shadow / diffuse modes are not available for shader programs (VX_DCR_B200_RT_KERNEL >= 2)");
    const bool machine = kernel == 0 && p.mode == 0 && !p.strict_order && rtu_machine_supports(a) && p.scene.n_tlas <= (1u << 27) && p.scene.n_bvh <= (1u << 27) &&
                         p.scene.n_tri <= (1u << 27) && env_u64("VORTEX_B200_RTU_MACHINE", 1) >= 2;
    // ... and the same records walked one ray per lane (k_rtu_direct): the default for fast-order frames
    const bool direct = !machine && kernel == 0 && p.mode == 0 && !p.strict_order && rtu_machine_supports(a) && p.scene.n_tlas <= (1u << 27) &&
                        p.scene.n_bvh <= (1u << 27) && p.scene.n_tri <= (1u << 27) && env_u64("VORTEX_B200_RTU_DIRECT", 1) != 0;
    if (sched_prepare_locked(1, a.dst_width, a.dst_height, true, &p.sched, /*cost_every_frame=*/!machine) != 0) return -1;
    int rc;
    if (machine || direct) {
      RtuMachine m{};
      m.scene = p.scene, m.arg = a, m.out = p.out, m.part = part_, m.camera = p.camera, m.counters = counters_, m.tile_counter = tile_counter_;
      m.stats = stats_on_, m.sched = p.sched;
      m.quorum_refill = (uint32_t)env_u64("VORTEX_B200_RTU_QUORUM_REFILL", 12), m.quorum_inner = (uint32_t)env_u64("VORTEX_B200_RTU_QUORUM_INNER", 8);
      m.chunk = 32;
      if (m.quorum_refill < 2 || m.quorum_refill > 32) m.quorum_refill = 12;
      if (m.quorum_inner < 1 || m.quorum_inner >= m.quorum_refill) m.quorum_inner = m.quorum_refill / 2;
      rc = machine ? launch_rtu_machine(m, prop_.multiProcessorCount, stream_) : launch_rtu_direct(m, prop_.multiProcessorCount, stream_);
    } else {
      rc = kernel >= 2 ? launch_rtu_program(p, kernel - 2, prop_.multiProcessorCount, stream_) : launch_rtu_render(p, prop_.multiProcessorCount, stream_);
    }
    if (rc < 0) {
      sched_active_ = -1, sched_timed_ = -1;   // no frame ran: nothing to read costs or a trial time from
      return fail("kernel launch failed: %s", cudaGetErrorString((cudaError_t)(-rc)));
    }
    launches_ += rc;
    return end_run_locked();
  }

  int trace_rtu_locked(const std::vector<uint8_t> &args, const uint64_t *rays_addr, uint64_t count, uint64_t hits_addr) {
    if (args.size() < sizeof(KernelArgB)) return fail("layout B needs a 216-byte kernel_arg_t");
    KernelArgB a;
    memcpy(&a, args.data(), sizeof(a));
    RtuTrace p{};
    uint64_t n = rays_addr ? count : (uint64_t)a.dst_width * a.dst_height;
    p.rays = nullptr;
    if (pipe_hits_) {
      p.hits = pipe_hits_, p.rays = pipe_rays_;
    } else {
      if (resolve(hits_addr, n * sizeof(HitA), 4, "hits", &p.hits)) return -1;
      if (rays_addr && resolve(*rays_addr, n * 24, 4, "rays", &p.rays)) return -1;
      last_out_ = Output{hits_addr, n * sizeof(HitA), 0, 0, false, true};
    }
    if (begin_run_locked() != 0) return -1;
    if (rtu_scene_locked(a, &p.scene)) return -1;
    CUDA_TRY(cudaEventRecord(ev_prep_, stream_));
    p.arg = a, p.count = n, p.counters = counters_, p.stats = stats_on_;
    p.camera = dcr_get(VX_DCR_B200_RT_CAMERA), p.strict_order = dcr_get(VX_DCR_B200_RT_ORDER);
    int rc;
    if (!p.strict_order && dcr_get(VX_DCR_B200_RT_KERNEL) != 1 && n < 0xFFFFFFE0ull && p.scene.n_tlas <= (1u << 27) && p.scene.n_bvh <= (1u << 27) &&
        p.scene.n_tri <= (1u << 27) && env_u64("VORTEX_B200_RTU_MACHINE", 1) != 0) {   // the ray state machine with rays as work items
      RtuMachine m{};
      m.scene = p.scene, m.arg = a, m.part = Partition{}, m.camera = p.camera, m.counters = counters_, m.tile_counter = tile_counter_;
      m.stats = stats_on_, m.rays = p.rays, m.hits = p.hits, m.count = n;
      m.quorum_refill = (uint32_t)env_u64("VORTEX_B200_RTU_TRACE_QUORUM_REFILL", 16);
      m.quorum_inner = (uint32_t)env_u64("VORTEX_B200_RTU_TRACE_QUORUM_INNER", 12);
      m.chunk = 64;
      if (m.quorum_refill < 2 || m.quorum_refill > 32) m.quorum_refill = 16;
      if (m.quorum_inner < 1 || m.quorum_inner >= m.quorum_refill) m.quorum_inner = m.quorum_refill / 2;
      rc = launch_rtu_machine(m, prop_.multiProcessorCount, stream_);
    } else {
      rc = launch_rtu_trace(p, prop_.multiProcessorCount, stream_);
    }
    if (rc < 0) {
      sched_active_ = -1, sched_timed_ = -1;   // no frame ran: nothing to read costs or a trial time from
      return fail("kernel launch failed: %s", cudaGetErrorString((cudaError_t)(-rc)));
    }
    launches_ += rc;
    return end_run_locked();
  }

  // ---- vx_b200_trace_rays_host: rays and hit records stay in host memory, chunks are pipelined over upload / compute / download streams ----
  static constexpr int kPipeSlots = 6;
  struct HostPipe {
    float *rays[kPipeSlots] = {};
    HitA *hits[kPipeSlots] = {};
    cudaEvent_t up[kPipeSlots] = {}, done[kPipeSlots] = {}, down[kPipeSlots] = {};
    cudaStream_t s_up = nullptr, s_down = nullptr;
    uint64_t cap = 0;   // rays per slot
  } pipe_;
  const float *pipe_rays_ = nullptr;   // staging slot of the chunk being launched (null outside trace_host)
  HitA *pipe_hits_ = nullptr;

  int pipe_init_locked(uint64_t chunk) {
    if (pipe_.cap >= chunk) return 0;
    pipe_free_locked();
    CUDA_TRY(cudaStreamCreateWithFlags(&pipe_.s_up, cudaStreamNonBlocking));
    CUDA_TRY(cudaStreamCreateWithFlags(&pipe_.s_down, cudaStreamNonBlocking));
    for (int i = 0; i < kPipeSlots; i++) {
      CUDA_TRY(cudaMalloc(&pipe_.rays[i], chunk * 24));
      CUDA_TRY(cudaMalloc(&pipe_.hits[i], chunk * sizeof(HitA)));
      CUDA_TRY(cudaEventCreateWithFlags(&pipe_.up[i], cudaEventDisableTiming));
      CUDA_TRY(cudaEventCreateWithFlags(&pipe_.done[i], cudaEventDisableTiming));
      CUDA_TRY(cudaEventCreateWithFlags(&pipe_.down[i], cudaEventDisableTiming));
    }
    pipe_.cap = chunk;
    for (int c = 1; c < kRunCtxs; c++) {   // extra run contexts: the kernels of consecutive chunks overlap (tail of one under the others)
      RunCtx &x = ctx_[c];
      if (x.stream) continue;
      CUDA_TRY(cudaStreamCreateWithFlags(&x.stream, cudaStreamNonBlocking));
      CUDA_TRY(cudaMalloc(&x.counters, sizeof(WorkCounters)));
      CUDA_TRY(cudaMalloc(&x.tile_counter, sizeof(uint32_t)));
      CUDA_TRY(cudaMallocHost(&x.h_counters, sizeof(WorkCounters)));
      CUDA_TRY(cudaEventCreate(&x.ev_begin));
      CUDA_TRY(cudaEventCreate(&x.ev_prep));
      CUDA_TRY(cudaEventCreate(&x.ev_end));
    }
    return 0;
  }
  void pipe_free_locked() {
    for (int i = 0; i < kPipeSlots; i++) {
      if (pipe_.rays[i]) cudaFree(pipe_.rays[i]);
      if (pipe_.hits[i]) cudaFree(pipe_.hits[i]);
      if (pipe_.up[i]) cudaEventDestroy(pipe_.up[i]);
      if (pipe_.done[i]) cudaEventDestroy(pipe_.done[i]);
      if (pipe_.down[i]) cudaEventDestroy(pipe_.down[i]);
    }
    if (pipe_.s_up) cudaStreamDestroy(pipe_.s_up);
    if (pipe_.s_down) cudaStreamDestroy(pipe_.s_down);
    pipe_ = HostPipe{};
  }
  static constexpr int kRunCtxs = 4;
  struct RunCtx {
    cudaStream_t stream = nullptr;
    WorkCounters *counters = nullptr, *h_counters = nullptr;
    uint32_t *tile_counter = nullptr;
    cudaEvent_t ev_begin = nullptr, ev_prep = nullptr, ev_end = nullptr;
    bool run_pending = false;
    uint64_t launches = 0, prep_launches = 0;
  } ctx_[kRunCtxs];      // [0] is only a parking slot for the device's own context while another one is loaded
  int cur_ctx_ = 0;
  void use_ctx_locked(int k) {
    if (k == cur_ctx_) return;
    ctx_[cur_ctx_] = RunCtx{stream_, counters_, h_counters_, tile_counter_, ev_begin_, ev_prep_, ev_end_, run_pending_, launches_, prep_launches_};
    const RunCtx &x = ctx_[k];
    stream_ = x.stream, counters_ = x.counters, h_counters_ = x.h_counters, tile_counter_ = x.tile_counter;
    ev_begin_ = x.ev_begin, ev_prep_ = x.ev_prep, ev_end_ = x.ev_end, run_pending_ = x.run_pending;
    launches_ = x.launches, prep_launches_ = x.prep_launches;
    cur_ctx_ = k;
  }
  void alt_free_locked() {
    use_ctx_locked(0);
    for (int c = 1; c < kRunCtxs; c++) {
      RunCtx &x = ctx_[c];
      if (x.stream) cudaStreamDestroy(x.stream);
      if (x.counters) cudaFree(x.counters);
      if (x.tile_counter) cudaFree(x.tile_counter);
      if (x.h_counters) cudaFreeHost(x.h_counters);
      if (x.ev_begin) cudaEventDestroy(x.ev_begin);
      if (x.ev_prep) cudaEventDestroy(x.ev_prep);
      if (x.ev_end) cudaEventDestroy(x.ev_end);
      x = RunCtx{};
    }
  }

  // ---- frame-to-frame schedule feedback (TileSchedule, rt_raycast.h) ----
  // A cycle of kSchedCycle frames of one (layout, image size, partition): frames 0 and 1 run in row-major order and record the
  // groups' costs, frame 2 runs in descending-cost order, and the faster of frames 1 and 2 (kernel time, 2 % margin) is kept
  // for the rest of the cycle.  While the order is in use layout B refreshes the costs every frame (a clock read per tile);
  // layout A records them with a slower build of its kernel, on the trial frames and every 16th.
  static constexpr uint32_t kSchedCycle = 64;
  struct SchedState {
    uint64_t key[5] = {0, 0, 0, 0, 0};
    bool valid = false, pending = false, use_order = false, identity_loaded = true;
    uint32_t group_shift = 0, n_groups = 0, lead = 0, frames = 0;
    float ms_identity = 0, ms_ordered = 0;
    uint32_t *d_order = nullptr, *d_cost = nullptr, *h_cost = nullptr, *h_order = nullptr;
    bool order_dirty = false;
    std::vector<uint32_t> order;      // descending cost, from the last measured frame
  } sched_[2];   // [0] layout A frames, [1] layout B frames
  int sched_active_ = -1;   // which state the run in flight records costs for
  int sched_timed_ = -1;    // which state wants this run's kernel time, and as what
  uint32_t sched_timed_phase_ = 0;
  void sched_free_locked() {
    for (auto &st : sched_) {
      if (st.d_order) cudaFree(st.d_order);
      if (st.d_cost) cudaFree(st.d_cost);
      if (st.h_cost) cudaFreeHost(st.h_cost);
      if (st.h_order) cudaFreeHost(st.h_order);
      st = SchedState{};
    }
  }
  // Tiles of this rank exactly as the kernels enumerate them (8x4 tiles inside this rank's bands).
  uint32_t rank_tiles(uint32_t W, uint32_t H) const {
    const uint32_t tcols = (W + 7) / 8, bands_total = (H + part_.band_rows - 1) / part_.band_rows;
    const uint32_t bands = bands_total > part_.rank ? (bands_total - part_.rank + part_.world - 1) / part_.world : 0;
    return bands * (part_.band_rows / 4) * tcols;
  }
  // h_order is page-locked (a pageable source would make the copy wait for the stream's earlier work) and is rewritten only
  // here, after the previous run has been finished, so the asynchronous copy never races a modification
  int sched_upload_locked(SchedState &st, bool identity) {
    for (uint32_t i = 0; i < st.n_groups; i++) st.h_order[i] = identity ? i : st.order[i];
    CUDA_TRY(cudaMemcpyAsync(st.d_order, st.h_order, st.n_groups * 4, cudaMemcpyHostToDevice, stream_));
    st.identity_loaded = identity, st.order_dirty = false;
    return 0;
  }
  int sched_prepare_locked(int which, uint32_t W, uint32_t H, bool eligible, TileSchedule *out, bool cost_every_frame = false) {
    *out = TileSchedule{};
    sched_active_ = -1, sched_timed_ = -1;
    const uint32_t ntiles = rank_tiles(W, H);
    if (!eligible || ntiles < 2048 || env_u64("VORTEX_B200_SCHEDULE", 1) == 0) return 0;
    SchedState &st = sched_[which];
    const uint64_t key[5] = {W, H, part_.rank, part_.world, part_.band_rows};
    if (!st.valid || memcmp(key, st.key, sizeof(key)) != 0) {
      st.group_shift = 4;   // about 256 groups of 2^k tiles
      while ((ntiles >> st.group_shift) > 256) st.group_shift++;
      st.n_groups = ntiles >> st.group_shift;
      st.lead = ntiles - (st.n_groups << st.group_shift);
      if (st.d_order) cudaFree(st.d_order);
      if (st.d_cost) cudaFree(st.d_cost);
      if (st.h_cost) cudaFreeHost(st.h_cost);
      if (st.h_order) cudaFreeHost(st.h_order);
      st.d_order = st.d_cost = st.h_cost = st.h_order = nullptr;
      CUDA_TRY(cudaMalloc(&st.d_order, st.n_groups * 4));
      CUDA_TRY(cudaMalloc(&st.d_cost, st.n_groups * 4));
      CUDA_TRY(cudaMallocHost(&st.h_cost, st.n_groups * 4));
      CUDA_TRY(cudaMallocHost(&st.h_order, st.n_groups * 4));
      st.order.resize(st.n_groups);
      for (uint32_t i = 0; i < st.n_groups; i++) st.order[i] = i;   // identity until a measured order exists
      memcpy(st.key, key, sizeof(key));
      st.valid = true, st.pending = false, st.use_order = false, st.frames = 0;
    }
    const uint32_t phase = st.frames % kSchedCycle;
    st.frames++;
    if (st.pending) {   // costs of the last measured frame (complete: every run is finished before the next begins)
      uint64_t sum = 0;
      for (uint32_t i = 0; i < st.n_groups; i++) sum += st.h_cost[i];
      if (getenv("VORTEX_B200_SCHED_DEBUG")) {
        fprintf(stderr, "[sched %d] phase=%u groups=%u shift=%u lead=%u use=%d ms identity %.4f ordered %.4f cost:", which, phase, st.n_groups,
                st.group_shift, st.lead, (int)st.use_order, st.ms_identity, st.ms_ordered);
        for (uint32_t i = 0; i < st.n_groups; i += 8) fprintf(stderr, " %u", st.h_cost[i]);
        fprintf(stderr, "\n");
      }
      for (uint32_t i = 0; i < st.n_groups; i++) st.order[i] = i;   // sum == 0 (nothing measured): back to the identity
      if (sum) {
        const uint32_t *c = st.h_cost;
        std::stable_sort(st.order.begin(), st.order.end(), [c](uint32_t a, uint32_t b) { return c[a] > c[b]; });
      }
      st.order_dirty = true;
      st.pending = false;
    }
    bool identity;
    if (phase <= 1) identity = true;                       // frame 0 warms up and records the costs, frame 1 is the timed row-major half
    else if (phase == 2) identity = false;                 // the timed descending-cost half
    else {
      if (phase == 3) st.use_order = st.ms_ordered > 0 && st.ms_ordered < 0.98f * st.ms_identity;
      identity = !st.use_order;
    }
    if (st.frames == 1 || identity != st.identity_loaded || (!identity && st.order_dirty)) {
      if (sched_upload_locked(st, identity) != 0) return -1;
    }
    const bool measure = phase < 3 || (st.use_order && (cost_every_frame || (phase & 15u) == 0));
    if (identity && !measure) return 0;                    // row-major order, nothing recorded: the kernels' plain path
    out->order = st.d_order, out->cost = nullptr;
    out->group_shift = st.group_shift, out->n_groups = st.n_groups, out->lead = st.lead;
    if (measure) {
      CUDA_TRY(cudaMemsetAsync(st.d_cost, 0, st.n_groups * 4, stream_));
      out->cost = st.d_cost;
      sched_active_ = which;
    }
    if (phase == 1 || phase == 2) sched_timed_ = which, sched_timed_phase_ = phase;
    return 0;
  }

  // coherence ordering scratch (grow-only)
  RayOrderBuffers order_{};
  uint64_t order_cap_ = 0;
  int ray_order_reserve_locked(uint64_t n) {
    if (order_cap_ >= n) return 0;
    ray_order_free_locked();
    for (int i = 0; i < 2; i++) {
      CUDA_TRY(cudaMalloc(&order_.keys[i], n * 4));
      CUDA_TRY(cudaMalloc(&order_.idx[i], n * 4));
    }
    order_.tmp_bytes = ray_order_tmp_bytes(n);
    CUDA_TRY(cudaMalloc(&order_.tmp, order_.tmp_bytes ? order_.tmp_bytes : 16));
    CUDA_TRY(cudaMalloc(&order_.jumps, 4));
    order_cap_ = n;
    return 0;
  }
  void ray_order_free_locked() {
    for (int i = 0; i < 2; i++) {
      if (order_.keys[i]) cudaFree(order_.keys[i]);
      if (order_.idx[i]) cudaFree(order_.idx[i]);
    }
    if (order_.tmp) cudaFree(order_.tmp);
    if (order_.jumps) cudaFree(order_.jumps);
    order_ = RayOrderBuffers{};
    order_cap_ = 0;
  }

  std::mutex mu_;
  int cuda_dev_ = -1;
  cudaDeviceProp prop_{};
  cudaStream_t stream_ = nullptr;
  uint8_t *arena_ = nullptr;
  AddressSpace space_;
  uint64_t generation_ = 0;
  std::unordered_map<uint64_t, std::vector<uint8_t>> small_shadow_;   // host copies of small uploads (argument blocks)
  std::unordered_map<uint32_t, uint32_t> dcrs_;
  Partition part_;
  // framebuffer of the image owner, mapped over NVLink (vx_b200_set_remote_framebuffer)
  void *remote_base_ = nullptr;
  uint32_t *remote_fb_ = nullptr;
  uint64_t remote_bytes_ = 0;
  uint8_t *peer_fb_arena_ = nullptr;   // same-process owner of the image (DeviceGroup, VORTEX_B200_FB=peer)
  // strip-pipelined readback
  static constexpr uint32_t kStrips = 8, kMaxStrips = 16;
  uint64_t rb_addr_ = 0, rb_bytes_ = 0;
  uint8_t *rb_host_ = nullptr;
  bool rb_issued_ = false, kend_valid_ = false, half_valid_ = false, strip_reverse_ = false;
  float strip_half_share_ = 0.5f;
  cudaStream_t aux_stream_ = nullptr, copy_stream_ = nullptr;
  cudaEvent_t ev_strip_[kMaxStrips + 1] = {}, ev_kend_ = nullptr, ev_half_ = nullptr;
  uint32_t *strip_counters_ = nullptr;
  int strips_init_locked() {
    if (aux_stream_) return 0;
    CUDA_TRY(cudaStreamCreateWithFlags(&aux_stream_, cudaStreamNonBlocking));
    CUDA_TRY(cudaStreamCreateWithFlags(&copy_stream_, cudaStreamNonBlocking));
    for (auto &e : ev_strip_) CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CUDA_TRY(cudaEventCreate(&ev_kend_));
    CUDA_TRY(cudaEventCreate(&ev_half_));
    CUDA_TRY(cudaMalloc(&strip_counters_, sizeof(uint32_t) * kMaxStrips));
    return 0;
  }
  void strips_free_locked() {
    if (!aux_stream_) return;
    cudaStreamDestroy(aux_stream_), cudaStreamDestroy(copy_stream_);
    for (auto &e : ev_strip_) cudaEventDestroy(e), e = nullptr;
    cudaEventDestroy(ev_kend_), cudaEventDestroy(ev_half_);
    cudaFree(strip_counters_);
    aux_stream_ = copy_stream_ = nullptr, ev_kend_ = ev_half_ = nullptr, strip_counters_ = nullptr;
  }
  uint8_t *stage_ = nullptr;           // page-locked ring for small uploads
  cudaEvent_t stage_ev_[kStageSlots] = {};
  int stage_next_ = 0;
  Output last_out_;
  // layout-A GPU triangle records
  GpuTri *gtri_ = nullptr;
  uint64_t gtri_cap_ = 0, gtri_slots_ = 0;
  uint64_t gtri_key_[4] = {0, 0, 0, 0};
  // layout-A 4-wide records (fast order of the state machine)
  GpuNode4 *wide_ = nullptr;
  uint32_t *wide_bases_ = nullptr, *wide_bases_host_ = nullptr;
  uint64_t wide_cap_ = 0, wide_bases_cap_ = 0, wide_key_[5] = {0};
  bool wide_valid_ = false;
  // layout-B decoded nodes (TLAS then BLAS) and edge-form triangles
  GpuNodeB *rtu_nodes_ = nullptr;
  GpuTri *rtu_tri_ = nullptr;
  uint64_t rtu_nodes_cap_ = 0, rtu_tri_cap_ = 0, rtu_key_[11] = {0};
  uint32_t *rtu_bases_ = nullptr, *rtu_bases_host_ = nullptr;
  uint64_t rtu_bases_cap_ = 0;
  bool rtu_valid_ = false;
  // run bookkeeping
  WorkCounters *counters_ = nullptr, *h_counters_ = nullptr;
  uint32_t *tile_counter_ = nullptr;
  cudaEvent_t ev_begin_ = nullptr, ev_prep_ = nullptr, ev_end_ = nullptr;
  bool run_pending_ = false;
  uint64_t launches_ = 0, prep_launches_ = 0;
  int stats_on_ = 0;
  vx_b200_stats_t stats_{};
};

#include "vx_group.inc"   // class DeviceGroup: what a vx_device_h points at (one or several Devices)

struct Buffer {   // what a vx_buffer_h points at (reference: callbacks.inc:14-18)
  DeviceGroup *device;
  uint64_t addr;
  uint64_t size;
};

}  // namespace vxb

using vxb::Buffer;
using Device = vxb::DeviceGroup;
using vxb::fail;

// =====================================================================================================
// C ABI (2): the runtime calls.  Argument checks follow runtime/common/callbacks.inc line by line.
// =====================================================================================================
static int b200_open(vx_device_h *hdevice) {
  if (!hdevice) return -1;
  auto d = new Device();
  if (d->init() != 0) {
    std::string keep = vxb::g_last_error;
    delete d;
    vxb::g_last_error = keep;
    return -1;
  }
  *hdevice = d;
  return 0;
}

VXB_EXPORT int vx_dev_close(vx_device_h hdevice) {
  if (!hdevice) return -1;
  if (getenv("VORTEX_B200_PERF")) vx_dump_perf(hdevice, stdout);
  delete (Device *)hdevice;
  return 0;
}

VXB_EXPORT int vx_dev_caps(vx_device_h hdevice, uint32_t caps_id, uint64_t *value) {
  if (!hdevice || !value) return -1;
  return ((Device *)hdevice)->get_caps(caps_id, value);
}

VXB_EXPORT int vx_mem_alloc(vx_device_h hdevice, uint64_t size, int flags, vx_buffer_h *hbuffer) {
  if (!hdevice || !hbuffer || size == 0) return -1;
  auto d = (Device *)hdevice;
  uint64_t addr;
  if (d->mem_alloc(size, flags, &addr) != 0) return -1;
  *hbuffer = new Buffer{d, addr, size};
  return 0;
}

VXB_EXPORT int vx_mem_reserve(vx_device_h hdevice, uint64_t address, uint64_t size, int flags, vx_buffer_h *hbuffer) {
  if (!hdevice || !hbuffer || size == 0) return -1;
  auto d = (Device *)hdevice;
  if (d->mem_reserve(address, size, flags) != 0) return -1;
  *hbuffer = new Buffer{d, address, size};
  return 0;
}

VXB_EXPORT int vx_mem_free(vx_buffer_h hbuffer) {
  if (!hbuffer) return 0;   // reference: freeing a null handle succeeds (callbacks.inc:100-101)
  auto b = (Buffer *)hbuffer;
  int err = b->device->mem_free(b->addr);
  delete b;
  return err;
}

VXB_EXPORT int vx_mem_access(vx_buffer_h hbuffer, uint64_t offset, uint64_t size, int flags) {
  if (!hbuffer) return -1;
  auto b = (Buffer *)hbuffer;
  if (offset + size > b->size) return -1;
  return b->device->mem_access(b->addr + offset, size, flags);
}

VXB_EXPORT int vx_mem_address(vx_buffer_h hbuffer, uint64_t *address) {
  if (!hbuffer || !address) return -1;
  *address = ((Buffer *)hbuffer)->addr;
  return 0;
}

VXB_EXPORT int vx_mem_info(vx_device_h hdevice, uint64_t *mem_free, uint64_t *mem_used) {
  if (!hdevice) return -1;
  return ((Device *)hdevice)->mem_info(mem_free, mem_used);
}

VXB_EXPORT int vx_copy_to_dev(vx_buffer_h hbuffer, const void *host_ptr, uint64_t dst_offset, uint64_t size) {
  if (!hbuffer || !host_ptr) return -1;
  auto b = (Buffer *)hbuffer;
  if (dst_offset + size > b->size) return -1;
  return b->device->upload(b->addr + dst_offset, host_ptr, size);
}

VXB_EXPORT int vx_copy_from_dev(void *host_ptr, vx_buffer_h hbuffer, uint64_t src_offset, uint64_t size) {
  if (!hbuffer || !host_ptr) return -1;
  auto b = (Buffer *)hbuffer;
  if (src_offset + size > b->size) return -1;
  return b->device->download(host_ptr, b->addr + src_offset, size);
}

VXB_EXPORT int vx_start(vx_device_h hdevice, vx_buffer_h hkernel, vx_buffer_h harguments) {
  if (!hdevice || !hkernel || !harguments) return -1;
  return ((Device *)hdevice)->start(((Buffer *)hkernel)->addr, ((Buffer *)harguments)->addr);
}

VXB_EXPORT int vx_ready_wait(vx_device_h hdevice, uint64_t timeout) {
  if (!hdevice) return -1;
  return ((Device *)hdevice)->ready_wait(timeout);
}

VXB_EXPORT int vx_dcr_read(vx_device_h hdevice, uint32_t addr, uint32_t *value) {
  if (!hdevice || !value) return -1;
  return ((Device *)hdevice)->dcr_read(addr, value);
}

VXB_EXPORT int vx_dcr_write(vx_device_h hdevice, uint32_t addr, uint32_t value) {
  if (!hdevice) return -1;
  return ((Device *)hdevice)->dcr_write(addr, value);
}

VXB_EXPORT int vx_mpm_query(vx_device_h hdevice, uint32_t addr, uint32_t core_id, uint64_t *value) {
  if (!hdevice || !value) return -1;
  return ((Device *)hdevice)->mpm_query(addr, core_id, value);
}

VXB_EXPORT int vx_dev_open(vx_device_h *hdevice) {
  // the stub opens the driver and then programs five DCRs (runtime/stub/vortex.cpp:25-49,84-91)
  vx_device_h h;
  if (b200_open(&h) != 0) return -1;
  vx_dcr_write(h, VX_DCR_BASE_STARTUP_ADDR0, 0x80000000u);
  vx_dcr_write(h, VX_DCR_BASE_STARTUP_ADDR1, 0);
  vx_dcr_write(h, VX_DCR_BASE_STARTUP_ARG0, 0);
  vx_dcr_write(h, VX_DCR_BASE_STARTUP_ARG1, 0);
  vx_dcr_write(h, VX_DCR_BASE_MPM_CLASS, 0);
  *hdevice = h;
  return 0;
}

// ---- upload helpers (reference: runtime/stub/utils.cpp:25-122) ----
VXB_EXPORT int vx_upload_kernel_bytes(vx_device_h hdevice, const void *content, uint64_t size, vx_buffer_h *hbuffer) {
  if (!hdevice || !content || size <= 8 || !hbuffer) return -1;
  if (size < 16) return -1;
  uint64_t vma[2];
  memcpy(vma, content, 16);   // image header: {min_vma, max_vma}
  uint64_t bin_size = size - 16, runtime_size = vma[1] - vma[0];
  if (vma[1] <= vma[0] || bin_size > runtime_size) return fail("malformed kernel image header");
  vx_buffer_h buf;
  if (vx_mem_reserve(hdevice, vma[0], runtime_size, 0, &buf) != 0) return -1;
  if (vx_mem_access(buf, 0, bin_size, VX_MEM_READ) != 0 || vx_mem_access(buf, bin_size, runtime_size - bin_size, VX_MEM_READ_WRITE) != 0 ||
      (bin_size && vx_copy_to_dev(buf, (const uint8_t *)content + 16, 0, bin_size) != 0)) {
    vx_mem_free(buf);
    return -1;
  }
  *hbuffer = buf;
  return 0;
}

static int read_file(const char *filename, std::vector<char> &content) {
  std::ifstream ifs(filename, std::ios::binary);
  if (!ifs) {
    fprintf(stderr, "Error: %s not found\n", filename);
    return -1;
  }
  ifs.seekg(0, ifs.end);
  auto size = ifs.tellg();
  content.resize((size_t)size);
  ifs.seekg(0, ifs.beg);
  ifs.read(content.data(), size);
  return 0;
}

VXB_EXPORT int vx_upload_kernel_file(vx_device_h hdevice, const char *filename, vx_buffer_h *hbuffer) {
  if (!hdevice || !filename || !hbuffer) return -1;
  std::vector<char> content;
  if (read_file(filename, content) != 0) return -1;
  return vx_upload_kernel_bytes(hdevice, content.data(), content.size(), hbuffer);
}

VXB_EXPORT int vx_upload_bytes(vx_device_h hdevice, const void *content, uint64_t size, vx_buffer_h *hbuffer) {
  if (!hdevice || !content || size == 0 || !hbuffer) return -1;
  vx_buffer_h buf;
  if (vx_mem_alloc(hdevice, size, VX_MEM_READ, &buf) != 0) return -1;
  if (vx_copy_to_dev(buf, content, 0, size) != 0) {
    vx_mem_free(buf);
    return -1;
  }
  *hbuffer = buf;
  return 0;
}

VXB_EXPORT int vx_upload_file(vx_device_h hdevice, const char *filename, vx_buffer_h *hbuffer) {
  if (!hdevice || !filename || !hbuffer) return -1;
  std::vector<char> content;
  if (read_file(filename, content) != 0) return -1;
  return vx_upload_bytes(hdevice, content.data(), content.size(), hbuffer);
}

VXB_EXPORT int vx_dump_perf(vx_device_h hdevice, FILE *stream) {
  if (!hdevice || !stream) return -1;
  vx_b200_stats_t s;
  ((Device *)hdevice)->get_stats(&s);
  double mrays = s.kernel_ms > 0 ? (double)s.rays / (s.kernel_ms * 1e3) : 0.0;
  fprintf(stream, "PERF: rays=%llu, kernel_ms=%.4f, Mrays/s=%.1f, launches=%llu, prep_ms=%.4f\n", (unsigned long long)s.rays, s.kernel_ms,
          mrays, (unsigned long long)s.kernel_launches, s.prep_ms);
  if (s.box_tests || s.tri_tests)
    fprintf(stream, "PERF: box_tests=%llu, tri_tests=%llu, blas_entered=%llu, shaded_hits=%llu\n", (unsigned long long)s.box_tests,
            (unsigned long long)s.tri_tests, (unsigned long long)s.blas_entered, (unsigned long long)s.shaded_hits);
  // the RT unit's own summary (sim/simx/rt_unit.cpp print_stats: SIMT efficiency, warps, per-thread latency), in
  // traversal steps instead of simulated cycles
  if (s.warp_steps)
    fprintf(stream, "PERF: rt: warps=%llu, warp_steps=%llu, thread_steps=%llu, simt_efficiency=%.4f, steps_per_ray=%.2f\n",
            (unsigned long long)s.warps, (unsigned long long)s.warp_steps, (unsigned long long)s.lane_steps,
            (double)s.lane_steps / (32.0 * (double)s.warp_steps), s.rays ? (double)s.lane_steps / (double)s.rays : 0.0);
  // RTUnit::print_stats' latency table (sim/simx/rt_unit.cpp:265-292): per warp status, what its lanes were doing
  uint64_t any = 0;
  for (int a = 0; a < 3; a++)
    for (int b = 0; b < 5; b++) any += s.latency_dist[a][b];
  if (any) {
    static const char *wn[3] = {"warp_fetching_work", "warp_shading", "warp_traversing"};
    static const char *rn[5] = {"awaiting_processing", "awaiting_scheduling", "awaiting_mf", "executing_op", "trace_complete"};
    for (int a = 0; a < 3; a++) {
      fprintf(stream, "PERF: rt: %s:", wn[a]);
      for (int b = 0; b < 5; b++) fprintf(stream, " %s=%.2f%%", rn[b], 100.0 * (double)s.latency_dist[a][b] / (double)any);
      fprintf(stream, "\n");
    }
    fprintf(stream, "PERF: rt: steps_per_ray_log2_hist:");
    for (int a = 0; a < 16; a++) fprintf(stream, " %llu", (unsigned long long)s.steps_hist[a]);
    fprintf(stream, "\n");
  }
  if (s.gpus > 1)
    fprintf(stream, "PERF: gpus=%u, group_ms=%.4f, kernel_ms fastest/slowest GPU = %.4f / %.4f\n", s.gpus, s.group_ms, s.kernel_ms_min, s.kernel_ms_max);
  return 0;
}

// =====================================================================================================
// C ABI (1): the dispatcher hook
// =====================================================================================================
VXB_EXPORT int vx_dev_init(callbacks_t *cb) {
  if (!cb) return -1;
  cb->dev_open = b200_open;   // the stub programs the start-up DCRs itself after this returns
  cb->dev_close = [](vx_device_h h) -> int {
    if (!h) return -1;
    delete (Device *)h;
    return 0;
  };
  cb->dev_caps = vx_dev_caps;
  cb->mem_alloc = vx_mem_alloc;
  cb->mem_reserve = vx_mem_reserve;
  cb->mem_free = vx_mem_free;
  cb->mem_access = vx_mem_access;
  cb->mem_address = vx_mem_address;
  cb->mem_info = vx_mem_info;
  cb->copy_to_dev = vx_copy_to_dev;
  cb->copy_from_dev = vx_copy_from_dev;
  cb->start = vx_start;
  cb->ready_wait = vx_ready_wait;
  cb->dcr_read = vx_dcr_read;
  cb->dcr_write = vx_dcr_write;
  cb->mpm_query = vx_mpm_query;
  return 0;
}

// =====================================================================================================
// C ABI (3): extensions
// =====================================================================================================
VXB_EXPORT int vx_b200_select_device(int cuda_device) {
  vxb::g_selected_device = cuda_device;
  return 0;
}
VXB_EXPORT int vx_b200_select_gpus(int n_gpus) {
  vxb::g_selected_gpus = n_gpus;
  return 0;
}
VXB_EXPORT int vx_b200_set_readback(vx_device_h h, vx_buffer_h hframebuffer, void *host_ptr, uint64_t bytes) {
  if (!h || !hframebuffer) return -1;
  auto b = (Buffer *)hframebuffer;
  if (host_ptr && bytes > b->size) return fail("vx_b200_set_readback: more bytes than the framebuffer holds");
  return ((Device *)h)->set_readback(b->addr, host_ptr, bytes);
}
VXB_EXPORT int vx_b200_flush_l2(vx_device_h h) {
  if (!h) return -1;
  return ((Device *)h)->flush_l2();
}
VXB_EXPORT int vx_b200_set_partition(vx_device_h h, uint32_t rank, uint32_t world, uint32_t band_rows) {
  if (!h) return -1;
  return ((Device *)h)->set_partition(rank, world, band_rows);
}
VXB_EXPORT int vx_b200_build_bvh(vx_device_h h, vx_buffer_h htri, uint64_t tri_first, uint64_t tri_count, vx_buffer_h hbvh,
                                 uint64_t node_first, vx_buffer_h htriIdx, uint64_t slot_first, uint64_t *nodes_used) {
  if (!h || !htri || !hbvh || !htriIdx) return -1;
  auto t = (Buffer *)htri, b = (Buffer *)hbvh, i = (Buffer *)htriIdx;
  if ((tri_first + tri_count) * sizeof(vxb::TriA) > t->size || (node_first + 2 * tri_count - 1) * sizeof(vxb::BvhNodeA) > b->size ||
      (slot_first + tri_count) * 4 > i->size)
    return fail("vx_b200_build_bvh: range outside a buffer");
  return ((Device *)h)->build_bvh(t->addr + tri_first * sizeof(vxb::TriA), tri_count, tri_first, b->addr + node_first * sizeof(vxb::BvhNodeA),
                                  i->addr + slot_first * 4, nodes_used);
}
VXB_EXPORT int vx_b200_build_bvh_static(vx_device_h h, vx_buffer_h htri, uint64_t tri_first, uint64_t tri_count, vx_buffer_h hbvh,
                                        uint64_t node_first, vx_buffer_h htriIdx, uint64_t slot_first, uint64_t *nodes_used) {
  if (!h || !htri || !hbvh || !htriIdx) return -1;
  auto t = (Buffer *)htri, b = (Buffer *)hbvh, i = (Buffer *)htriIdx;
  if (tri_count > vxb::kBuildStaticMaxTris)
    return fail("vx_b200_build_bvh_static: %llu triangles exceed the static scratch array's limit of %u (use vx_b200_build_bvh)",
                (unsigned long long)tri_count, vxb::kBuildStaticMaxTris);
  if (tri_count == 0 || (tri_first + tri_count) * sizeof(vxb::TriA) > t->size ||
      (node_first + 2 * tri_count - 1) * sizeof(vxb::BvhNodeA) > b->size || (slot_first + tri_count) * 4 > i->size)
    return fail("vx_b200_build_bvh_static: range outside a buffer");
  return ((Device *)h)->build_bvh_static(t->addr + tri_first * sizeof(vxb::TriA), tri_count, tri_first,
                                         b->addr + node_first * sizeof(vxb::BvhNodeA), i->addr + slot_first * 4, nodes_used);
}
VXB_EXPORT int vx_b200_build_qbvh(vx_device_h h, vx_buffer_h htri, uint64_t tri_first, uint64_t tri_count, vx_buffer_h htriEx,
                                  vx_buffer_h hqbvh, uint64_t node_first, uint64_t *nodes_used, float *root_box6) {
  if (!h || !htri || !hqbvh) return -1;
  auto t = (Buffer *)htri, q = (Buffer *)hqbvh, x = (Buffer *)htriEx;
  if ((tri_first + tri_count) * sizeof(vxb::TriA) > t->size || (x && (tri_first + tri_count) * sizeof(vxb::TriExB) > x->size) ||
      node_first * sizeof(vxb::QNodeB) >= q->size)
    return fail("vx_b200_build_qbvh: range outside a buffer");
  uint64_t capacity = q->size / sizeof(vxb::QNodeB) - node_first;
  return ((Device *)h)->build_qbvh(t->addr + tri_first * sizeof(vxb::TriA), tri_count, tri_first,
                                   x ? x->addr + tri_first * sizeof(vxb::TriExB) : ~0ull, q->addr + node_first * sizeof(vxb::QNodeB), capacity,
                                   nodes_used, root_box6);
}
VXB_EXPORT int vx_b200_export_buffer(vx_buffer_h hbuffer, void *ipc_handle_64, uint64_t *offset) {
  if (!hbuffer || !ipc_handle_64 || !offset) return -1;
  auto b = (Buffer *)hbuffer;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t");
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaSetDevice(b->device->cuda_device());
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, b->device->arena());
  if (e != cudaSuccess) return fail("cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
  memcpy(ipc_handle_64, &h, 64);
  *offset = b->addr;
  return 0;
}
VXB_EXPORT int vx_b200_set_remote_framebuffer(vx_device_h h, const void *ipc_handle_64, uint64_t offset, uint64_t bytes) {
  if (!h) return -1;
  return ((Device *)h)->set_remote_framebuffer(ipc_handle_64, offset, bytes);
}
VXB_EXPORT int vx_b200_mark_written(vx_buffer_h hbuffer) {
  if (!hbuffer) return -1;
  auto b = (Buffer *)hbuffer;
  return b->device->mark_written(b->addr);
}
VXB_EXPORT int vx_b200_probe_gather(vx_device_h h, uint64_t buffer_bytes, uint32_t record_bytes, double *gbs) {
  if (!h || !gbs) return -1;
  return ((Device *)h)->probe_gather(buffer_bytes, record_bytes, gbs);
}
VXB_EXPORT int vx_b200_device_ptr(vx_buffer_h hbuffer, uint64_t *cuda_ptr) {
  if (!hbuffer || !cuda_ptr) return -1;
  auto b = (Buffer *)hbuffer;
  if (b->device->make_resident(b->addr) != 0) return -1;   // several GPUs: the pieces of a kernel output are gathered on GPU 0 first
  *cuda_ptr = (uint64_t)(uintptr_t)(b->device->arena() + b->addr);
  return 0;
}
VXB_EXPORT int vx_b200_stream(vx_device_h h, uint64_t *cuda_stream) {
  if (!h || !cuda_stream) return -1;
  *cuda_stream = (uint64_t)(uintptr_t)((Device *)h)->stream();
  return 0;
}
VXB_EXPORT int vx_b200_last_kernel_ms(vx_device_h h, float *ms) {
  if (!h || !ms) return -1;
  *ms = ((Device *)h)->last_ms();
  return 0;
}
VXB_EXPORT int vx_b200_enable_stats(vx_device_h h, int enable) {
  if (!h) return -1;
  ((Device *)h)->enable_stats(enable);
  return 0;
}
VXB_EXPORT int vx_b200_get_stats(vx_device_h h, vx_b200_stats_t *stats) {
  if (!h || !stats) return -1;
  return ((Device *)h)->get_stats(stats);
}
VXB_EXPORT int vx_b200_trace_primary(vx_device_h h, vx_buffer_h harguments, vx_buffer_h hhits) {
  if (!h || !harguments || !hhits) return -1;
  return ((Device *)h)->trace(((Buffer *)harguments)->addr, nullptr, 0, ((Buffer *)hhits)->addr);
}
VXB_EXPORT int vx_b200_trace_rays(vx_device_h h, vx_buffer_h harguments, vx_buffer_h hrays, uint64_t count, vx_buffer_h hhits) {
  if (!h || !harguments || !hrays || !hhits) return -1;
  if (count == 0) return 0;
  uint64_t ra = ((Buffer *)hrays)->addr;
  return ((Device *)h)->trace(((Buffer *)harguments)->addr, &ra, count, ((Buffer *)hhits)->addr);
}
VXB_EXPORT int vx_b200_trace_rays_host(vx_device_h h, vx_buffer_h harguments, const void *host_rays, uint64_t count, void *host_hits) {
  if (!h || !harguments || !host_rays || !host_hits) return -1;
  if (count == 0) return 0;
  return ((Device *)h)->trace_host(((Buffer *)harguments)->addr, host_rays, count, host_hits);
}
VXB_EXPORT const char *vx_b200_last_error(void) { return vxb::g_last_error.c_str(); }
