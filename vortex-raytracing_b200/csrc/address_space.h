// address_space.h — the Vortex device address space as this backend hands it out.
//
// A Vortex host stores the addresses returned by vx_mem_address() into kernel_arg_t (and, on the RT-unit
// path, truncates them to 32 bits for DCR writes: tests/regression/raytracing/tracer.cpp:252-256), so they
// must look like simx's: offsets in [USER_BASE_ADDR, GLOBAL_MEM_SIZE) = [0x10000, 4 GiB), never raw CUDA
// pointers (reference: runtime/simx/vortex.cpp:52,208-246 over sim/common/mem_alloc.h:72-152).
// Here one contiguous CUDA allocation backs the whole space: device pointer = arena_base + address.
#pragma once
#include <cstdint>
#include <map>

namespace vxb {

class AddressSpace {
public:
  struct Block {
    uint64_t size;       // bytes as requested by the caller
    uint64_t span;       // bytes occupied (aligned)
    uint64_t write_gen;  // generation of the last upload into this block (0 = never written)
    int flags;
  };

  static constexpr uint64_t kUserBase = 0x10000;   // USER_BASE_ADDR (hw/VX_config.toml:160)
  static constexpr uint64_t kBlockAlign = 64;      // CACHE_BLOCK_SIZE (runtime/common/common.h:27)
  static constexpr uint64_t kPageAlign = 4096;     // MEM_PAGE_SIZE
  static constexpr uint64_t kPlaceAlign = 256;     // start alignment of every block (>= 16 B for 128-bit loads)

  explicit AddressSpace(uint64_t capacity) : capacity_(capacity) {}

  uint64_t capacity() const { return capacity_; }
  uint64_t allocated() const { return allocated_; }
  uint64_t free_bytes() const { return capacity_ - kUserBase - allocated_; }

  // first fit over the gaps between live blocks
  int allocate(uint64_t size, int flags, uint64_t *addr) {
    if (size == 0 || addr == nullptr) return -1;
    uint64_t span = align_up(size, kPlaceAlign);
    uint64_t cursor = kUserBase;
    for (auto &kv : blocks_) {
      if (kv.first >= cursor && kv.first - cursor >= span) break;
      uint64_t end = align_up(kv.first + kv.second.span, kPlaceAlign);
      if (end > cursor) cursor = end;
    }
    if (cursor + span > capacity_ || cursor + span < cursor) return -1;
    blocks_[cursor] = Block{size, span, 0, flags};
    allocated_ += span;
    *addr = cursor;
    return 0;
  }

  // fixed-address reservation (kernel images at 0x8000_0000 etc.: runtime/stub/utils.cpp:37)
  int reserve(uint64_t addr, uint64_t size, int flags) {
    if (size == 0) return -1;
    uint64_t span = align_up(size, kPageAlign);
    if (addr + span > capacity_ || addr + span < addr) return -1;
    auto it = blocks_.lower_bound(addr);
    if (it != blocks_.end() && it->first < addr + span) return -1;
    if (it != blocks_.begin()) {
      auto pv = std::prev(it);
      if (pv->first + pv->second.span > addr) return -1;
    }
    blocks_[addr] = Block{size, span, 0, flags};
    allocated_ += span;
    return 0;
  }

  int release(uint64_t addr) {
    auto it = blocks_.find(addr);
    if (it == blocks_.end()) return -1;
    allocated_ -= it->second.span;
    blocks_.erase(it);
    return 0;
  }

  // block containing [addr, addr+size); returns its start in *start
  Block *find(uint64_t addr, uint64_t size, uint64_t *start) {
    auto it = blocks_.upper_bound(addr);
    if (it == blocks_.begin()) return nullptr;
    --it;
    if (addr + size > it->first + it->second.span || addr + size < addr) return nullptr;
    if (start) *start = it->first;
    return &it->second;
  }

  static uint64_t align_up(uint64_t v, uint64_t a) { return (v + a - 1) / a * a; }

private:
  uint64_t capacity_;
  uint64_t allocated_ = 0;
  std::map<uint64_t, Block> blocks_;
};

}  // namespace vxb
