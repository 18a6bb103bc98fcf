// tests/emu/emu_device.inl — TEST INFRASTRUCTURE ONLY: host stand-ins for the shared-memory addressing, mbarrier and bulk-copy
// (TMA) wrappers of rdf_fusion_b200/csrc/rf_device.cuh, included by that header (inside namespace rf::dev) when it is compiled
// by g++ for tests/emu.  A bulk copy completes at once (memcpy) and then signals its mbarrier, which is a legal — the most
// eager — schedule of the asynchronous engine; an mbarrier is {arrivals pending, bytes pending, phase} behind one mutex.
struct EmuMbar {
  uint32_t count, pending;
  int64_t tx;
  uint32_t phase;
};
static_assert(sizeof(EmuMbar) <= 24, "");
inline std::mutex& emu_mbar_mutex() {
  static std::mutex m;
  return m;
}
inline std::map<const void*, EmuMbar>& emu_mbars() {
  static std::map<const void*, EmuMbar> m;
  return m;
}
inline void emu_mbar_step(EmuMbar& b) {
  if (b.pending == 0 && b.tx == 0) {
    b.phase ^= 1;
    b.pending = b.count;
  }
}
// "shared-memory address": a 32-bit offset from an anchor inside the same library image (static storage is within 2 GB)
inline char* emu_smem_anchor() {
  static char anchor;
  return &anchor;
}
inline uint32_t smem_u32(const void* p) { return (uint32_t)(int32_t)((const char*)p - emu_smem_anchor()); }
inline uint32_t smem_ld32(uint32_t a) { return *(const uint32_t*)(emu_smem_anchor() + (int32_t)a); }
inline void smem_st32(uint32_t a, uint32_t v) { *(uint32_t*)(emu_smem_anchor() + (int32_t)a) = v; }

inline void mbar_init(uint64_t* bar, uint32_t count) {
  std::lock_guard<std::mutex> g(emu_mbar_mutex());
  emu_mbars()[bar] = EmuMbar{count, count, 0, 0};
}
inline void mbar_fence_init() {}
inline void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  std::lock_guard<std::mutex> g(emu_mbar_mutex());
  EmuMbar& b = emu_mbars().at(bar);
  b.tx += bytes;
  b.pending--;
  emu_mbar_step(b);
}
inline void mbar_arrive(uint64_t* bar) {
  std::lock_guard<std::mutex> g(emu_mbar_mutex());
  EmuMbar& b = emu_mbars().at(bar);
  b.pending--;
  emu_mbar_step(b);
}
inline void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (true) {
    {
      std::lock_guard<std::mutex> g(emu_mbar_mutex());
      if (emu_mbars().at(bar).phase != parity) return;  // the phase with this parity has completed
    }
    std::this_thread::yield();
  }
}
inline void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  std::memcpy(smem_dst, gsrc, bytes);
  std::lock_guard<std::mutex> g(emu_mbar_mutex());
  EmuMbar& b = emu_mbars().at(bar);
  b.tx -= bytes;
  emu_mbar_step(b);
}
inline void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes) { std::memcpy(gdst, smem_src, bytes); }
inline void bulk_prefetch_l2(const void*, uint32_t) {}
inline void bulk_commit() {}
template <int N>
inline void bulk_wait_read() {}
template <int N>
inline void bulk_wait() {}
inline void fence_proxy_async() {}
