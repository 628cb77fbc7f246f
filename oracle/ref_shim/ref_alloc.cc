// CPU ORACLE support (test infrastructure, not product code) -- see ref_alloc.h.
// DistributeOctTree sorts (size, ExtractorNode*) pairs (ORBextractor.cc:680), i.e. ties between nodes holding the same number of
// points are broken by heap ADDRESS, which under glibc malloc depends on the allocation history of the process.
// Replacing the global operator new inside the reference libraries (linked -Bsymbolic, so only their own allocations are
// affected) with a monotone bump arena makes the tie "later created = higher address": the declared, reproducible rule.
// operator delete is a no-op for arena memory; arenas are recycled whole.
#include "ref_alloc.h"

#include <sys/mman.h>

#include <atomic>
#include <cstdlib>
#include <new>

namespace refalloc {
const size_t kArenaBytes = (size_t)1 << 30;  // virtual; pages are committed on first touch only
const int kMaxArenas = 512;
struct Arena { char* base; std::atomic<int> busy; };
static Arena g_arenas[kMaxArenas];
static std::atomic<int> g_narenas{0};
static thread_local char* tl_cur = nullptr;
static thread_local char* tl_end = nullptr;
static thread_local bool tl_overflow = false;

Arena* acquire() {
  const int n = g_narenas.load();
  for (int i = 0; i < n && i < kMaxArenas; i++) {
    int expected = 0;
    if (g_arenas[i].base && g_arenas[i].busy.compare_exchange_strong(expected, 1)) return &g_arenas[i];
  }
  const int i = g_narenas.fetch_add(1);
  if (i >= kMaxArenas) { g_narenas.fetch_sub(1); return nullptr; }
  void* p = mmap(nullptr, kArenaBytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
  if (p == MAP_FAILED) return nullptr;
  g_arenas[i].busy.store(1);
  g_arenas[i].base = (char*)p;
  return &g_arenas[i];
}
void release(Arena* a) { if (a) a->busy.store(0); }
void bind(Arena* a) { tl_cur = a->base; tl_end = a->base + kArenaBytes; tl_overflow = false; }
void unbind() { tl_cur = tl_end = nullptr; }
bool overflowed() { return tl_overflow; }
static inline bool in_arena(void* p) {
  const int n = g_narenas.load(std::memory_order_relaxed);
  for (int i = 0; i < n && i < kMaxArenas; i++) {
    char* b = g_arenas[i].base;
    if (b && (char*)p >= b && (char*)p < b + kArenaBytes) return true;
  }
  return false;
}
static inline void* alloc(size_t n) {
  if (tl_cur) {
    const size_t a = (n + 15) & ~(size_t)15;
    if (tl_cur + a <= tl_end) { void* p = tl_cur; tl_cur += a; return p; }
    tl_overflow = true;
  }
  void* p = malloc(n ? n : 1);
  if (!p) throw std::bad_alloc();
  return p;
}
}  // namespace refalloc

void* operator new(size_t n) { return refalloc::alloc(n); }
void* operator new[](size_t n) { return refalloc::alloc(n); }
void operator delete(void* p) noexcept { if (p && !refalloc::in_arena(p)) free(p); }
void operator delete[](void* p) noexcept { operator delete(p); }
void operator delete(void* p, size_t) noexcept { operator delete(p); }
void operator delete[](void* p, size_t) noexcept { operator delete(p); }
