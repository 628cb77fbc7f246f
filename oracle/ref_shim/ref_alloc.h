// CPU ORACLE support (test infrastructure, not product code).
// Node allocator of the reference builds (see ref_alloc.cc): while an arena is bound to the calling thread, every operator new
// is served from it by a monotone bump pointer, so "created later = higher address" -- the reproducible form of the
// (size, ExtractorNode*) tie of DistributeOctTree's sort (ORBextractor.cc:680).
#pragma once
namespace refalloc {
struct Arena;
Arena* acquire();          // a free arena, cursor reset; nullptr when none can be mapped
void release(Arena* a);    // memory handed out from it must no longer be in use
void bind(Arena* a);       // operator new on THIS thread now bumps from `a`
void unbind();
bool overflowed();         // this thread ran out of arena since the last bind (results must be discarded)
struct Scope {             // acquire + bind for the calling thread, unbind + release on exit
  Arena* a;
  Scope() : a(acquire()) { if (a) bind(a); }
  ~Scope() { unbind(); if (a) release(a); }
  bool ok() const { return a != nullptr; }
};
}  // namespace refalloc
