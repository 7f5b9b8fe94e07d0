// memtune.h — keep freed model-sized buffers inside the process.
//
// A push or a pull of a large model allocates a handful of model-sized buffers (receive body,
// decoded tensors, encoded request).  glibc serves every allocation above 32 MiB with its own
// mmap and gives it back with munmap, so each of them is first-touched page by page on every
// iteration — on this class of machine ~0.3 s per 436 MB buffer, more than the copy into it
// (0.07 s warm).  With mmap disabled for malloc and trimming deferred, freed buffers stay in
// the heap and the next iteration reuses warm pages: a 109 M-parameter push + pull goes from
// ~3.5 s to ~2.0 s per iteration (profiles/r2/host_plane.md).  The price is that the process
// keeps its peak footprint; PSB_RETAIN_MEMORY=0 restores the default behaviour.
#pragma once

#include <malloc.h>

#include <climits>
#include <cstdlib>
#include <cstring>

namespace psb {

inline void retain_freed_memory() {
  const char* e = std::getenv("PSB_RETAIN_MEMORY");
  if (e && std::strcmp(e, "0") == 0) return;
  mallopt(M_MMAP_MAX, 0);              // large blocks come from the heap (brk; mmap on failure)
  mallopt(M_TRIM_THRESHOLD, INT_MAX);  // do not hand the top of the heap back after each free
  mallopt(M_TOP_PAD, 1 << 30);         // grow in big steps; a trim keeps this much
}

}  // namespace psb
