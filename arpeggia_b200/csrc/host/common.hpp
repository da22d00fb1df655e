// Shared internals of the host side (not part of the ABI).
#pragma once
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <string_view>
#include <utility>
#include <vector>

#include "arpeggia_cuda.h"
#include "arpeggia_host.h"

// Tracing: NVTX ranges around every phase of the path (header-only NVTX v3: no link dependency, a no-op unless a
// tool -- nsys, ncu --nvtx -- is attached).  Ranges are named "arp:<phase>" (device path) / "arph:<phase>" (host side).
#include <nvtx3/nvToolsExt.h>

namespace arp {

struct NvtxRange {
  bool open = true;
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  void end() {
    if (open) nvtxRangePop();
    open = false;
  }
  ~NvtxRange() { end(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};

struct Error {
  int code;
  std::string msg;
};

// thread-local message behind arph_last_error() / arp_last_error(NULL)
std::string& last_error();

inline std::string fixed_field(const char* p, size_t cap) {
  size_t n = 0;
  while (n < cap && p[n] != '\0') ++n;
  return std::string(p, n);
}

// ---- geometry (geometry.cpp) ----------------------------------------------------------------------
struct Plane {
  double c[3];
  double n[3];
};
// center_and_normal, residues.rs:270-298.  pts = 3*n doubles (xyz interleaved). false when n < 3.
bool fit_plane(const double* pts, int64_t n, Plane& out);
double plane_point_dist(const Plane& p, const double pt[3]);           // residues.rs:31-35
double plane_point_angle(const Plane& p, const double pt[3]);          // residues.rs:45-54
double plane_dihedral(const double n1[3], const double n2[3]);         // residues.rs:65-75

// Exact cosine-domain images of every angle comparison on the path.  The reference compares
// acos(c).to_degrees() (optionally folded to <= 90) with a constant; acos is monotone, so each comparison
// is equivalent to comparing c with a boundary double, found here once by bisection with the HOST libm.
// The device then reproduces the host's decisions bit for bit without evaluating acos.
struct AngleBounds {
  // plain angle A(c) = deg(acos(c)), decreasing in c:  A >= T  <=>  c <= ge[T]  (and c >= -1)
  double a_ge_90, a_ge_130, a_ge_60;
  // A <= 120  <=>  c >= a_le_120  (and c <= 1)
  double a_le_120;
  // folded angle F(c) (residues.rs:49-53):  F <= T  <=>  c >= pos[T] (c in [0,1])  or  c <= neg[T] (c in [-1,0))
  double f_le_30_pos, f_le_30_neg;
  double f_le_60_pos, f_le_60_neg;
  // F < T
  double f_lt_30_pos, f_lt_30_neg;
  double f_lt_60_pos, f_lt_60_neg;
};
const AngleBounds& angle_bounds();

// std::vector whose resize() leaves new elements uninitialised (trivial types only): the big per-atom arrays are resized
// once and then written by every host thread, so the pages are first touched -- and faulted in -- by the threads that
// fill them instead of by one thread zero-filling 100 MB.
template <class T>
struct NoInitAlloc : std::allocator<T> {
  template <class U>
  struct rebind {
    using other = NoInitAlloc<U>;
  };
  NoInitAlloc() = default;
  template <class U>
  NoInitAlloc(const NoInitAlloc<U>&) {}
  template <class U>
  void construct(U* p) noexcept {
    ::new (static_cast<void*>(p)) U;
  }
  template <class U, class... Args>
  void construct(U* p, Args&&... args) {
    ::new (static_cast<void*>(p)) U(std::forward<Args>(args)...);
  }
};
template <class T>
using RawVec = std::vector<T, NoInitAlloc<T>>;

// the text of a file: mapped read-only when it is a regular file, else read into a string
struct FileText {
  std::string owned;
  void* map = nullptr;
  size_t map_len = 0;
  std::string_view view;
  bool open(const char* path) {
    const int fd = ::open(path, O_RDONLY);
    if (fd < 0) return false;
    struct stat st;
    if (::fstat(fd, &st) == 0 && S_ISREG(st.st_mode) && st.st_size > 0) {
      void* m = ::mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
      if (m != MAP_FAILED) {
        map = m;
        map_len = (size_t)st.st_size;
        ::madvise(m, map_len, MADV_WILLNEED);
        view = std::string_view(static_cast<const char*>(m), map_len);
        ::close(fd);
        return true;
      }
    }
    char chunk[1 << 16];
    ssize_t got;
    while ((got = ::read(fd, chunk, sizeof chunk)) > 0) owned.append(chunk, (size_t)got);
    ::close(fd);
    if (got < 0) return false;
    view = owned;
    return true;
  }
  ~FileText() {
    if (map) ::munmap(map, map_len);
  }
  FileText() = default;
  FileText(const FileText&) = delete;
  FileText& operator=(const FileText&) = delete;
};

// ARPH_TIMING=1: the phases of the packer and of the side-table builder on stderr (a development aid)
struct PhaseClock {
  bool on = std::getenv("ARPH_TIMING") != nullptr;
  std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now();
  void mark(const char* what) {
    if (!on) return;
    const auto now = std::chrono::steady_clock::now();
    std::fprintf(stderr, "[arph] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(now - t).count());
    t = now;
  }
};

}  // namespace arp
