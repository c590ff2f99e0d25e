// Shared-memory process group for CPU ranks on one host.
//
// The reference gets its communicator from an MPI C library through mpi4py
// (mpipy.py:5, :208-210) and uses exactly three collectives on host buffers:
// Scatter (mpipy.py:236-241), Gather (mpipy.py:121-127) and a commented-out
// Barrier (mpipy.py:92-93).  No MPI exists on the target image, so this file is
// the native stand-in: rank/size from the launcher, a POSIX shared-memory
// segment as the transport, and barrier / bcast / scatter / gather / allgather /
// allreduce built on it.  Every wait is bounded: a peer that died or stalled
// turns into error MTB_ERR_TIMEOUT (or MTB_ERR_ABORTED once any rank has raised
// the abort word) instead of the silent hang an MPI_Gather would give
// (SURVEY section 5 "Failure detection").
//
// Layout of the segment:
//   Header | slot[0] | slot[1] | ... | slot[world-1]
// Every collective moves data through the slots in chunks of slot_bytes and is
// bracketed by counter barriers.  Counters only ever increase, so there is no
// flag-reset race between consecutive collectives.
#include <atomic>
#include <cerrno>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "../../include/mtb200_host.h"

namespace {

constexpr uint64_t kMagic = 0x4d54423230304831ull;  // "MTB200H1"

struct alignas(64) Header {
  std::atomic<uint64_t> magic;
  std::atomic<uint32_t> world;
  std::atomic<uint32_t> attached;
  uint64_t slot_bytes;
  alignas(64) std::atomic<uint64_t> arrive;   // total barrier arrivals so far
  alignas(64) std::atomic<uint32_t> abort_word;
  alignas(64) std::atomic<uint64_t> heartbeat[64];
};

struct Comm {
  int rank = 0, world = 1;
  uint64_t slot_bytes = 0;
  size_t map_bytes = 0;
  Header* hdr = nullptr;
  uint8_t* slots = nullptr;
  uint64_t epoch = 0;          // barriers completed by this rank
  int timeout_ms = 30000;
  std::string name;
  bool owner = false;
};

inline uint8_t* slot(Comm* c, int r) { return c->slots + (size_t)r * c->slot_bytes; }

int wait_barrier(Comm* c) {
  c->epoch += 1;
  c->hdr->heartbeat[c->rank].fetch_add(1, std::memory_order_relaxed);
  c->hdr->arrive.fetch_add(1, std::memory_order_acq_rel);
  const uint64_t target = c->epoch * (uint64_t)c->world;
  auto t0 = std::chrono::steady_clock::now();
  uint32_t spins = 0;
  while (c->hdr->arrive.load(std::memory_order_acquire) < target) {
    if (c->hdr->abort_word.load(std::memory_order_acquire) != 0) return MTB_ERR_ABORTED;
    if (++spins > 2000) {
      std::this_thread::sleep_for(std::chrono::microseconds(50));
      auto ms = std::chrono::duration_cast<std::chrono::milliseconds>(
                    std::chrono::steady_clock::now() - t0).count();
      if (ms > c->timeout_ms) {
        c->hdr->abort_word.store(1u + (uint32_t)c->rank, std::memory_order_release);
        return MTB_ERR_TIMEOUT;
      }
    }
  }
  return MTB_OK;
}

#define CHECK_BAR(c) do { int _e = wait_barrier(c); if (_e != MTB_OK) return _e; } while (0)

}  // namespace

extern "C" {

int mtb_shm_open(const char* name, int rank, int world, uint64_t slot_bytes, int timeout_ms,
                 void** out) {
  if (!name || !out || world < 1 || world > 64 || rank < 0 || rank >= world) return MTB_ERR_ARG;
  slot_bytes = (slot_bytes + 63) / 64 * 64;
  Comm* c = new Comm();
  c->rank = rank; c->world = world; c->slot_bytes = slot_bytes; c->timeout_ms = timeout_ms;
  c->name = name;
  c->map_bytes = sizeof(Header) + (size_t)world * slot_bytes;
  int fd = -1;
  if (rank == 0) {
    shm_unlink(name);
    fd = shm_open(name, O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0) { delete c; return MTB_ERR_SYS; }
    if (ftruncate(fd, (off_t)c->map_bytes) != 0) { close(fd); shm_unlink(name); delete c; return MTB_ERR_SYS; }
    c->owner = true;
  } else {
    auto t0 = std::chrono::steady_clock::now();
    for (;;) {
      fd = shm_open(name, O_RDWR, 0600);
      if (fd >= 0) {
        struct stat st;
        if (fstat(fd, &st) == 0 && (size_t)st.st_size >= c->map_bytes) break;
        close(fd); fd = -1;
      }
      std::this_thread::sleep_for(std::chrono::milliseconds(2));
      auto ms = std::chrono::duration_cast<std::chrono::milliseconds>(
                    std::chrono::steady_clock::now() - t0).count();
      if (ms > timeout_ms) { delete c; return MTB_ERR_TIMEOUT; }
    }
  }
  void* p = mmap(nullptr, c->map_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (p == MAP_FAILED) { delete c; return MTB_ERR_SYS; }
  c->hdr = reinterpret_cast<Header*>(p);
  c->slots = reinterpret_cast<uint8_t*>(p) + sizeof(Header);
  if (rank == 0) {
    // ftruncate zero-fills, so every counter starts at 0; publish magic last.
    c->hdr->slot_bytes = slot_bytes;
    c->hdr->world.store((uint32_t)world, std::memory_order_relaxed);
    c->hdr->magic.store(kMagic, std::memory_order_release);
  } else {
    auto t0 = std::chrono::steady_clock::now();
    while (c->hdr->magic.load(std::memory_order_acquire) != kMagic) {
      std::this_thread::sleep_for(std::chrono::milliseconds(1));
      auto ms = std::chrono::duration_cast<std::chrono::milliseconds>(
                    std::chrono::steady_clock::now() - t0).count();
      if (ms > timeout_ms) { munmap(p, c->map_bytes); delete c; return MTB_ERR_TIMEOUT; }
    }
    if (c->hdr->world.load() != (uint32_t)world || c->hdr->slot_bytes != slot_bytes) {
      munmap(p, c->map_bytes); delete c; return MTB_ERR_ARG;
    }
  }
  c->hdr->attached.fetch_add(1);
  *out = c;
  int e = wait_barrier(c);      // everyone mapped before anyone proceeds / unlinks
  if (e == MTB_OK && rank == 0) shm_unlink(name);   // segment lives until last unmap
  return e;
}

int mtb_shm_close(void* h) {
  Comm* c = static_cast<Comm*>(h);
  if (!c) return MTB_ERR_ARG;
  munmap(c->hdr, c->map_bytes);
  delete c;
  return MTB_OK;
}

int mtb_shm_rank(void* h) { return static_cast<Comm*>(h)->rank; }
int mtb_shm_size(void* h) { return static_cast<Comm*>(h)->world; }
void mtb_shm_set_timeout(void* h, int ms) { static_cast<Comm*>(h)->timeout_ms = ms; }
void mtb_shm_abort(void* h) {
  Comm* c = static_cast<Comm*>(h);
  c->hdr->abort_word.store(1u + (uint32_t)c->rank, std::memory_order_release);
}
uint64_t mtb_shm_heartbeat(void* h, int rank) {
  return static_cast<Comm*>(h)->hdr->heartbeat[rank].load(std::memory_order_relaxed);
}

int mtb_shm_barrier(void* h) { return wait_barrier(static_cast<Comm*>(h)); }

// root's buffer -> everyone
int mtb_shm_bcast(void* h, void* buf, uint64_t bytes, int root) {
  Comm* c = static_cast<Comm*>(h);
  uint8_t* b = static_cast<uint8_t*>(buf);
  for (uint64_t off = 0; off < bytes; off += c->slot_bytes) {
    uint64_t n = bytes - off < c->slot_bytes ? bytes - off : c->slot_bytes;
    if (c->rank == root) memcpy(slot(c, root), b + off, n);
    CHECK_BAR(c);
    if (c->rank != root) memcpy(b + off, slot(c, root), n);
    CHECK_BAR(c);
  }
  return MTB_OK;
}

// send (root only, world*bytes) -> recv (bytes each): MPI_Scatter, mpipy.py:236-241
int mtb_shm_scatter(void* h, const void* send, void* recv, uint64_t bytes, int root) {
  Comm* c = static_cast<Comm*>(h);
  const uint8_t* s = static_cast<const uint8_t*>(send);
  uint8_t* r = static_cast<uint8_t*>(recv);
  for (uint64_t off = 0; off < bytes; off += c->slot_bytes) {
    uint64_t n = bytes - off < c->slot_bytes ? bytes - off : c->slot_bytes;
    if (c->rank == root)
      for (int p = 0; p < c->world; ++p) memcpy(slot(c, p), s + (uint64_t)p * bytes + off, n);
    CHECK_BAR(c);
    memcpy(r + off, slot(c, c->rank), n);
    CHECK_BAR(c);
  }
  return MTB_OK;
}

// send (bytes each) -> recv (root only, world*bytes): MPI_Gather, mpipy.py:121-127
int mtb_shm_gather(void* h, const void* send, void* recv, uint64_t bytes, int root) {
  Comm* c = static_cast<Comm*>(h);
  const uint8_t* s = static_cast<const uint8_t*>(send);
  uint8_t* r = static_cast<uint8_t*>(recv);
  for (uint64_t off = 0; off < bytes; off += c->slot_bytes) {
    uint64_t n = bytes - off < c->slot_bytes ? bytes - off : c->slot_bytes;
    memcpy(slot(c, c->rank), s + off, n);
    CHECK_BAR(c);
    if (c->rank == root)
      for (int p = 0; p < c->world; ++p) memcpy(r + (uint64_t)p * bytes + off, slot(c, p), n);
    CHECK_BAR(c);
  }
  return MTB_OK;
}

int mtb_shm_allgather(void* h, const void* send, void* recv, uint64_t bytes) {
  Comm* c = static_cast<Comm*>(h);
  const uint8_t* s = static_cast<const uint8_t*>(send);
  uint8_t* r = static_cast<uint8_t*>(recv);
  for (uint64_t off = 0; off < bytes; off += c->slot_bytes) {
    uint64_t n = bytes - off < c->slot_bytes ? bytes - off : c->slot_bytes;
    memcpy(slot(c, c->rank), s + off, n);
    CHECK_BAR(c);
    for (int p = 0; p < c->world; ++p) memcpy(r + (uint64_t)p * bytes + off, slot(c, p), n);
    CHECK_BAR(c);
  }
  return MTB_OK;
}

// In-place fp32 allreduce, sum or average.  Two-shot inside the segment: rank r
// reduces slice r of every slot (always in rank order 0..N-1, so every rank
// ends with bit-identical results), writes it back into ALL slots' slice r...
// which here is simply its own slice of slot[r]; peers then read slice p from
// slot[p].  This is the host-memory analogue of the device two-shot kernel.
int mtb_shm_allreduce_f32(void* h, float* buf, uint64_t count, int average) {
  Comm* c = static_cast<Comm*>(h);
  const uint64_t cap = c->slot_bytes / sizeof(float);
  const float scale = average ? 1.0f / (float)c->world : 1.0f;
  for (uint64_t off = 0; off < count; off += cap) {
    uint64_t n = count - off < cap ? count - off : cap;
    float* mine = reinterpret_cast<float*>(slot(c, c->rank));
    memcpy(mine, buf + off, n * sizeof(float));
    CHECK_BAR(c);
    uint64_t per = (n + c->world - 1) / c->world;
    uint64_t lo = per * (uint64_t)c->rank, hi = lo + per;
    if (lo > n) lo = n;
    if (hi > n) hi = n;
    // reduce into a private temp first: slot[rank][lo:hi) is still being read by nobody
    // else (peers only read their own slice range), so accumulate in place in rank order.
    for (uint64_t i = lo; i < hi; ++i) {
      float acc = 0.f;
      for (int p = 0; p < c->world; ++p) acc += reinterpret_cast<float*>(slot(c, p))[i];
      mine[i] = acc * scale;   // slot[rank] slice [lo,hi) now holds the result
    }
    CHECK_BAR(c);
    for (int p = 0; p < c->world; ++p) {
      uint64_t plo = per * (uint64_t)p, phi = plo + per;
      if (plo > n) plo = n;
      if (phi > n) phi = n;
      if (phi > plo)
        memcpy(buf + off + plo, reinterpret_cast<float*>(slot(c, p)) + plo, (phi - plo) * sizeof(float));
    }
    CHECK_BAR(c);
  }
  return MTB_OK;
}

}  // extern "C"
