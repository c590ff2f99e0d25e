// IDX dataset reader and helpers (native data path).
//
// The reference parses MNIST through convolutional.extract_data/extract_labels
// (imported at mpipy.py:12, called at mpipy.py:215-218), which gunzip the IDX
// files, skip the header and normalise pixels to [-0.5, 0.5].  This is the C++
// equivalent: zlib's gz* API reads both plain and gzip'd files transparently.
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include <zlib.h>

#include "../../include/mtb200_host.h"

namespace {

uint32_t be32(const unsigned char* p) {
  return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | (uint32_t)p[3];
}

int read_exact(gzFile f, void* dst, uint64_t n) {
  uint8_t* d = static_cast<uint8_t*>(dst);
  while (n > 0) {
    unsigned chunk = n > (1u << 30) ? (1u << 30) : (unsigned)n;
    int got = gzread(f, d, chunk);
    if (got <= 0) return MTB_ERR_FORMAT;
    d += got;
    n -= (uint64_t)got;
  }
  return MTB_OK;
}

// software CRC32C (Castagnoli), table driven
uint32_t g_table[8][256];
bool g_table_ready = false;
void init_table() {
  for (uint32_t i = 0; i < 256; ++i) {
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ 0x82F63B78u : (c >> 1);
    g_table[0][i] = c;
  }
  for (uint32_t i = 0; i < 256; ++i)
    for (int t = 1; t < 8; ++t)
      g_table[t][i] = (g_table[t - 1][i] >> 8) ^ g_table[0][g_table[t - 1][i] & 0xff];
  g_table_ready = true;
}

}  // namespace

extern "C" {

void mtb_normalize_u8(const uint8_t* in, float* out, int64_t n) {
  unsigned hw = std::thread::hardware_concurrency();
  int nt = (n < (1 << 20)) ? 1 : (int)(hw ? (hw > 16 ? 16 : hw) : 4);
  auto work = [&](int64_t lo, int64_t hi) {
    for (int64_t i = lo; i < hi; ++i) out[i] = ((float)in[i] - 127.5f) / 255.0f;
  };
  if (nt == 1) { work(0, n); return; }
  std::vector<std::thread> th;
  int64_t per = (n + nt - 1) / nt;
  for (int t = 0; t < nt; ++t) {
    int64_t lo = per * t, hi = lo + per > n ? n : lo + per;
    if (lo < hi) th.emplace_back(work, lo, hi);
  }
  for (auto& t : th) t.join();
}

int mtb_idx_read_images(const char* path, int64_t n, int64_t rows, int64_t cols, float* out_f32,
                        uint8_t* out_u8) {
  gzFile f = gzopen(path, "rb");
  if (!f) return MTB_ERR_SYS;
  unsigned char hdr[16];
  int e = read_exact(f, hdr, 16);
  if (e != MTB_OK || be32(hdr) != 2051u || (int64_t)be32(hdr + 4) < n ||
      (int64_t)be32(hdr + 8) != rows || (int64_t)be32(hdr + 12) != cols) {
    gzclose(f);
    return MTB_ERR_FORMAT;
  }
  uint64_t total = (uint64_t)n * rows * cols;
  std::vector<uint8_t> tmp;
  uint8_t* raw = out_u8;
  if (!raw) { tmp.resize(total); raw = tmp.data(); }
  e = read_exact(f, raw, total);
  gzclose(f);
  if (e != MTB_OK) return e;
  if (out_f32) mtb_normalize_u8(raw, out_f32, (int64_t)total);
  return MTB_OK;
}

int mtb_idx_read_labels(const char* path, int64_t n, int64_t* out_i64) {
  gzFile f = gzopen(path, "rb");
  if (!f) return MTB_ERR_SYS;
  unsigned char hdr[8];
  int e = read_exact(f, hdr, 8);
  if (e != MTB_OK || be32(hdr) != 2049u || (int64_t)be32(hdr + 4) < n) {
    gzclose(f);
    return MTB_ERR_FORMAT;
  }
  std::vector<uint8_t> tmp((size_t)n);
  e = read_exact(f, tmp.data(), (uint64_t)n);
  gzclose(f);
  if (e != MTB_OK) return e;
  for (int64_t i = 0; i < n; ++i) out_i64[i] = tmp[(size_t)i];
  return MTB_OK;
}

uint32_t mtb_crc32c(const void* data, uint64_t n, uint32_t init) {
  if (!g_table_ready) init_table();
  const uint8_t* p = static_cast<const uint8_t*>(data);
  uint32_t c = ~init;
  while (n >= 8) {
    uint32_t lo, hi;
    memcpy(&lo, p, 4);
    memcpy(&hi, p + 4, 4);
    lo ^= c;
    c = g_table[7][lo & 0xff] ^ g_table[6][(lo >> 8) & 0xff] ^ g_table[5][(lo >> 16) & 0xff] ^
        g_table[4][lo >> 24] ^ g_table[3][hi & 0xff] ^ g_table[2][(hi >> 8) & 0xff] ^
        g_table[1][(hi >> 16) & 0xff] ^ g_table[0][hi >> 24];
    p += 8;
    n -= 8;
  }
  while (n--) c = g_table[0][(c ^ *p++) & 0xff] ^ (c >> 8);
  return ~c;
}

}  // extern "C"
