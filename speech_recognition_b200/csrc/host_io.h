// Host-side byte work next to the front-end: RIFF/WAVE header walk for the
// 16-bit PCM ingest and the CRC-32C of TFRecord framing (the reference's feature cache,
// run/make_tfrecord.py:39-58 / data.py:64-79).  No CUDA in this header.
#pragma once
#include <cstdint>
#include <cstring>

namespace b200mel {

struct WavInfo {
  int32_t format = 0;  // 1 = PCM, 0xFFFE = extensible (sub-format read from the extension)
  int32_t channels = 0;
  int32_t sample_rate = 0;
  int32_t bits = 0;
  int64_t data_offset = 0;
  int64_t data_bytes = 0;
};

inline uint32_t rd_u32(const uint8_t* p) { return p[0] | (p[1] << 8) | (p[2] << 16) | (static_cast<uint32_t>(p[3]) << 24); }
inline uint16_t rd_u16(const uint8_t* p) { return static_cast<uint16_t>(p[0] | (p[1] << 8)); }

// Returns 0 on success; 1 = not RIFF/WAVE, 2 = no fmt chunk before data, 3 = no data chunk, 4 = truncated fmt.
// A data chunk whose declared size runs past the end of the file (streamed writers leave 0 / 0xFFFFFFFF) is
// clipped to the bytes present.
inline int wav_parse(const uint8_t* d, int64_t n, WavInfo* out) {
  if (n < 12 || std::memcmp(d, "RIFF", 4) != 0 || std::memcmp(d + 8, "WAVE", 4) != 0) return 1;
  int64_t pos = 12;
  bool have_fmt = false;
  WavInfo w;
  while (pos + 8 <= n) {
    const uint8_t* c = d + pos;
    int64_t size = rd_u32(c + 4);
    int64_t body = pos + 8;
    if (std::memcmp(c, "fmt ", 4) == 0) {
      if (size < 16 || body + 16 > n) return 4;
      w.format = rd_u16(d + body);
      w.channels = rd_u16(d + body + 2);
      w.sample_rate = static_cast<int32_t>(rd_u32(d + body + 4));
      w.bits = rd_u16(d + body + 14);
      if (w.format == 0xFFFE && size >= 26 && body + 26 <= n) w.format = rd_u16(d + body + 24);
      have_fmt = true;
    } else if (std::memcmp(c, "data", 4) == 0) {
      if (!have_fmt) return 2;
      w.data_offset = body;
      w.data_bytes = (size == 0xFFFFFFFFll || body + size > n) ? n - body : size;
      *out = w;
      return 0;
    }
    pos = body + size + (size & 1);
  }
  return 3;
}

// CRC-32C (Castagnoli, reflected polynomial 0x82F63B78), slicing-by-8.
struct Crc32cTables {
  uint32_t t[8][256];
  Crc32cTables() {
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t c = i;
      for (int k = 0; k < 8; ++k) c = (c >> 1) ^ ((c & 1) ? 0x82F63B78u : 0u);
      t[0][i] = c;
    }
    for (uint32_t i = 0; i < 256; ++i)
      for (int s = 1; s < 8; ++s) t[s][i] = (t[s - 1][i] >> 8) ^ t[0][t[s - 1][i] & 0xFF];
  }
};

inline uint32_t crc32c(const uint8_t* p, int64_t n, uint32_t crc = 0) {
  static const Crc32cTables T;
  uint32_t c = ~crc;
  while (n > 0 && (reinterpret_cast<uintptr_t>(p) & 7)) {
    c = (c >> 8) ^ T.t[0][(c ^ *p++) & 0xFF];
    --n;
  }
  while (n >= 8) {
    uint64_t v;
    std::memcpy(&v, p, 8);
    v ^= c;
    c = T.t[7][v & 0xFF] ^ T.t[6][(v >> 8) & 0xFF] ^ T.t[5][(v >> 16) & 0xFF] ^ T.t[4][(v >> 24) & 0xFF] ^
        T.t[3][(v >> 32) & 0xFF] ^ T.t[2][(v >> 40) & 0xFF] ^ T.t[1][(v >> 48) & 0xFF] ^ T.t[0][(v >> 56) & 0xFF];
    p += 8;
    n -= 8;
  }
  while (n-- > 0) c = (c >> 8) ^ T.t[0][(c ^ *p++) & 0xFF];
  return ~c;
}

// TFRecord's masked CRC: rotate right by 15 and add a constant.
inline uint32_t tfrecord_masked_crc(const uint8_t* p, int64_t n) {
  uint32_t c = crc32c(p, n);
  return ((c >> 15) | (c << 17)) + 0xA282EAD8u;
}

}  // namespace b200mel
