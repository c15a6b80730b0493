// nprand.hpp — bit-exact NumPy ``RandomState`` (MT19937 as modified by RandomKit) so that searcher sampling is
// reproducible from ``reproducibility.experiment_seed`` and matches NumPy draw for draw.
// Parity: reference master/pkg/nprand/nprand.go:1-168 (seeding, Bits32/64, byte stream, masked rejection
// sampling for bounded ints, 53-bit unit interval).
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>

namespace det {

class NpRand {
 public:
  explicit NpRand(uint32_t seed = 0) { this->seed(seed); }

  void seed(uint32_t s) {
    for (int pos = 0; pos < kN; ++pos) {
      key_[pos] = s;
      s = 1812433253u * (s ^ (s >> 30)) + static_cast<uint32_t>(pos) + 1u;
    }
    pos_ = kN;
  }

  uint32_t bits32() {
    uint32_t y;
    if (pos_ == kN) {
      int i = 0;
      for (; i < kN - kM; ++i) {
        y = (key_[i] & 0x80000000u) | (key_[i + 1] & 0x7fffffffu);
        key_[i] = key_[i + kM] ^ (y >> 1) ^ ((0u - (y & 1u)) & 0x9908b0dfu);
      }
      for (; i < kN - 1; ++i) {
        y = (key_[i] & 0x80000000u) | (key_[i + 1] & 0x7fffffffu);
        key_[i] = key_[i + (kM - kN)] ^ (y >> 1) ^ ((0u - (y & 1u)) & 0x9908b0dfu);
      }
      y = (key_[kN - 1] & 0x80000000u) | (key_[0] & 0x7fffffffu);
      key_[kN - 1] = key_[kM - 1] ^ (y >> 1) ^ ((0u - (y & 1u)) & 0x9908b0dfu);
      pos_ = 0;
    }
    y = key_[pos_++];
    y ^= y >> 11;
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= y >> 18;
    return y;
  }

  uint64_t bits64() {
    const uint64_t upper = static_cast<uint64_t>(bits32()) << 32;
    return upper | static_cast<uint64_t>(bits32());
  }

  // Byte stream: each 32-bit draw yields 4 bytes, least-significant first; leftovers are discarded per call.
  void read(uint8_t* p, size_t n) {
    int left = 0;
    uint32_t val = 0;
    for (size_t i = 0; i < n; ++i) {
      if (left == 0) { val = bits32(); left = 4; }
      p[i] = static_cast<uint8_t>(val);
      val >>= 8;
      --left;
    }
  }

  // Uniform integer in [0, limit] by masked rejection.
  uint64_t bits_limit(uint64_t limit) {
    if (limit == 0) return 0;
    uint64_t mask = limit;
    mask |= mask >> 1; mask |= mask >> 2; mask |= mask >> 4; mask |= mask >> 8; mask |= mask >> 16; mask |= mask >> 32;
    if (limit <= 0xffffffffull) {
      for (;;) {
        const uint64_t v = static_cast<uint64_t>(bits32()) & mask;
        if (v <= limit) return v;
      }
    }
    for (;;) {
      const uint64_t v = bits64() & mask;
      if (v <= limit) return v;
    }
  }

  int64_t int64_range(int64_t low, int64_t high) {
    if (high <= low) throw std::invalid_argument("nprand: high <= low");
    return low + static_cast<int64_t>(bits_limit(static_cast<uint64_t>(high) - static_cast<uint64_t>(low) - 1));
  }
  int64_t int64n(int64_t n) {
    if (n < 0) throw std::invalid_argument("nprand: n < 0");
    return static_cast<int64_t>(bits_limit(static_cast<uint64_t>(n) - 1));
  }
  int intn(int n) { return static_cast<int>(int64n(n)); }

  double unit_interval() {
    const double a = static_cast<double>(bits32() >> 5);
    const double b = static_cast<double>(bits32() >> 6);
    return (a * 67108864.0 + b) / 9007199254740992.0;
  }
  double uniform(double low, double high) {
    if (high <= low) throw std::invalid_argument("nprand: high <= low");
    return low + (high - low) * unit_interval();
  }

  // RFC 4122 version-4 UUID drawn from the byte stream (how trial request ids are minted).
  std::string uuid4() {
    uint8_t u[16];
    read(u, 16);
    u[6] = (u[6] & 0x0f) | 0x40;
    u[8] = (u[8] & 0x3f) | 0x80;
    static const char* hex = "0123456789abcdef";
    std::string s;
    for (int i = 0; i < 16; ++i) {
      if (i == 4 || i == 6 || i == 8 || i == 10) s.push_back('-');
      s.push_back(hex[u[i] >> 4]);
      s.push_back(hex[u[i] & 15]);
    }
    return s;
  }

 private:
  static constexpr int kN = 624, kM = 397;
  uint32_t key_[kN];
  int pos_ = kN;
};

}  // namespace det
