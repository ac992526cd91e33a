// oracle/ncd_hashes.cpp — CPU statements of the digests.  TEST INFRASTRUCTURE (see
// ncd_oracle.cpp for who may load it).
//
// The reference computes no hash anywhere (controller.go:539,600,725,795 are
// reflect.DeepEqual calls), so hash VALUES have no reference parity to pin: "parity
// unpinned" for NCD-H64, which is defined here and in DESIGN.md.  SHA-256 follows FIPS 180-4
// and XXH64 follows the xxHash specification; both are pinned by their published known
// answers in tests/test_oracle_hashes.py ("abc" -> ba7816bf...f20015ad;
// XXH64("", 0) = ef46db3751d8e999).
#include <stdint.h>
#include <string.h>

namespace {

inline uint64_t fmix64(uint64_t x) {
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdull;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ull;
  x ^= x >> 33;
  return x;
}
inline uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
inline uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }
inline uint32_t rotr32(uint32_t x, int r) { return (x >> r) | (x << (32 - r)); }
inline uint64_t rd64(const uint8_t* p) {
  uint64_t v = 0;
  for (int i = 7; i >= 0; --i) v = (v << 8) | p[i];
  return v;
}
inline uint32_t rd32(const uint8_t* p) { return p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); }

}  // namespace

extern "C" {

// ------------------------------------------------------------------ NCD-H64
// The record is read as 16-byte vectors (zero padded), grouped in tiles of 256 vectors.
// Vector v of a tile, words d0 d1 (little endian), keys k0 = sm(C + 2v), k1 = sm(C + 2v + 1):
//   A += lo32(d0 ^ k0) * hi32(d0 ^ k0) + d1 ;  B += lo32(d1 ^ k1) * hi32(d1 ^ k1) + d0
// tile t gives T = fmix(fmix(A + sm(2t)) + rotl(B, 31) + sm(2t + 1)); digest =
// fmix(sum T + len * 0x9E3779B185EBCA87).  All arithmetic mod 2^64.
uint64_t oracle_h64(const uint8_t* data, uint64_t len) {
  const uint64_t C = 0x6e63642d68363400ull;
  uint64_t sum = 0;
  const uint64_t nvec = (len + 15) / 16;
  for (uint64_t t = 0; t * 256 < nvec; ++t) {
    uint64_t A = 0, B = 0;
    for (uint64_t v = 0; v < 256 && t * 256 + v < nvec; ++v) {
      uint8_t buf[16] = {0};
      const uint64_t off = (t * 256 + v) * 16;
      const uint64_t nb = len - off < 16 ? len - off : 16;
      memcpy(buf, data + off, nb);
      const uint64_t d0 = rd64(buf), d1 = rd64(buf + 8);
      const uint64_t x0 = d0 ^ splitmix64(C + 2 * v), x1 = d1 ^ splitmix64(C + 2 * v + 1);
      A += (x0 & 0xffffffffull) * (x0 >> 32) + d1;
      B += (x1 & 0xffffffffull) * (x1 >> 32) + d0;
    }
    sum += fmix64(fmix64(A + splitmix64(2 * t)) + rotl64(B, 31) + splitmix64(2 * t + 1));
  }
  return fmix64(sum + len * 0x9E3779B185EBCA87ull);
}

// ------------------------------------------------------------------ SHA-256 (FIPS 180-4)
static const uint32_t K[64] = {
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5,
    0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174,
    0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da,
    0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967,
    0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
    0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070,
    0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3,
    0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};

static void sha_block(uint32_t h[8], const uint8_t* p) {
  uint32_t w[64];
  for (int i = 0; i < 16; ++i) w[i] = ((uint32_t)p[4 * i] << 24) | (p[4 * i + 1] << 16) | (p[4 * i + 2] << 8) | p[4 * i + 3];
  for (int i = 16; i < 64; ++i) {
    const uint32_t s0 = rotr32(w[i - 15], 7) ^ rotr32(w[i - 15], 18) ^ (w[i - 15] >> 3);
    const uint32_t s1 = rotr32(w[i - 2], 17) ^ rotr32(w[i - 2], 19) ^ (w[i - 2] >> 10);
    w[i] = w[i - 16] + s0 + w[i - 7] + s1;
  }
  uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
  for (int i = 0; i < 64; ++i) {
    const uint32_t t1 = hh + (rotr32(e, 6) ^ rotr32(e, 11) ^ rotr32(e, 25)) + ((e & f) ^ (~e & g)) + K[i] + w[i];
    const uint32_t t2 = (rotr32(a, 2) ^ rotr32(a, 13) ^ rotr32(a, 22)) + ((a & b) ^ (a & c) ^ (b & c));
    hh = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
  }
  h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
}

void oracle_sha256(const uint8_t* data, uint64_t len, uint8_t out[32]) {
  uint32_t h[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};
  uint64_t i = 0;
  for (; i + 64 <= len; i += 64) sha_block(h, data + i);
  uint8_t tail[128] = {0};
  const uint64_t rem = len - i;
  memcpy(tail, data + i, rem);
  tail[rem] = 0x80;
  const uint64_t tl = rem < 56 ? 64 : 128;
  const uint64_t bits = len * 8;
  for (int k = 0; k < 8; ++k) tail[tl - 1 - k] = (uint8_t)(bits >> (8 * k));
  sha_block(h, tail);
  if (tl == 128) sha_block(h, tail + 64);
  for (int k = 0; k < 8; ++k) {
    out[4 * k] = (uint8_t)(h[k] >> 24);
    out[4 * k + 1] = (uint8_t)(h[k] >> 16);
    out[4 * k + 2] = (uint8_t)(h[k] >> 8);
    out[4 * k + 3] = (uint8_t)h[k];
  }
}

// ------------------------------------------------------------------ XXH64 (xxHash spec)
uint64_t oracle_xxh64(const uint8_t* p, uint64_t len, uint64_t seed) {
  const uint64_t P1 = 11400714785074694791ull, P2 = 14029467366897019727ull, P3 = 1609587929392839161ull,
                 P4 = 9650029242287828579ull, P5 = 2870177450012600261ull;
  auto round = [&](uint64_t acc, uint64_t in) { return rotl64(acc + in * P2, 31) * P1; };
  auto merge = [&](uint64_t acc, uint64_t v) { return (acc ^ round(0, v)) * P1 + P4; };
  const uint8_t* end = p + len;
  uint64_t h;
  if (len >= 32) {
    uint64_t v1 = seed + P1 + P2, v2 = seed + P2, v3 = seed, v4 = seed - P1;
    do {
      v1 = round(v1, rd64(p));
      v2 = round(v2, rd64(p + 8));
      v3 = round(v3, rd64(p + 16));
      v4 = round(v4, rd64(p + 24));
      p += 32;
    } while (p + 32 <= end);
    h = rotl64(v1, 1) + rotl64(v2, 7) + rotl64(v3, 12) + rotl64(v4, 18);
    h = merge(h, v1);
    h = merge(h, v2);
    h = merge(h, v3);
    h = merge(h, v4);
  } else {
    h = seed + P5;
  }
  h += len;
  while (p + 8 <= end) {
    h ^= round(0, rd64(p));
    h = rotl64(h, 27) * P1 + P4;
    p += 8;
  }
  if (p + 4 <= end) {
    h ^= (uint64_t)rd32(p) * P1;
    h = rotl64(h, 23) * P2 + P3;
    p += 4;
  }
  while (p < end) {
    h ^= (*p) * P5;
    h = rotl64(h, 11) * P1;
    ++p;
  }
  h ^= h >> 33;
  h *= P2;
  h ^= h >> 29;
  h *= P3;
  h ^= h >> 32;
  return h;
}

// Digests of every record of an arena laid out as in include/ncdrift.h (payload, rec_off, rec_len).
void oracle_h64_arena(const uint8_t* payload, const uint64_t* rec_off, const uint32_t* rec_len, uint64_t n,
                      uint64_t* out) {
  for (uint64_t i = 0; i < n; ++i) out[i] = oracle_h64(payload + rec_off[i], rec_len[i]);
}
void oracle_sha256_arena(const uint8_t* payload, const uint64_t* rec_off, const uint32_t* rec_len, uint64_t n,
                         uint8_t* out) {
  for (uint64_t i = 0; i < n; ++i) oracle_sha256(payload + rec_off[i], rec_len[i], out + 32 * i);
}

}  // extern "C"
