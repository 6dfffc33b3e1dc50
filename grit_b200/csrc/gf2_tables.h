// CRC-32C (Castagnoli, reflected 0x82F63B78) as linear algebra over GF(2).
//
// adv(n, s): the CRC register s after n further zero bytes.  It is linear in s,
// so any "advance by n bytes" is a 32x32 bit matrix, stored here as four
// 256-entry byte tables: Op[k][b] = adv(n, b << 8k), adv(n, s) = XOR_k Op[k][byte k of s].
//
// The page kernels read a 4 KiB page as 8-lane groups of a warp: lane l of the
// group loads the 16 bytes at 128*j + 16*l for j = 0..31.  Word slot i of lane l
// therefore sees words 32j + 4l + i, 128 bytes apart, and runs the chain
//     c = w(j=0);  c = adv(128, c) ^ w(j)        (tables A, conflict-free copies)
// Four chains fold into one with adv(4, .) (tables T) and the lane's position
// in the page is applied once with adv(116 - 16l, .) (tables W[l]); the eight
// lane values XOR to the register of the whole page.  With init/xor-out
// 0xFFFFFFFF the CRC of a 4096-byte page is that register ^ crc32c(4096 zero bytes).
//
// Reference: the CRC definition is Go's hash/crc32 (Castagnoli), named by
// BASELINE.json north_star; /root/reference itself computes no CRC on this path.
#pragma once
#include <cstdint>

namespace grit {

struct Gf2Tables {
  uint32_t A[4][256];     // adv(128, .)
  uint32_t T[4][256];     // adv(4, .)
  uint32_t W[8][4][256];  // adv(116 - 16*l, .), l = 0..7
};

constexpr uint32_t kCrcZeroPage = 0x98F94189u;  // crc32c of 4096 zero bytes

inline uint32_t crc_byte_step(uint32_t s) {
  for (int k = 0; k < 8; k++) s = (s & 1) ? (s >> 1) ^ 0x82F63B78u : s >> 1;
  return s;
}

inline void build_gf2_tables(Gf2Tables* t) {
  uint32_t byte_tab[256];
  for (uint32_t b = 0; b < 256; b++) byte_tab[b] = crc_byte_step(b);
  auto adv = [&](int n, uint32_t s) {
    for (int i = 0; i < n; i++) s = byte_tab[s & 0xFF] ^ (s >> 8);
    return s;
  };
  for (int k = 0; k < 4; k++)
    for (uint32_t b = 0; b < 256; b++) {
      t->A[k][b] = adv(128, b << (8 * k));
      t->T[k][b] = adv(4, b << (8 * k));
      for (int l = 0; l < 8; l++) t->W[l][k][b] = adv(116 - 16 * l, b << (8 * k));
    }
}

}  // namespace grit
