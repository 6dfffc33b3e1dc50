// Host-side helpers of the C ABI that touch no GPU: CRC-32C of small metadata
// blobs and the completion sentinel.
#include <fcntl.h>
#include <unistd.h>

#include <cstring>
#include <mutex>
#include <string>

#include "../../include/gritpack.h"

namespace {

// slice-by-8 tables for the Castagnoli polynomial (reflected 0x82F63B78)
uint32_t g_tab[8][256];
std::once_flag g_once;

void init_tab() {
  for (uint32_t i = 0; i < 256; i++) {
    uint32_t c = i;
    for (int k = 0; k < 8; k++) c = (c & 1) ? (c >> 1) ^ 0x82F63B78u : c >> 1;
    g_tab[0][i] = c;
  }
  for (uint32_t i = 0; i < 256; i++)
    for (int t = 1; t < 8; t++) g_tab[t][i] = g_tab[0][g_tab[t - 1][i] & 0xFF] ^ (g_tab[t - 1][i] >> 8);
}

}  // namespace

extern "C" {

// Go: crc32.Update(crc, castagnoliTable, data).  The reference has no call site;
// north_star names the polynomial.  Used for manifest checksums only.
uint32_t grit_crc32c(uint32_t crc, const void* data, size_t len) {
  std::call_once(g_once, init_tab);
  const unsigned char* p = static_cast<const unsigned char*>(data);
  uint32_t c = ~crc;
  while (len >= 8) {
    uint32_t lo, hi;
    memcpy(&lo, p, 4);
    memcpy(&hi, p + 4, 4);
    lo ^= c;
    c = g_tab[7][lo & 0xFF] ^ g_tab[6][(lo >> 8) & 0xFF] ^ g_tab[5][(lo >> 16) & 0xFF] ^ g_tab[4][lo >> 24] ^
        g_tab[3][hi & 0xFF] ^ g_tab[2][(hi >> 8) & 0xFF] ^ g_tab[1][(hi >> 16) & 0xFF] ^ g_tab[0][hi >> 24];
    p += 8;
    len -= 8;
  }
  while (len--) c = g_tab[0][(c ^ *p++) & 0xFF] ^ (c >> 8);
  return ~c;
}

// CreateSentinelFile, pkg/gritagent/copy/copy.go:92-102: os.Create + WriteString.
int grit_write_sentinel(const char* dir, const char* name) {
  if (!dir || !name) return GRIT_EINVAL;
  const std::string path = std::string(dir) + "/" + name;
  const int fd = open(path.c_str(), O_WRONLY | O_CREAT | O_TRUNC | O_CLOEXEC, 0666);
  if (fd < 0) return GRIT_EIO;
  static const char payload[] = "Transfer Completed\n";
  const ssize_t n = write(fd, payload, sizeof payload - 1);
  const int rc = close(fd);
  return (n == (ssize_t)(sizeof payload - 1) && rc == 0) ? GRIT_OK : GRIT_EIO;
}

}  // extern "C"
