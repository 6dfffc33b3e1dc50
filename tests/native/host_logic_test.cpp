// Host logic of the directory packer, exercised without a GPU: the I/O thread pool, the manifest codec, the cutting
// of the page stream into file pieces, the striped page store.  These live in an anonymous namespace of
// grit_b200/csrc/dirpack.cpp, so this file includes that translation unit; tests/test_host_logic.py builds it with
// nvcc (the headers pull in cuda_runtime.h; nothing here calls CUDA) and runs it in the "not gpu" suite.
#include "../../grit_b200/csrc/dirpack.cpp"

#include <cstdlib>
#include <random>

namespace {

int g_failed = 0;
#define CHECK(cond)                                                      \
  do {                                                                   \
    if (!(cond)) {                                                       \
      fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond);  \
      g_failed++;                                                        \
    }                                                                    \
  } while (0)

void test_pool() {
  for (unsigned threads : {1u, 2u, 5u, 16u}) {
    IoPool pool(threads);
    std::atomic<long> total{0};
    auto hammer = [&](unsigned seed) {
      std::mt19937 rng(seed);
      for (int round = 0; round < 300; round++) {
        const size_t n = rng() % 40;
        std::vector<std::atomic<int>> hit(n);
        for (auto& h : hit) h = 0;
        pool.for_each(n, [&](size_t i) {
          hit[i]++;
          total++;
        });
        for (size_t i = 0; i < n; i++) CHECK(hit[i] == 1);  // every index once, all done when for_each returns
      }
    };
    std::thread a(hammer, 1u), b(hammer, 2u);  // two callers at once, as the reader and the writer side are
    hammer(3u);
    a.join();
    b.join();
    CHECK(total > 0);
  }
}

Manifest sample_manifest() {
  Manifest m;
  m.entries.push_back({0, 0755, 0, 0, "main"});
  m.entries.push_back({0, 0700, 0, 0, "main/checkpoint"});
  m.entries.push_back({1, 0444, 0, 0, "main/checkpoint/empty.img"});
  m.entries.push_back({1, 0600, 77, 0, "main/checkpoint/inventory.img"});
  m.entries.push_back({1, 0600, 3 * 4096, 1, "main/checkpoint/pages-1.img"});
  m.entries.push_back({1, 0644, 4097, 4, "main/rootfs-diff.tar"});
  m.n_pages = 6;
  m.n_slots = 5;
  m.bytes_logical = 77 + 3 * 4096 + 4097;
  m.pagemap_crc = 0x12345678;
  m.hashes_crc = 0x9abcdef0;
  m.n_shards = 8;
  m.stripe_slots = 256;
  m.parent = "../ckpt-0";
  m.peers = {"/pv/ns/ckpt-rank0", "/pv/ns/ckpt-rank1"};
  return m;
}

void test_manifest(const std::string& tmp) {
  const Manifest m = sample_manifest();
  const std::string enc = encode_manifest(m);
  std::string msg;
  CHECK(write_file(tmp + "/" + kManifest, enc.data(), enc.size(), &msg) == GRIT_OK);
  CHECK(has_manifest(tmp));
  Manifest back;
  CHECK(load_manifest(tmp, &back, &msg) == GRIT_OK);
  CHECK(back.n_pages == m.n_pages && back.n_slots == m.n_slots && back.bytes_logical == m.bytes_logical);
  CHECK(back.pagemap_crc == m.pagemap_crc && back.hashes_crc == m.hashes_crc);
  CHECK(back.n_shards == m.n_shards && back.stripe_slots == m.stripe_slots);
  CHECK(back.parent == m.parent && back.peers == m.peers && back.entries.size() == m.entries.size());
  for (size_t i = 0; i < m.entries.size() && i < back.entries.size(); i++) {
    const Entry &x = m.entries[i], &y = back.entries[i];
    CHECK(x.type == y.type && x.mode == y.mode && x.size == y.size && x.path == y.path);
    if (x.type == 1) CHECK(x.first_page == y.first_page);
  }
  CHECK(encode_manifest(back) == enc);  // the codec is its own inverse
  // every single-byte change is caught by the checksum
  for (size_t off = 0; off < enc.size(); off++) {
    std::string bad = enc;
    bad[off] ^= 0x20;
    CHECK(write_file(tmp + "/" + kManifest, bad.data(), bad.size(), &msg) == GRIT_OK);
    Manifest junk;
    CHECK(load_manifest(tmp, &junk, &msg) == GRIT_EFORMAT);
  }
  // entries that do not tile the image are refused even when the checksum is right
  Manifest gap = m;
  gap.entries[4].first_page = 2;
  const std::string genc = encode_manifest(gap);
  CHECK(write_file(tmp + "/" + kManifest, genc.data(), genc.size(), &msg) == GRIT_OK);
  Manifest junk;
  CHECK(load_manifest(tmp, &junk, &msg) == GRIT_EFORMAT);
  unlink((tmp + "/" + kManifest).c_str());
  CHECK(!has_manifest(tmp));
}

void test_paths() {
  for (const char* ok : {"a", "a/b", "main/checkpoint/pages-1.img", "..a", "a..", "a/..b/c", "...", "a/.hidden"})
    CHECK(safe_relative(ok));
  for (const char* bad : {"", "/", "/etc/passwd", "..", "../x", "a/../b", "a/..", ".", "./a", "a/./b", "a//b", "a/"})
    CHECK(!safe_relative(bad));
}

// the page stream a list of files stands for, and the same stream rebuilt from pieces + zero_tails
void test_pieces() {
  std::mt19937 rng(7);
  const std::vector<uint64_t> sizes = {0, 77, 4096, 4097, 1, 3 * 4096 + 5, 0, 9 * 1048576 + 123, 8192, 4095};
  std::vector<Entry> entries;
  std::vector<std::string> contents;
  uint64_t pages = 0;
  for (size_t i = 0; i < sizes.size(); i++) {
    entries.push_back({1, 0644, sizes[i], pages, "f" + std::to_string(i)});
    pages += entries.back().pages();
    std::string c(sizes[i], '\0');
    for (auto& ch : c) ch = (char)(1 + rng() % 255);  // never zero, so missing bytes show
    contents.push_back(c);
  }
  std::vector<const Entry*> files;
  for (const auto& e : entries)
    if (e.size) files.push_back(&e);
  std::string stream(pages * kPage, '\0');
  for (size_t i = 0; i < entries.size(); i++) memcpy(&stream[entries[i].first_page * kPage], contents[i].data(), sizes[i]);

  for (uint64_t chunk : std::vector<uint64_t>{1, 2, 3, 256, 1000, pages})
    for (uint64_t max_piece : std::vector<uint64_t>{4096, 1ull << 20, 8ull << 20}) {
      std::string rebuilt(pages * kPage, '\x55');
      for (uint64_t p0 = 0; p0 < pages; p0 += chunk) {
        const uint64_t p1 = std::min(pages, p0 + chunk);
        std::vector<Piece> ps;
        char* buf = &rebuilt[p0 * kPage];
        pieces_for(files, p0, p1, buf, max_piece, &ps);
        zero_tails(files, p0, p1, buf);
        for (const auto& p : ps) {
          CHECK(p.len > 0 && p.len <= max_piece && p.file_off + p.len <= p.e->size);
          CHECK(p.buf >= buf && p.buf + p.len <= buf + (p1 - p0) * kPage);
          memcpy(p.buf, contents[p.e - entries.data()].data() + p.file_off, p.len);  // what read_piece would do
        }
      }
      CHECK(rebuilt == stream);
    }
  CHECK(piece_bytes_for(1024, 8) == (1u << 20) && piece_bytes_for(4096, 8) == (2u << 20));
  CHECK(piece_bytes_for(16384, 1) == (8u << 20) && piece_bytes_for(1, 64) == (1u << 20));
  Opts o;
  CHECK(o.chunk_for(1) * kPage == (4u << 20) && o.chunk_for(262144) * kPage == (8u << 20));
  CHECK(o.chunk_for(3921026) * kPage == (16u << 20) && o.chunk_for(1ull << 26) * kPage == (16u << 20));
}

void test_page_store(const std::string& tmp) {
  std::mt19937 rng(11);
  IoPool pool(6);
  for (uint32_t stripe : {256u, 700u, 2048u}) {
    const uint64_t n_slots = 5 * 8 * stripe / 2 + 123;  // several rounds over the shard files and a ragged tail
    std::string data(n_slots * kPage, '\0');
    for (size_t i = 0; i < data.size(); i += 509) data[i] = (char)(rng() | 1);
    std::string msg;
    {
      PageStore w;
      w.dir = tmp;
      w.stripe = stripe;
      CHECK(w.open_all(true, &msg) == GRIT_OK);
      for (uint64_t s = 0; s < n_slots;) {  // written in ragged chunks, as compaction hands them out
        const uint64_t n = std::min<uint64_t>(n_slots - s, 1 + rng() % 3000);
        CHECK(w.io(s, n, &data[s * kPage], &pool, true) == 0);
        s += n;
      }
      CHECK(w.sync_all() == 0);
    }
    PageStore r;
    r.dir = tmp;
    r.stripe = stripe;
    CHECK(r.open_all(false, &msg) == GRIT_OK);
    CHECK(r.check_sizes(n_slots, &msg) == GRIT_OK);
    CHECK(r.check_sizes(n_slots + 1, &msg) == GRIT_EFORMAT);
    uint64_t sum = 0;
    for (uint32_t i = 0; i < r.k; i++) sum += r.shard_bytes(i, n_slots);
    CHECK(sum == n_slots * kPage);
    std::string back(data.size(), '\x33');
    for (uint64_t s = 0; s < n_slots;) {
      const uint64_t n = std::min<uint64_t>(n_slots - s, 1 + rng() % 5000);
      CHECK(r.io(s, n, &back[s * kPage], &pool, false) == 0);
      s += n;
    }
    CHECK(back == data);
    // the layout DESIGN.md states: slot s is in stripe t = s / stripe, stripe number t / k of file t % k
    const uint64_t s = 1 * stripe * 8 + 5 * stripe + 17, t = s / stripe;  // second round over the files, sixth file
    std::string page(kPage, '\0');
    CHECK(pread(r.fds[t % r.k], &page[0], kPage, (off_t)(((t / r.k) * stripe + s % stripe) * kPage)) == (ssize_t)kPage);
    CHECK(memcmp(page.data(), &data[s * kPage], kPage) == 0);
  }
}

}  // namespace

int main(int argc, char** argv) {
  if (argc < 2) return fprintf(stderr, "usage: %s <scratch dir>\n", argv[0]), 2;
  const std::string tmp = argv[1];
  test_pool();
  test_manifest(tmp);
  test_paths();
  test_pieces();
  test_page_store(tmp);
  if (g_failed) return fprintf(stderr, "%d check(s) failed\n", g_failed), 1;
  printf("host logic ok\n");
  return 0;
}
