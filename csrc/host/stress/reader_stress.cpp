// Sanitizer stress harness for the host input runtime (record_io.cpp / example_scan.cpp).
// Built by tests/test_native_reader.py with -fsanitize=thread and -fsanitize=address,undefined
// and run as a plain executable: the reader threads, the bounded queue and the early-close
// path are driven hard enough that a missing lock or a use-after-free shows up as a sanitizer
// report (SURVEY.md 5.2 asks for a TSAN build of the C++ runtime; the reference has none).
//
// usage: reader_stress <scratch dir> [iterations]
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../../include/kfb200_host.h"

namespace {

uint64_t g_rng = 0x9E3779B97F4A7C15ull;
uint32_t rnd() {
  g_rng ^= g_rng << 13;
  g_rng ^= g_rng >> 7;
  g_rng ^= g_rng << 17;
  return static_cast<uint32_t>(g_rng >> 16);
}

// one TFRecord frame: u64 length, masked crc of the length, payload, masked crc of the payload
void write_record(FILE* f, const std::vector<uint8_t>& rec) {
  uint64_t len = rec.size();
  uint32_t c = kfb_masked_crc32c(&len, 8);
  std::fwrite(&len, 8, 1, f);
  std::fwrite(&c, 4, 1, f);
  if (len) std::fwrite(rec.data(), 1, len, f);
  c = kfb_masked_crc32c(rec.data(), len);
  std::fwrite(&c, 4, 1, f);
}

#define REQUIRE(cond, ...)                    \
  do {                                        \
    if (!(cond)) {                            \
      std::fprintf(stderr, "FAIL: " __VA_ARGS__); \
      std::fprintf(stderr, "\n");             \
      std::exit(1);                           \
    }                                         \
  } while (0)

}  // namespace

int main(int argc, char** argv) {
  REQUIRE(argc >= 2, "usage: reader_stress <dir> [iterations]");
  const std::string dir = argv[1];
  const int iters = argc > 2 ? std::atoi(argv[2]) : 20;
  const int nfiles = 6, per_file = 50;
  std::vector<std::string> paths;
  uint64_t total_bytes = 0;
  for (int i = 0; i < nfiles; ++i) {
    paths.push_back(dir + "/stress-" + std::to_string(i));
    FILE* f = std::fopen(paths.back().c_str(), "wb");
    REQUIRE(f != nullptr, "cannot create %s", paths.back().c_str());
    for (int r = 0; r < per_file; ++r) {
      std::vector<uint8_t> rec(rnd() % 2000);
      for (auto& b : rec) b = static_cast<uint8_t>(rnd());
      total_bytes += rec.size();
      write_record(f, rec);
    }
    std::fclose(f);
  }
  std::vector<const char*> cpaths;
  for (auto& p : paths) cpaths.push_back(p.c_str());

  for (int it = 0; it < iters; ++it) {
    const int threads = 1 + it % 4, capacity = 1 + (it * 7) % 9;
    // (a) full pass: every record arrives exactly once, CRC-checked
    void* h = kfb_rr_open(cpaths.data(), nfiles, threads, capacity, 1, 0, it & 1, it);
    REQUIRE(h != nullptr, "open: %s", kfb_host_last_error());
    const uint8_t* data;
    uint64_t len, count = 0, bytes = 0;
    int rc;
    while ((rc = kfb_rr_next(h, &data, &len)) == 1) {
      ++count;
      bytes += len;
      if (len) bytes += data[len - 1] & 0;  // touch the buffer the producer thread filled
    }
    REQUIRE(rc == 0, "next: %s", kfb_host_last_error());
    REQUIRE(count == static_cast<uint64_t>(nfiles * per_file) && bytes == total_bytes,
            "pass %d: %llu records / %llu bytes", it, (unsigned long long)count,
            (unsigned long long)bytes);
    kfb_rr_close(h);
    // (a2) deterministic interleave: same count, and the same order on a second pass
    {
      std::vector<uint64_t> order[2];
      for (int pass = 0; pass < 2; ++pass) {
        void* hd = kfb_rr_open2(cpaths.data(), nfiles, threads, capacity, 1, 0, 0, it, 1);
        REQUIRE(hd != nullptr, "open2: %s", kfb_host_last_error());
        while ((rc = kfb_rr_next(hd, &data, &len)) == 1) order[pass].push_back(len);
        REQUIRE(rc == 0, "next (deterministic): %s", kfb_host_last_error());
        kfb_rr_close(hd);
      }
      REQUIRE(order[0].size() == static_cast<size_t>(nfiles * per_file) && order[0] == order[1],
              "deterministic interleave differs between passes");
      // endless, closed early
      void* hd = kfb_rr_open2(cpaths.data(), nfiles, threads, capacity, 1, 1, 0, it, 1);
      const int take2 = static_cast<int>(rnd() % 300);
      for (int k = 0; k < take2; ++k) REQUIRE(kfb_rr_next(hd, &data, &len) == 1, "repeat next (det)");
      kfb_rr_close(hd);
    }
    // (b) endless repeat, closed while the readers are blocked on a full queue or mid-file
    h = kfb_rr_open(cpaths.data(), nfiles, threads, capacity, 1, 1, 1, it + 100);
    REQUIRE(h != nullptr, "open: %s", kfb_host_last_error());
    const int take = static_cast<int>(rnd() % 400);
    for (int k = 0; k < take; ++k) REQUIRE(kfb_rr_next(h, &data, &len) == 1, "repeat next");
    kfb_rr_close(h);
    // (c) two independent readers consumed from two threads at once
    void* h1 = kfb_rr_open(cpaths.data(), nfiles, 2, 4, 1, 0, 0, 1);
    void* h2 = kfb_rr_open(cpaths.data(), nfiles, 3, 2, 0, 0, 1, 2);
    REQUIRE(h1 && h2, "open pair");
    uint64_t n1 = 0, n2 = 0;
    std::thread t1([&] { const uint8_t* d; uint64_t l; while (kfb_rr_next(h1, &d, &l) == 1) ++n1; });
    std::thread t2([&] { const uint8_t* d; uint64_t l; while (kfb_rr_next(h2, &d, &l) == 1) ++n2; });
    t1.join();
    t2.join();
    REQUIRE(n1 == n2 && n1 == static_cast<uint64_t>(nfiles * per_file), "pair counts");
    kfb_rr_close(h1);
    kfb_rr_close(h2);
  }
  // (d) a corrupt file: the error is reported after the intact records before it
  {
    FILE* f = std::fopen(paths[0].c_str(), "r+b");
    REQUIRE(f != nullptr, "reopen");
    std::fseek(f, -6, SEEK_END);
    std::fputc(0x5A, f);
    std::fputc(0xA5, f);
    std::fclose(f);
    void* h = kfb_rr_open(cpaths.data(), nfiles, 3, 4, 1, 0, 0, 0);
    const uint8_t* data;
    uint64_t len;
    int rc;
    while ((rc = kfb_rr_next(h, &data, &len)) == 1) {
    }
    REQUIRE(rc == -1 && std::strstr(kfb_host_last_error(), "corrupt") != nullptr,
            "corrupt file not reported (rc=%d)", rc);
    kfb_rr_close(h);
  }
  // (e) the tf.Example scanner on hostile input: random bytes, and a well-formed message cut at
  // every length - it may find nothing, but must never read outside [data, data + len) (the
  // buffer is heap-allocated at its exact size so ASAN sees an overrun)
  {
    // Example{features{feature{"k": bytes_list{"abc"}}, feature{"n": int64_list{1, 300}}}}
    const uint8_t good[] = {0x0A, 0x1C, 0x0A, 0x0C, 0x0A, 0x01, 'k',  0x12, 0x07, 0x0A,
                            0x05, 0x0A, 0x03, 'a',  'b',  'c',  0x0A, 0x0C, 0x0A, 0x01,
                            'n',  0x12, 0x07, 0x1A, 0x05, 0x0A, 0x03, 0x01, 0xAC, 0x02};
    const size_t good_len = sizeof(good);
    const uint8_t* out;
    uint64_t out_len;
    double nums[4];
    REQUIRE(kfb_example_get_bytes(good, good_len, "k", &out, &out_len) == 1 && out_len == 3 &&
                std::memcmp(out, "abc", 3) == 0, "scanner: bytes feature");
    REQUIRE(kfb_example_get_numbers(good, good_len, "n", nums, 4) == 2 && nums[0] == 1 &&
                nums[1] == 300, "scanner: int64 feature");
    REQUIRE(kfb_example_get_numbers(good, good_len, "missing", nums, 4) == -1, "scanner: missing");
    for (size_t cut = 0; cut <= good_len; ++cut) {
      std::vector<uint8_t> buf(good, good + cut);
      kfb_example_get_bytes(buf.data(), cut, "k", &out, &out_len);
      kfb_example_get_numbers(buf.data(), cut, "n", nums, 4);
    }
    for (int round = 0; round < 20000; ++round) {
      std::vector<uint8_t> buf(rnd() % 64);
      for (auto& b : buf) b = static_cast<uint8_t>(rnd());
      if (round & 1) {  // keep a valid prefix so the parser gets deep before the garbage
        const size_t keep = std::min(buf.size(), static_cast<size_t>(rnd() % good_len));
        if (keep) std::memcpy(buf.data(), good, keep);
      }
      kfb_example_get_bytes(buf.data(), buf.size(), "k", &out, &out_len);
      kfb_example_get_numbers(buf.data(), buf.size(), "n", nums, 2);
    }
  }
  // (f) image transform: random crops / resizes / windows of every method from several threads
  // at once (it keeps thread-local scratch); source and destination are heap blocks of exactly
  // the size the geometry implies, so an out-of-range tap is an ASAN report
  {
    auto worker = [](uint64_t seed) {
      uint64_t st = seed * 0x9E3779B97F4A7C15ull + 1;
      auto r = [&st](uint32_t n) {
        st ^= st << 13;
        st ^= st >> 7;
        st ^= st << 17;
        return static_cast<int>((st >> 20) % n);
      };
      for (int it = 0; it < 300; ++it) {
        kfb_image_op op{};
        op.ih = 1 + r(60);
        op.iw = 1 + r(60);
        op.src_stride = op.iw * 3 + r(8);
        op.ch = 1 + r(op.ih);
        op.cw = 1 + r(op.iw);
        op.top = r(op.ih - op.ch + 1);
        op.left = r(op.iw - op.cw + 1);
        op.vh = 1 + r(70);
        op.vw = 1 + r(70);
        op.oh = 1 + r(op.vh);
        op.ow = 1 + r(op.vw);
        op.oy = r(op.vh - op.oh + 1);
        op.ox = r(op.vw - op.ow + 1);
        op.method = it % 5;
        if (op.method == 4) {  // crop_and_resize window, partly outside the image now and then
          op.box[0] = -0.1f + 0.01f * r(60);
          op.box[1] = -0.1f + 0.01f * r(60);
          op.box[2] = op.box[0] + 0.01f * (1 + r(70));
          op.box[3] = op.box[1] + 0.01f * (1 + r(70));
          op.per_channel = 1;
          for (int k = 0; k < 3; ++k) {
            op.scale3[k] = 0.01f;
            op.bias3[k] = -1.f;
          }
          op.no_clamp = r(2);
        }
        op.flip = r(2);
        op.distort_order = r(3);
        op.brightness = 5.f;
        op.saturation = 1.2f;
        op.hue = 0.15f;
        op.contrast = 0.8f;
        op.scale = 1.f / 127.5f;
        op.bias = -1.f;
        op.out_f16 = r(2);
        // the last row needs only iw*3 bytes: allocate exactly that
        std::vector<uint8_t> src(static_cast<size_t>(op.ih - 1) * op.src_stride + op.iw * 3);
        for (auto& b : src) b = static_cast<uint8_t>(r(256));
        std::vector<uint8_t> out(static_cast<size_t>(op.oh) * op.ow * 3 * (op.out_f16 ? 2 : 4));
        REQUIRE(kfb_image_transform(src.data(), &op, out.data()) == 0, "image transform failed");
      }
      kfb_image_op bad{};
      uint8_t px[3] = {0, 0, 0};
      float o[3];
      REQUIRE(kfb_image_transform(px, &bad, o) == -1, "bad geometry accepted");
    };
    std::vector<std::thread> ts;
    for (int t = 0; t < 4; ++t) ts.emplace_back(worker, static_cast<uint64_t>(t + 1));
    for (auto& t : ts) t.join();
  }
  std::printf("reader_stress ok (%d iterations)\n", iters);
  return 0;
}
