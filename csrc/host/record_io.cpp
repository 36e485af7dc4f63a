// Host-side input runtime (plain C++17, no CUDA): TFRecord streaming with CRC32C validation on
// background threads, and a tf.Example field scanner, so the Python input pipeline only sees
// ready byte ranges. Counterpart of the pieces of TensorFlow's C++ input runtime the reference
// drives through RecordInput / tf.data (T/preprocessing.py:358-399, 601-609).
//
// Threading model: `num_threads` readers each own a disjoint subset of the files (file i goes
// to reader i % num_threads, like a parallel interleave with cycle length = num_threads). Two
// ways to merge their output:
//   * sloppy (kfb_rr_open): all readers push into one bounded queue, the consumer pops whatever
//     came first - the order depends on thread timing
//     (datasets_sloppy_parallel_interleave=true in the reference, T/preprocessing.py:516-519);
//   * deterministic (kfb_rr_open2, deterministic=1): every reader has its own bounded queue and
//     the consumer takes one record from each live reader in turn - tf.data's interleave with
//     block length 1 - so the record order is a function of the file list alone.
// With one reader both give the file order. A reader that hits a corrupt record stores the
// error and stops; the consumer sees it after draining what was queued before it.
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <deque>
#include <mutex>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include "../../include/kfb200_host.h"

#if defined(__SSE4_2__)
#include <nmmintrin.h>
#endif

namespace {

// ---- CRC32C (Castagnoli), hardware instruction when compiled for it, slicing table otherwise
uint32_t g_table[256];
std::once_flag g_table_once;

[[maybe_unused]] void init_table() {
  for (uint32_t i = 0; i < 256; ++i) {
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ 0x82F63B78u : c >> 1;
    g_table[i] = c;
  }
}

uint32_t crc32c(const uint8_t* p, size_t n) {
  uint32_t c = 0xFFFFFFFFu;
#if defined(__SSE4_2__)
  while (n >= 8) {
    uint64_t v;
    std::memcpy(&v, p, 8);
    c = static_cast<uint32_t>(_mm_crc32_u64(c, v));
    p += 8;
    n -= 8;
  }
  while (n--) c = _mm_crc32_u8(c, *p++);
#else
  std::call_once(g_table_once, init_table);
  while (n--) c = g_table[(c ^ *p++) & 0xFF] ^ (c >> 8);
#endif
  return c ^ 0xFFFFFFFFu;
}

inline uint32_t masked(uint32_t crc) { return ((crc >> 15) | (crc << 17)) + 0xA282EAD8u; }

struct Reader {
  std::vector<std::string> files;
  int num_threads = 1, capacity = 64;
  bool check_crc = true, repeat = false, shuffle = false;
  uint64_t seed = 0;

  std::mutex mu;
  std::condition_variable not_empty, not_full;
  std::deque<std::vector<uint8_t>> queue;               // sloppy mode
  bool deterministic = false;
  std::vector<std::deque<std::vector<uint8_t>>> queues;  // deterministic mode: one per reader
  std::vector<char> finished;                            // reader t has returned
  int cursor = 0;                                        // reader the consumer takes from next
  int per_reader_capacity = 1;
  int live_readers = 0;
  bool stop = false;
  std::string error;
  std::vector<std::thread> threads;
  std::vector<uint8_t> current;  // record handed to the consumer (valid until the next pop)

  void fail(const std::string& msg) {
    std::lock_guard<std::mutex> lk(mu);
    if (error.empty()) error = msg;
  }

  bool push(int tid, std::vector<uint8_t>&& rec) {
    std::unique_lock<std::mutex> lk(mu);
    if (deterministic) {
      auto& q = queues[tid];
      not_full.wait(lk, [&] { return stop || static_cast<int>(q.size()) < per_reader_capacity; });
      if (stop) return false;
      q.emplace_back(std::move(rec));
      not_empty.notify_all();  // the consumer may be waiting for exactly this reader
      return true;
    }
    not_full.wait(lk, [&] { return stop || static_cast<int>(queue.size()) < capacity; });
    if (stop) return false;
    queue.emplace_back(std::move(rec));
    not_empty.notify_one();
    return true;
  }

  // returns false on error / stop
  bool read_file(int tid, const std::string& path) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) {
      fail("cannot open " + path);
      return false;
    }
    bool ok = true;
    for (;;) {
      uint8_t head[12];
      size_t got = std::fread(head, 1, 12, f);
      if (got == 0) break;  // clean end of file
      if (got != 12) {
        fail("truncated record header in " + path);
        ok = false;
        break;
      }
      uint64_t len;
      uint32_t len_crc;
      std::memcpy(&len, head, 8);
      std::memcpy(&len_crc, head + 8, 4);
      if (check_crc && masked(crc32c(head, 8)) != len_crc) {
        fail("corrupt record length in " + path);
        ok = false;
        break;
      }
      if (len > (1ull << 32)) {
        fail("implausible record length in " + path);
        ok = false;
        break;
      }
      std::vector<uint8_t> rec(len);
      uint32_t data_crc;
      if (std::fread(rec.data(), 1, len, f) != len || std::fread(&data_crc, 1, 4, f) != 4) {
        fail("truncated record in " + path);
        ok = false;
        break;
      }
      if (check_crc && masked(crc32c(rec.data(), len)) != data_crc) {
        fail("corrupt record data in " + path);
        ok = false;
        break;
      }
      if (!push(tid, std::move(rec))) {
        ok = false;
        break;
      }
    }
    std::fclose(f);
    return ok;
  }

  void run(int tid) {
    std::vector<std::string> mine;
    for (size_t i = tid; i < files.size(); i += num_threads) mine.push_back(files[i]);
    std::mt19937_64 rng(seed * 1000003ull + tid);
    bool ok = true;
    do {
      if (shuffle) std::shuffle(mine.begin(), mine.end(), rng);
      for (const auto& p : mine) {
        ok = read_file(tid, p);
        if (!ok) break;
      }
    } while (ok && repeat && !mine.empty());
    std::lock_guard<std::mutex> lk(mu);
    --live_readers;
    if (deterministic) finished[tid] = 1;
    not_empty.notify_all();
  }
};

thread_local std::string g_last_error;

}  // namespace

extern "C" {

const char* kfb_host_last_error() { return g_last_error.c_str(); }

uint32_t kfb_crc32c(const void* data, uint64_t len) {
  return crc32c(static_cast<const uint8_t*>(data), len);
}

uint32_t kfb_masked_crc32c(const void* data, uint64_t len) {
  return masked(crc32c(static_cast<const uint8_t*>(data), len));
}

void* kfb_rr_open2(const char* const* paths, int npaths, int num_threads, int capacity,
                   int check_crc, int repeat, int shuffle_files, uint64_t seed,
                   int deterministic) {
  if (npaths <= 0 || paths == nullptr) {
    g_last_error = "record reader: no files";
    return nullptr;
  }
  auto* r = new Reader();
  for (int i = 0; i < npaths; ++i) r->files.emplace_back(paths[i]);
  r->num_threads = num_threads < 1 ? 1 : (num_threads > npaths ? npaths : num_threads);
  r->capacity = capacity < 1 ? 1 : capacity;
  r->check_crc = check_crc != 0;
  r->repeat = repeat != 0;
  r->shuffle = shuffle_files != 0;
  r->seed = seed;
  r->live_readers = r->num_threads;
  r->deterministic = deterministic != 0;
  if (r->deterministic) {
    r->queues.resize(r->num_threads);
    r->finished.assign(r->num_threads, 0);
    r->per_reader_capacity = std::max(1, r->capacity / r->num_threads);
  }
  for (int t = 0; t < r->num_threads; ++t) r->threads.emplace_back([r, t] { r->run(t); });
  return r;
}

void* kfb_rr_open(const char* const* paths, int npaths, int num_threads, int capacity,
                  int check_crc, int repeat, int shuffle_files, uint64_t seed) {
  return kfb_rr_open2(paths, npaths, num_threads, capacity, check_crc, repeat, shuffle_files,
                      seed, 0);
}

// 1: *data / *len describe the next record (valid until the next call on this handle);
// 0: end of data; -1: error (kfb_host_last_error()).
int kfb_rr_next(void* handle, const uint8_t** data, uint64_t* len) {
  auto* r = static_cast<Reader*>(handle);
  std::unique_lock<std::mutex> lk(r->mu);
  if (r->deterministic) {
    const int n = r->num_threads;
    for (int tried = 0; tried < n;) {
      const int t = r->cursor;
      r->not_empty.wait(lk, [&] { return !r->queues[t].empty() || r->finished[t]; });
      if (!r->queues[t].empty()) {
        r->current = std::move(r->queues[t].front());
        r->queues[t].pop_front();
        r->cursor = (t + 1) % n;
        r->not_full.notify_all();
        *data = r->current.data();
        *len = r->current.size();
        return 1;
      }
      // reader t is exhausted (or failed): leave it out of the rotation from now on
      r->cursor = (t + 1) % n;
      ++tried;
      if (!r->error.empty()) break;
    }
    if (!r->error.empty()) {
      g_last_error = r->error;
      return -1;
    }
    return 0;
  }
  r->not_empty.wait(lk, [&] { return !r->queue.empty() || r->live_readers == 0; });
  if (r->queue.empty()) {
    if (!r->error.empty()) {
      g_last_error = r->error;
      return -1;
    }
    return 0;
  }
  r->current = std::move(r->queue.front());
  r->queue.pop_front();
  r->not_full.notify_one();
  *data = r->current.data();
  *len = r->current.size();
  return 1;
}

void kfb_rr_close(void* handle) {
  auto* r = static_cast<Reader*>(handle);
  {
    std::lock_guard<std::mutex> lk(r->mu);
    r->stop = true;
  }
  r->not_full.notify_all();
  for (auto& t : r->threads) t.join();
  delete r;
}

}  // extern "C"
