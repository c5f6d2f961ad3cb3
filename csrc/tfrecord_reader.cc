// Native input pipeline: TFRecord framing + masked CRC-32C + a minimal tf.train.Example walk +
// a multi-threaded random shuffle queue with TensorFlow's capacity / min_after_dequeue semantics.
//
// Restates, without TensorFlow, what the reference builds from library ops (image_input.py:98-143):
//   string_input_producer(filenames)            -> shuffled, endless file list shared by the readers
//   TFRecordReader + parse_single_example       -> record framing, feature "image_raw" (bytes)
//   decode_raw(float64) + reshape + cast(f32)   -> 64x64x3 doubles (98,304 B) to float32
//   shuffle_batch(num_threads=16, capacity=min_after_dequeue+3*batch, min_after_dequeue=10776)
// Exposed through a C ABI (bound with ctypes in data/native_reader.py).
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <random>
#include <string>
#include <thread>
#include <vector>

namespace {

// ---- CRC-32C -----------------------------------------------------------------------------------------
uint32_t g_table[8][256];
std::once_flag g_table_once;
void init_table() {
  for (uint32_t i = 0; i < 256; ++i) {
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c >> 1) ^ ((c & 1) ? 0x82F63B78u : 0u);
    g_table[0][i] = c;
  }
  for (int t = 1; t < 8; ++t)
    for (uint32_t i = 0; i < 256; ++i) g_table[t][i] = (g_table[t - 1][i] >> 8) ^ g_table[0][g_table[t - 1][i] & 0xFF];
}
uint32_t crc32c(const uint8_t* p, size_t n, uint32_t crc) {
  std::call_once(g_table_once, init_table);
  crc ^= 0xFFFFFFFFu;
#if defined(__SSE4_2__)
  while (n >= 8) { uint64_t v; memcpy(&v, p, 8); crc = (uint32_t)__builtin_ia32_crc32di(crc, v); p += 8; n -= 8; }
  while (n--) crc = __builtin_ia32_crc32qi(crc, *p++);
#else
  while (n >= 8) {
    uint32_t lo, hi; memcpy(&lo, p, 4); memcpy(&hi, p + 4, 4); lo ^= crc;
    crc = g_table[7][lo & 0xFF] ^ g_table[6][(lo >> 8) & 0xFF] ^ g_table[5][(lo >> 16) & 0xFF] ^ g_table[4][lo >> 24] ^
          g_table[3][hi & 0xFF] ^ g_table[2][(hi >> 8) & 0xFF] ^ g_table[1][(hi >> 16) & 0xFF] ^ g_table[0][hi >> 24];
    p += 8; n -= 8;
  }
  while (n--) crc = (crc >> 8) ^ g_table[0][(crc ^ *p++) & 0xFF];
#endif
  return crc ^ 0xFFFFFFFFu;
}
uint32_t masked(uint32_t c) { return ((c >> 15) | (c << 17)) + 0xA282EAD8u; }

// ---- protobuf walk -------------------------------------------------------------------------------------
bool read_varint(const uint8_t*& p, const uint8_t* end, uint64_t* out) {
  uint64_t r = 0; int shift = 0;
  while (p < end) {
    uint8_t b = *p++;
    r |= (uint64_t)(b & 0x7F) << shift;
    if (!(b & 0x80)) { *out = r; return true; }
    shift += 7;
    if (shift > 63) return false;
  }
  return false;
}
struct Span { const uint8_t* p; size_t n; };
// finds the first length-delimited field `field` at this level; other fields are skipped
template <class F>
bool for_each_ld(Span s, uint32_t field, F&& fn) {
  const uint8_t* p = s.p; const uint8_t* end = s.p + s.n;
  while (p < end) {
    uint64_t key; if (!read_varint(p, end, &key)) return false;
    uint32_t f = (uint32_t)(key >> 3), w = (uint32_t)(key & 7);
    if (w == 0) { uint64_t v; if (!read_varint(p, end, &v)) return false; }
    else if (w == 1) { p += 8; }
    else if (w == 5) { p += 4; }
    else if (w == 2) {
      uint64_t len; if (!read_varint(p, end, &len)) return false;
      if (len > (uint64_t)(end - p)) return false;
      if (f == field && fn(Span{p, (size_t)len})) return true;
      p += len;
    } else return false;
  }
  return false;
}
// Example.features(1).feature(1: map entry{key=1,value=2: Feature{bytes_list=1{value=1}}})
bool find_bytes_feature(Span example, const std::string& name, Span* out) {
  return for_each_ld(example, 1, [&](Span features) {
    return for_each_ld(features, 1, [&](Span entry) {
      Span key{nullptr, 0}, val{nullptr, 0};
      for_each_ld(entry, 1, [&](Span k) { key = k; return true; });
      for_each_ld(entry, 2, [&](Span v) { val = v; return true; });
      if (!key.p || key.n != name.size() || memcmp(key.p, name.data(), key.n) != 0 || !val.p) return false;
      return for_each_ld(val, 1, [&](Span bytes_list) {
        return for_each_ld(bytes_list, 1, [&](Span v) { *out = v; return true; });
      });
    });
  });
}

// ---- shuffle reader -------------------------------------------------------------------------------------
struct Reader {
  std::vector<std::string> files;
  int elems;  // h*w*c
  size_t capacity, min_after;
  bool shuffle;
  std::vector<std::thread> threads;
  std::mutex mu;
  std::condition_variable not_full, not_empty;
  std::vector<uint8_t> pool;  // capacity records of rec_bytes() each: float32, or the raw float64 of the file
  bool raw_f64 = false;
  size_t rec_bytes() const { return (size_t)elems * (raw_f64 ? 8 : 4); }
  size_t size = 0;
  std::mt19937_64 rng;
  std::atomic<bool> stop{false};
  std::atomic<int> live_threads{0};
  std::string error;
  // shared endless file list (string_input_producer): index into a per-epoch shuffled order
  std::mutex file_mu;
  std::vector<int> order;
  size_t next_file = 0;
  std::mt19937_64 file_rng;
  long long records_read = 0;
  bool verify_crc = true;
  int epochs_limit = 0;  // 0 = endless
  int epoch = 0;

  int take_file() {
    std::lock_guard<std::mutex> l(file_mu);
    if (next_file >= order.size()) {
      if (!order.empty()) {  // end of an epoch (the very first call only fills the list)
        if (epochs_limit && epoch + 1 >= epochs_limit) return -1;
        ++epoch;
      }
      order.resize(files.size());
      for (size_t i = 0; i < order.size(); ++i) order[i] = (int)i;
      if (shuffle) std::shuffle(order.begin(), order.end(), file_rng);
      next_file = 0;
    }
    return order[next_file++];
  }

  void fail(const std::string& msg) {
    std::lock_guard<std::mutex> l(mu);
    if (error.empty()) error = msg;
    stop = true;
    not_empty.notify_all(); not_full.notify_all();
  }

  void push(const double* src) {
    std::unique_lock<std::mutex> l(mu);
    not_full.wait(l, [&] { return size < capacity || stop.load(); });
    if (stop) return;
    if (raw_f64) {  // records stay float64: the cast happens on the GPU (dcgan_decode_f64_f32)
      memcpy(pool.data() + size * rec_bytes(), src, rec_bytes());
    } else {
      float* dst = reinterpret_cast<float*>(pool.data() + size * rec_bytes());
      for (int i = 0; i < elems; ++i) dst[i] = (float)src[i];
    }
    ++size; ++records_read;
    l.unlock();
    not_empty.notify_one();
  }

  void worker() {
    std::vector<uint8_t> buf;
    std::vector<double> tmp(elems);
    while (!stop) {
      int fi = take_file();
      if (fi < 0) break;
      FILE* f = fopen(files[fi].c_str(), "rb");
      if (!f) { fail("cannot open " + files[fi]); break; }
      while (!stop) {
        uint8_t hdr[12];
        size_t got = fread(hdr, 1, 12, f);
        if (got == 0) break;
        if (got != 12) { fail(files[fi] + ": truncated record header"); break; }
        uint64_t len; uint32_t hcrc; memcpy(&len, hdr, 8); memcpy(&hcrc, hdr + 8, 4);
        if (verify_crc && masked(crc32c(hdr, 8, 0)) != hcrc) { fail(files[fi] + ": corrupted record length"); break; }
        buf.resize(len + 4);
        if (fread(buf.data(), 1, len + 4, f) != len + 4) { fail(files[fi] + ": truncated record"); break; }
        uint32_t dcrc; memcpy(&dcrc, buf.data() + len, 4);
        if (verify_crc && masked(crc32c(buf.data(), len, 0)) != dcrc) { fail(files[fi] + ": corrupted record data"); break; }
        Span raw{nullptr, 0};
        if (!find_bytes_feature(Span{buf.data(), (size_t)len}, "image_raw", &raw)) { fail(files[fi] + ": no image_raw feature"); break; }
        if (raw.n != (size_t)elems * 8) { fail(files[fi] + ": image_raw has " + std::to_string(raw.n) + " bytes"); break; }
        memcpy(tmp.data(), raw.p, raw.n);  // decode_raw(float64): unaligned-safe copy
        push(tmp.data());
      }
      fclose(f);
    }
    if (--live_threads == 0) { std::lock_guard<std::mutex> l(mu); not_empty.notify_all(); }
  }

  // RandomShuffleQueue.dequeue_many(n): waits until size - n >= min_after_dequeue (or the producers are done)
  int dequeue(int n, void* out_v) {
    uint8_t* out = static_cast<uint8_t*>(out_v);
    const size_t rb = rec_bytes();
    std::unique_lock<std::mutex> l(mu);
    for (int i = 0; i < n; ++i) {
      not_empty.wait(l, [&] {
        return stop.load() || size > min_after || (live_threads.load() == 0 && size > 0) ||
               (live_threads.load() == 0 && size == 0);
      });
      if (!error.empty()) return -1;
      if (size == 0) return i;  // producers finished (finite epochs)
      size_t idx = shuffle ? (size_t)(rng() % size) : 0;
      uint8_t* src = pool.data() + idx * rb;
      memcpy(out + (size_t)i * rb, src, rb);
      if (shuffle) {
        if (idx != size - 1) memcpy(src, pool.data() + (size - 1) * rb, rb);
      } else {
        memmove(pool.data(), pool.data() + rb, (size - 1) * rb);
      }
      --size;
      not_full.notify_one();
    }
    return n;
  }
};

thread_local std::string g_last_error;

}  // namespace

extern "C" {

uint32_t dcgan_io_crc32c(const uint8_t* data, size_t n, uint32_t crc) { return crc32c(data, n, crc); }

const char* dcgan_io_last_error(void) { return g_last_error.c_str(); }

void* dcgan_io_reader_create2(const char** files, int nfiles, int h, int w, int c, long capacity, long min_after,
                              int threads, unsigned long long seed, int shuffle, int epochs, int verify_crc,
                              int raw_f64);
void* dcgan_io_reader_create(const char** files, int nfiles, int h, int w, int c, long capacity, long min_after,
                             int threads, unsigned long long seed, int shuffle, int epochs, int verify_crc) {
  return dcgan_io_reader_create2(files, nfiles, h, w, c, capacity, min_after, threads, seed, shuffle, epochs,
                                 verify_crc, 0);
}

// raw_f64 != 0: the queue holds the records as stored (float64); dequeue copies 8 bytes per element
void* dcgan_io_reader_create2(const char** files, int nfiles, int h, int w, int c, long capacity, long min_after,
                              int threads, unsigned long long seed, int shuffle, int epochs, int verify_crc,
                              int raw_f64) {
  if (nfiles <= 0 || capacity <= 0 || min_after < 0 || min_after >= capacity || threads <= 0) {
    g_last_error = "bad reader arguments";
    return nullptr;
  }
  Reader* r = new Reader();
  for (int i = 0; i < nfiles; ++i) r->files.emplace_back(files[i]);
  r->elems = h * w * c;
  r->raw_f64 = raw_f64 != 0;
  r->capacity = (size_t)capacity; r->min_after = (size_t)min_after; r->shuffle = shuffle != 0;
  r->pool.resize((size_t)capacity * r->rec_bytes());
  r->rng.seed(seed); r->file_rng.seed(seed ^ 0x9E3779B97F4A7C15ull);
  r->epochs_limit = epochs; r->verify_crc = verify_crc != 0;
  r->live_threads = threads;
  for (int t = 0; t < threads; ++t) r->threads.emplace_back([r] { r->worker(); });
  return r;
}

int dcgan_io_reader_dequeue(void* handle, int n, void* out) {
  Reader* r = static_cast<Reader*>(handle);
  int got = r->dequeue(n, out);
  if (got < 0) g_last_error = r->error;
  return got;
}

long dcgan_io_reader_size(void* handle) {
  Reader* r = static_cast<Reader*>(handle);
  std::lock_guard<std::mutex> l(r->mu);
  return (long)r->size;
}

long long dcgan_io_reader_records_read(void* handle) {
  Reader* r = static_cast<Reader*>(handle);
  std::lock_guard<std::mutex> l(r->mu);
  return r->records_read;
}

void dcgan_io_reader_destroy(void* handle) {
  Reader* r = static_cast<Reader*>(handle);
  {
    std::lock_guard<std::mutex> l(r->mu);
    r->stop = true;
  }
  r->not_full.notify_all(); r->not_empty.notify_all();
  for (auto& t : r->threads) t.join();
  delete r;
}

}  // extern "C"
