// Native input pipeline for real data: the reference's queue-runner pipeline (mnist_input.py:44-73) as C++ threads.
//
//   filename queue (shuffled, infinite epochs)  ->  N reader threads: TFRecord framing (length + masked CRC-32C of both
//   the length and the payload), tf.train.Example parsing (`image_raw`: bytes, `label`: int64)  ->  RandomShuffleQueue
//   (capacity, min_after_dequeue; dequeue = uniformly random element, blocks while size <= min_after_dequeue)  ->
//   dtm_loader_next() copies one batch straight into the caller's (pinned) buffers: uint8 pixels [B][image_bytes] and
//   int32 labels [B] - the packed layout NativeEngine.stage_packed moves to the device with one H2D copy.
//
// Files are mapped read-only for the loader's lifetime and the queue holds views into the mappings, so a pixel is copied
// once on its way from the page cache to the batch buffer.
//
// Python keeps an equivalent implementation (data/readers.py) as the portable path and as the oracle for the tests.
// Pure host code: no CUDA calls, usable on machines without a GPU.
#include <fcntl.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include "../common/host_util.h"

extern "C" uint32_t dtm_crc32c(const char* data, size_t n, uint32_t crc);

namespace {

inline uint32_t masked_crc(const uint8_t* p, size_t n) {
  const uint32_t c = dtm_crc32c(reinterpret_cast<const char*>(p), n, 0);
  return ((c >> 15) | (c << 17)) + 0xa282ead8u;
}

// ---- minimal protobuf reader -----------------------------------------------------------------------------------
struct Cursor {
  const uint8_t* p;
  const uint8_t* end;
  bool ok = true;
  uint64_t varint() {
    if (p < end && !(*p & 0x80)) return *p++;            // tags and small lengths: one byte
    uint64_t v = 0;
    for (int shift = 0; shift < 64 && p < end; shift += 7) {
      const uint8_t b = *p++;
      v |= static_cast<uint64_t>(b & 0x7F) << shift;
      if (!(b & 0x80)) return v;
    }
    ok = false;
    return 0;
  }
  // reads a length-delimited field body
  Cursor sub() {
    const uint64_t n = varint();
    if (!ok || n > static_cast<uint64_t>(end - p)) { ok = false; return Cursor{p, p, false}; }
    Cursor c{p, p + n, true};
    p += n;
    return c;
  }
  void skip(uint32_t wire) {
    switch (wire) {
      case 0: varint(); break;
      case 1: if (end - p >= 8) p += 8; else ok = false; break;
      case 2: sub(); break;
      case 5: if (end - p >= 4) p += 4; else ok = false; break;
      default: ok = false;
    }
  }
};

// tf.train.Example { features(1) { map<string, Feature> feature(1) } } ; Feature { bytes_list(1){value(1)} | float_list(2) |
// int64_list(3){value(1), packed or not} }.  Extracts `image_raw` and `label`.
// `label_pos` receives the address of the varint the label was read from (the last one seen).
bool parse_example(const uint8_t* rec, size_t len, const uint8_t** img, size_t* img_len, int64_t* label,
                   const uint8_t** label_pos = nullptr) {
  bool have_img = false, have_label = false;
  const uint8_t* lpos = nullptr;
  Cursor ex{rec, rec + len};
  while (ex.ok && ex.p < ex.end) {
    const uint64_t tag = ex.varint();
    if ((tag >> 3) != 1 || (tag & 7) != 2) { ex.skip(tag & 7); continue; }
    Cursor feats = ex.sub();
    while (feats.ok && feats.p < feats.end) {
      const uint64_t t2 = feats.varint();
      if ((t2 >> 3) != 1 || (t2 & 7) != 2) { feats.skip(t2 & 7); continue; }
      Cursor entry = feats.sub();
      const uint8_t* key = nullptr;
      size_t key_len = 0;
      Cursor value{nullptr, nullptr, false};
      while (entry.ok && entry.p < entry.end) {
        const uint64_t t3 = entry.varint();
        if ((t3 & 7) != 2) { entry.skip(t3 & 7); continue; }
        Cursor body = entry.sub();
        if ((t3 >> 3) == 1) { key = body.p; key_len = static_cast<size_t>(body.end - body.p); }
        else if ((t3 >> 3) == 2) value = body;
      }
      if (!entry.ok) return false;
      if (!value.ok) continue;
      if (key_len == 9 && memcmp(key, "image_raw", 9) == 0) {
        while (value.ok && value.p < value.end) {
          const uint64_t t4 = value.varint();
          if ((t4 >> 3) == 1 && (t4 & 7) == 2) {          // bytes_list
            Cursor bl = value.sub();
            while (bl.ok && bl.p < bl.end) {
              const uint64_t t5 = bl.varint();
              if ((t5 >> 3) == 1 && (t5 & 7) == 2) {
                Cursor b = bl.sub();
                if (!bl.ok) return false;
                *img = b.p;
                *img_len = static_cast<size_t>(b.end - b.p);
                have_img = true;
              } else {
                bl.skip(t5 & 7);
              }
            }
            if (!bl.ok) return false;
          } else {
            value.skip(t4 & 7);
          }
        }
      } else if (key_len == 5 && memcmp(key, "label", 5) == 0) {
        while (value.ok && value.p < value.end) {
          const uint64_t t4 = value.varint();
          if ((t4 >> 3) == 3 && (t4 & 7) == 2) {          // int64_list
            Cursor il = value.sub();
            while (il.ok && il.p < il.end) {
              const uint64_t t5 = il.varint();
              if ((t5 >> 3) == 1 && (t5 & 7) == 0) {        // unpacked value
                lpos = il.p;
                *label = static_cast<int64_t>(il.varint());
                have_label = true;
              } else if ((t5 >> 3) == 1 && (t5 & 7) == 2) { // packed values
                Cursor pk = il.sub();
                if (pk.ok && pk.p < pk.end) { lpos = pk.p; *label = static_cast<int64_t>(pk.varint()); have_label = pk.ok; }
              } else {
                il.skip(t5 & 7);
              }
            }
            if (!il.ok) return false;
          } else {
            value.skip(t4 & 7);
          }
        }
      }
      if (!value.ok) return false;
    }
    if (!feats.ok) return false;
  }
  if (label_pos) *label_pos = lpos;
  return ex.ok && have_img && have_label;
}

// Records written by one writer share their whole wire structure: same fields in the same order with the same lengths,
// only the pixel bytes and the (one-byte varint) label differ.  A reader thread remembers the first record it parsed in
// full; a later record of the same length whose every byte outside the pixels and the label value equals the remembered
// one has, byte for byte, the same parse - its pixels and label are read at the remembered offsets (a 90-byte memcmp
// instead of ~60 varints).  Anything else takes the general parser above.
struct RecordShape {
  bool valid = false;
  size_t len = 0, img_off = 0, img_len = 0, label_off = 0;
  std::vector<uint8_t> bytes;

  void remember(const uint8_t* rec, size_t n, const uint8_t* img, size_t il, const uint8_t* label_pos) {
    valid = false;
    if (!label_pos || (*label_pos & 0x80)) return;                  // multi-byte label varint: its length is structure
    const size_t lo = static_cast<size_t>(label_pos - rec);
    const size_t io = static_cast<size_t>(img - rec);
    if (lo >= io && lo < io + il) return;
    len = n; img_off = io; img_len = il; label_off = lo;
    bytes.assign(rec, rec + n);
    valid = true;
  }
  static bool same(const uint8_t* a, const uint8_t* b, size_t lo, size_t hi, size_t skip) {   // [lo, hi) without `skip`
    if (skip >= lo && skip < hi) return memcmp(a + lo, b + lo, skip - lo) == 0 && memcmp(a + skip + 1, b + skip + 1, hi - skip - 1) == 0;
    return memcmp(a + lo, b + lo, hi - lo) == 0;
  }
  bool match(const uint8_t* rec, size_t n) const {
    return valid && n == len && !(rec[label_off] & 0x80) && same(rec, bytes.data(), 0, img_off, label_off) &&
           same(rec, bytes.data(), img_off + img_len, len, label_off);
  }
};

// ---- the loader ------------------------------------------------------------------------------------------------
struct Loader {
  std::vector<std::string> files;
  int batch = 0, image_bytes = 0, capacity = 0, min_after = 0;
  bool shuffle = true;
  uint64_t seed = 0;

  // The shuffle queue holds VIEWS: a queue element is the address of a record's pixels inside its file's mapping (verified
  // by the payload CRC, immutable, mapped for the loader's lifetime) plus the label - 16 bytes instead of a pixel copy.
  // Pixels are copied exactly once, from the page cache into the caller's batch buffer.
  struct Entry {
    const uint8_t* pixels;
    int32_t label;
  };
  std::vector<Entry> filled;          // queue order; size() <= capacity
  size_t head = 0;                    // unshuffled mode dequeues from the front: elements before `head` are gone
  std::mutex mu;
  std::condition_variable not_full, enough;
  std::atomic<bool> stop{false};
  std::string error;                  // first producer error (guarded by mu)
  std::vector<std::thread> threads;
  std::mt19937_64 consumer_rng;
  uint64_t records_read = 0;

  void fail(const std::string& msg) {
    std::lock_guard<std::mutex> lk(mu);
    if (error.empty()) error = msg;
    enough.notify_all();
  }

  // ONE record stream for all reader threads, like the reference's single TFRecordReader behind its filename queue
  // (mnist_input.py:57-71): under `read_mu` a thread only walks the record headers of the current file (length + its CRC,
  // 12 bytes per record) and takes the next CHUNK_RECORDS records as (offset, length) pairs into the file's read-only
  // mapping; payload CRC and Example parsing run outside every lock, straight from the page cache - every record is
  // delivered once per epoch, whichever thread asks.
  struct Mapping {
    const uint8_t* base = nullptr;
    size_t size = 0;
    bool mapped = false;
    std::vector<uint8_t> owned;          // files that cannot be mapped (pipes, odd file systems) are read whole instead
    std::string path;
    ~Mapping() {
      if (mapped) munmap(const_cast<uint8_t*>(base), size);
    }
  };
  std::mutex read_mu;
  std::map<std::string, std::shared_ptr<Mapping>> mapped_files;     // for the loader's lifetime: pages are faulted in once
  std::shared_ptr<Mapping> cur;
  size_t cur_off = 0;
  std::vector<std::string> order;
  size_t next_file = 0;
  uint64_t epoch_records = 0;
  bool first_epoch = true;
  std::mt19937_64 file_rng;

  // A run of consecutive records of ONE file.  Taking records one by one made the threads hand the reader lock to each
  // other for every 800 bytes (two threads were slower than one); a chunk amortises it.
  struct Chunk {
    std::shared_ptr<Mapping> file;                       // keeps the mapping alive while the records are in use
    std::vector<std::pair<size_t, uint64_t>> recs;     // (offset of the payload in the file, payload length)
  };
  static constexpr int CHUNK_RECORDS = 128;

  static std::shared_ptr<Mapping> open_mapping(const std::string& path, std::string* why) {
    auto m = std::make_shared<Mapping>();
    m->path = path;
    const int fd = open(path.c_str(), O_RDONLY | O_CLOEXEC);
    if (fd < 0) { *why = "Failed to find file: " + path; return nullptr; }
    struct stat st;
    if (fstat(fd, &st) == 0 && S_ISREG(st.st_mode) && st.st_size > 0) {
      void* p = mmap(nullptr, static_cast<size_t>(st.st_size), PROT_READ, MAP_PRIVATE, fd, 0);
      if (p != MAP_FAILED) {
        m->base = static_cast<const uint8_t*>(p);
        m->size = static_cast<size_t>(st.st_size);
        m->mapped = true;
      }
    }
    if (!m->mapped) {
      uint8_t buf[1 << 16];
      for (;;) {
        const ssize_t got = read(fd, buf, sizeof buf);
        if (got < 0) { close(fd); *why = "read error in " + path; return nullptr; }
        if (got == 0) break;
        m->owned.insert(m->owned.end(), buf, buf + got);
      }
      m->base = m->owned.data();
      m->size = m->owned.size();
    }
    close(fd);
    return m;
  }

  // false: stop requested or error recorded
  bool next_chunk(Chunk& c) {
    std::lock_guard<std::mutex> rl(read_mu);
    c.file.reset();
    c.recs.clear();
    for (;;) {
      if (stop.load()) return false;
      if (!cur) {
        if (!c.recs.empty()) return true;          // a chunk never spans two files
        if (next_file == order.size()) {          // epoch boundary
          if (!first_epoch && epoch_records == 0) { fail("input files contain no records"); return false; }
          first_epoch = false;
          order = files;
          if (shuffle) std::shuffle(order.begin(), order.end(), file_rng);
          next_file = 0;
          epoch_records = 0;
        }
        const std::string& path = order[next_file++];
        auto it = mapped_files.find(path);
        if (it != mapped_files.end()) {
          cur = it->second;
        } else {
          std::string why;
          cur = open_mapping(path, &why);
          if (!cur) { fail(why); return false; }
          mapped_files[path] = cur;               // queue elements point into it: it lives as long as the loader
        }
        cur_off = 0;
      }
      c.file = cur;
      if (cur_off == cur->size) { cur.reset(); continue; }        // clean end of file
      auto bad = [&](const std::string& what) { const std::string path = cur->path; cur.reset(); fail(what + path); return false; };
      if (cur->size - cur_off < 12) return bad("truncated TFRecord header in ");
      const uint8_t* hdr = cur->base + cur_off;
      uint64_t len;
      uint32_t len_crc;
      memcpy(&len, hdr, 8);
      memcpy(&len_crc, hdr + 8, 4);
      if (masked_crc(hdr, 8) != len_crc) return bad("corrupt TFRecord length in ");
      if (len > (1ull << 30)) return bad("implausible TFRecord length in ");
      if (cur->size - cur_off - 12 < len + 4) return bad("truncated TFRecord in ");
      c.recs.emplace_back(cur_off + 12, len);
      cur_off += 12 + len + 4;
      // the header walk is a pointer chase through memory; records are mostly of one size, so the header four records on
      // is very likely at four times this stride
      if (cur->size - cur_off > 4 * (16 + len)) __builtin_prefetch(cur->base + cur_off + 4 * (16 + len));
      ++epoch_records;
      if (static_cast<int>(c.recs.size()) >= CHUNK_RECORDS) return true;
    }
  }

  // Shuffle-queue hand-off: verified records enter the queue as views, as many as there is room for per lock round.
  void produce() {
    Chunk c;
    RecordShape shape;
    std::vector<Entry> ents;
    while (next_chunk(c)) {
      ents.clear();
      for (const auto& rec : c.recs) {       // payload CRC, Example parsing and validation: outside every lock
        const uint8_t* payload = c.file->base + rec.first;
        uint32_t data_crc;
        memcpy(&data_crc, payload + rec.second, 4);
        if (masked_crc(payload, rec.second) != data_crc) { fail("corrupt TFRecord payload in " + c.file->path); return; }
        const uint8_t* img = nullptr;
        size_t img_len = 0;
        int64_t label = 0;
        if (shape.match(payload, rec.second)) {
          img = payload + shape.img_off;
          img_len = shape.img_len;
          label = payload[shape.label_off];
        } else {
          const uint8_t* label_pos = nullptr;
          if (!parse_example(payload, rec.second, &img, &img_len, &label, &label_pos)) {
            fail("record without image_raw/label features in " + c.file->path);
            return;
          }
          shape.remember(payload, rec.second, img, img_len, label_pos);
        }
        if (static_cast<int>(img_len) != image_bytes) {
          fail("record holds " + std::to_string(img_len) + " image bytes, expected " + std::to_string(image_bytes));
          return;
        }
        ents.push_back(Entry{img, static_cast<int32_t>(label)});
      }
      size_t done = 0;
      while (done < ents.size()) {
        std::unique_lock<std::mutex> lk(mu);
        not_full.wait(lk, [&] { return stop.load() || queued() < static_cast<size_t>(capacity); });
        if (stop.load()) return;
        const size_t k = std::min(static_cast<size_t>(capacity) - queued(), ents.size() - done);
        filled.insert(filled.end(), ents.begin() + done, ents.begin() + done + k);
        records_read += k;
        done += k;
        if (queued() > static_cast<size_t>(min_after)) enough.notify_one();
      }
    }
  }

  size_t queued() const { return filled.size() - head; }

  // Batch ring (dtm_loader_ring_*): one more thread dequeues whole batches into caller-owned (pinned) buffers ahead of
  // the trainer, which then only swaps buffer indices.
  struct Ring {
    bool running = false;
    std::vector<uint8_t*> buffers;
    uint64_t label_offset = 0;
    std::deque<int> free, ready;
    std::mutex mu;
    std::condition_variable cv;
    std::string error;
    std::thread worker;
  } ring;
};

thread_local std::string g_loader_error;

}  // namespace

extern "C" void* dtm_loader_create(const char* const* files, int n_files, int batch, int image_bytes, int num_threads,
                                   uint64_t seed, int shuffle, int min_after_dequeue) {
  if (!files || n_files <= 0 || batch <= 0 || image_bytes <= 0) { dtm::dtm_fail("loader: bad arguments"); return nullptr; }
  for (int i = 0; i < n_files; ++i) {
    FILE* f = fopen(files[i], "rb");
    if (!f) {
      g_loader_error = std::string("Failed to find file: ") + files[i];   // the reference's ValueError text (mnist_input.py:55)
      dtm::dtm_fail(g_loader_error.c_str());
      return nullptr;
    }
    fclose(f);
  }
  Loader* L = new Loader();
  L->files.assign(files, files + n_files);
  L->batch = batch;
  L->image_bytes = image_bytes;
  L->shuffle = shuffle != 0;
  L->min_after = L->shuffle ? (min_after_dequeue < 0 ? 0 : min_after_dequeue) : 0;
  L->capacity = (min_after_dequeue < 0 ? 0 : min_after_dequeue) + 3 * batch;        // mnist_input.py:70
  L->seed = seed;
  L->filled.reserve(static_cast<size_t>(L->capacity) + L->batch);
  L->consumer_rng.seed(seed);
  // without shuffling the order must be the file order: one reader
  const int nthreads = L->shuffle ? (num_threads < 1 ? 1 : num_threads) : 1;
  L->file_rng.seed(seed + 1);
  for (int t = 0; t < nthreads; ++t) L->threads.emplace_back([L] { L->produce(); });
  return L;
}

// dequeue_many(batch): every element drawn uniformly from the queue while it holds more than min_after_dequeue elements
// (views only under the lock; their places are free for the producers at once), then the pixels are copied out.
// Returns an empty string, or the error that ended the stream.
static std::string dequeue_batch(Loader* L, uint8_t* pixels_out, int32_t* labels_out) {
  std::vector<Loader::Entry> picked(L->batch);
  {
    std::unique_lock<std::mutex> lk(L->mu);
    for (int i = 0; i < L->batch; ++i) {
      if (static_cast<int>(L->queued()) <= L->min_after) {
        L->not_full.notify_all();                  // about to sleep: the producers may use the places taken so far
        L->enough.wait(lk, [&] { return L->stop.load() || !L->error.empty() || static_cast<int>(L->queued()) > L->min_after; });
      }
      if (!L->error.empty()) return L->error;
      if (L->stop.load()) return "loader is shutting down";
      if (L->shuffle) {
        const size_t j = static_cast<size_t>(L->consumer_rng() % L->filled.size());
        picked[i] = L->filled[j];
        L->filled[j] = L->filled.back();
        L->filled.pop_back();
      } else {
        picked[i] = L->filled[L->head++];
        if (L->head == L->filled.size()) { L->filled.clear(); L->head = 0; }
      }
    }
    if (L->head > static_cast<size_t>(L->capacity)) {         // unshuffled: drop the consumed front now and then
      L->filled.erase(L->filled.begin(), L->filled.begin() + L->head);
      L->head = 0;
    }
  }
  L->not_full.notify_all();
  const size_t ib = static_cast<size_t>(L->image_bytes);
  constexpr int AHEAD = 6;                            // picks are random addresses: start their cache misses early
  for (int i = 0; i < L->batch; ++i) {
    if (i + AHEAD < L->batch) {
      const uint8_t* nxt = picked[i + AHEAD].pixels;
      for (size_t o = 0; o < ib; o += 64) __builtin_prefetch(nxt + o);
    }
    memcpy(pixels_out + static_cast<size_t>(i) * ib, picked[i].pixels, ib);
    labels_out[i] = picked[i].label;
  }
  return std::string();
}

extern "C" int dtm_loader_next(void* handle, uint8_t* pixels_out, int32_t* labels_out) {
  Loader* L = static_cast<Loader*>(handle);
  if (!L || !pixels_out || !labels_out) return dtm::dtm_fail("loader: null argument");
  if (L->ring.running) return dtm::dtm_fail("loader: dtm_loader_next after dtm_loader_ring_start (the ring's thread owns the queue's consumer side)");
  g_loader_error = dequeue_batch(L, pixels_out, labels_out);
  return g_loader_error.empty() ? 0 : dtm::dtm_fail(g_loader_error.c_str());
}

// ---- batch ring: a thread of the loader assembles whole batches ahead of the trainer ------------------------------------
extern "C" int dtm_loader_ring_start(void* handle, uint8_t* const* buffers, int n_buffers, uint64_t label_offset) {
  Loader* L = static_cast<Loader*>(handle);
  if (!L || !buffers || n_buffers < 2) return dtm::dtm_fail("loader ring: needs a loader and at least two buffers");
  if (L->ring.running) return dtm::dtm_fail("loader ring: already started");
  if (label_offset < static_cast<uint64_t>(L->batch) * L->image_bytes || label_offset % 4) {
    return dtm::dtm_fail("loader ring: labels must follow the pixels at a 4-byte aligned offset");
  }
  Loader::Ring& R = L->ring;
  R.buffers.assign(buffers, buffers + n_buffers);
  R.label_offset = label_offset;
  for (int i = 0; i < n_buffers; ++i) R.free.push_back(i);
  R.running = true;
  R.worker = std::thread([L] {
    Loader::Ring& R = L->ring;
    for (;;) {
      int idx;
      {
        std::unique_lock<std::mutex> lk(R.mu);
        R.cv.wait(lk, [&] { return L->stop.load() || !R.free.empty(); });
        if (L->stop.load()) return;
        idx = R.free.front();
        R.free.pop_front();
      }
      uint8_t* buf = R.buffers[idx];
      std::string err = dequeue_batch(L, buf, reinterpret_cast<int32_t*>(buf + R.label_offset));
      std::lock_guard<std::mutex> lk(R.mu);
      if (!err.empty()) {
        R.error = err;
        R.cv.notify_all();
        return;
      }
      R.ready.push_back(idx);
      R.cv.notify_all();
    }
  });
  return 0;
}

extern "C" int dtm_loader_ring_acquire(void* handle) {
  Loader* L = static_cast<Loader*>(handle);
  if (!L || !L->ring.running) { dtm::dtm_fail("loader ring: not started"); return -1; }
  Loader::Ring& R = L->ring;
  std::unique_lock<std::mutex> lk(R.mu);
  R.cv.wait(lk, [&] { return !R.ready.empty() || !R.error.empty() || L->stop.load(); });
  if (!R.ready.empty()) {                         // batches assembled before an error are still delivered
    const int idx = R.ready.front();
    R.ready.pop_front();
    return idx;
  }
  g_loader_error = R.error.empty() ? "loader is shutting down" : R.error;
  dtm::dtm_fail(g_loader_error.c_str());
  return -1;
}

extern "C" int dtm_loader_ring_release(void* handle, int index) {
  Loader* L = static_cast<Loader*>(handle);
  if (!L || !L->ring.running) return dtm::dtm_fail("loader ring: not started");
  Loader::Ring& R = L->ring;
  if (index < 0 || index >= static_cast<int>(R.buffers.size())) return dtm::dtm_fail("loader ring: bad buffer index");
  {
    std::lock_guard<std::mutex> lk(R.mu);
    R.free.push_back(index);
  }
  R.cv.notify_all();
  return 0;
}

extern "C" uint64_t dtm_loader_records_read(void* handle) {
  Loader* L = static_cast<Loader*>(handle);
  if (!L) return 0;
  std::lock_guard<std::mutex> lk(L->mu);
  return L->records_read;
}

extern "C" void dtm_loader_destroy(void* handle) {
  Loader* L = static_cast<Loader*>(handle);
  if (!L) return;
  {
    std::lock_guard<std::mutex> lk(L->mu);
    L->stop.store(true);
  }
  L->not_full.notify_all();
  L->enough.notify_all();
  if (L->ring.running) {
    { std::lock_guard<std::mutex> lk(L->ring.mu); }
    L->ring.cv.notify_all();
    L->ring.worker.join();
  }
  for (auto& t : L->threads) t.join();
  delete L;
}
