// Native input pipeline for real data: the reference's queue-runner pipeline (mnist_input.py:44-73) as C++ threads.
//
//   filename queue (shuffled, infinite epochs)  ->  N reader threads: TFRecord framing (length + masked CRC-32C of both
//   the length and the payload), tf.train.Example parsing (`image_raw`: bytes, `label`: int64)  ->  RandomShuffleQueue
//   (capacity, min_after_dequeue; dequeue = uniformly random element, blocks while size <= min_after_dequeue)  ->
//   dtm_loader_next() copies one batch straight into the caller's (pinned) buffers: uint8 pixels [B][image_bytes] and
//   int32 labels [B] - the packed layout NativeEngine.stage_packed moves to the device with one H2D copy.
//
// Python keeps an equivalent implementation (data/readers.py) as the portable path and as the oracle for the tests.
// Pure host code: no CUDA calls, usable on machines without a GPU.
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <condition_variable>
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
bool parse_example(const uint8_t* rec, size_t len, const uint8_t** img, size_t* img_len, int64_t* label) {
  bool have_img = false, have_label = false;
  Cursor ex{rec, rec + len};
  while (ex.ok && ex.p < ex.end) {
    const uint64_t tag = ex.varint();
    if ((tag >> 3) != 1 || (tag & 7) != 2) { ex.skip(tag & 7); continue; }
    Cursor feats = ex.sub();
    while (feats.ok && feats.p < feats.end) {
      const uint64_t t2 = feats.varint();
      if ((t2 >> 3) != 1 || (t2 & 7) != 2) { feats.skip(t2 & 7); continue; }
      Cursor entry = feats.sub();
      std::string key;
      Cursor value{nullptr, nullptr, false};
      while (entry.ok && entry.p < entry.end) {
        const uint64_t t3 = entry.varint();
        if ((t3 & 7) != 2) { entry.skip(t3 & 7); continue; }
        Cursor body = entry.sub();
        if ((t3 >> 3) == 1) key.assign(reinterpret_cast<const char*>(body.p), body.end - body.p);
        else if ((t3 >> 3) == 2) value = body;
      }
      if (!entry.ok) return false;
      if (!value.ok) continue;
      if (key == "image_raw") {
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
      } else if (key == "label") {
        while (value.ok && value.p < value.end) {
          const uint64_t t4 = value.varint();
          if ((t4 >> 3) == 3 && (t4 & 7) == 2) {          // int64_list
            Cursor il = value.sub();
            while (il.ok && il.p < il.end) {
              const uint64_t t5 = il.varint();
              if ((t5 >> 3) == 1 && (t5 & 7) == 0) {        // unpacked value
                *label = static_cast<int64_t>(il.varint());
                have_label = true;
              } else if ((t5 >> 3) == 1 && (t5 & 7) == 2) { // packed values
                Cursor pk = il.sub();
                if (pk.ok && pk.p < pk.end) { *label = static_cast<int64_t>(pk.varint()); have_label = pk.ok; }
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
  return ex.ok && have_img && have_label;
}

// ---- the loader ------------------------------------------------------------------------------------------------
struct Loader {
  std::vector<std::string> files;
  int batch = 0, image_bytes = 0, capacity = 0, min_after = 0;
  bool shuffle = true;
  uint64_t seed = 0;

  std::vector<uint8_t> slab;          // capacity x image_bytes
  std::vector<int32_t> labels;        // capacity
  std::vector<int> filled;            // slots holding a record, in queue order
  std::vector<int> free_slots;
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
  // (mnist_input.py:57-71): a thread takes the next raw record under `read_mu` (framing + both CRCs), then parses the Example
  // and fills the shuffle queue on its own - every record is delivered once per epoch, whichever thread asks.
  std::mutex read_mu;
  FILE* cur = nullptr;
  std::string cur_path;
  std::vector<std::string> order;
  size_t next_file = 0;
  uint64_t epoch_records = 0;
  bool first_epoch = true;
  std::mt19937_64 file_rng;

  // A chunk of consecutive raw records (payload + its CRC word each) of ONE file.  Taking records one by one made the
  // threads hand the reader lock to each other for every 800 bytes (two threads were slower than one); a chunk amortises it.
  struct Chunk {
    std::vector<uint8_t> bytes;
    std::vector<std::pair<size_t, uint64_t>> recs;     // (offset of the payload in `bytes`, payload length)
    std::string path;
  };
  static constexpr int CHUNK_RECORDS = 128;

  // false: stop requested or error recorded
  bool next_chunk(Chunk& c) {
    std::lock_guard<std::mutex> rl(read_mu);
    c.bytes.clear();
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
        cur_path = order[next_file++];
        cur = fopen(cur_path.c_str(), "rb");
        if (!cur) { fail("Failed to find file: " + cur_path); return false; }
        setvbuf(cur, nullptr, _IOFBF, 1 << 20);
      }
      c.path = cur_path;
      uint8_t hdr[12];
      const size_t got = fread(hdr, 1, 12, cur);
      if (got == 0) { fclose(cur); cur = nullptr; continue; }     // clean end of file
      auto bad = [&](const std::string& what) { fclose(cur); cur = nullptr; fail(what + cur_path); return false; };
      if (got != 12) return bad("truncated TFRecord header in ");
      uint64_t len;
      uint32_t len_crc;
      memcpy(&len, hdr, 8);
      memcpy(&len_crc, hdr + 8, 4);
      if (masked_crc(hdr, 8) != len_crc) return bad("corrupt TFRecord length in ");
      if (len > (1ull << 30)) return bad("implausible TFRecord length in ");
      const size_t off = c.bytes.size();
      c.bytes.resize(off + len + 4);
      if (fread(c.bytes.data() + off, 1, len + 4, cur) != len + 4) return bad("truncated TFRecord in ");
      c.recs.emplace_back(off, len);
      ++epoch_records;
      if (static_cast<int>(c.recs.size()) >= CHUNK_RECORDS) return true;
    }
  }

  // Shuffle-queue hand-off: the mutex only guards the slot lists; pixels are copied in and out of a slot while nobody else
  // can name it (a producer owns a slot between taking it off the free list and appending it to `filled`).
  void produce() {
    Chunk c;
    std::vector<const uint8_t*> imgs;
    std::vector<int32_t> labs;
    std::vector<int> slots;
    while (next_chunk(c)) {
      imgs.clear();
      labs.clear();
      for (const auto& rec : c.recs) {       // payload CRC, Example parsing and validation: outside every lock
        const uint8_t* payload = c.bytes.data() + rec.first;
        uint32_t data_crc;
        memcpy(&data_crc, payload + rec.second, 4);
        if (masked_crc(payload, rec.second) != data_crc) { fail("corrupt TFRecord payload in " + c.path); return; }
        const uint8_t* img = nullptr;
        size_t img_len = 0;
        int64_t label = 0;
        if (!parse_example(payload, rec.second, &img, &img_len, &label)) { fail("record without image_raw/label features in " + c.path); return; }
        if (static_cast<int>(img_len) != image_bytes) {
          fail("record holds " + std::to_string(img_len) + " image bytes, expected " + std::to_string(image_bytes));
          return;
        }
        imgs.push_back(img);
        labs.push_back(static_cast<int32_t>(label));
      }
      size_t done = 0;
      while (done < imgs.size()) {           // as many free slots as there are, at least one, per lock round
        slots.clear();
        {
          std::unique_lock<std::mutex> lk(mu);
          not_full.wait(lk, [&] { return stop.load() || !free_slots.empty(); });
          if (stop.load()) return;
          while (!free_slots.empty() && done + slots.size() < imgs.size()) {
            slots.push_back(free_slots.back());
            free_slots.pop_back();
          }
        }
        for (size_t k = 0; k < slots.size(); ++k) {
          memcpy(slab.data() + static_cast<size_t>(slots[k]) * image_bytes, imgs[done + k], image_bytes);
          labels[slots[k]] = labs[done + k];
        }
        {
          std::lock_guard<std::mutex> lk(mu);
          for (int sl : slots) filled.push_back(sl);
          records_read += slots.size();
          if (static_cast<int>(filled.size()) > min_after) enough.notify_one();
        }
        done += slots.size();
      }
    }
  }

  ~Loader() {
    if (cur) fclose(cur);
  }
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
  L->slab.resize(static_cast<size_t>(L->capacity) * image_bytes);
  L->labels.resize(L->capacity);
  L->free_slots.resize(L->capacity);
  for (int i = 0; i < L->capacity; ++i) L->free_slots[i] = L->capacity - 1 - i;
  L->consumer_rng.seed(seed);
  // without shuffling the order must be the file order: one reader
  const int nthreads = L->shuffle ? (num_threads < 1 ? 1 : num_threads) : 1;
  L->file_rng.seed(seed + 1);
  for (int t = 0; t < nthreads; ++t) L->threads.emplace_back([L] { L->produce(); });
  return L;
}

extern "C" int dtm_loader_next(void* handle, uint8_t* pixels_out, int32_t* labels_out) {
  Loader* L = static_cast<Loader*>(handle);
  if (!L || !pixels_out || !labels_out) return dtm::dtm_fail("loader: null argument");
  // dequeue_many(batch): every element drawn uniformly from the queue while it holds more than min_after_dequeue elements
  // (slot indices only under the lock), then the pixels are copied out and the slots go back to the free list
  std::vector<int> picked(L->batch);
  {
    std::unique_lock<std::mutex> lk(L->mu);
    for (int i = 0; i < L->batch; ++i) {
      L->enough.wait(lk, [&] { return !L->error.empty() || static_cast<int>(L->filled.size()) > L->min_after; });
      if (!L->error.empty()) {
        g_loader_error = L->error;
        for (int k = 0; k < i; ++k) L->free_slots.push_back(picked[k]);
        return dtm::dtm_fail(g_loader_error.c_str());
      }
      size_t j = 0;
      if (L->shuffle) j = static_cast<size_t>(L->consumer_rng() % L->filled.size());
      picked[i] = L->filled[j];
      if (L->shuffle) {
        L->filled[j] = L->filled.back();
        L->filled.pop_back();
      } else {
        L->filled.erase(L->filled.begin());
      }
    }
  }
  for (int i = 0; i < L->batch; ++i) {
    memcpy(pixels_out + static_cast<size_t>(i) * L->image_bytes, L->slab.data() + static_cast<size_t>(picked[i]) * L->image_bytes,
           L->image_bytes);
    labels_out[i] = L->labels[picked[i]];
  }
  {
    std::lock_guard<std::mutex> lk(L->mu);
    for (int i = 0; i < L->batch; ++i) L->free_slots.push_back(picked[i]);
  }
  L->not_full.notify_all();
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
  for (auto& t : L->threads) t.join();
  delete L;
}
