// Fuzz driver for the tf.train.Example parser and the remembered-record-shape fast path of csrc/io/record_loader.cpp, built
// by tools/sanitize_host.sh under AddressSanitizer + UBSan.  Every input lives in a heap buffer of exactly its size, so any
// read past the record is reported.  Inputs: random bytes, valid MNIST examples, and valid examples with random byte flips,
// truncations and extensions.  A mutated record that the fast path accepts must decode to what the general parser says.
#include "../csrc/io/record_loader.cpp"

namespace dtm { int dtm_fail(const char* m) { fprintf(stderr, "fail: %s\n", m); return -1; } }

static void put_varint(std::vector<uint8_t>& v, uint64_t x) {
  while (x >= 0x80) { v.push_back(static_cast<uint8_t>(x) | 0x80); x >>= 7; }
  v.push_back(static_cast<uint8_t>(x));
}
static void put_bytes_field(std::vector<uint8_t>& v, int field, const std::vector<uint8_t>& b) {
  put_varint(v, (field << 3) | 2);
  put_varint(v, b.size());
  v.insert(v.end(), b.begin(), b.end());
}
static std::vector<uint8_t> example(const std::vector<uint8_t>& img, uint64_t label) {
  std::vector<uint8_t> bl, f1, e1, il, pk, f2, e2, feats, ex;
  put_bytes_field(bl, 1, img);
  put_bytes_field(f1, 1, bl);
  put_bytes_field(e1, 1, std::vector<uint8_t>{'i', 'm', 'a', 'g', 'e', '_', 'r', 'a', 'w'});
  put_bytes_field(e1, 2, f1);
  put_varint(pk, label);
  put_bytes_field(il, 1, pk);
  put_bytes_field(f2, 3, il);
  put_bytes_field(e2, 1, std::vector<uint8_t>{'l', 'a', 'b', 'e', 'l'});
  put_bytes_field(e2, 2, f2);
  put_bytes_field(feats, 1, e1);
  put_bytes_field(feats, 1, e2);
  put_bytes_field(ex, 1, feats);
  return ex;
}

int main(int argc, char** argv) {
  const long iters = argc > 1 ? atol(argv[1]) : 300000;
  std::mt19937_64 rng(12345);
  RecordShape shape;
  long parsed = 0, fast = 0, rejected = 0;
  for (long it = 0; it < iters; ++it) {
    std::vector<uint8_t> rec;
    const int kind = static_cast<int>(rng() % 4);
    if (kind == 0) {
      rec.resize(rng() % 96);
      for (auto& b : rec) b = static_cast<uint8_t>(rng());
    } else {
      std::vector<uint8_t> img(16);
      for (auto& b : img) b = static_cast<uint8_t>(rng());
      const uint64_t label = (rng() % 3 == 0) ? rng() : rng() % 10;
      rec = example(img, label);
      if (kind >= 2) {
        const int flips = 1 + static_cast<int>(rng() % 3);
        for (int f = 0; f < flips && !rec.empty(); ++f) rec[rng() % rec.size()] ^= static_cast<uint8_t>(1u << (rng() % 8));
      }
      if (kind == 3 && !rec.empty()) {
        if (rng() % 2) rec.resize(rng() % rec.size()); else rec.push_back(static_cast<uint8_t>(rng()));
      }
    }
    // exact-size heap copy: ASan flags any access beyond it
    uint8_t* buf = static_cast<uint8_t*>(malloc(rec.size() ? rec.size() : 1));
    if (!rec.empty()) memcpy(buf, rec.data(), rec.size());
    const uint8_t* img = nullptr;
    size_t img_len = 0;
    int64_t label = 0;
    const uint8_t* label_pos = nullptr;
    const bool ok = parse_example(buf, rec.size(), &img, &img_len, &label, &label_pos);
    if (ok) {
      ++parsed;
      if (img < buf || img + img_len > buf + rec.size()) { fprintf(stderr, "image outside the record\n"); return 1; }
      if (shape.match(buf, rec.size())) {
        ++fast;
        if (buf + shape.img_off != img || shape.img_len != img_len || static_cast<int64_t>(buf[shape.label_off]) != label) {
          fprintf(stderr, "fast path disagrees with the parser at iteration %ld\n", it);
          return 2;
        }
      }
      shape.remember(buf, rec.size(), img, img_len, label_pos);
    } else {
      ++rejected;
      if (shape.match(buf, rec.size())) { fprintf(stderr, "fast path accepts a record the parser rejects (iteration %ld)\n", it); return 3; }
    }
    free(buf);
  }
  printf("fuzz: %ld inputs, %ld parsed (%ld also matched the remembered shape), %ld rejected\n", iters, parsed, fast, rejected);
  return 0;
}
