// Native I/O for the framework (C ABI, no TensorFlow dependency):
//   * crc32c (Castagnoli) + TF/LevelDB masking
//   * TFRecord reader / writer (length + masked crc framing)
//   * tf.train.Example field extraction (hand-rolled protobuf wire decoding)
//   * threaded shuffling record prefetcher (tf.train.shuffle_batch's queue runners)
//   * TensorBundle (Saver V2) writer / reader: <prefix>.index SSTable + <prefix>.data-XXXXX-of-YYYYY shards
//   * TensorBoard event-file writer (scalar summaries)
//
// The reference delegates all of this to the TF1 runtime: MonitoredTrainingSession(checkpoint_dir=...) at
// /root/reference/distribute_train.py:169-180, TFRecordReader / parse_single_example / shuffle_batch at
// /root/reference/distribute_input.py:54-106, tf.summary.scalar at /root/reference/distribute_tower.py:139-143.
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <map>
#include <mutex>
#include <random>
#include <string>
#include <thread>
#include <vector>

namespace {

// ------------------------------------------------------------------------------------------------ crc32c
uint32_t g_crc_table[8][256];
std::once_flag g_crc_once;
void crc_init() {
  const uint32_t poly = 0x82f63b78u;
  for (uint32_t i = 0; i < 256; ++i) {
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ poly : (c >> 1);
    g_crc_table[0][i] = c;
  }
  for (uint32_t i = 0; i < 256; ++i)
    for (int t = 1; t < 8; ++t) g_crc_table[t][i] = (g_crc_table[t - 1][i] >> 8) ^ g_crc_table[0][g_crc_table[t - 1][i] & 0xff];
}
uint32_t crc32c_extend(uint32_t crc, const uint8_t* p, size_t n) {
  std::call_once(g_crc_once, crc_init);
  uint32_t c = crc ^ 0xffffffffu;
  while (n >= 8) {
    uint32_t lo, hi;
    memcpy(&lo, p, 4);
    memcpy(&hi, p + 4, 4);
    lo ^= c;
    c = g_crc_table[7][lo & 0xff] ^ g_crc_table[6][(lo >> 8) & 0xff] ^ g_crc_table[5][(lo >> 16) & 0xff] ^
        g_crc_table[4][lo >> 24] ^ g_crc_table[3][hi & 0xff] ^ g_crc_table[2][(hi >> 8) & 0xff] ^
        g_crc_table[1][(hi >> 16) & 0xff] ^ g_crc_table[0][hi >> 24];
    p += 8;
    n -= 8;
  }
  while (n--) c = g_crc_table[0][(c ^ *p++) & 0xff] ^ (c >> 8);
  return c ^ 0xffffffffu;
}
inline uint32_t crc_mask(uint32_t crc) { return ((crc >> 15) | (crc << 17)) + 0xa282ead8u; }
inline uint32_t crc_unmask(uint32_t m) {
  uint32_t rot = m - 0xa282ead8u;
  return (rot >> 17) | (rot << 15);
}

// ------------------------------------------------------------------------------------------------ varints / protobuf
void put_varint(std::string* s, uint64_t v) {
  while (v >= 0x80) {
    s->push_back(static_cast<char>(v | 0x80));
    v >>= 7;
  }
  s->push_back(static_cast<char>(v));
}
bool get_varint(const uint8_t*& p, const uint8_t* end, uint64_t* v) {
  uint64_t r = 0;
  for (int shift = 0; shift <= 63 && p < end; shift += 7) {
    uint8_t b = *p++;
    r |= static_cast<uint64_t>(b & 0x7f) << shift;
    if (!(b & 0x80)) {
      *v = r;
      return true;
    }
  }
  return false;
}
void put_fixed32(std::string* s, uint32_t v) { s->append(reinterpret_cast<const char*>(&v), 4); }
void put_fixed64(std::string* s, uint64_t v) { s->append(reinterpret_cast<const char*>(&v), 8); }
void put_tag(std::string* s, int field, int wire) { put_varint(s, (static_cast<uint64_t>(field) << 3) | wire); }
void put_len_field(std::string* s, int field, const std::string& payload) {
  put_tag(s, field, 2);
  put_varint(s, payload.size());
  s->append(payload);
}
// Iterate over the fields of one message. Returns false on malformed input.
struct PbField {
  int field, wire;
  uint64_t varint;          // wire 0 / 1 / 5 (fixed widths zero-extended)
  const uint8_t* data;      // wire 2
  size_t len;
};
bool pb_next(const uint8_t*& p, const uint8_t* end, PbField* f) {
  uint64_t tag;
  if (!get_varint(p, end, &tag)) return false;
  f->field = static_cast<int>(tag >> 3);
  f->wire = static_cast<int>(tag & 7);
  f->data = nullptr;
  f->len = 0;
  f->varint = 0;
  switch (f->wire) {
    case 0:
      return get_varint(p, end, &f->varint);
    case 1:
      if (end - p < 8) return false;
      memcpy(&f->varint, p, 8);
      p += 8;
      return true;
    case 5: {
      if (end - p < 4) return false;
      uint32_t v;
      memcpy(&v, p, 4);
      f->varint = v;
      p += 4;
      return true;
    }
    case 2: {
      uint64_t n;
      if (!get_varint(p, end, &n) || static_cast<uint64_t>(end - p) < n) return false;
      f->data = p;
      f->len = n;
      p += n;
      return true;
    }
    default:
      return false;
  }
}

// ------------------------------------------------------------------------------------------------ TFRecord
struct RecordReader {
  FILE* f = nullptr;
  std::vector<uint8_t> buf;
  bool check_crc = true;
};
// 1 = record read, 0 = clean EOF, -1 = truncated / io error, -2 = crc mismatch
int read_record(RecordReader* r) {
  uint8_t hdr[12];
  size_t n = fread(hdr, 1, 12, r->f);
  if (n == 0) return 0;
  if (n != 12) return -1;
  uint64_t len;
  uint32_t len_crc;
  memcpy(&len, hdr, 8);
  memcpy(&len_crc, hdr + 8, 4);
  if (r->check_crc && crc_mask(crc32c_extend(0, hdr, 8)) != len_crc) return -2;
  if (len > (1ull << 34)) return -1;
  r->buf.resize(len);
  if (len && fread(r->buf.data(), 1, len, r->f) != len) return -1;
  uint32_t data_crc;
  if (fread(&data_crc, 1, 4, r->f) != 4) return -1;
  if (r->check_crc && crc_mask(crc32c_extend(0, r->buf.data(), len)) != data_crc) return -2;
  return 1;
}
bool write_record(FILE* f, const void* data, uint64_t len) {
  uint8_t hdr[12];
  memcpy(hdr, &len, 8);
  uint32_t c = crc_mask(crc32c_extend(0, hdr, 8));
  memcpy(hdr + 8, &c, 4);
  uint32_t dc = crc_mask(crc32c_extend(0, static_cast<const uint8_t*>(data), len));
  return fwrite(hdr, 1, 12, f) == 12 && (len == 0 || fwrite(data, 1, len, f) == len) && fwrite(&dc, 1, 4, f) == 4;
}

// ------------------------------------------------------------------------------------------------ Example decoding
// Example{1: Features{1: map<string, Feature>}}; map entry {1: key, 2: Feature};
// Feature{1: BytesList{1: bytes*}, 2: FloatList{1: packed float*}, 3: Int64List{1: packed varint*}}
bool find_feature(const uint8_t* buf, size_t len, const char* name, int* kind, const uint8_t** list, size_t* list_len) {
  const size_t name_len = strlen(name);
  const uint8_t *p = buf, *end = buf + len;
  PbField f;
  while (p < end) {
    if (!pb_next(p, end, &f)) return false;
    if (f.field != 1 || f.wire != 2) continue;
    const uint8_t *q = f.data, *qend = f.data + f.len;  // Features
    PbField e;
    while (q < qend) {
      if (!pb_next(q, qend, &e)) return false;
      if (e.field != 1 || e.wire != 2) continue;
      const uint8_t *m = e.data, *mend = e.data + e.len;  // map entry
      PbField kv;
      const uint8_t* key = nullptr;
      size_t key_len = 0;
      const uint8_t* feat = nullptr;
      size_t feat_len = 0;
      while (m < mend) {
        if (!pb_next(m, mend, &kv)) return false;
        if (kv.field == 1 && kv.wire == 2) {
          key = kv.data;
          key_len = kv.len;
        } else if (kv.field == 2 && kv.wire == 2) {
          feat = kv.data;
          feat_len = kv.len;
        }
      }
      if (key && key_len == name_len && memcmp(key, name, name_len) == 0) {
        *kind = 0;
        *list = nullptr;
        *list_len = 0;
        if (!feat) return true;
        const uint8_t *t = feat, *tend = feat + feat_len;
        PbField l;
        while (t < tend) {
          if (!pb_next(t, tend, &l)) return false;
          if (l.wire == 2 && l.field >= 1 && l.field <= 3) {
            *kind = l.field;
            *list = l.data;
            *list_len = l.len;
          }
        }
        return true;
      }
    }
  }
  return false;
}

// ------------------------------------------------------------------------------------------------ prefetcher
struct Prefetcher {
  std::vector<std::string> paths;
  bool shuffle_files, loop;
  size_t capacity;
  std::mt19937_64 rng;
  std::mutex mu;
  std::condition_variable cv_not_full, cv_not_empty;
  std::vector<std::vector<uint8_t>> pool;
  std::vector<uint8_t> current;
  std::thread worker;
  std::atomic<bool> stop{false};
  bool done = false;
  int error = 0;
  bool shuffle_records = true;

  void run() {
    std::vector<size_t> order(paths.size());
    for (size_t i = 0; i < order.size(); ++i) order[i] = i;
    std::mt19937_64 frng(rng());
    do {
      if (shuffle_files) std::shuffle(order.begin(), order.end(), frng);
      for (size_t idx : order) {
        RecordReader r;
        r.f = fopen(paths[idx].c_str(), "rb");
        if (!r.f) {
          std::lock_guard<std::mutex> lk(mu);
          error = -3;
          continue;
        }
        int rc;
        while (!stop.load() && (rc = read_record(&r)) == 1) {
          std::unique_lock<std::mutex> lk(mu);
          cv_not_full.wait(lk, [&] { return pool.size() < capacity || stop.load(); });
          if (stop.load()) break;
          pool.emplace_back(std::move(r.buf));
          r.buf.clear();
          cv_not_empty.notify_one();
        }
        if (rc < 0) {
          std::lock_guard<std::mutex> lk(mu);
          error = rc;
        }
        fclose(r.f);
        if (stop.load()) break;
      }
    } while (loop && !stop.load() && error == 0 && !paths.empty());
    std::lock_guard<std::mutex> lk(mu);
    done = true;
    cv_not_empty.notify_all();
  }
};

// ------------------------------------------------------------------------------------------------ SSTable (LevelDB table format, uncompressed)
struct BlockBuilder {
  std::string buf;
  std::vector<uint32_t> restarts{0};
  std::string last_key;
  int counter = 0;
  int interval;
  explicit BlockBuilder(int restart_interval) : interval(restart_interval) {}
  bool empty() const { return buf.empty(); }
  void add(const std::string& key, const std::string& value) {
    size_t shared = 0;
    if (counter < interval) {
      const size_t n = std::min(last_key.size(), key.size());
      while (shared < n && last_key[shared] == key[shared]) ++shared;
    } else {
      restarts.push_back(static_cast<uint32_t>(buf.size()));
      counter = 0;
    }
    put_varint(&buf, shared);
    put_varint(&buf, key.size() - shared);
    put_varint(&buf, value.size());
    buf.append(key, shared, std::string::npos);
    buf.append(value);
    last_key = key;
    ++counter;
  }
  std::string finish() {
    std::string out = buf;
    for (uint32_t r : restarts) put_fixed32(&out, r);
    put_fixed32(&out, static_cast<uint32_t>(restarts.size()));
    return out;
  }
  void reset() {
    buf.clear();
    restarts.assign(1, 0);
    last_key.clear();
    counter = 0;
  }
};
struct BlockHandle {
  uint64_t offset = 0, size = 0;
  std::string encode() const {
    std::string s;
    put_varint(&s, offset);
    put_varint(&s, size);
    return s;
  }
};
const uint64_t kTableMagic = 0xdb4775248b80fb57ull;

bool write_block(FILE* f, const std::string& contents, uint64_t* offset, BlockHandle* h) {
  h->offset = *offset;
  h->size = contents.size();
  uint8_t trailer[5];
  trailer[0] = 0;  // no compression
  uint32_t crc = crc32c_extend(crc32c_extend(0, reinterpret_cast<const uint8_t*>(contents.data()), contents.size()), trailer, 1);
  crc = crc_mask(crc);
  memcpy(trailer + 1, &crc, 4);
  if (fwrite(contents.data(), 1, contents.size(), f) != contents.size()) return false;
  if (fwrite(trailer, 1, 5, f) != 5) return false;
  *offset += contents.size() + 5;
  return true;
}
bool write_table(const std::string& path, const std::map<std::string, std::string>& kv, size_t block_size) {
  FILE* f = fopen(path.c_str(), "wb");
  if (!f) return false;
  uint64_t offset = 0;
  BlockBuilder data(16), index(1);
  bool ok = true;
  std::string pending_last_key;
  auto flush = [&]() {
    if (data.empty()) return;
    BlockHandle h;
    ok = ok && write_block(f, data.finish(), &offset, &h);
    index.add(pending_last_key, h.encode());
    data.reset();
  };
  for (const auto& e : kv) {
    data.add(e.first, e.second);
    pending_last_key = e.first;
    if (data.buf.size() >= block_size) flush();
  }
  flush();
  BlockBuilder meta(16);
  BlockHandle meta_h, index_h;
  ok = ok && write_block(f, meta.finish(), &offset, &meta_h);
  ok = ok && write_block(f, index.finish(), &offset, &index_h);
  std::string footer = meta_h.encode() + index_h.encode();
  footer.resize(40, '\0');
  put_fixed64(&footer, kTableMagic);
  ok = ok && fwrite(footer.data(), 1, footer.size(), f) == footer.size();
  ok = (fclose(f) == 0) && ok;
  return ok;
}
bool read_file(const std::string& path, std::string* out) {
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END);
  long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  out->resize(n);
  bool ok = n == 0 || fread(&(*out)[0], 1, n, f) == static_cast<size_t>(n);
  fclose(f);
  return ok;
}
bool read_block(const std::string& file, uint64_t off, uint64_t size, std::string* contents, bool verify) {
  if (off + size + 5 > file.size()) return false;
  if (file[off + size] != 0) return false;  // compressed blocks are not produced by bundles
  if (verify) {
    uint32_t stored;
    memcpy(&stored, file.data() + off + size + 1, 4);
    uint32_t crc = crc32c_extend(0, reinterpret_cast<const uint8_t*>(file.data()) + off, size + 1);
    if (crc_unmask(stored) != crc) return false;
  }
  contents->assign(file, off, size);
  return true;
}
bool parse_block(const std::string& b, std::vector<std::pair<std::string, std::string>>* out) {
  if (b.size() < 4) return false;
  uint32_t nrestarts;
  memcpy(&nrestarts, b.data() + b.size() - 4, 4);
  if (static_cast<uint64_t>(nrestarts) * 4 + 4 > b.size()) return false;
  const uint8_t* p = reinterpret_cast<const uint8_t*>(b.data());
  const uint8_t* end = p + b.size() - 4 - nrestarts * 4;
  std::string key;
  while (p < end) {
    uint64_t shared, non_shared, vlen;
    if (!get_varint(p, end, &shared) || !get_varint(p, end, &non_shared) || !get_varint(p, end, &vlen)) return false;
    if (shared > key.size() || static_cast<uint64_t>(end - p) < non_shared + vlen) return false;
    key.resize(shared);
    key.append(reinterpret_cast<const char*>(p), non_shared);
    p += non_shared;
    out->emplace_back(key, std::string(reinterpret_cast<const char*>(p), vlen));
    p += vlen;
  }
  return true;
}
bool read_table(const std::string& path, std::vector<std::pair<std::string, std::string>>* out) {
  std::string file;
  if (!read_file(path, &file) || file.size() < 48) return false;
  uint64_t magic;
  memcpy(&magic, file.data() + file.size() - 8, 8);
  if (magic != kTableMagic) return false;
  const uint8_t* p = reinterpret_cast<const uint8_t*>(file.data()) + file.size() - 48;
  const uint8_t* end = p + 40;
  uint64_t mo, ms, io, is;
  if (!get_varint(p, end, &mo) || !get_varint(p, end, &ms) || !get_varint(p, end, &io) || !get_varint(p, end, &is)) return false;
  std::string index_block;
  if (!read_block(file, io, is, &index_block, true)) return false;
  std::vector<std::pair<std::string, std::string>> index_entries;
  if (!parse_block(index_block, &index_entries)) return false;
  for (const auto& ie : index_entries) {
    const uint8_t* q = reinterpret_cast<const uint8_t*>(ie.second.data());
    const uint8_t* qend = q + ie.second.size();
    uint64_t off, size;
    if (!get_varint(q, qend, &off) || !get_varint(q, qend, &size)) return false;
    std::string blk;
    if (!read_block(file, off, size, &blk, true)) return false;
    if (!parse_block(blk, out)) return false;
  }
  return true;
}

// ------------------------------------------------------------------------------------------------ TensorBundle
struct BundleEntry {
  int dtype = 0;
  std::vector<long long> shape;
  int shard = 0;
  long long offset = 0, size = 0;
  uint32_t crc = 0;  // masked
};
std::string encode_entry(const BundleEntry& e) {
  std::string s;
  put_tag(&s, 1, 0);
  put_varint(&s, e.dtype);
  std::string shape;
  for (long long d : e.shape) {
    std::string dim;
    put_tag(&dim, 1, 0);
    put_varint(&dim, static_cast<uint64_t>(d));
    put_len_field(&shape, 2, dim);
  }
  put_len_field(&s, 2, shape);
  if (e.shard) {
    put_tag(&s, 3, 0);
    put_varint(&s, e.shard);
  }
  if (e.offset) {
    put_tag(&s, 4, 0);
    put_varint(&s, e.offset);
  }
  put_tag(&s, 5, 0);
  put_varint(&s, e.size);
  put_tag(&s, 6, 5);
  put_fixed32(&s, e.crc);
  return s;
}
bool decode_entry(const std::string& v, BundleEntry* e) {
  const uint8_t* p = reinterpret_cast<const uint8_t*>(v.data());
  const uint8_t* end = p + v.size();
  PbField f;
  while (p < end) {
    if (!pb_next(p, end, &f)) return false;
    switch (f.field) {
      case 1: e->dtype = static_cast<int>(f.varint); break;
      case 2: {
        const uint8_t *q = f.data, *qend = f.data + f.len;
        PbField d;
        while (q < qend) {
          if (!pb_next(q, qend, &d)) return false;
          if (d.field == 2 && d.wire == 2) {
            const uint8_t *r = d.data, *rend = d.data + d.len;
            PbField s;
            long long size = 0;
            while (r < rend) {
              if (!pb_next(r, rend, &s)) return false;
              if (s.field == 1) size = static_cast<long long>(s.varint);
            }
            e->shape.push_back(size);
          }
        }
        break;
      }
      case 3: e->shard = static_cast<int>(f.varint); break;
      case 4: e->offset = static_cast<long long>(f.varint); break;
      case 5: e->size = static_cast<long long>(f.varint); break;
      case 6: e->crc = static_cast<uint32_t>(f.varint); break;
      default: break;
    }
  }
  return true;
}
std::string shard_name(const std::string& prefix, int shard, int num_shards) {
  char buf[64];
  snprintf(buf, sizeof(buf), ".data-%05d-of-%05d", shard, num_shards);
  return prefix + buf;
}
struct BundleWriter {
  std::string prefix;
  int num_shards;
  std::vector<FILE*> files;
  std::vector<long long> offsets;
  std::map<std::string, std::string> index;
  bool ok = true;
};
struct BundleReader {
  std::string prefix;
  int num_shards = 1;
  std::vector<std::pair<std::string, BundleEntry>> entries;
};

struct EventWriter {
  FILE* f = nullptr;
};

}  // namespace

extern "C" {

int dtf_io_version(void) { return 2; }

unsigned int dtf_crc32c(const void* data, unsigned long long n) { return crc32c_extend(0, static_cast<const uint8_t*>(data), n); }
unsigned int dtf_crc32c_masked(const void* data, unsigned long long n) {
  return crc_mask(crc32c_extend(0, static_cast<const uint8_t*>(data), n));
}

// ---- TFRecord
void* dtf_tfrecord_open(const char* path, int check_crc) {
  FILE* f = fopen(path, "rb");
  if (!f) return nullptr;
  RecordReader* r = new RecordReader();
  r->f = f;
  r->check_crc = check_crc != 0;
  return r;
}
int dtf_tfrecord_next(void* h, const unsigned char** data, unsigned long long* len) {
  RecordReader* r = static_cast<RecordReader*>(h);
  int rc = read_record(r);
  if (rc == 1) {
    *data = r->buf.data();
    *len = r->buf.size();
  }
  return rc;
}
void dtf_tfrecord_close(void* h) {
  RecordReader* r = static_cast<RecordReader*>(h);
  if (r->f) fclose(r->f);
  delete r;
}
void* dtf_tfrecord_writer_open(const char* path) { return fopen(path, "wb"); }
int dtf_tfrecord_write(void* h, const void* data, unsigned long long len) {
  return write_record(static_cast<FILE*>(h), data, len) ? 0 : -1;
}
int dtf_tfrecord_writer_close(void* h) { return fclose(static_cast<FILE*>(h)); }

// ---- Example access. kind: 1 bytes, 2 float, 3 int64. Returns number of values, -1 if the feature is absent.
int dtf_example_count(const void* buf, unsigned long long len, const char* name, int* kind_out) {
  int kind;
  const uint8_t* list;
  size_t list_len;
  if (!find_feature(static_cast<const uint8_t*>(buf), len, name, &kind, &list, &list_len)) return -1;
  *kind_out = kind;
  if (kind == 0) return 0;
  const uint8_t *p = list, *end = list + list_len;
  PbField f;
  int n = 0;
  while (p < end) {
    if (!pb_next(p, end, &f)) return -2;
    if (f.field != 1) continue;
    if (kind == 1) ++n;
    else if (kind == 2) n += (f.wire == 2) ? static_cast<int>(f.len / 4) : 1;
    else {
      if (f.wire == 2) {
        const uint8_t *q = f.data, *qend = f.data + f.len;
        uint64_t v;
        while (q < qend && get_varint(q, qend, &v)) ++n;
      } else ++n;
    }
  }
  return n;
}
// bytes value #idx: offset / length relative to buf
int dtf_example_bytes(const void* buf, unsigned long long len, const char* name, int idx, unsigned long long* off,
                      unsigned long long* vlen) {
  int kind;
  const uint8_t* list;
  size_t list_len;
  if (!find_feature(static_cast<const uint8_t*>(buf), len, name, &kind, &list, &list_len) || kind != 1) return -1;
  const uint8_t *p = list, *end = list + list_len;
  PbField f;
  int n = 0;
  while (p < end) {
    if (!pb_next(p, end, &f)) return -2;
    if (f.field == 1 && f.wire == 2) {
      if (n == idx) {
        *off = f.data - static_cast<const uint8_t*>(buf);
        *vlen = f.len;
        return 0;
      }
      ++n;
    }
  }
  return -1;
}
int dtf_example_floats(const void* buf, unsigned long long len, const char* name, float* out, int max_n) {
  int kind;
  const uint8_t* list;
  size_t list_len;
  if (!find_feature(static_cast<const uint8_t*>(buf), len, name, &kind, &list, &list_len) || kind != 2) return -1;
  const uint8_t *p = list, *end = list + list_len;
  PbField f;
  int n = 0;
  while (p < end) {
    if (!pb_next(p, end, &f)) return -2;
    if (f.field != 1) continue;
    if (f.wire == 2) {
      for (size_t i = 0; i + 4 <= f.len && n < max_n; i += 4) memcpy(&out[n++], f.data + i, 4);
    } else if (f.wire == 5 && n < max_n) {
      uint32_t u = static_cast<uint32_t>(f.varint);
      memcpy(&out[n++], &u, 4);
    }
  }
  return n;
}
int dtf_example_int64s(const void* buf, unsigned long long len, const char* name, long long* out, int max_n) {
  int kind;
  const uint8_t* list;
  size_t list_len;
  if (!find_feature(static_cast<const uint8_t*>(buf), len, name, &kind, &list, &list_len) || kind != 3) return -1;
  const uint8_t *p = list, *end = list + list_len;
  PbField f;
  int n = 0;
  while (p < end) {
    if (!pb_next(p, end, &f)) return -2;
    if (f.field != 1) continue;
    if (f.wire == 2) {
      const uint8_t *q = f.data, *qend = f.data + f.len;
      uint64_t v;
      while (q < qend && n < max_n && get_varint(q, qend, &v)) out[n++] = static_cast<long long>(v);
    } else if (f.wire == 0 && n < max_n) {
      out[n++] = static_cast<long long>(f.varint);
    }
  }
  return n;
}

// ---- shuffling prefetcher. paths: '\n'-separated. capacity in records.
void* dtf_prefetch_open(const char* paths_nl, int shuffle_files, int shuffle_records, int loop,
                        unsigned long long capacity, unsigned long long seed) {
  Prefetcher* p = new Prefetcher();
  std::string s(paths_nl);
  size_t pos = 0;
  while (pos <= s.size()) {
    size_t nl = s.find('\n', pos);
    if (nl == std::string::npos) nl = s.size();
    if (nl > pos) p->paths.emplace_back(s.substr(pos, nl - pos));
    pos = nl + 1;
  }
  p->shuffle_files = shuffle_files != 0;
  p->shuffle_records = shuffle_records != 0;
  p->loop = loop != 0;
  p->capacity = capacity ? capacity : 1;
  p->rng.seed(seed);
  p->worker = std::thread([p] { p->run(); });
  return p;
}
// 1 = record (valid until the next call), 0 = end of data, <0 = error
int dtf_prefetch_next(void* h, const unsigned char** data, unsigned long long* len) {
  Prefetcher* p = static_cast<Prefetcher*>(h);
  std::unique_lock<std::mutex> lk(p->mu);
  p->cv_not_empty.wait(lk, [&] { return !p->pool.empty() || p->done; });
  if (p->pool.empty()) return p->error ? p->error : 0;
  size_t idx = 0;
  if (p->shuffle_records) idx = std::uniform_int_distribution<size_t>(0, p->pool.size() - 1)(p->rng);
  if (p->shuffle_records) std::swap(p->pool[idx], p->pool.back());
  if (p->shuffle_records) {
    p->current = std::move(p->pool.back());
    p->pool.pop_back();
  } else {
    p->current = std::move(p->pool.front());
    p->pool.erase(p->pool.begin());
  }
  p->cv_not_full.notify_one();
  *data = p->current.data();
  *len = p->current.size();
  return 1;
}
void dtf_prefetch_close(void* h) {
  Prefetcher* p = static_cast<Prefetcher*>(h);
  p->stop.store(true);
  {
    std::lock_guard<std::mutex> lk(p->mu);
    p->cv_not_full.notify_all();
  }
  if (p->worker.joinable()) p->worker.join();
  delete p;
}

// ---- TensorBundle writer
void* dtf_bundle_writer_open(const char* prefix, int num_shards) {
  BundleWriter* w = new BundleWriter();
  w->prefix = prefix;
  w->num_shards = num_shards < 1 ? 1 : num_shards;
  for (int s = 0; s < w->num_shards; ++s) {
    FILE* f = fopen((shard_name(w->prefix, s, w->num_shards) + ".tempstate").c_str(), "wb");
    if (!f) w->ok = false;
    w->files.push_back(f);
    w->offsets.push_back(0);
  }
  return w;
}
int dtf_bundle_writer_add(void* h, const char* name, int dtype, const long long* shape, int rank, const void* data,
                          unsigned long long nbytes, int shard) {
  BundleWriter* w = static_cast<BundleWriter*>(h);
  if (!w->ok || shard < 0 || shard >= w->num_shards || !w->files[shard]) return -1;
  BundleEntry e;
  e.dtype = dtype;
  e.shape.assign(shape, shape + rank);
  e.shard = shard;
  e.offset = w->offsets[shard];
  e.size = static_cast<long long>(nbytes);
  e.crc = crc_mask(crc32c_extend(0, static_cast<const uint8_t*>(data), nbytes));
  if (nbytes && fwrite(data, 1, nbytes, w->files[shard]) != nbytes) {
    w->ok = false;
    return -1;
  }
  w->offsets[shard] += nbytes;
  w->index[name] = encode_entry(e);
  return 0;
}
int dtf_bundle_writer_finish(void* h) {
  BundleWriter* w = static_cast<BundleWriter*>(h);
  bool ok = w->ok;
  for (int s = 0; s < w->num_shards; ++s) {
    if (w->files[s]) ok = (fclose(w->files[s]) == 0) && ok;
    const std::string fin = shard_name(w->prefix, s, w->num_shards);
    ok = (rename((fin + ".tempstate").c_str(), fin.c_str()) == 0) && ok;
  }
  // header: BundleHeaderProto{1: num_shards, 2: endianness (0 = little, default omitted), 3: VersionDef{1: producer=1}}
  std::string header;
  put_tag(&header, 1, 0);
  put_varint(&header, w->num_shards);
  std::string version;
  put_tag(&version, 1, 0);
  put_varint(&version, 1);
  put_len_field(&header, 3, version);
  w->index[""] = header;
  const std::string tmp = w->prefix + ".index.tempstate";
  ok = write_table(tmp, w->index, 256 * 1024) && ok;
  ok = (rename(tmp.c_str(), (w->prefix + ".index").c_str()) == 0) && ok;
  delete w;
  return ok ? 0 : -1;
}

// ---- TensorBundle reader
void* dtf_bundle_reader_open(const char* prefix) {
  std::vector<std::pair<std::string, std::string>> kv;
  if (!read_table(std::string(prefix) + ".index", &kv)) return nullptr;
  BundleReader* r = new BundleReader();
  r->prefix = prefix;
  for (const auto& e : kv) {
    if (e.first.empty()) {
      const uint8_t* p = reinterpret_cast<const uint8_t*>(e.second.data());
      const uint8_t* end = p + e.second.size();
      PbField f;
      while (p < end && pb_next(p, end, &f))
        if (f.field == 1) r->num_shards = static_cast<int>(f.varint);
      continue;
    }
    BundleEntry be;
    if (!decode_entry(e.second, &be)) {
      delete r;
      return nullptr;
    }
    r->entries.emplace_back(e.first, be);
  }
  return r;
}
int dtf_bundle_reader_num_tensors(void* h) { return static_cast<int>(static_cast<BundleReader*>(h)->entries.size()); }
int dtf_bundle_reader_num_shards(void* h) { return static_cast<BundleReader*>(h)->num_shards; }
int dtf_bundle_reader_entry(void* h, int i, char* name_out, int name_cap, int* dtype, long long* shape8, int* rank,
                            long long* nbytes, int* shard) {
  BundleReader* r = static_cast<BundleReader*>(h);
  if (i < 0 || i >= static_cast<int>(r->entries.size())) return -1;
  const auto& e = r->entries[i];
  if (static_cast<int>(e.first.size()) + 1 > name_cap || e.second.shape.size() > 8) return -2;
  memcpy(name_out, e.first.c_str(), e.first.size() + 1);
  *dtype = e.second.dtype;
  *rank = static_cast<int>(e.second.shape.size());
  for (size_t d = 0; d < e.second.shape.size(); ++d) shape8[d] = e.second.shape[d];
  *nbytes = e.second.size;
  *shard = e.second.shard;
  return 0;
}
// 0 ok, -1 not found, -2 size mismatch, -3 io error, -4 crc mismatch
int dtf_bundle_reader_read(void* h, const char* name, void* dst, unsigned long long nbytes) {
  BundleReader* r = static_cast<BundleReader*>(h);
  for (const auto& e : r->entries) {
    if (e.first != name) continue;
    if (static_cast<unsigned long long>(e.second.size) != nbytes) return -2;
    FILE* f = fopen(shard_name(r->prefix, e.second.shard, r->num_shards).c_str(), "rb");
    if (!f) return -3;
    bool ok = fseek(f, e.second.offset, SEEK_SET) == 0 && (nbytes == 0 || fread(dst, 1, nbytes, f) == nbytes);
    fclose(f);
    if (!ok) return -3;
    if (crc_mask(crc32c_extend(0, static_cast<const uint8_t*>(dst), nbytes)) != e.second.crc) return -4;
    return 0;
  }
  return -1;
}
void dtf_bundle_reader_close(void* h) { delete static_cast<BundleReader*>(h); }

// ---- TensorBoard events: Event{1: wall_time double, 2: step int64, 3: file_version string, 5: Summary}
//      Summary{1: Value{1: tag, 2: simple_value float}}
void* dtf_events_open(const char* path, double wall_time) {
  FILE* f = fopen(path, "wb");
  if (!f) return nullptr;
  EventWriter* w = new EventWriter();
  w->f = f;
  std::string ev;
  put_tag(&ev, 1, 1);
  uint64_t bits;
  memcpy(&bits, &wall_time, 8);
  put_fixed64(&ev, bits);
  put_len_field(&ev, 3, "brain.Event:2");
  write_record(f, ev.data(), ev.size());
  fflush(f);
  return w;
}
int dtf_events_scalar(void* h, const char* tag, float value, long long step, double wall_time) {
  EventWriter* w = static_cast<EventWriter*>(h);
  std::string val;
  put_len_field(&val, 1, tag);
  put_tag(&val, 2, 5);
  uint32_t fb;
  memcpy(&fb, &value, 4);
  put_fixed32(&val, fb);
  std::string summary;
  put_len_field(&summary, 1, val);
  std::string ev;
  put_tag(&ev, 1, 1);
  uint64_t bits;
  memcpy(&bits, &wall_time, 8);
  put_fixed64(&ev, bits);
  put_tag(&ev, 2, 0);
  put_varint(&ev, static_cast<uint64_t>(step));
  put_len_field(&ev, 5, summary);
  return write_record(w->f, ev.data(), ev.size()) ? 0 : -1;
}
int dtf_events_flush(void* h) { return fflush(static_cast<EventWriter*>(h)->f); }
int dtf_events_close(void* h) {
  EventWriter* w = static_cast<EventWriter*>(h);
  int rc = fclose(w->f);
  delete w;
  return rc;
}

}  // extern "C"
