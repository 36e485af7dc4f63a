// tf.train.Example field scanner (protobuf wire format, no protobuf dependency): finds one
// feature by key inside a serialized Example and hands back its bytes value in place or its
// numeric values converted to double. Used by the native input path to pull `image/encoded`,
// `image/class/label` and the bounding boxes out of ImageNet / COCO records without building
// Python objects for the other fields (T/preprocessing.py:27-98 parses the same records).
#include <cstdint>
#include <cstring>

#include "../../include/kfb200_host.h"

namespace {

struct Span {
  const uint8_t* p;
  const uint8_t* end;
};

bool varint(Span& s, uint64_t* out) {
  uint64_t v = 0;
  int shift = 0;
  while (s.p < s.end && shift < 70) {
    const uint8_t b = *s.p++;
    v |= static_cast<uint64_t>(b & 0x7F) << shift;
    if (!(b & 0x80)) {
      *out = v;
      return true;
    }
    shift += 7;
  }
  return false;
}

// Reads one field; for length-delimited fields `val` is the payload span.
bool field(Span& s, uint32_t* num, uint32_t* wire, Span* val, uint64_t* scalar) {
  uint64_t key;
  if (!varint(s, &key)) return false;
  *num = static_cast<uint32_t>(key >> 3);
  *wire = static_cast<uint32_t>(key & 7);
  switch (*wire) {
    case 0:
      return varint(s, scalar);
    case 1:
      if (s.end - s.p < 8) return false;
      std::memcpy(scalar, s.p, 8);
      s.p += 8;
      return true;
    case 5: {
      if (s.end - s.p < 4) return false;
      uint32_t v;
      std::memcpy(&v, s.p, 4);
      *scalar = v;
      s.p += 4;
      return true;
    }
    case 2: {
      uint64_t n;
      if (!varint(s, &n) || static_cast<uint64_t>(s.end - s.p) < n) return false;
      val->p = s.p;
      val->end = s.p + n;
      s.p += n;
      return true;
    }
    default:
      return false;
  }
}

// Locates Feature `key` in Example `data`; returns its payload span (the Feature message).
bool find_feature(const uint8_t* data, uint64_t len, const char* key, Span* feature) {
  const size_t klen = std::strlen(key);
  Span ex{data, data + len};
  uint32_t num, wire;
  Span features{nullptr, nullptr};
  uint64_t scalar;
  while (ex.p < ex.end) {
    Span v{nullptr, nullptr};
    if (!field(ex, &num, &wire, &v, &scalar)) return false;
    if (num != 1 || wire != 2) continue;  // Example.features
    features = v;
    while (features.p < features.end) {
      Span entry{nullptr, nullptr};
      if (!field(features, &num, &wire, &entry, &scalar)) return false;
      if (num != 1 || wire != 2) continue;  // map entry
      Span k{nullptr, nullptr}, f{nullptr, nullptr};
      Span e = entry;
      while (e.p < e.end) {
        Span v2{nullptr, nullptr};
        if (!field(e, &num, &wire, &v2, &scalar)) return false;
        if (wire != 2) continue;
        if (num == 1) k = v2;
        else if (num == 2) f = v2;
      }
      if (k.p != nullptr && static_cast<size_t>(k.end - k.p) == klen &&
          std::memcmp(k.p, key, klen) == 0) {
        *feature = f;
        return f.p != nullptr || true;
      }
    }
  }
  return false;
}

}  // namespace

extern "C" {

// First element of a BytesList feature, in place. 1 found, 0 missing / not bytes.
int kfb_example_get_bytes(const uint8_t* data, uint64_t len, const char* key, const uint8_t** out,
                          uint64_t* out_len) {
  Span f{nullptr, nullptr};
  if (!find_feature(data, len, key, &f) || f.p == nullptr) return 0;
  uint32_t num, wire;
  uint64_t scalar;
  while (f.p < f.end) {
    Span list{nullptr, nullptr};
    if (!field(f, &num, &wire, &list, &scalar)) return 0;
    if (num != 1 || wire != 2) continue;  // Feature.bytes_list
    while (list.p < list.end) {
      Span v{nullptr, nullptr};
      if (!field(list, &num, &wire, &v, &scalar)) return 0;
      if (num == 1 && wire == 2) {
        *out = v.p;
        *out_len = static_cast<uint64_t>(v.end - v.p);
        return 1;
      }
    }
  }
  return 0;
}

// Values of a FloatList / Int64List feature as doubles. Returns the number of values the
// feature holds (at most `cap` are written), or -1 if the key is missing.
int kfb_example_get_numbers(const uint8_t* data, uint64_t len, const char* key, double* out,
                            int cap) {
  Span f{nullptr, nullptr};
  if (!find_feature(data, len, key, &f)) return -1;
  if (f.p == nullptr) return 0;
  int count = 0;
  uint32_t num, wire;
  uint64_t scalar;
  while (f.p < f.end) {
    Span list{nullptr, nullptr};
    if (!field(f, &num, &wire, &list, &scalar)) return count;
    if (wire != 2 || (num != 2 && num != 3)) continue;
    const bool is_float = num == 2;
    while (list.p < list.end) {
      Span v{nullptr, nullptr};
      uint32_t n2, w2;
      if (!field(list, &n2, &w2, &v, &scalar)) return count;
      if (n2 != 1) continue;
      if (w2 == 2) {  // packed
        if (is_float) {
          for (const uint8_t* p = v.p; p + 4 <= v.end; p += 4) {
            float x;
            std::memcpy(&x, p, 4);
            if (count < cap) out[count] = x;
            ++count;
          }
        } else {
          Span pv = v;
          uint64_t x;
          while (pv.p < pv.end && varint(pv, &x)) {
            if (count < cap) out[count] = static_cast<double>(static_cast<int64_t>(x));
            ++count;
          }
        }
      } else if (is_float && w2 == 5) {
        float x;
        const uint32_t bits = static_cast<uint32_t>(scalar);
        std::memcpy(&x, &bits, 4);
        if (count < cap) out[count] = x;
        ++count;
      } else if (!is_float && w2 == 0) {
        if (count < cap) out[count] = static_cast<double>(static_cast<int64_t>(scalar));
        ++count;
      }
    }
  }
  return count;
}

}  // extern "C"
