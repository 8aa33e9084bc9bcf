// N2: the batch wire format without TensorFlow (host code; no CUDA in this file).
//
// The reference stores one tf.train.Example per BATCH in TFRecord shards
// (/root/reference/compress_data/utils.py:35-42 create_example; compress_data/main.py:97-125 writer;
// tensorflow_codes/run.py:40-66 reader): features "positive_sample" int64 [B*3], "negative_sample" int64 [B*N],
// "subsampling_weight" float [B], "mode" int64 [B].  This file restates the two public formats involved:
//   TFRecord framing  : u64 length | u32 masked_crc32c(length) | data | u32 masked_crc32c(data)   (little endian),
//                       masked(c) = ((c >> 15) | (c << 17)) + 0xa282ead8, CRC-32C (Castagnoli, reflected 0x82F63B78)
//   protobuf wire     : Example{1: Features{1: map<string, Feature>}}, Feature{1: BytesList, 2: FloatList, 3: Int64List},
//                       FloatList{1: packed float}, Int64List{1: packed varint}  (tensorflow/core/example/*.proto)
// The parser accepts map entries in any order and packed or unpacked repeated scalars; the writer emits the map in
// sorted key order with packed lists (what protobuf's deterministic serialisation emits -- tests compare bytes).
#include <string.h>

#include "kge_internal.h"

namespace kge {
namespace {

uint32_t g_crc_table[8][256];
bool g_crc_ready = false;

void crc_init() {
    for (uint32_t i = 0; i < 256; ++i) {
        uint32_t c = i;
        for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ 0x82F63B78u : c >> 1;
        g_crc_table[0][i] = c;
    }
    for (uint32_t i = 0; i < 256; ++i)
        for (int t = 1; t < 8; ++t) g_crc_table[t][i] = (g_crc_table[t - 1][i] >> 8) ^ g_crc_table[0][g_crc_table[t - 1][i] & 0xFF];
    g_crc_ready = true;
}

uint32_t crc32c(const uint8_t* p, size_t n) {
    if (!g_crc_ready) crc_init();
    uint32_t c = 0xFFFFFFFFu;
    while (n >= 8) {  // slicing-by-8
        uint32_t lo, hi;
        memcpy(&lo, p, 4);
        memcpy(&hi, p + 4, 4);
        lo ^= c;
        c = g_crc_table[7][lo & 0xFF] ^ g_crc_table[6][(lo >> 8) & 0xFF] ^ g_crc_table[5][(lo >> 16) & 0xFF] ^
            g_crc_table[4][lo >> 24] ^ g_crc_table[3][hi & 0xFF] ^ g_crc_table[2][(hi >> 8) & 0xFF] ^
            g_crc_table[1][(hi >> 16) & 0xFF] ^ g_crc_table[0][hi >> 24];
        p += 8;
        n -= 8;
    }
    while (n--) c = (c >> 8) ^ g_crc_table[0][(c ^ *p++) & 0xFF];
    return c ^ 0xFFFFFFFFu;
}

inline uint32_t mask_crc(uint32_t c) { return ((c >> 15) | (c << 17)) + 0xa282ead8u; }

// ---- protobuf wire helpers ----------------------------------------------------------------------
struct Cur {
    const uint8_t* p;
    const uint8_t* end;
    bool ok;
};
uint64_t rd_varint(Cur& c) {
    uint64_t v = 0;
    for (int shift = 0; shift < 70; shift += 7) {
        if (c.p >= c.end) { c.ok = false; return 0; }
        const uint8_t b = *c.p++;
        v |= (uint64_t)(b & 0x7F) << (shift < 64 ? shift : 63);
        if (!(b & 0x80)) return v;
    }
    c.ok = false;
    return 0;
}
Cur rd_len(Cur& c) {  // length-delimited payload as a sub-cursor
    const uint64_t n = rd_varint(c);
    if (!c.ok || n > (uint64_t)(c.end - c.p)) { c.ok = false; return Cur{c.p, c.p, false}; }
    Cur s{c.p, c.p + n, true};
    c.p += n;
    return s;
}
void skip_field(Cur& c, uint32_t wt) {
    switch (wt) {
        case 0: rd_varint(c); break;
        case 1: if (c.end - c.p < 8) c.ok = false; else c.p += 8; break;
        case 2: rd_len(c); break;
        case 5: if (c.end - c.p < 4) c.ok = false; else c.p += 4; break;
        default: c.ok = false;
    }
}

size_t varint_size(uint64_t v) {
    size_t n = 1;
    while (v >= 0x80) { v >>= 7; ++n; }
    return n;
}
uint8_t* wr_varint(uint8_t* o, uint64_t v) {
    while (v >= 0x80) { *o++ = (uint8_t)(v | 0x80); v >>= 7; }
    *o++ = (uint8_t)v;
    return o;
}

struct Sink {
    int64_t* i64;
    float* f32;
    size_t cap, n;
};

// Feature payload -> values appended to the sink (int64 list or float list)
bool parse_feature(Cur f, Sink& s) {
    while (f.ok && f.p < f.end) {
        const uint64_t tag = rd_varint(f);
        const uint32_t field = (uint32_t)(tag >> 3), wt = (uint32_t)(tag & 7);
        if (wt != 2 || (field != 2 && field != 3)) { skip_field(f, wt); continue; }
        Cur list = rd_len(f);  // FloatList / Int64List
        while (list.ok && list.p < list.end) {
            const uint64_t t2 = rd_varint(list);
            const uint32_t fld = (uint32_t)(t2 >> 3), w2 = (uint32_t)(t2 & 7);
            if (fld != 1) { skip_field(list, w2); continue; }
            if (field == 3) {  // int64
                if (!s.i64) return false;
                if (w2 == 2) {
                    Cur pk = rd_len(list);
                    while (pk.ok && pk.p < pk.end) {
                        const uint64_t v = rd_varint(pk);
                        if (s.n >= s.cap) return false;
                        s.i64[s.n++] = (int64_t)v;
                    }
                    if (!pk.ok) return false;
                } else if (w2 == 0) {
                    const uint64_t v = rd_varint(list);
                    if (s.n >= s.cap) return false;
                    s.i64[s.n++] = (int64_t)v;
                } else return false;
            } else {  // float
                if (!s.f32) return false;
                if (w2 == 2) {
                    Cur pk = rd_len(list);
                    if (!pk.ok || ((pk.end - pk.p) & 3)) return false;
                    const size_t cnt = (size_t)(pk.end - pk.p) / 4;
                    if (s.n + cnt > s.cap) return false;
                    memcpy(s.f32 + s.n, pk.p, cnt * 4);
                    s.n += cnt;
                } else if (w2 == 5) {
                    if (list.end - list.p < 4 || s.n >= s.cap) return false;
                    memcpy(s.f32 + s.n++, list.p, 4);
                    list.p += 4;
                } else return false;
            }
        }
        if (!list.ok) return false;
    }
    return f.ok;
}

}  // namespace
}  // namespace kge

using namespace kge;

extern "C" {

uint32_t kge_crc32c(const void* data, size_t n) { return crc32c(static_cast<const uint8_t*>(data), n); }

// One TFRecord frame at `offset` of a shard held in memory.  0 = ok, 1 = clean end of file, negative = error.
int kge_tfrecord_next(const uint8_t* file, size_t len, size_t offset, size_t* data_offset, size_t* data_len,
                      size_t* next_offset, int verify_crc) {
    if (!file || !data_offset || !data_len || !next_offset) return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    if (offset == len) return 1;
    if (offset > len || len - offset < 12) return fail(KGE_ERR_INVALID_ARGUMENT, "truncated TFRecord header at %zu", offset);
    uint64_t n;
    uint32_t crc_len;
    memcpy(&n, file + offset, 8);
    memcpy(&crc_len, file + offset + 8, 4);
    if (verify_crc && mask_crc(crc32c(file + offset, 8)) != crc_len)
        return fail(KGE_ERR_INVALID_ARGUMENT, "TFRecord length CRC mismatch at %zu", offset);
    if (n > len - offset - 12 || len - offset - 12 - n < 4)
        return fail(KGE_ERR_INVALID_ARGUMENT, "truncated TFRecord payload at %zu", offset);
    const uint8_t* data = file + offset + 12;
    if (verify_crc) {
        uint32_t crc_data;
        memcpy(&crc_data, data + n, 4);
        if (mask_crc(crc32c(data, (size_t)n)) != crc_data)
            return fail(KGE_ERR_INVALID_ARGUMENT, "TFRecord data CRC mismatch at %zu", offset);
    }
    *data_offset = offset + 12;
    *data_len = (size_t)n;
    *next_offset = offset + 12 + (size_t)n + 4;
    return 0;
}

// frame = header + data + footer around `data` (out needs data_len + 16 bytes); returns bytes written
size_t kge_tfrecord_frame(const uint8_t* data, size_t data_len, uint8_t* out) {
    const uint64_t n = data_len;
    memcpy(out, &n, 8);
    const uint32_t c1 = mask_crc(crc32c(out, 8));
    memcpy(out + 8, &c1, 4);
    memcpy(out + 12, data, data_len);
    const uint32_t c2 = mask_crc(crc32c(data, data_len));
    memcpy(out + 12 + data_len, &c2, 4);
    return data_len + 16;
}

// Serialized tf.train.Example -> the four features of one batch record (run.py:40-53).  counts[4] receives the
// number of values of positive_sample, negative_sample, subsampling_weight, mode.
int kge_record_parse_example(const uint8_t* buf, size_t len, int64_t* positive, size_t cap_positive, int64_t* negative,
                             size_t cap_negative, float* weight, size_t cap_weight, int64_t* mode, size_t cap_mode,
                             size_t* counts) {
    if (!buf || !counts) return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    Sink sinks[4] = {{positive, nullptr, cap_positive, 0}, {negative, nullptr, cap_negative, 0},
                     {nullptr, weight, cap_weight, 0}, {mode, nullptr, cap_mode, 0}};
    static const char* const names[4] = {"positive_sample", "negative_sample", "subsampling_weight", "mode"};
    Cur ex{buf, buf + len, true};
    while (ex.ok && ex.p < ex.end) {
        const uint64_t tag = rd_varint(ex);
        if ((tag >> 3) != 1 || (tag & 7) != 2) { skip_field(ex, (uint32_t)(tag & 7)); continue; }
        Cur feats = rd_len(ex);  // Features
        while (feats.ok && feats.p < feats.end) {
            const uint64_t t1 = rd_varint(feats);
            if ((t1 >> 3) != 1 || (t1 & 7) != 2) { skip_field(feats, (uint32_t)(t1 & 7)); continue; }
            Cur entry = rd_len(feats);  // map entry {1: key, 2: Feature}
            const uint8_t* key = nullptr;
            size_t key_len = 0;
            Cur value{nullptr, nullptr, false};
            while (entry.ok && entry.p < entry.end) {
                const uint64_t t2 = rd_varint(entry);
                if ((t2 & 7) != 2) { skip_field(entry, (uint32_t)(t2 & 7)); continue; }
                Cur payload = rd_len(entry);
                if ((t2 >> 3) == 1) { key = payload.p; key_len = (size_t)(payload.end - payload.p); }
                else if ((t2 >> 3) == 2) value = payload;
            }
            if (!entry.ok) return fail(KGE_ERR_INVALID_ARGUMENT, "malformed feature map entry");
            for (int i = 0; i < 4; ++i)
                if (key && key_len == strlen(names[i]) && memcmp(key, names[i], key_len) == 0 && value.ok) {
                    if (!parse_feature(value, sinks[i]))
                        return fail(KGE_ERR_INVALID_ARGUMENT, "feature %s: malformed or larger than the buffer", names[i]);
                }
        }
        if (!feats.ok) return fail(KGE_ERR_INVALID_ARGUMENT, "malformed Features message");
    }
    if (!ex.ok) return fail(KGE_ERR_INVALID_ARGUMENT, "malformed Example message");
    for (int i = 0; i < 4; ++i) counts[i] = sinks[i].n;
    return KGE_OK;
}

// The four features -> serialized tf.train.Example (create_example, utils.py:35-42).  out == NULL: size query.
size_t kge_record_build_example(const int64_t* positive, size_t n_positive, const int64_t* negative, size_t n_negative,
                                const float* weight, size_t n_weight, const int64_t* mode, size_t n_mode, uint8_t* out,
                                size_t cap) {
    // sorted key order: mode, negative_sample, positive_sample, subsampling_weight
    struct F { const char* name; const int64_t* i; const float* f; size_t n; };
    const F fs[4] = {{"mode", mode, nullptr, n_mode}, {"negative_sample", negative, nullptr, n_negative},
                     {"positive_sample", positive, nullptr, n_positive}, {"subsampling_weight", nullptr, weight, n_weight}};
    size_t packed[4], list[4], feature[4], entry[4], features_len = 0;
    for (int k = 0; k < 4; ++k) {
        packed[k] = 0;
        if (fs[k].i) for (size_t j = 0; j < fs[k].n; ++j) packed[k] += varint_size((uint64_t)fs[k].i[j]);
        else packed[k] = 4 * fs[k].n;
        list[k] = fs[k].n ? 1 + varint_size(packed[k]) + packed[k] : 0;  // empty repeated field: nothing emitted
        feature[k] = 1 + varint_size(list[k]) + list[k];
        const size_t klen = strlen(fs[k].name);
        entry[k] = 1 + varint_size(klen) + klen + 1 + varint_size(feature[k]) + feature[k];
        features_len += 1 + varint_size(entry[k]) + entry[k];
    }
    const size_t total = 1 + varint_size(features_len) + features_len;
    if (!out) return total;
    if (cap < total) return 0;
    uint8_t* o = out;
    *o++ = 0x0A;  // Example.features
    o = wr_varint(o, features_len);
    for (int k = 0; k < 4; ++k) {
        *o++ = 0x0A;  // Features.feature map entry
        o = wr_varint(o, entry[k]);
        const size_t klen = strlen(fs[k].name);
        *o++ = 0x0A;  // key
        o = wr_varint(o, klen);
        memcpy(o, fs[k].name, klen);
        o += klen;
        *o++ = 0x12;  // value: Feature
        o = wr_varint(o, feature[k]);
        *o++ = fs[k].i ? 0x1A : 0x12;  // Feature.int64_list (3) / float_list (2)
        o = wr_varint(o, list[k]);
        if (fs[k].n) {
            *o++ = 0x0A;  // packed value
            o = wr_varint(o, packed[k]);
            if (fs[k].i) for (size_t j = 0; j < fs[k].n; ++j) o = wr_varint(o, (uint64_t)fs[k].i[j]);
            else { memcpy(o, fs[k].f, 4 * fs[k].n); o += 4 * fs[k].n; }
        }
    }
    return (size_t)(o - out);
}

}  // extern "C"
