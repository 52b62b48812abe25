// parquet.cu -- Parquet decode on the GPU into K0's typed columns (fixture
// notebooks/featurestore/aws/s3/data/telco-delta/*.snappy.parquet, written by Spark 2.4 / parquet-mr 1.10).
//
// Replaces, for flat schemas, the Parquet scan behind `spark.read.format("delta" | "parquet").load(...)` that feeds the
// same groupBy / join / statistics operators as the CSV scan of ComputeFeatures.scala:104.  Host: the footer and the page
// headers (Thrift compact protocol) are walked once -- a few dozen bytes per page -- and only the byte ranges of the asked
// column's chunks are copied to the device; a chunk's dictionary page (one per chunk, a single sequential Snappy stream of up
// to a megabyte) is unpacked by the host while the device works on the data pages, and uploaded before the last kernel.  Device, one warp per page (pages are independent: the grid is the page list):
//   k_pq_snappy   Snappy block decompression: every lane reads the tag (a broadcast load), literals and back references are
//                 moved by all 32 lanes (long literals as aligned words funnel-shifted out of the input)
//   k_pq_decode   A. definition levels (RLE / bit-packed hybrid, bit width 1: OPTIONAL columns) -> a byte per row;
//                 B. dictionary indices (hybrid, PLAIN_DICTIONARY / RLE_DICTIONARY) -> a word per non-null value -- lane 0
//                 reads the run headers, the values of a run are produced by all lanes
//   k_pq_place    (eight warps per page) 256 rows at a time: ballot of the levels -> validity words, prefix popcount -> which
//                 packed value belongs to the row -> PLAIN value or dictionary entry stored in the row's slot (INT32, INT64,
//                 FLOAT, DOUBLE); the dependent loads (index -> dictionary entry) of eight warps overlap.
//                 INT96 timestamps become int64 microseconds since the epoch, PLAIN BOOLEAN bits become int32 0 / 1.
// Declined (FSX_E_UNSUPPORTED; read those with the host's Parquet reader): nested schemas, PLAIN-stored BYTE_ARRAY and
// FIXED_LEN_BYTE_ARRAY columns, data page v2, codecs other than UNCOMPRESSED and SNAPPY, DELTA_* and RLE value encodings.
#include <string>
#include <unordered_map>
#include <vector>

#include "common.cuh"

namespace {

// ---- Thrift compact protocol (just enough to walk Parquet's metadata) ------------------------------------------------
struct TReader {
    const uint8_t *p, *end;
    bool ok = true;
    uint8_t byte() { if (p >= end) { ok = false; return 0; } return *p++; }
    uint64_t varint() {
        uint64_t v = 0;
        for (int shift = 0; shift < 64; shift += 7) {
            const uint8_t b = byte();
            v |= (uint64_t)(b & 0x7F) << shift;
            if (!(b & 0x80)) break;
        }
        return v;
    }
    int64_t zigzag() { const uint64_t v = varint(); return (int64_t)(v >> 1) ^ -(int64_t)(v & 1); }
    std::string binary() {
        const uint64_t n = varint();
        if (!ok || n > (uint64_t)(end - p)) { ok = false; return std::string(); }
        std::string s((const char *)p, (size_t)n);
        p += n;
        return s;
    }
    // field header: returns the wire type (0 = STOP) and updates the running field id
    int field(int16_t &id) {
        const uint8_t b = byte();
        if (b == 0) return 0;
        const int delta = b >> 4, type = b & 0x0F;
        if (delta) id = (int16_t)(id + delta); else id = (int16_t)zigzag();
        return type;
    }
    void list_header(int &size, int &etype) {
        const uint8_t b = byte();
        size = b >> 4; etype = b & 0x0F;
        if (size == 15) size = (int)varint();
    }
    int depth = 0;                                               // containers inside containers being skipped
    void skip(int type) {
        if (type >= 9 && ++depth > 64) { ok = false; return; }    // a byte per nesting level would otherwise buy a stack frame each
        skip_value(type);
        if (type >= 9) depth--;
    }
    void skip_value(int type) {
        switch (type) {
        case 1: case 2: break;                                   // bool encoded in the field header
        case 3: byte(); break;
        case 4: case 5: case 6: zigzag(); break;
        case 7: for (int i = 0; i < 8; i++) byte(); break;
        case 8: binary(); break;
        case 9: case 10: { int n, et; list_header(n, et); for (int i = 0; i < n && ok; i++) skip_elem(et); break; }
        case 11: { const int n = (int)varint(); if (n) { const uint8_t kv = byte(); for (int i = 0; i < n && ok; i++) { skip_elem(kv >> 4); skip_elem(kv & 0x0F); } } break; }
        case 12: { int16_t id = 0; for (int t; ok && (t = field(id)) != 0;) skip(t); break; }
        default: ok = false;
        }
    }
    void skip_elem(int et) { if (et == 1 || et == 2) byte(); else skip(et); }   // booleans inside containers take a byte
};

enum { PQ_BOOLEAN = 0, PQ_INT32 = 1, PQ_INT64 = 2, PQ_INT96 = 3, PQ_FLOAT = 4, PQ_DOUBLE = 5, PQ_BYTE_ARRAY = 6, PQ_FLBA = 7 };
enum { ENC_PLAIN = 0, ENC_PLAIN_DICT = 2, ENC_RLE = 3, ENC_BIT_PACKED = 4, ENC_RLE_DICT = 8 };
enum { PAGE_DATA = 0, PAGE_INDEX = 1, PAGE_DICT = 2, PAGE_DATA_V2 = 3 };

struct PqColumnSchema { std::string name; int type = -1; int repetition = 0; };
struct PqChunk { int type = -1, codec = 0; int64_t num_values = 0, total_compressed = 0, data_page_offset = 0, dict_page_offset = 0; std::vector<std::string> path; };
struct PqRowGroup { int64_t num_rows = 0; std::vector<PqChunk> chunks; };
struct PqFile { std::vector<PqColumnSchema> cols; std::vector<PqRowGroup> groups; int64_t num_rows = 0; bool nested = false; };

bool parse_chunk_meta(TReader &r, PqChunk &c) {
    int16_t id = 0;
    for (int t; r.ok && (t = r.field(id)) != 0;) {
        switch (id) {
        case 1: c.type = (int)r.zigzag(); break;
        case 3: { int n, et; r.list_header(n, et); for (int i = 0; i < n && r.ok; i++) c.path.push_back(r.binary()); break; }
        case 4: c.codec = (int)r.zigzag(); break;
        case 5: c.num_values = r.zigzag(); break;
        case 7: c.total_compressed = r.zigzag(); break;
        case 9: c.data_page_offset = r.zigzag(); break;
        case 11: c.dict_page_offset = r.zigzag(); break;
        default: r.skip(t);
        }
    }
    return r.ok;
}

bool parse_footer(const uint8_t *file, int64_t n, PqFile &f, std::string &err) {
    if (n < 12 || memcmp(file, "PAR1", 4) != 0 || memcmp(file + n - 4, "PAR1", 4) != 0) { err = "not a Parquet file (magic bytes)"; return false; }
    uint32_t flen;
    memcpy(&flen, file + n - 8, 4);
    if ((int64_t)flen + 12 > n) { err = "footer length beyond the file"; return false; }
    TReader r{file + n - 8 - flen, file + n - 8};
    int16_t id = 0;
    std::vector<int> children;
    for (int t; r.ok && (t = r.field(id)) != 0;) {
        if (id == 2 && t == 9) {                                 // schema: the root first, then the leaves of a flat schema
            int cnt, et;
            r.list_header(cnt, et);
            for (int i = 0; i < cnt && r.ok; i++) {
                PqColumnSchema s;
                int nchildren = 0;
                int16_t sid = 0;
                for (int st; r.ok && (st = r.field(sid)) != 0;) {
                    switch (sid) {
                    case 1: s.type = (int)r.zigzag(); break;
                    case 3: s.repetition = (int)r.zigzag(); break;
                    case 4: s.name = r.binary(); break;
                    case 5: nchildren = (int)r.zigzag(); break;
                    default: r.skip(st);
                    }
                }
                if (i == 0) continue;                            // the root group
                if (nchildren > 0 || s.repetition == 2) f.nested = true;
                f.cols.push_back(s);
            }
        } else if (id == 3) f.num_rows = r.zigzag();
        else if (id == 4 && t == 9) {
            int cnt, et;
            r.list_header(cnt, et);
            for (int i = 0; i < cnt && r.ok; i++) {
                PqRowGroup g;
                int16_t gid = 0;
                for (int gt; r.ok && (gt = r.field(gid)) != 0;) {
                    if (gid == 1 && gt == 9) {
                        int nc, cet;
                        r.list_header(nc, cet);
                        for (int k = 0; k < nc && r.ok; k++) {
                            PqChunk c;
                            int16_t cid = 0;
                            for (int ct; r.ok && (ct = r.field(cid)) != 0;) {
                                if (cid == 3 && ct == 12) parse_chunk_meta(r, c);
                                else r.skip(ct);
                            }
                            g.chunks.push_back(c);
                        }
                    } else if (gid == 3) g.num_rows = r.zigzag();
                    else r.skip(gt);
                }
                f.groups.push_back(g);
            }
        } else r.skip(t);
    }
    if (!r.ok) { err = "malformed footer"; return false; }
    return true;
}

struct PqPage { int type = 0, encoding = 0, def_encoding = ENC_RLE; int32_t num_values = 0, uncompressed = 0, compressed = 0; int64_t payload = 0; };

bool parse_page_header(TReader &r, PqPage &pg) {
    int16_t id = 0;
    for (int t; r.ok && (t = r.field(id)) != 0;) {
        if (id == 1) pg.type = (int)r.zigzag();
        else if (id == 2) pg.uncompressed = (int32_t)r.zigzag();
        else if (id == 3) pg.compressed = (int32_t)r.zigzag();
        else if ((id == 5 || id == 7) && t == 12) {
            int16_t hid = 0;
            for (int ht; r.ok && (ht = r.field(hid)) != 0;) {
                if (hid == 1) pg.num_values = (int32_t)r.zigzag();
                else if (hid == 2) pg.encoding = (int)r.zigzag();
                else if (hid == 3 && id == 5) pg.def_encoding = (int)r.zigzag();
                else r.skip(ht);
            }
        } else r.skip(t);
    }
    return r.ok;
}

// ---- device ------------------------------------------------------------------------------------------------------------
struct PqPageDev {
    long long src;            // offset of the (compressed) payload in the device image of the column's chunks
    long long dst;            // offset of the uncompressed payload in the page buffer
    int compressed, uncompressed;
    int codec;
    int type, encoding, num_values;
    long long row0;           // first row of this page in the output column
    long long dict;           // page-buffer offset of the chunk's dictionary values (-1: none)
    int dict_count;
};

__device__ __forceinline__ unsigned pq_varint(const uint8_t *&p, const uint8_t *end) {
    unsigned v = 0;
    for (int shift = 0; shift < 35 && p < end; shift += 7) {
        const uint8_t b = *p++;
        v |= (unsigned)(b & 0x7F) << shift;
        if (!(b & 0x80)) break;
    }
    return v;
}

constexpr int kLanes = 32;
constexpr int kThreads = 8 * kLanes;     // k_pq_place: eight warps per page

// one warp per page.  Every lane reads the tag bytes itself (one broadcast load, no shuffle, no divergence); the bytes of a
// literal or a back reference are moved by all lanes.  Byte i of a back reference of offset `off` is byte (i mod off) of the
// `off` bytes before it, so a reference that overlaps its own output needs no ordering between the lanes.
// (Measured and dropped: a shared-memory window of the last 4-64 KB of output for the back references to read from, which
// costs resident pages per SM and gains nothing; and a 16-byte register window of the input so that a tag needs no load,
// whose shifts cost more than the L1 hit they replace (20.0 against 17.0 ms on the 64 M-row id column); and requesting the
// next element's tag bytes before the current element's bytes are moved (17.6 ms: no gain).  What a warp waits for is the
// previous element's store becoming visible to the next element's source load -- in such data every back reference
// reads the bytes written just before it.)
__global__ void __launch_bounds__(kLanes) k_pq_snappy(const uint8_t *__restrict__ file, uint8_t *__restrict__ buf, const PqPageDev *__restrict__ pages,
                                                      unsigned *__restrict__ bad) {
    const PqPageDev pg = pages[blockIdx.x];
    const int lane = threadIdx.x;
    const uint8_t *src = file + pg.src, *send = src + pg.compressed;
    uint8_t *dst = buf + pg.dst;
    if (pg.codec == 0) {                                          // UNCOMPRESSED: a plain copy
        for (int i = lane; i < pg.compressed; i += kLanes) dst[i] = src[i];
        return;
    }
    const uint8_t *p = src;
    const unsigned total = pq_varint(p, send);
    if (total != (unsigned)pg.uncompressed) { if (lane == 0) *bad = 1u; return; }
    unsigned out = 0;
    while (p < send && out < total) {
        // the image is padded: the four bytes after a tag are always readable (they may belong to the next element)
        const unsigned tag = p[0], b1 = p[1], b2 = p[2], b3 = p[3], b4 = p[4];
        const unsigned kind = tag & 3u;
        unsigned len, off = 0, adv;
        if (kind == 0) {                                          // literal
            len = tag >> 2;
            adv = 1;
            if (len >= 60) {
                const unsigned nb = len - 59;
                const unsigned w = b1 | (b2 << 8) | (b3 << 16) | (b4 << 24);
                len = nb == 4 ? w : (w & ((1u << (8 * nb)) - 1u));
                adv += nb;
            }
            len += 1;
            if (len == 0 || len > total - out || (size_t)(send - p) < (size_t)adv + len) { if (lane == 0) *bad = 1u; return; }
            p += adv;
            if (len < 64) {                                       // short literal: a byte per lane
                for (unsigned i = lane; i < len; i += kLanes) dst[out + i] = p[i];
            } else {                                              // long literal: aligned words of the output, each funnel-shifted out of two aligned words of the input
                const unsigned head = (4u - (out & 3u)) & 3u;
                for (unsigned i = lane; i < head; i += kLanes) dst[out + i] = p[i];
                const uint8_t *s0 = p + head;
                const unsigned sa = (unsigned)((size_t)s0 & 3u), words = (len - head) >> 2, o0 = out + head;
                const unsigned *sw = (const unsigned *)(s0 - sa);
                for (unsigned j = lane; j < words; j += kLanes) *(unsigned *)(dst + o0 + 4u * j) = __funnelshift_r(sw[j], sw[j + 1], sa * 8u);
                for (unsigned i = head + 4u * words + lane; i < len; i += kLanes) dst[out + i] = p[i];
            }
            p += len;
        } else {
            if (kind == 1) { len = ((tag >> 2) & 7u) + 4; off = ((tag >> 5) << 8) | b1; adv = 2; }
            else if (kind == 2) { len = (tag >> 2) + 1; off = b1 | (b2 << 8); adv = 3; }
            else { len = (tag >> 2) + 1; off = b1 | (b2 << 8) | (b3 << 16) | (b4 << 24); adv = 5; }
            if (off == 0 || off > out || len > total - out || (size_t)(send - p) < adv) { if (lane == 0) *bad = 1u; return; }
            p += adv;
            const unsigned from = out - off;                      // every source byte lies before `out`: reads never see this element's writes
            if (off >= len) { for (unsigned i = lane; i < len; i += kLanes) dst[out + i] = dst[from + i]; }
            else { for (unsigned i = lane; i < len; i += kLanes) dst[out + i] = dst[from + i % off]; }
        }
        __syncwarp();                                             // the next element may read what this one wrote
        out += len;
    }
    if (out != total && lane == 0) *bad = 1u;
}

// RLE / bit-packed hybrid stream -> `count` values in dst.  Lane 0 reads each run header and broadcasts it; the values of a
// run are produced by all lanes (a bit-packed value i sits at bit i * width of the run's bytes).  Uniform return value.
template <typename T>
__device__ bool pq_hybrid_expand(const uint8_t *p, const uint8_t *end, int width, int count, T *__restrict__ dst, int lane) {
    const unsigned mask = width >= 32 ? 0xFFFFFFFFu : ((1u << width) - 1u);
    if (width == 0) {                                             // a one-entry dictionary: every index is 0
        for (int i = lane; i < count; i += kLanes) dst[i] = (T)0;
        return true;
    }
    int pos = 0;
    while (pos < count) {
        unsigned h = 0, v = 0, adv = 0;
        if (lane == 0) {
            const uint8_t *q = p;
            if (q >= end) h = 0xFFFFFFFFu;
            else {
                h = pq_varint(q, end);
                if (!(h & 1u)) { const int nb = (width + 7) / 8; for (int k = 0; k < nb && q < end; k++) v |= (unsigned)(*q++) << (8 * k); }
                adv = (unsigned)(q - p);
            }
        }
        h = __shfl_sync(0xffffffffu, h, 0);
        v = __shfl_sync(0xffffffffu, v, 0);
        adv = __shfl_sync(0xffffffffu, adv, 0);
        if (h == 0xFFFFFFFFu || (h >> 1) == 0) return false;     // stream ends early / empty run
        p += adv;
        if (h & 1u) {                                             // bit-packed: (h >> 1) groups of 8 values
            const long long run = (long long)(h >> 1) * 8;
            const int take = run < (long long)(count - pos) ? (int)run : count - pos;
#pragma unroll 4
            for (int i = lane; i < take; i += kLanes) {
                const unsigned long long bit = (unsigned long long)i * (unsigned)width;
                const uint8_t *q = p + (bit >> 3);
                unsigned long long w = 0;
                for (int k = 0; k < 5; k++) if (q + k < end) w |= (unsigned long long)q[k] << (8 * k);
                dst[pos + i] = (T)((unsigned)(w >> (bit & 7)) & mask);
            }
            p += ((unsigned long long)run * (unsigned)width + 7) >> 3;
            pos += take;
        } else {                                                  // (h >> 1) copies of v
            const unsigned run = h >> 1;
            const int take = run < (unsigned)(count - pos) ? (int)run : count - pos;
            for (int i = lane; i < take; i += kLanes) dst[pos + i] = (T)v;
            pos += take;
        }
    }
    return true;
}

// one warp per DATA page:  A. definition levels -> one byte per row;  B. dictionary indices -> one word per non-null value
// (k_pq_place then puts the values into the rows)
__global__ void __launch_bounds__(kLanes) k_pq_decode(const uint8_t *__restrict__ buf, const PqPageDev *__restrict__ pages, int optional, int elem,
                                                      uint8_t *__restrict__ level_bytes, unsigned *__restrict__ index_words, unsigned *__restrict__ bad) {
    const PqPageDev pg = pages[blockIdx.x];
    if (pg.type != PAGE_DATA) return;
    const int lane = threadIdx.x;
    const uint8_t *p = buf + pg.dst, *pend = p + pg.uncompressed;
    const int n = pg.num_values;
    uint8_t *lvl = level_bytes + pg.row0;
    unsigned *idx = index_words + pg.row0;
    int nn = n;                                                   // non-null values of the page
    if (optional) {
        if (p + 4 > pend) { if (lane == 0) *bad = 1u; return; }
        const unsigned len = (unsigned)p[0] | ((unsigned)p[1] << 8) | ((unsigned)p[2] << 16) | ((unsigned)p[3] << 24);
        const uint8_t *levels = p + 4;
        if (len > (unsigned)(pend - levels)) { if (lane == 0) *bad = 1u; return; }
        if (!pq_hybrid_expand<uint8_t>(levels, levels + len, 1, n, lvl, lane)) { if (lane == 0) *bad = 1u; return; }
        p = levels + len;
        __syncwarp();
        int mine = 0;
        for (int i = lane; i < n; i += kLanes) mine += lvl[i] != 0;
        nn = (int)__reduce_add_sync(0xffffffffu, (unsigned)mine);
    }
    const bool dict = pg.encoding == ENC_PLAIN_DICT || pg.encoding == ENC_RLE_DICT;
    if (dict) {
        if (p >= pend || pg.dict < 0) { if (lane == 0) *bad = 1u; return; }
        const int width = *p;
        if (width > 32 || !pq_hybrid_expand<unsigned>(p + 1, pend, width, nn, idx, lane)) { if (lane == 0) *bad = 1u; return; }
        __syncwarp();
    } else if (pg.encoding != ENC_PLAIN) { if (lane == 0) *bad = 2u; return; }
    else if (p + (elem == 1 ? ((size_t)nn + 7) / 8 : (size_t)nn * elem) > pend) { if (lane == 0) *bad = 1u; return; }
}

// kThreads threads per DATA page, a tile of kThreads rows at a time: ballot of the levels -> validity words; the rank of a
// non-null row among the page's values = values before the tile + those of the lower warps + those of the lower lanes;
// the row's value is the rank-th PLAIN value or the dictionary entry the rank-th index names.
__global__ void __launch_bounds__(kThreads) k_pq_place(const uint8_t *__restrict__ buf, const PqPageDev *__restrict__ pages, int optional, int elem,
                                                       const uint8_t *__restrict__ level_bytes, const unsigned *__restrict__ index_words,
                                                       uint8_t *__restrict__ out, unsigned *__restrict__ valid_words, unsigned *__restrict__ bad) {
    constexpr int kWarps = kThreads / kLanes;
    __shared__ int warp_total[2][kWarps];
    const PqPageDev pg = pages[blockIdx.x];
    if (pg.type != PAGE_DATA) return;
    const int tid = threadIdx.x, lane = tid % kLanes, warp = tid / kLanes;
    const uint8_t *p = buf + pg.dst, *pend = p + pg.uncompressed;
    const int n = pg.num_values;
    if (optional) {                                               // k_pq_decode has checked these bounds
        if (p + 4 > pend) return;
        const unsigned len = (unsigned)p[0] | ((unsigned)p[1] << 8) | ((unsigned)p[2] << 16) | ((unsigned)p[3] << 24);
        if (len > (unsigned)(pend - p - 4)) return;
        p += 4 + len;
    }
    const bool dict = pg.encoding == ENC_PLAIN_DICT || pg.encoding == ENC_RLE_DICT;
    if (!dict && pg.encoding != ENC_PLAIN) return;
    if (dict && pg.dict < 0) return;
    const uint8_t *lvl = level_bytes + pg.row0;
    const unsigned *idx = index_words + pg.row0;
    const uint8_t *dict_values = buf + (pg.dict < 0 ? 0 : pg.dict);
    int running = 0, parity = 0;
    for (int base = 0; base < n; base += kThreads, parity ^= 1) {
        const int i = base + tid;
        const bool valid = i < n && (!optional || lvl[i] != 0);
        const unsigned ballot = __ballot_sync(0xffffffffu, valid);
        if (lane == 0) {
            warp_total[parity][warp] = __popc(ballot);
            if (ballot) {                                         // the warp's 32 rows may straddle two validity words
                const long long row = pg.row0 + base + warp * kLanes;
                const int shift = (int)(row & 31);
                atomicOr(&valid_words[row >> 5], ballot << shift);
                if (shift && (ballot >> (32 - shift))) atomicOr(&valid_words[(row >> 5) + 1], ballot >> (32 - shift));
            }
        }
        __syncthreads();
        int before = running, tile = 0;
#pragma unroll
        for (int w = 0; w < kWarps; w++) {
            const int t = warp_total[parity][w];
            if (w < warp) before += t;
            tile += t;
        }
        running += tile;
        if (!valid) continue;
        const int rank = before + __popc(ballot & ((1u << lane) - 1u));
        const long long row = pg.row0 + i;
        if (elem == 1) {                                          // BOOLEAN, PLAIN: bit `rank` of the page's value bytes
            const uint8_t *src = p + (rank >> 3);
            if (dict || src >= pend) { *bad = 1u; continue; }
            ((unsigned *)out)[row] = (*src >> (rank & 7)) & 1u;
            continue;
        }
        const uint8_t *src;
        if (dict) {
            const unsigned d = idx[rank];
            if (d >= (unsigned)pg.dict_count) { *bad = 1u; continue; }
            src = dict_values + (size_t)d * elem;
        } else {
            src = p + (size_t)rank * elem;
            if (src + elem > pend) { *bad = 1u; continue; }
        }
        unsigned long long v = 0;
        for (int k = 0; k < 8 && k < elem; k++) v |= (unsigned long long)src[k] << (8 * k);
        if (elem == 12) {                                         // INT96: v = nanoseconds of the day, then the Julian day
            const long long day = (long long)(int)((unsigned)src[8] | ((unsigned)src[9] << 8) | ((unsigned)src[10] << 16) | ((unsigned)src[11] << 24));
            v = (unsigned long long)(day - 2440588LL) * 86400000000ULL + (unsigned long long)((long long)v / 1000LL);   // wraps like Java's long for absurd days
        }
        if (elem >= 8) ((unsigned long long *)out)[row] = v; else ((unsigned *)out)[row] = (unsigned)v;
    }
}

// bytes per stored value: 4 / 8 for the numeric types, 12 for INT96 (Spark 2.4's timestamps: nanoseconds of the day + Julian day,
// decoded to int64 microseconds since the epoch as Spark's TimestampType holds them), 1 standing for BOOLEAN (one BIT per
// value in a PLAIN page, decoded to int32 0 / 1); 0 = not decoded here
int pq_elem(int type) {
    return type == PQ_INT32 || type == PQ_FLOAT ? 4 : (type == PQ_INT64 || type == PQ_DOUBLE ? 8 : type == PQ_INT96 ? 12 : type == PQ_BOOLEAN ? 1 : 0);
}
int pq_out_elem(int elem) { return elem == 12 ? 8 : elem == 1 ? 4 : elem; }
int32_t pq_dtype(int type) {
    return type == PQ_INT32 || type == PQ_BOOLEAN ? FSX_I32 : type == PQ_INT64 || type == PQ_INT96 ? FSX_I64 : type == PQ_FLOAT ? FSX_F32 : FSX_F64;
}
const char *pq_type_name(int t) {
    static const char *names[] = {"boolean", "int32", "int64", "int96", "float", "double", "byte_array", "fixed_len_byte_array"};
    return t >= 0 && t < 8 ? names[t] : "?";
}

// Walks the page headers of column `ci` in every row group: one PqPageDev per dictionary / data page, laid out back to back
// (16-byte aligned) in the page buffer.  Returns FSX_OK or the status to fail with (message in err).
struct PqRange { const uint8_t *file; long long file_pos, len, image_pos; };   // a column chunk: where it is in its file, where in the device image

// What the device needs for one column of one or several files: the page list (rows numbered through all files in order),
// the chunks to copy, and the sizes of the device image / page buffer.
struct PqHostDict { const uint8_t *src; int compressed, uncompressed, codec; long long buf_pos; };   // a dictionary page the host unpacks
struct PqPlan {
    std::vector<PqPageDev> pages;
    std::vector<PqRange> ranges;
    std::vector<PqHostDict> host_dicts;
    long long image_len = 0, bufpos = 0, nrows = 0;
};

// Snappy block format on the host.  Dictionary pages come through here: one per column chunk, at most about a megabyte, and
// a single sequential stream -- on the device it would be one warp working alone long after the data pages are done.
bool snappy_decode_host(const uint8_t *p, size_t n, size_t expect, std::vector<uint8_t> &out) {
    const uint8_t *end = p + n;
    size_t total = 0;
    for (int shift = 0; p < end; shift += 7) { const uint8_t b = *p++; total |= (size_t)(b & 0x7F) << shift; if (!(b & 0x80)) break; if (shift > 28) return false; }
    if (total != expect) return false;
    out.resize(total);
    uint8_t *o = out.data();
    size_t at = 0;
    while (p < end && at < total) {
        const unsigned tag = *p++, kind = tag & 3u;
        size_t len, off = 0;
        if (kind == 0) {
            len = tag >> 2;
            if (len >= 60) { const unsigned nb = (unsigned)len - 59; if ((size_t)(end - p) < nb) return false; len = 0; for (unsigned k = 0; k < nb; k++) len |= (size_t)(*p++) << (8 * k); }
            len += 1;
            if ((size_t)(end - p) < len || len > total - at) return false;
            memcpy(o + at, p, len);
            p += len;
            at += len;
            continue;
        }
        if (kind == 1) { if (end - p < 1) return false; len = ((tag >> 2) & 7u) + 4; off = ((size_t)(tag >> 5) << 8) | *p++; }
        else if (kind == 2) { if (end - p < 2) return false; len = (tag >> 2) + 1; off = (size_t)p[0] | ((size_t)p[1] << 8); p += 2; }
        else { if (end - p < 4) return false; len = (tag >> 2) + 1; off = (size_t)p[0] | ((size_t)p[1] << 8) | ((size_t)p[2] << 16) | ((size_t)p[3] << 24); p += 4; }
        if (off == 0 || off > at || len > total - at) return false;
        if (off >= len) memcpy(o + at, o + at - off, len);
        else for (size_t i = 0; i < len; i++) o[at + i] = o[at + i - off];        // overlapping: byte by byte
        at += len;
    }
    return at == total;
}

// String (BYTE_ARRAY) columns: the distinct strings of the file, merged over the chunks' dictionary pages in order of first
// appearance; each chunk's dictionary becomes a table of int32 codes that the device gathers from like any int32 dictionary.
struct PqStrings {
    std::vector<std::string> distinct;
    std::unordered_map<std::string, int32_t> code_of;
    struct Upload { long long buf_pos; std::vector<int32_t> codes; };
    std::vector<Upload> uploads;
};

int32_t walk_pages(const uint8_t *file, int64_t nbytes, const PqFile &f, int ci, bool optional, int elem, PqPlan &plan, std::string &err,
                   PqStrings *strings = nullptr) {
    std::vector<PqPageDev> &pages = plan.pages;
    std::vector<PqRange> &ranges = plan.ranges;
    long long &image_len = plan.image_len, &bufpos = plan.bufpos;
    char msg[256];
#define PQ_REQUIRE(cond, code, ...) do { if (!(cond)) { snprintf(msg, sizeof msg, __VA_ARGS__); err = msg; return (code); } } while (0)
    long long row0 = plan.nrows;                                  // a second file continues where the first one ended
    for (const PqRowGroup &g : f.groups) {
        PQ_REQUIRE((size_t)ci < g.chunks.size(), FSX_E_INVALID, "row group without column %d", ci);
        const PqChunk &c = g.chunks[(size_t)ci];
        PQ_REQUIRE(c.codec == 0 || c.codec == 1, FSX_E_UNSUPPORTED, "compression codec %d (UNCOMPRESSED and SNAPPY are decoded)", c.codec);
        int64_t pos = c.dict_page_offset > 0 && c.dict_page_offset < c.data_page_offset ? c.dict_page_offset : c.data_page_offset;
        const int64_t chunk_end = pos + c.total_compressed;
        PQ_REQUIRE(pos > 0 && c.total_compressed >= 0 && chunk_end <= nbytes, FSX_E_INVALID, "column chunk outside the file");
        const PqRange range{file, pos, c.total_compressed, image_len};
        ranges.push_back(range);
        image_len += (c.total_compressed + 15) & ~15LL;
        long long dict = -1, values_seen = 0;
        int dict_count = 0;
        const long long group_row0 = row0;
        while (pos < chunk_end && values_seen < c.num_values) {
            TReader r{file + pos, file + chunk_end};
            PqPage pg;
            PQ_REQUIRE(parse_page_header(r, pg), FSX_E_INVALID, "malformed page header");
            pg.payload = r.p - file;
            PQ_REQUIRE(pg.compressed >= 0 && pg.uncompressed >= 0 && pg.payload + pg.compressed <= chunk_end, FSX_E_INVALID, "page beyond its chunk");
            // an UNCOMPRESSED page is copied as it is: its two sizes must agree, or the copy would leave the chunk / the page's slot
            PQ_REQUIRE(c.codec != 0 || pg.compressed == pg.uncompressed, FSX_E_INVALID, "uncompressed page with sizes %d / %d", pg.compressed, pg.uncompressed);
            PQ_REQUIRE(pg.type != PAGE_DATA_V2, FSX_E_UNSUPPORTED, "data page v2");
            if (pg.type == PAGE_DATA || pg.type == PAGE_DICT) {
                PqPageDev d;
                d.src = range.image_pos + (pg.payload - range.file_pos); d.dst = bufpos; d.compressed = pg.compressed; d.uncompressed = pg.uncompressed; d.codec = c.codec;
                d.type = pg.type; d.encoding = pg.encoding; d.num_values = pg.num_values; d.row0 = group_row0 + values_seen; d.dict = dict; d.dict_count = dict_count;
                if (pg.type == PAGE_DICT && strings) {             // decoded here; the device sees a table of codes in its place
                    PQ_REQUIRE(pg.encoding == ENC_PLAIN || pg.encoding == ENC_PLAIN_DICT, FSX_E_UNSUPPORTED, "dictionary page encoding %d", pg.encoding);
                    PQ_REQUIRE(pg.num_values >= 0, FSX_E_INVALID, "dictionary page with %d values", pg.num_values);
                    std::vector<uint8_t> plain;
                    const uint8_t *q = file + pg.payload, *qend = q + pg.compressed;
                    if (c.codec == 1) {
                        PQ_REQUIRE(snappy_decode_host(q, (size_t)pg.compressed, (size_t)pg.uncompressed, plain), FSX_E_DATA, "corrupt dictionary page (Snappy stream)");
                        q = plain.data(); qend = q + plain.size();
                    }
                    PqStrings::Upload up;
                    up.buf_pos = bufpos;
                    for (int k = 0; k < pg.num_values; k++) {       // PLAIN BYTE_ARRAY: 4-byte length + bytes
                        PQ_REQUIRE(qend - q >= 4, FSX_E_DATA, "dictionary page shorter than its %d values", pg.num_values);
                        uint32_t len;
                        memcpy(&len, q, 4);
                        q += 4;
                        PQ_REQUIRE((size_t)(qend - q) >= len, FSX_E_DATA, "dictionary entry beyond its page");
                        std::string v((const char *)q, len);
                        q += len;
                        auto it = strings->code_of.find(v);
                        if (it == strings->code_of.end()) {
                            it = strings->code_of.emplace(v, (int32_t)strings->distinct.size()).first;
                            strings->distinct.push_back(std::move(v));
                        }
                        up.codes.push_back(it->second);
                    }
                    dict = bufpos; dict_count = pg.num_values;
                    bufpos += ((long long)pg.num_values * 4 + 15) & ~15LL;
                    strings->uploads.push_back(std::move(up));
                    pos = pg.payload + pg.compressed;
                    continue;
                }
                if (pg.type == PAGE_DICT) {                        // unpacked by the host while the device works on the data pages
                    PQ_REQUIRE(pg.encoding == ENC_PLAIN || pg.encoding == ENC_PLAIN_DICT, FSX_E_UNSUPPORTED, "dictionary page encoding %d", pg.encoding);
                    PQ_REQUIRE(pg.num_values >= 0 && (long long)pg.num_values * elem <= pg.uncompressed, FSX_E_INVALID, "dictionary page shorter than its %d values", pg.num_values);
                    dict = bufpos; dict_count = pg.num_values;
                    plan.host_dicts.push_back(PqHostDict{file + pg.payload, pg.compressed, pg.uncompressed, c.codec, bufpos});
                    bufpos += ((long long)pg.uncompressed + 15) & ~15LL;
                    pos = pg.payload + pg.compressed;
                    continue;
                } else {
                    PQ_REQUIRE(!strings || pg.encoding == ENC_PLAIN_DICT || pg.encoding == ENC_RLE_DICT, FSX_E_UNSUPPORTED,
                               "string values stored PLAIN (the writer gave up on the dictionary): read this column on the host");
                    PQ_REQUIRE(pg.encoding == ENC_PLAIN || pg.encoding == ENC_PLAIN_DICT || pg.encoding == ENC_RLE_DICT, FSX_E_UNSUPPORTED,
                               "value encoding %d (PLAIN and dictionary encodings are decoded)", pg.encoding);
                    PQ_REQUIRE(!optional || pg.def_encoding == ENC_RLE, FSX_E_UNSUPPORTED, "definition level encoding %d", pg.def_encoding);
                    PQ_REQUIRE(pg.num_values >= 0 && values_seen + pg.num_values <= c.num_values, FSX_E_INVALID, "more page values than the chunk declares");
                    values_seen += pg.num_values;
                }
                pages.push_back(d);
                bufpos += ((long long)pg.uncompressed + 15) & ~15LL;
            }
            pos = pg.payload + pg.compressed;
        }
        PQ_REQUIRE(values_seen == c.num_values && values_seen == g.num_rows, FSX_E_INVALID, "pages hold %lld values, the row group %lld rows",
                   values_seen, (long long)g.num_rows);
        row0 += g.num_rows;
    }
    plan.nrows = row0;
    return FSX_OK;
#undef PQ_REQUIRE
}

// Footers of all files + the page walk of `column` in each: the plan, the column's physical type and whether it is OPTIONAL
// (the same in every file).  strings != nullptr asks for a BYTE_ARRAY column as dictionary codes.
int32_t plan_column(const void *const *files, const int64_t *nbytes, int32_t nfiles, const char *column, PqPlan &plan, int &type, int &optional,
                    PqStrings *strings, std::string &err) {
    char msg[256];
    type = -1; optional = 0;
    for (int32_t k = 0; k < nfiles; k++) {
        PqFile f;
        const uint8_t *file = (const uint8_t *)files[k];
        if (!file || !parse_footer(file, nbytes[k], f, err)) { if (err.empty()) err = "NULL file"; snprintf(msg, sizeof msg, "file %d: %s", k, err.c_str()); err = msg; return FSX_E_INVALID; }
        if (f.nested) { snprintf(msg, sizeof msg, "file %d: nested or repeated schema", k); err = msg; return FSX_E_UNSUPPORTED; }
        int ci = -1;
        for (size_t i = 0; i < f.cols.size(); i++) if (f.cols[i].name == column) ci = (int)i;
        if (ci < 0) { snprintf(msg, sizeof msg, "file %d has no column '%s'", k, column); err = msg; return FSX_E_INVALID; }
        const int t = f.cols[(size_t)ci].type, opt = f.cols[(size_t)ci].repetition == 1;
        if (k == 0) { type = t; optional = opt; }
        else if (t != type || opt != optional) { snprintf(msg, sizeof msg, "file %d stores '%s' as %s%s, file 0 as %s%s", k, column, pq_type_name(t), opt ? "?" : "", pq_type_name(type), optional ? "?" : ""); err = msg; return FSX_E_UNSUPPORTED; }
        if (strings) {
            if (t != PQ_BYTE_ARRAY) { snprintf(msg, sizeof msg, "column '%s' is %s, not byte_array", column, pq_type_name(t)); err = msg; return FSX_E_INVALID; }
        } else if (pq_elem(t) == 0) {
            snprintf(msg, sizeof msg, "column '%s' is %s; int32 / int64 / float / double / int96 / boolean columns are decoded as values, byte_array through fsx_parquet_decode_strings", column, pq_type_name(t)); err = msg; return FSX_E_UNSUPPORTED;
        }
        std::string werr;
        const int32_t rc = walk_pages(file, nbytes[k], f, ci, opt != 0, strings ? 4 : pq_elem(t), plan, werr, strings);
        if (rc != FSX_OK) { snprintf(msg, sizeof msg, "file %d, column '%s': %s", k, column, werr.c_str()); err = msg; return rc; }
    }
    return FSX_OK;
}

// ---- device driver (CUDA runtime calls from here on; tests/pq_emulation.py compiles everything above for the host) ----
// file image + page list -> one K0 column (three launches, whatever the page count)
int32_t decode_on_device(fsx_ctx *ctx, const PqPlan &plan, int optional, int elem, int32_t dtype, const PqStrings *strings, const char *column, fsx_col **out) {
    const std::vector<PqPageDev> &pages = plan.pages;
    const std::vector<PqRange> &ranges = plan.ranges;
    const long long image_len = plan.image_len, bufpos = plan.bufpos, nrows = plan.nrows;
    fsx_col *col = nullptr;
    FSX_CHECK(fsx_col_new(ctx, dtype, nrows, true, &col));
    std::vector<void *> tmp;
    auto cleanup = [&]() { for (void *p : tmp) fsx_dev_free(ctx, p); };
    auto fail = [&](int32_t rc) { cleanup(); fsx_col_release(ctx, col); return rc; };
    uint8_t *d_file = nullptr, *d_buf = nullptr, *d_levels = nullptr;
    unsigned *d_index = nullptr;
    PqPageDev *d_pages = nullptr;
    unsigned *d_bad = nullptr;
    int32_t rc = fsx_dev_alloc(ctx, (size_t)image_len + 16, (void **)&d_file);
    if (rc == FSX_OK) { tmp.push_back(d_file); rc = fsx_dev_alloc(ctx, (size_t)bufpos + 64, (void **)&d_buf); }
    if (rc == FSX_OK) { tmp.push_back(d_buf); rc = fsx_dev_alloc(ctx, sizeof(PqPageDev) * (pages.size() + 1), (void **)&d_pages); }
    if (rc == FSX_OK) { tmp.push_back(d_pages); rc = fsx_dev_alloc(ctx, (size_t)nrows + 64, (void **)&d_levels); }
    if (rc == FSX_OK) { tmp.push_back(d_levels); rc = fsx_dev_alloc(ctx, (size_t)nrows * 4 + 64, (void **)&d_index); }
    if (rc == FSX_OK) { tmp.push_back(d_index); rc = fsx_dev_alloc(ctx, 16, (void **)&d_bad); }
    if (rc != FSX_OK) return fail(rc);
    tmp.push_back(d_bad);
    cudaError_t e = cudaSuccess;
    std::vector<std::vector<uint8_t>> unpacked;                   // kept until the stream has drained
    bool host_bad = false;
    for (const PqRange &r : ranges) {
        if (e == cudaSuccess && r.len) e = cudaMemcpyAsync(d_file + r.image_pos, r.file + r.file_pos, (size_t)r.len, cudaMemcpyHostToDevice, ctx->stream);
        ctx->timing.h2d_bytes += r.len;
    }
    if (strings) for (const PqStrings::Upload &u : strings->uploads)
        if (e == cudaSuccess && !u.codes.empty()) e = cudaMemcpyAsync(d_buf + u.buf_pos, u.codes.data(), u.codes.size() * 4, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess && !pages.empty()) e = cudaMemcpyAsync(d_pages, pages.data(), sizeof(PqPageDev) * pages.size(), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(d_bad, 0, 16, ctx->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(col->validity, 0, (size_t)((nrows + 127) / 128) * 16, ctx->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(col->data, 0, (size_t)nrows * pq_out_elem(elem), ctx->stream);
    if (e == cudaSuccess && !pages.empty()) {
        k_pq_snappy<<<(unsigned)pages.size(), kLanes, 0, ctx->stream>>>(d_file, d_buf, d_pages, d_bad);
        k_pq_decode<<<(unsigned)pages.size(), kLanes, 0, ctx->stream>>>(d_buf, d_pages, optional, elem, d_levels, d_index, d_bad);
        e = cudaGetLastError();
        // the chunks' dictionary pages: unpacked here while the two kernels above run, needed only by the placement
        for (const PqHostDict &hd : plan.host_dicts) {
            if (e != cudaSuccess) break;
            const uint8_t *bytes = hd.src;
            if (hd.codec == 1) {
                unpacked.emplace_back();
                if (!snappy_decode_host(hd.src, (size_t)hd.compressed, (size_t)hd.uncompressed, unpacked.back())) { host_bad = true; break; }
                bytes = unpacked.back().data();
            }
            if (hd.uncompressed > 0) e = cudaMemcpyAsync(d_buf + hd.buf_pos, bytes, (size_t)hd.uncompressed, cudaMemcpyHostToDevice, ctx->stream);
            ctx->timing.h2d_bytes += hd.uncompressed;
        }
        if (e == cudaSuccess && !host_bad) k_pq_place<<<(unsigned)pages.size(), kThreads, 0, ctx->stream>>>(d_buf, d_pages, optional, elem, d_levels, d_index, (uint8_t *)col->data, (unsigned *)col->validity, d_bad);
        ctx->launches += 3;
        if (e == cudaSuccess) e = cudaGetLastError();
    }
    unsigned bad = 0;
    if (e == cudaSuccess) e = cudaMemcpyAsync(&bad, d_bad, 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) { fail(FSX_E_CUDA); FSX_CUDA(ctx, e); }
    if (host_bad) bad = 1;
    if (bad) return fail(fsx_fail(ctx, bad == 2 ? FSX_E_UNSUPPORTED : FSX_E_DATA, "fsx_parquet_decode: column '%s': %s", column,
                                  bad == 2 ? "an encoding the device decoder does not handle" : "corrupt page data (Snappy stream, levels or dictionary indices)"));
    cleanup();
    if (!optional) { fsx_dev_free(ctx, col->validity); col->validity = nullptr; }       // REQUIRED column: no nulls
    *out = col;
    return FSX_OK;
}

} // namespace

// "name:type?,..." (a trailing ? marks OPTIONAL) and the row count, from the footer alone.  Host code, no GPU needed.
extern "C" int32_t fsx_parquet_schema(const void *file_bytes, int64_t nbytes, char *out, size_t outlen, int64_t *nrows_out) {
    PqFile f;
    std::string err;
    if (!file_bytes || !parse_footer((const uint8_t *)file_bytes, nbytes, f, err)) return fsx_fail(nullptr, FSX_E_INVALID, "fsx_parquet_schema: %s", err.empty() ? "NULL file" : err.c_str());
    std::string s;
    for (size_t i = 0; i < f.cols.size(); i++) {
        if (i) s += ",";
        s += f.cols[i].name + ":" + pq_type_name(f.cols[i].type) + (f.cols[i].repetition == 1 ? "?" : "");
    }
    if (out && outlen) { strncpy(out, s.c_str(), outlen - 1); out[outlen - 1] = 0; }
    if (nrows_out) *nrows_out = f.num_rows;
    return FSX_OK;
}

extern "C" int32_t fsx_parquet_decode_column_files(fsx_ctx *ctx, const void *const *files, const int64_t *nbytes, int32_t nfiles, const char *column,
                                                   fsx_col **out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, files && nbytes && nfiles > 0 && column && out, FSX_E_INVALID, "fsx_parquet_decode_column: NULL argument or no file");
    *out = nullptr;
    PqPlan plan;
    std::string err;
    int type = -1, optional = 0;
    const int32_t prc = plan_column(files, nbytes, nfiles, column, plan, type, optional, nullptr, err);
    FSX_REQUIRE(ctx, prc == FSX_OK, prc, "fsx_parquet_decode_column: %s", err.c_str());
    fsx_col *col = nullptr;
    FSX_CHECK(decode_on_device(ctx, plan, optional, pq_elem(type), pq_dtype(type), nullptr, column, &col));
    *out = col;
    FSX_END();
}

extern "C" int32_t fsx_parquet_decode_column(fsx_ctx *ctx, const void *file_bytes, int64_t nbytes, const char *column, fsx_col **out) {
    return fsx_parquet_decode_column_files(ctx, &file_bytes, &nbytes, 1, column, out);
}

// A dictionary-encoded string column as int32 codes + its dictionary (see fsx.h).
extern "C" int32_t fsx_parquet_decode_strings_files(fsx_ctx *ctx, const void *const *files, const int64_t *nbytes, int32_t nfiles, const char *column,
                                                    fsx_col **codes_out, char *dict_bytes, int64_t dict_bytes_cap, int64_t *dict_offsets, int32_t dict_cap,
                                                    int32_t *ndistinct_out, int64_t *dict_bytes_needed) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, files && nbytes && nfiles > 0 && column && codes_out && ndistinct_out && dict_bytes_needed, FSX_E_INVALID,
                "fsx_parquet_decode_strings: NULL argument or no file");
    *codes_out = nullptr;
    PqPlan plan;
    PqStrings strings;
    std::string err;
    int type = -1, optional = 0;
    const int32_t prc = plan_column(files, nbytes, nfiles, column, plan, type, optional, &strings, err);
    FSX_REQUIRE(ctx, prc == FSX_OK, prc, "fsx_parquet_decode_strings: %s", err.c_str());
    int64_t bytes = 0;
    for (const std::string &v : strings.distinct) bytes += (int64_t)v.size();
    *ndistinct_out = (int32_t)strings.distinct.size();
    *dict_bytes_needed = bytes;
    FSX_REQUIRE(ctx, (int64_t)strings.distinct.size() < (int64_t)dict_cap + (strings.distinct.empty() ? 1 : 0) && bytes <= dict_bytes_cap &&
                         (strings.distinct.empty() || (dict_bytes && dict_offsets)), FSX_E_INVALID,
                "fsx_parquet_decode_strings: the dictionary of '%s' has %zu entries / %lld bytes; the buffers hold %d offsets / %lld bytes (sizes returned: call again)",
                column, strings.distinct.size(), (long long)bytes, dict_cap, (long long)dict_bytes_cap);
    fsx_col *col = nullptr;
    FSX_CHECK(decode_on_device(ctx, plan, optional, 4, FSX_I32, &strings, column, &col));
    int64_t at = 0;
    for (size_t i = 0; i < strings.distinct.size(); i++) {
        dict_offsets[i] = at;
        memcpy(dict_bytes + at, strings.distinct[i].data(), strings.distinct[i].size());
        at += (int64_t)strings.distinct[i].size();
    }
    if (!strings.distinct.empty()) dict_offsets[strings.distinct.size()] = at;
    *codes_out = col;
    FSX_END();
}

extern "C" int32_t fsx_parquet_decode_strings(fsx_ctx *ctx, const void *file_bytes, int64_t nbytes, const char *column, fsx_col **codes_out,
                                              char *dict_bytes, int64_t dict_bytes_cap, int64_t *dict_offsets, int32_t dict_cap,
                                              int32_t *ndistinct_out, int64_t *dict_bytes_needed) {
    return fsx_parquet_decode_strings_files(ctx, &file_bytes, &nbytes, 1, column, codes_out, dict_bytes, dict_bytes_cap, dict_offsets, dict_cap, ndistinct_out,
                                            dict_bytes_needed);
}
