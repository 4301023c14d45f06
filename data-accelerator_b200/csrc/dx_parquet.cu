// dx_parquet.cu — the on-disk form of state tables: Parquet files under <location>/A/ and <location>/B/ plus
// <location>/metadata.info (active=A|B, standby=B|A), as StateTableHandler keeps them (datax/handler/StateTableHandler.scala:26-52
// metadata, :70-83 `STORED AS PARQUET LOCATION '<location><suffix>/'`, :99-115 INSERT OVERWRITE standby + flip, :117-125 persist).
// Host code only (a restart happens once; nothing here is on the per-batch path).
//
// Writer: one file, one row group, one v1 data page per column, PLAIN, uncompressed, every column OPTIONAL with RLE/bit-packed
// definition levels; long -> INT64, double -> DOUBLE, boolean -> BOOLEAN, string -> BYTE_ARRAY (UTF8), timestamp -> INT96 (Julian day
// + nanoseconds of the day: what Spark 2.4 writes and expects by default, spark.sql.parquet.outputTimestampType = INT96).
// Reader: what Spark 2.4's writer produces for such tables — v1 data pages, PLAIN or dictionary encoded (PLAIN_DICTIONARY /
// RLE_DICTIONARY indices in the RLE/bit-packed hybrid), definition levels for OPTIONAL columns, UNCOMPRESSED or SNAPPY pages, any
// number of row groups and of part files in the directory.  The format is Apache Parquet's (parquet-format 2.x: thrift compact
// FileMetaData / PageHeader); nothing of it is in /root/reference, the reference reaches it through spark-sql.
#include "dx_common.h"
#include "dx_parquet.h"
#include <algorithm>
#include <dirent.h>
#include <fstream>
#include <sstream>
#include <sys/stat.h>

namespace {
struct PqError : std::runtime_error { using std::runtime_error::runtime_error; };

// ---- thrift compact protocol -------------------------------------------------------------------------------------------
enum { T_STOP = 0, T_TRUE = 1, T_FALSE = 2, T_BYTE = 3, T_I16 = 4, T_I32 = 5, T_I64 = 6, T_DOUBLE = 7, T_BINARY = 8, T_LIST = 9, T_SET = 10, T_MAP = 11, T_STRUCT = 12 };
struct TWriter {
    std::string out; std::vector<int> last;
    void byte(uint8_t b) { out.push_back((char)b); }
    void varint(uint64_t v) { while (v >= 0x80) { byte((uint8_t)(v | 0x80)); v >>= 7; } byte((uint8_t)v); }
    void zz(int64_t v) { varint(((uint64_t)v << 1) ^ (uint64_t)(v >> 63)); }
    void begin() { last.push_back(0); }
    void end() { byte(T_STOP); last.pop_back(); }
    void field(int id, int type) {
        const int d = id - last.back();
        if (d > 0 && d <= 15) byte((uint8_t)((d << 4) | type)); else { byte((uint8_t)type); zz(id); }
        last.back() = id;
    }
    void i32(int id, int64_t v) { field(id, T_I32); zz(v); }
    void i64(int id, int64_t v) { field(id, T_I64); zz(v); }
    void str(int id, const std::string& s) { field(id, T_BINARY); varint(s.size()); out += s; }
    void list(int id, int elem, size_t n) { field(id, T_LIST); if (n < 15) byte((uint8_t)((n << 4) | elem)); else { byte((uint8_t)(0xF0 | elem)); varint(n); } }
};
struct TReader {
    const uint8_t* p; const uint8_t* e;
    uint8_t byte() { if (p >= e) throw PqError("truncated thrift data"); return *p++; }
    uint64_t varint() { uint64_t v = 0; int sh = 0; for (;;) { const uint8_t b = byte(); v |= (uint64_t)(b & 0x7f) << sh; if (!(b & 0x80)) return v; sh += 7; if (sh > 63) throw PqError("bad varint"); } }
    int64_t zz() { const uint64_t v = varint(); return (int64_t)(v >> 1) ^ -(int64_t)(v & 1); }
    std::string str() { const uint64_t n = varint(); if ((uint64_t)(e - p) < n) throw PqError("truncated thrift string"); std::string s((const char*)p, (size_t)n); p += n; return s; }
    // field header: false at STOP
    bool field(int* last, int* id, int* type) {
        const uint8_t b = byte();
        if (b == T_STOP) return false;
        *type = b & 15;
        const int d = b >> 4;
        *id = d ? *last + d : (int)zz();
        *last = *id;
        return true;
    }
    void list_header(int* elem, size_t* n) { const uint8_t b = byte(); *elem = b & 15; *n = b >> 4; if (*n == 15) *n = (size_t)varint(); }
    void skip(int type) {
        switch (type) {
        case T_TRUE: case T_FALSE: return;
        case T_BYTE: byte(); return;
        case T_I16: case T_I32: case T_I64: varint(); return;
        case T_DOUBLE: if (e - p < 8) throw PqError("truncated"); p += 8; return;
        case T_BINARY: str(); return;
        case T_LIST: case T_SET: { int el; size_t n; list_header(&el, &n); for (size_t i = 0; i < n; i++) skip(el == T_TRUE || el == T_FALSE ? T_BYTE : el); return; }
        case T_MAP: { const uint64_t n = varint(); if (!n) return; const uint8_t kv = byte(); for (uint64_t i = 0; i < n; i++) { skip(kv >> 4); skip(kv & 15); } return; }
        case T_STRUCT: { int last = 0, id, ty; while (field(&last, &id, &ty)) skip(ty); return; }
        default: throw PqError("unknown thrift type");
        }
    }
};

enum { PQ_BOOLEAN = 0, PQ_INT32 = 1, PQ_INT64 = 2, PQ_INT96 = 3, PQ_FLOAT = 4, PQ_DOUBLE = 5, PQ_BYTE_ARRAY = 6, PQ_FIXED = 7 };
enum { ENC_PLAIN = 0, ENC_PLAIN_DICT = 2, ENC_RLE = 3, ENC_BIT_PACKED = 4, ENC_RLE_DICT = 8 };
const long long US_PER_DAY = 86400000000ll;
const long long JULIAN_EPOCH = 2440588;

void put_le(std::string& s, uint64_t v, int n) { for (int i = 0; i < n; i++) s.push_back((char)((v >> (8 * i)) & 0xff)); }
uint64_t get_le(const uint8_t* p, int n) { uint64_t v = 0; for (int i = 0; i < n; i++) v |= (uint64_t)p[i] << (8 * i); return v; }

// ---- snappy (raw format) -----------------------------------------------------------------------------------------------
std::string snappy_uncompress(const uint8_t* p, size_t n) {
    const uint8_t* e = p + n;
    uint64_t len = 0; int sh = 0;
    for (;;) { if (p >= e) throw PqError("snappy: truncated"); const uint8_t b = *p++; len |= (uint64_t)(b & 0x7f) << sh; if (!(b & 0x80)) break; sh += 7; }
    std::string out; out.reserve((size_t)len);
    while (p < e) {
        const uint8_t tag = *p++;
        if ((tag & 3) == 0) {
            size_t l = (tag >> 2) + 1;
            if (l > 60) { const int nb = (int)l - 60; if (e - p < nb) throw PqError("snappy: truncated"); l = (size_t)get_le(p, nb) + 1; p += nb; }
            if ((size_t)(e - p) < l) throw PqError("snappy: truncated literal");
            out.append((const char*)p, l); p += l;
        } else {
            size_t l, off;
            if ((tag & 3) == 1) { if (p >= e) throw PqError("snappy: truncated"); l = 4 + ((tag >> 2) & 7); off = ((size_t)(tag >> 5) << 8) | *p++; }
            else if ((tag & 3) == 2) { if (e - p < 2) throw PqError("snappy: truncated"); l = (tag >> 2) + 1; off = (size_t)get_le(p, 2); p += 2; }
            else { if (e - p < 4) throw PqError("snappy: truncated"); l = (tag >> 2) + 1; off = (size_t)get_le(p, 4); p += 4; }
            if (off == 0 || off > out.size()) throw PqError("snappy: bad offset");
            const size_t from = out.size() - off;
            for (size_t i = 0; i < l; i++) out.push_back(out[from + i]);
        }
    }
    if (out.size() != len) throw PqError("snappy: length mismatch");
    return out;
}

// ---- RLE / bit-packed hybrid -------------------------------------------------------------------------------------------
void hybrid_decode(const uint8_t* p, const uint8_t* e, int width, size_t n, std::vector<uint32_t>* out) {
    out->clear(); out->reserve(n);
    if (width == 0) { out->assign(n, 0); return; }
    TReader r{p, e};
    const int vb = (width + 7) / 8;
    while (out->size() < n) {
        if (r.p >= r.e) throw PqError("levels / indices run out");
        const uint64_t h = r.varint();
        if (h & 1) {
            const size_t groups = (size_t)(h >> 1), bytes = groups * width;
            if ((size_t)(r.e - r.p) < bytes && (size_t)(r.e - r.p) * 8 < (n - out->size()) * (size_t)width) throw PqError("bit-packed run truncated");
            uint64_t acc = 0; int have = 0; const uint8_t* q = r.p;
            for (size_t i = 0; i < groups * 8 && out->size() < n; i++) {
                while (have < width) { acc |= (uint64_t)(q < r.e ? *q : 0) << have; q++; have += 8; }
                out->push_back((uint32_t)(acc & ((1ull << width) - 1))); acc >>= width; have -= width;
            }
            r.p += std::min(bytes, (size_t)(r.e - r.p));
        } else {
            const size_t cnt = (size_t)(h >> 1);
            if (r.e - r.p < vb) throw PqError("RLE run truncated");
            const uint32_t v = (uint32_t)get_le(r.p, vb); r.p += vb;
            for (size_t i = 0; i < cnt && out->size() < n; i++) out->push_back(v);
        }
    }
}
std::string hybrid_encode_bits(const std::vector<uint8_t>& bits) {      // width 1, bit-packed runs
    std::string s;
    const size_t groups = (bits.size() + 7) / 8;
    TWriter w; w.varint((groups << 1) | 1); s = w.out;
    for (size_t g = 0; g < groups; g++) { uint8_t b = 0; for (int k = 0; k < 8; k++) { const size_t i = g * 8 + k; if (i < bits.size() && bits[i]) b |= 1u << k; } s.push_back((char)b); }
    return s;
}

std::string read_file(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw PqError("cannot open '" + path + "'");
    std::ostringstream ss; ss << f.rdbuf();
    return ss.str();
}
bool is_dir(const std::string& p) { struct stat st; return stat(p.c_str(), &st) == 0 && S_ISDIR(st.st_mode); }
bool exists(const std::string& p) { struct stat st; return stat(p.c_str(), &st) == 0; }
void mkdirs(const std::string& p) {
    std::string cur;
    for (size_t i = 0; i <= p.size(); i++) {
        if (i == p.size() || p[i] == '/') { if (!cur.empty() && !exists(cur)) mkdir(cur.c_str(), 0777); }
        if (i < p.size()) cur.push_back(p[i]);
    }
}

struct SchemaEl { int type = -1, rep = 0, children = 0, converted = -1; std::string name; };
struct ColChunk { int type = 0, codec = 0; int64_t num_values = 0, data_off = -1, dict_off = -1, total_compressed = 0; std::vector<std::string> path; };

// one part file appended to the columns of `t`
void read_one(const std::string& path, dx_parquet* t) {
    const std::string file = read_file(path);
    const uint8_t* base = (const uint8_t*)file.data();
    if (file.size() < 12 || memcmp(base, "PAR1", 4) || memcmp(base + file.size() - 4, "PAR1", 4)) throw PqError("'" + path + "' is not a Parquet file");
    const uint32_t flen = (uint32_t)get_le(base + file.size() - 8, 4);
    if ((size_t)flen + 12 > file.size()) throw PqError("bad footer length in '" + path + "'");
    TReader r{base + file.size() - 8 - flen, base + file.size() - 8};
    std::vector<SchemaEl> schema; std::vector<std::vector<ColChunk>> groups; std::vector<int64_t> group_rows;
    int last = 0, id, ty;
    while (r.field(&last, &id, &ty)) {
        if (id == 2 && ty == T_LIST) {
            int el; size_t n; r.list_header(&el, &n);
            for (size_t i = 0; i < n; i++) {
                SchemaEl se; int l2 = 0, i2, t2;
                while (r.field(&l2, &i2, &t2)) {
                    if (i2 == 1) se.type = (int)r.zz(); else if (i2 == 3) se.rep = (int)r.zz(); else if (i2 == 4) se.name = r.str();
                    else if (i2 == 5) se.children = (int)r.zz(); else if (i2 == 6) se.converted = (int)r.zz(); else r.skip(t2);
                }
                schema.push_back(se);
            }
        } else if (id == 4 && ty == T_LIST) {
            int el; size_t n; r.list_header(&el, &n);
            for (size_t g = 0; g < n; g++) {
                std::vector<ColChunk> cols; int64_t rows = 0; int l2 = 0, i2, t2;
                while (r.field(&l2, &i2, &t2)) {
                    if (i2 == 1 && t2 == T_LIST) {
                        int e3; size_t n3; r.list_header(&e3, &n3);
                        for (size_t c = 0; c < n3; c++) {
                            ColChunk cc; int l3 = 0, i3, t3;
                            while (r.field(&l3, &i3, &t3)) {
                                if (i3 == 3 && t3 == T_STRUCT) {
                                    int l4 = 0, i4, t4;
                                    while (r.field(&l4, &i4, &t4)) {
                                        if (i4 == 1) cc.type = (int)r.zz(); else if (i4 == 4) cc.codec = (int)r.zz(); else if (i4 == 5) cc.num_values = r.zz();
                                        else if (i4 == 7) cc.total_compressed = r.zz(); else if (i4 == 9) cc.data_off = r.zz(); else if (i4 == 11) cc.dict_off = r.zz();
                                        else if (i4 == 3 && t4 == T_LIST) { int e5; size_t n5; r.list_header(&e5, &n5); for (size_t k = 0; k < n5; k++) cc.path.push_back(r.str()); }
                                        else r.skip(t4);
                                    }
                                } else r.skip(t3);
                            }
                            cols.push_back(cc);
                        }
                    } else if (i2 == 3) rows = r.zz(); else r.skip(t2);
                }
                groups.push_back(cols); group_rows.push_back(rows);
            }
        } else r.skip(ty);
    }
    if (schema.empty()) throw PqError("no schema in '" + path + "'");
    // flat schema only: root + leaves
    std::vector<SchemaEl> leaves(schema.begin() + 1, schema.end());
    for (auto& se : leaves) if (se.children) throw PqError("nested column '" + se.name + "' in '" + path + "': state tables are flat");
    if (t->cols.empty()) {
        for (auto& se : leaves) {
            dx_parquet::Col c; c.name = se.name;
            switch (se.type) {
            case PQ_INT64: case PQ_INT32: c.type = DX_T_I64; break;
            case PQ_DOUBLE: case PQ_FLOAT: c.type = DX_T_F64; break;
            case PQ_BOOLEAN: c.type = DX_T_BOOL; break;
            case PQ_INT96: c.type = DX_T_TS; break;
            case PQ_BYTE_ARRAY: c.type = DX_T_STR; c.offsets.push_back(0); break;
            default: throw PqError("column '" + se.name + "': unsupported Parquet type");
            }
            if (se.type == PQ_INT64 && (se.converted == 9 || se.converted == 10)) c.type = DX_T_TS;     // TIMESTAMP_MILLIS / MICROS
            c.pq_type = se.type; c.converted = se.converted;
            t->cols.push_back(c);
        }
    } else if (t->cols.size() != leaves.size()) throw PqError("'" + path + "' has a different number of columns than the files before it");
    for (size_t g = 0; g < groups.size(); g++) {
        if (groups[g].size() != leaves.size()) throw PqError("row group with missing columns in '" + path + "'");
        for (size_t c = 0; c < leaves.size(); c++) {
            const ColChunk& cc = groups[g][c]; dx_parquet::Col& col = t->cols[c];
            const bool optional = leaves[c].rep == 1;
            int64_t at = cc.dict_off > 0 && cc.dict_off < cc.data_off ? cc.dict_off : cc.data_off;
            int64_t seen = 0;
            std::string dict_plain; std::vector<std::string> dict_str; std::vector<uint64_t> dict_num; bool have_dict = false;
            while (seen < cc.num_values) {
                if (at < 4 || (size_t)at >= file.size()) throw PqError("page offset out of the file in '" + path + "'");
                TReader ph{base + at, base + file.size()};
                int ptype = -1; int64_t usize = 0, csize = 0, nvals = 0; int enc = 0, def_enc = ENC_RLE; int l2 = 0, i2, t2;
                while (ph.field(&l2, &i2, &t2)) {
                    if (i2 == 1) ptype = (int)ph.zz(); else if (i2 == 2) usize = ph.zz(); else if (i2 == 3) csize = ph.zz();
                    else if ((i2 == 5 || i2 == 7) && t2 == T_STRUCT) {
                        int l3 = 0, i3, t3;
                        while (ph.field(&l3, &i3, &t3)) { if (i3 == 1) nvals = ph.zz(); else if (i3 == 2) enc = (int)ph.zz(); else if (i3 == 3 && i2 == 5) def_enc = (int)ph.zz(); else ph.skip(t3); }
                    } else ph.skip(t2);
                }
                const uint8_t* body = ph.p;
                if ((size_t)(ph.e - body) < (size_t)csize) throw PqError("page body truncated in '" + path + "'");
                at = (int64_t)(body - base) + csize;
                std::string page;
                if (cc.codec == 0) page.assign((const char*)body, (size_t)csize);
                else if (cc.codec == 1) page = snappy_uncompress(body, (size_t)csize);
                else throw PqError("column '" + col.name + "': compression codec " + std::to_string(cc.codec) + " is not supported (UNCOMPRESSED and SNAPPY are)");
                if ((int64_t)page.size() != usize) throw PqError("page size mismatch in '" + path + "'");
                const uint8_t* q = (const uint8_t*)page.data(); const uint8_t* qe = q + page.size();
                auto plain_values = [&](const uint8_t*& v, const uint8_t* ve, size_t n, std::vector<uint64_t>* nums, std::vector<std::string>* strs) {
                    for (size_t i = 0; i < n; i++) {
                        switch (col.pq_type) {
                        case PQ_INT64: case PQ_DOUBLE: if (ve - v < 8) throw PqError("values truncated"); nums->push_back(get_le(v, 8)); v += 8; break;
                        case PQ_INT32: if (ve - v < 4) throw PqError("values truncated"); nums->push_back((uint64_t)(int64_t)(int32_t)get_le(v, 4)); v += 4; break;
                        case PQ_FLOAT: { if (ve - v < 4) throw PqError("values truncated"); uint32_t b = (uint32_t)get_le(v, 4); float f; memcpy(&f, &b, 4); double d = f; uint64_t u; memcpy(&u, &d, 8); nums->push_back(u); v += 4; break; }
                        case PQ_INT96: { if (ve - v < 12) throw PqError("values truncated"); const uint64_t nanos = get_le(v, 8); const int64_t jd = (int64_t)(int32_t)get_le(v + 8, 4);
                                         nums->push_back((uint64_t)((jd - JULIAN_EPOCH) * US_PER_DAY + (int64_t)(nanos / 1000))); v += 12; break; }
                        case PQ_BYTE_ARRAY: { if (ve - v < 4) throw PqError("values truncated"); const uint32_t l = (uint32_t)get_le(v, 4); v += 4; if ((size_t)(ve - v) < l) throw PqError("values truncated"); strs->emplace_back((const char*)v, l); v += l; break; }
                        default: throw PqError("unsupported value type");
                        }
                    }
                };
                if (ptype == 2) {       // dictionary page: PLAIN values
                    dict_num.clear(); dict_str.clear();
                    plain_values(q, qe, (size_t)nvals, &dict_num, &dict_str);
                    have_dict = true;
                    continue;
                }
                if (ptype != 0) throw PqError("page type " + std::to_string(ptype) + " in '" + path + "' (v1 data pages are read)");
                std::vector<uint32_t> defs;
                size_t present = (size_t)nvals;
                if (optional) {
                    if (def_enc != ENC_RLE) throw PqError("definition levels not RLE-encoded");
                    if (qe - q < 4) throw PqError("page truncated");
                    const uint32_t dl = (uint32_t)get_le(q, 4); q += 4;
                    if ((size_t)(qe - q) < dl) throw PqError("definition levels truncated");
                    hybrid_decode(q, q + dl, 1, (size_t)nvals, &defs); q += dl;
                    present = 0; for (uint32_t d : defs) present += d;
                }
                std::vector<uint64_t> nums; std::vector<std::string> strs;
                if (enc == ENC_PLAIN) {
                    if (col.pq_type == PQ_BOOLEAN) { for (size_t i = 0; i < present; i++) { if (q + i / 8 >= qe) throw PqError("values truncated"); nums.push_back((q[i / 8] >> (i % 8)) & 1); } }
                    else plain_values(q, qe, present, &nums, &strs);
                } else if (enc == ENC_PLAIN_DICT || enc == ENC_RLE_DICT) {
                    if (!have_dict) throw PqError("dictionary-encoded page without a dictionary page");
                    if (q >= qe && present) throw PqError("page truncated");
                    const int bw = present || q < qe ? *q++ : 0;
                    std::vector<uint32_t> idx; hybrid_decode(q, qe, bw, present, &idx);
                    for (uint32_t k : idx) {
                        if (col.pq_type == PQ_BYTE_ARRAY) { if (k >= dict_str.size()) throw PqError("dictionary index out of range"); strs.push_back(dict_str[k]); }
                        else { if (k >= dict_num.size()) throw PqError("dictionary index out of range"); nums.push_back(dict_num[k]); }
                    }
                } else throw PqError("column '" + col.name + "': encoding " + std::to_string(enc) + " is not supported");
                size_t vi = 0;
                for (size_t i = 0; i < (size_t)nvals; i++) {
                    const bool ok = !optional || defs[i];
                    col.valid.push_back(ok ? 1 : 0);
                    if (col.type == DX_T_STR) { if (ok) col.bytes += strs[vi]; col.offsets.push_back((int64_t)col.bytes.size()); }
                    else {
                        uint64_t v = ok ? nums[vi] : 0;
                        if (ok && col.pq_type == PQ_INT64 && col.converted == 9) v = (uint64_t)((int64_t)v * 1000);      // TIMESTAMP_MILLIS -> microseconds
                        col.values.push_back((int64_t)v);
                    }
                    if (ok) vi++;
                }
                seen += nvals;
            }
        }
        t->n_rows += group_rows[g];
    }
}
}  // namespace

int dx_parquet_read_impl(const std::string& path, dx_parquet** out, std::string* err) {
    try {
        std::unique_ptr<dx_parquet> t(new dx_parquet());
        if (is_dir(path)) {
            std::vector<std::string> files;
            if (DIR* d = opendir(path.c_str())) { while (dirent* e = readdir(d)) { const std::string n = e->d_name; if (n.empty() || n[0] == '.' || n[0] == '_') continue; if (!is_dir(path + "/" + n)) files.push_back(n); } closedir(d); }
            std::sort(files.begin(), files.end());
            for (auto& f : files) read_one(path + "/" + f, t.get());
        } else read_one(path, t.get());
        for (auto& c : t->cols) if ((int64_t)c.valid.size() != t->n_rows) throw PqError("column '" + c.name + "' has " + std::to_string(c.valid.size()) + " values for " + std::to_string(t->n_rows) + " rows");
        *out = t.release();
        return DX_OK;
    } catch (std::exception& e) { *err = e.what(); return DX_ERR_PARSE; }
}

int dx_parquet_write_impl(const std::string& path, const std::vector<std::string>& names, const std::vector<int>& types, int64_t n_rows,
                          const void* const* data, const uint8_t* const* valid, const uint8_t* const* string_bytes, std::string* err) {
    try {
        std::string file = "PAR1";
        struct Chunk { int pq; int64_t off, size; };
        std::vector<Chunk> chunks;
        for (size_t c = 0; c < names.size(); c++) {
            int pq;
            switch (types[c]) {
            case DX_T_I64: pq = PQ_INT64; break; case DX_T_F64: pq = PQ_DOUBLE; break; case DX_T_BOOL: pq = PQ_BOOLEAN; break;
            case DX_T_TS: pq = PQ_INT96; break; case DX_T_STR: pq = PQ_BYTE_ARRAY; break;
            default: throw PqError("column '" + names[c] + "': type cannot be stored");
            }
            std::vector<uint8_t> defs((size_t)n_rows);
            for (int64_t r = 0; r < n_rows; r++) defs[(size_t)r] = (!valid || !valid[c] || valid[c][r]) ? 1 : 0;
            std::string levels = hybrid_encode_bits(defs), vals;
            const int64_t* v64 = reinterpret_cast<const int64_t*>(data[c]);
            uint8_t bacc = 0; int bn = 0;
            for (int64_t r = 0; r < n_rows; r++) {
                if (!defs[(size_t)r]) continue;
                switch (pq) {
                case PQ_INT64: case PQ_DOUBLE: put_le(vals, (uint64_t)v64[r], 8); break;
                case PQ_BOOLEAN: if (v64[r]) bacc |= 1u << bn; if (++bn == 8) { vals.push_back((char)bacc); bacc = 0; bn = 0; } break;
                case PQ_INT96: {
                    const long long us = v64[r];
                    long long day = us / US_PER_DAY, rem = us % US_PER_DAY; if (rem < 0) { rem += US_PER_DAY; day--; }
                    put_le(vals, (uint64_t)(rem * 1000), 8); put_le(vals, (uint64_t)(uint32_t)(int32_t)(day + JULIAN_EPOCH), 4); break;
                }
                case PQ_BYTE_ARRAY: { const int64_t a = v64[r], b = v64[r + 1]; put_le(vals, (uint64_t)(b - a), 4); vals.append(reinterpret_cast<const char*>(string_bytes[c]) + a, (size_t)(b - a)); break; }
                }
            }
            if (pq == PQ_BOOLEAN && bn) vals.push_back((char)bacc);
            std::string body; put_le(body, levels.size(), 4); body += levels; body += vals;
            TWriter ph; ph.begin();
            ph.i32(1, 0); ph.i32(2, (int64_t)body.size()); ph.i32(3, (int64_t)body.size());
            ph.field(5, T_STRUCT); ph.begin(); ph.i32(1, n_rows); ph.i32(2, ENC_PLAIN); ph.i32(3, ENC_RLE); ph.i32(4, ENC_RLE); ph.end();
            ph.end();
            chunks.push_back({pq, (int64_t)file.size(), (int64_t)(ph.out.size() + body.size())});
            file += ph.out; file += body;
        }
        TWriter fm; fm.begin();
        fm.i32(1, 1);
        fm.list(2, T_STRUCT, names.size() + 1);
        fm.begin(); fm.str(4, "spark_schema"); fm.i32(5, (int64_t)names.size()); fm.end();
        for (size_t c = 0; c < names.size(); c++) {
            fm.begin(); fm.i32(1, chunks[c].pq); fm.i32(3, 1); fm.str(4, names[c]); if (chunks[c].pq == PQ_BYTE_ARRAY) fm.i32(6, 0); fm.end();
        }
        fm.i64(3, n_rows);
        fm.list(4, T_STRUCT, 1);
        fm.begin();
        fm.list(1, T_STRUCT, names.size());
        int64_t total = 0;
        for (size_t c = 0; c < names.size(); c++) {
            fm.begin(); fm.i64(2, chunks[c].off);
            fm.field(3, T_STRUCT); fm.begin();
            fm.i32(1, chunks[c].pq); fm.list(2, T_I32, 2); fm.zz(ENC_PLAIN); fm.zz(ENC_RLE);
            fm.list(3, T_BINARY, 1); fm.varint(names[c].size()); fm.out += names[c];
            fm.i32(4, 0); fm.i64(5, n_rows); fm.i64(6, chunks[c].size); fm.i64(7, chunks[c].size); fm.i64(9, chunks[c].off);
            fm.end(); fm.end();
            total += chunks[c].size;
        }
        fm.i64(2, total); fm.i64(3, n_rows);
        fm.end();
        fm.str(6, "dxgpu (parquet subset of Spark 2.4 state tables)");
        fm.end();
        file += fm.out; put_le(file, fm.out.size(), 4); file += "PAR1";
        const size_t slash = path.rfind('/');
        if (slash != std::string::npos) mkdirs(path.substr(0, slash));
        const std::string tmp = path + ".tmp";
        { std::ofstream f(tmp, std::ios::binary | std::ios::trunc); if (!f) throw PqError("cannot write '" + tmp + "'"); f.write(file.data(), (std::streamsize)file.size()); if (!f) throw PqError("short write to '" + tmp + "'"); }
        if (rename(tmp.c_str(), path.c_str())) throw PqError("cannot rename '" + tmp + "'");
        return DX_OK;
    } catch (std::exception& e) { *err = e.what(); return DX_ERR_STATE; }
}

// metadata.info (StateTableHandler.scala:26-52): key=value lines; absent file = active A, standby B
int dx_state_metadata_read(const std::string& location, std::string* active, std::string* standby, std::string* err) {
    *active = "A"; *standby = "B";
    const std::string f = location + "/metadata.info";
    if (!exists(f)) return DX_OK;
    try {
        std::istringstream ss(read_file(f)); std::string line;
        while (std::getline(ss, line)) {
            if (!line.empty() && line.back() == '\r') line.pop_back();
            if (line.empty()) continue;
            const size_t pos = line.find('=');
            if (pos == std::string::npos || pos == 0) { *err = "Invalid content in '" + f + "': '" + line + "'"; return DX_ERR_PARSE; }
            if (line.substr(0, pos) == "active") *active = line.substr(pos + 1); else if (line.substr(0, pos) == "standby") *standby = line.substr(pos + 1);
        }
        return DX_OK;
    } catch (std::exception& e) { *err = e.what(); return DX_ERR_PARSE; }
}
int dx_state_metadata_write(const std::string& location, const std::string& active, const std::string& standby, std::string* err) {
    mkdirs(location);
    const std::string f = location + "/metadata.info", tmp = f + ".tmp";
    { std::ofstream o(tmp, std::ios::trunc); if (!o) { *err = "cannot write '" + tmp + "'"; return DX_ERR_STATE; } o << "active=" << active << "\nstandby=" << standby; }
    if (rename(tmp.c_str(), f.c_str())) { *err = "cannot rename '" + tmp + "'"; return DX_ERR_STATE; }
    return DX_OK;
}
void dx_clear_directory_files(const std::string& dir) {
    if (DIR* d = opendir(dir.c_str())) {
        std::vector<std::string> files;
        while (dirent* e = readdir(d)) { const std::string n = e->d_name; if (n != "." && n != ".." && !is_dir(dir + "/" + n)) files.push_back(n); }
        closedir(d);
        for (auto& f : files) remove((dir + "/" + f).c_str());
    }
}
