// dx_json.h — minimal host-side JSON DOM, used only for the flow artefacts (Spark StructType
// schema JSON read by SchemaFile.loadInputSchema, reference datax/input/SchemaFile.scala:22-34).
// The event hot path never touches this; events are decoded on the GPU.
#pragma once
#include <string>
#include <vector>
#include <map>
#include <memory>
#include <stdexcept>

struct JVal {
    enum Kind { NUL, BOOL, NUM, STR, ARR, OBJ } kind = NUL;
    bool b = false;
    double num = 0;
    std::string str;
    std::vector<JVal> arr;
    std::vector<std::pair<std::string, JVal>> obj;
    const JVal* get(const std::string& k) const {
        for (auto& kv : obj) if (kv.first == k) return &kv.second;
        return nullptr;
    }
};

class JParser {
    const char* p; const char* e; int depth = 0;
    struct Nest { int& d; explicit Nest(int& x) : d(x) { d++; } ~Nest() { d--; } };
    void ws() { while (p < e && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) p++; }
    [[noreturn]] void fail(const char* m) { throw std::runtime_error(std::string("schema JSON: ") + m); }
    std::string str() {
        std::string out; p++;
        while (true) {
            if (p >= e) fail("unterminated string");
            char c = *p++;
            if (c == '"') break;
            if (c != '\\') { out.push_back(c); continue; }
            if (p >= e) fail("bad escape");
            char x = *p++;
            switch (x) {
            case 'n': out.push_back('\n'); break;
            case 't': out.push_back('\t'); break;
            case 'r': out.push_back('\r'); break;
            case 'b': out.push_back('\b'); break;
            case 'f': out.push_back('\f'); break;
            case 'u': {
                if (e - p < 4) fail("bad \\u");
                unsigned cp = (unsigned)std::stoul(std::string(p, p + 4), nullptr, 16); p += 4;
                if (cp < 0x80) out.push_back((char)cp);
                else if (cp < 0x800) { out.push_back((char)(0xC0 | (cp >> 6))); out.push_back((char)(0x80 | (cp & 63))); }
                else { out.push_back((char)(0xE0 | (cp >> 12))); out.push_back((char)(0x80 | ((cp >> 6) & 63))); out.push_back((char)(0x80 | (cp & 63))); }
                break;
            }
            default: out.push_back(x);
            }
        }
        return out;
    }
public:
    JParser(const char* s, size_t n) : p(s), e(s + n) {}
    JVal value() {
        ws();
        if (p >= e) fail("unexpected end");
        JVal v;
        Nest nest(depth);
        if (depth > 512) fail("nested deeper than 512 levels");
        if (*p == '{') {
            v.kind = JVal::OBJ; p++; ws();
            if (p < e && *p == '}') { p++; return v; }
            while (true) {
                ws(); if (p >= e || *p != '"') fail("expected key");
                std::string k = str(); ws();
                if (p >= e || *p != ':') fail("expected ':'");
                p++;
                v.obj.emplace_back(k, value()); ws();
                if (p < e && *p == ',') { p++; continue; }
                if (p < e && *p == '}') { p++; break; }
                fail("expected ',' or '}'");
            }
        } else if (*p == '[') {
            v.kind = JVal::ARR; p++; ws();
            if (p < e && *p == ']') { p++; return v; }
            while (true) {
                v.arr.push_back(value()); ws();
                if (p < e && *p == ',') { p++; continue; }
                if (p < e && *p == ']') { p++; break; }
                fail("expected ',' or ']'");
            }
        } else if (*p == '"') { v.kind = JVal::STR; v.str = str(); }
        else if (*p == 't' && e - p >= 4) { v.kind = JVal::BOOL; v.b = true; p += 4; }
        else if (*p == 'f' && e - p >= 5) { v.kind = JVal::BOOL; v.b = false; p += 5; }
        else if (*p == 'n' && e - p >= 4) { v.kind = JVal::NUL; p += 4; }
        else {
            const char* s = p;
            while (p < e && (*p == '-' || *p == '+' || *p == '.' || *p == 'e' || *p == 'E' || (*p >= '0' && *p <= '9'))) p++;
            if (s == p) fail("unexpected character");
            v.kind = JVal::NUM; v.num = std::stod(std::string(s, p));
        }
        return v;
    }
};

static inline JVal dx_parse_json(const std::string& s) {
    size_t off = 0;
    if (s.size() >= 3 && (unsigned char)s[0] == 0xEF && (unsigned char)s[1] == 0xBB && (unsigned char)s[2] == 0xBF) off = 3; // BOM
    JParser jp(s.data() + off, s.size() - off);
    return jp.value();
}
