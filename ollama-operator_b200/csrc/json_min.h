// json_min.h -- a JSON reader / string escaper just large enough for Ollama manifests and API requests
// (objects, arrays, strings with \u escapes, numbers, literals).  Header only; used by store.cpp and the HTTP front end.
#pragma once
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <utility>
#include <vector>

namespace b2 {

struct Json {
    enum T { NUL, BOOL, NUM, STR, ARR, OBJ } t = NUL;
    double num = 0;
    bool b = false;
    std::string str;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;
    const Json* get(const char* key) const {
        for (const auto& kv : obj)
            if (kv.first == key) return &kv.second;
        return nullptr;
    }
};
struct JsonParser {
    const char* p;
    const char* end;
    bool ok = true;
    void ws() { while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) p++; }
    bool lit(const char* s) {
        const size_t n = strlen(s);
        if ((size_t)(end - p) >= n && memcmp(p, s, n) == 0) { p += n; return true; }
        return false;
    }
    std::string string() {
        std::string s;
        if (p >= end || *p != '"') { ok = false; return s; }
        for (p++; p < end && *p != '"'; p++) {
            if (*p != '\\') { s.push_back(*p); continue; }
            if (++p >= end) break;
            switch (*p) {
            case 'n': s.push_back('\n'); break;
            case 't': s.push_back('\t'); break;
            case 'r': s.push_back('\r'); break;
            case 'b': s.push_back('\b'); break;
            case 'f': s.push_back('\f'); break;
            case 'u': {
                if (end - p < 5) { ok = false; return s; }
                unsigned cp = (unsigned)strtoul(std::string(p + 1, 4).c_str(), nullptr, 16);
                p += 4;
                if (cp < 0x80) s.push_back((char)cp);
                else if (cp < 0x800) { s.push_back((char)(0xC0 | (cp >> 6))); s.push_back((char)(0x80 | (cp & 0x3F))); }
                else { s.push_back((char)(0xE0 | (cp >> 12))); s.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); s.push_back((char)(0x80 | (cp & 0x3F))); }
            } break;
            default: s.push_back(*p);
            }
        }
        if (p >= end) { ok = false; return s; }
        p++;
        return s;
    }
    Json value(int depth = 0) {
        Json j;
        ws();
        if (!ok || p >= end || depth > 32) { ok = false; return j; }
        if (*p == '{') {
            j.t = Json::OBJ;
            p++; ws();
            if (p < end && *p == '}') { p++; return j; }
            while (ok) {
                ws();
                std::string k = string();
                ws();
                if (!ok || p >= end || *p != ':') { ok = false; break; }
                p++;
                j.obj.emplace_back(std::move(k), value(depth + 1));
                ws();
                if (p < end && *p == ',') { p++; continue; }
                if (p < end && *p == '}') { p++; break; }
                ok = false;
            }
        } else if (*p == '[') {
            j.t = Json::ARR;
            p++; ws();
            if (p < end && *p == ']') { p++; return j; }
            while (ok) {
                j.arr.push_back(value(depth + 1));
                ws();
                if (p < end && *p == ',') { p++; continue; }
                if (p < end && *p == ']') { p++; break; }
                ok = false;
            }
        } else if (*p == '"') {
            j.t = Json::STR;
            j.str = string();
        } else if (lit("true")) { j.t = Json::BOOL; j.b = true; }
        else if (lit("false")) { j.t = Json::BOOL; }
        else if (lit("null")) {}
        else {
            char* e = nullptr;
            j.num = strtod(p, &e);
            if (e == p || e > end) ok = false;
            else { j.t = Json::NUM; p = e; }
        }
        return j;
    }
};


inline Json json_parse(const std::string& text, bool* ok) {
    JsonParser jp{text.data(), text.data() + text.size()};
    Json j = jp.value();
    jp.ws();
    if (ok) *ok = jp.ok && jp.p == jp.end;
    return j;
}

inline std::string json_escape(const std::string& s) {
    std::string o;
    o.reserve(s.size() + 8);
    for (unsigned char c : s) {
        switch (c) {
        case '"': o += "\\\""; break;
        case '\\': o += "\\\\"; break;
        case '\n': o += "\\n"; break;
        case '\r': o += "\\r"; break;
        case '\t': o += "\\t"; break;
        default:
            if (c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); o += b; }
            else o.push_back((char)c);
        }
    }
    return o;
}

}  // namespace b2
