// gguf.hpp — GGUF v2/v3 metadata parser (host, C++).  The north star asks for a Go parser; no Go
// toolchain exists on this image, so the parser lives behind the C ABI where it can be tested.
// Container spec cross-checked against gguf-py 0.19.0 (gguf/constants.py:10-12 magic / version /
// default alignment 32; :4157-4170 value types) — the writer the test fixtures are produced with.
#pragma once
#include <cstdint>
#include <cstring>
#include <map>
#include <string>
#include <vector>

namespace b200 {

enum GgufValueType : uint32_t { GV_U8, GV_I8, GV_U16, GV_I16, GV_U32, GV_I32, GV_F32, GV_BOOL, GV_STR, GV_ARR, GV_U64, GV_I64, GV_F64 };

struct GgufValue {
    uint32_t type = GV_U32;
    double num = 0;                       // every scalar numeric / bool
    std::string str;                      // GV_STR
    uint32_t arr_type = 0;                // GV_ARR
    std::vector<double> arr_num;
    std::vector<std::string> arr_str;
};

struct GgufTensor {
    std::string name;
    int type = 0; int n_dims = 0; int64_t ne[4] = {1, 1, 1, 1};
    uint64_t offset = 0;                  // relative to data section
    uint64_t nbytes = 0;
    int64_t n_elems() const { return ne[0] * ne[1] * ne[2] * ne[3]; }
};

struct GgufMeta {
    uint32_t version = 0;
    uint64_t alignment = 32;
    uint64_t data_offset = 0;             // absolute file offset of the tensor data section
    uint64_t data_bytes = 0;              // bytes of tensor data (end of last tensor)
    std::map<std::string, GgufValue> kv;
    std::vector<GgufTensor> tensors;

    const GgufValue* find(const std::string& k) const { auto it = kv.find(k); return it == kv.end() ? nullptr : &it->second; }
    double num(const std::string& k, double d) const { auto v = find(k); return v && v->type != GV_STR && v->type != GV_ARR ? v->num : d; }
    std::string str(const std::string& k, const std::string& d = "") const { auto v = find(k); return v && v->type == GV_STR ? v->str : d; }
    const GgufTensor* tensor(const std::string& n) const { for (auto& t : tensors) if (t.name == n) return &t; return nullptr; }
};

inline int gguf_block_elems(int t) { switch (t) { case 0: case 1: case 30: return 1; case 8: return 32; case 12: case 13: case 14: return 256; default: return 0; } }
inline int gguf_block_bytes(int t) { switch (t) { case 0: return 4; case 1: case 30: return 2; case 8: return 34; case 12: return 144; case 13: return 176; case 14: return 210; default: return 0; } }

// Parses the header from `buf` (the first n bytes of the file).  Returns "" on success, "need more"
// if the buffer ends inside the header (caller retries with more bytes), else an error message.
inline std::string gguf_parse(const uint8_t* buf, size_t n, GgufMeta& out) {
    struct R {
        const uint8_t* p; const uint8_t* e; bool more = false;
        bool take(void* dst, size_t k) { if ((size_t)(e - p) < k) { more = true; return false; } memcpy(dst, p, k); p += k; return true; }
        bool str(std::string& s) { uint64_t len; if (!take(&len, 8)) return false; if ((uint64_t)(e - p) < len) { more = true; return false; } if (len > (1u << 28)) return false; s.assign((const char*)p, len); p += len; return true; }
    } r{buf, buf + n};
    static const int sz[] = {1, 1, 2, 2, 4, 4, 4, 1, 0, 0, 8, 8, 8};
    auto scalar = [&](uint32_t t, double& v) -> bool {
        uint8_t raw[8] = {0}; if (!r.take(raw, sz[t])) return false;
        switch (t) {
            case GV_U8: case GV_BOOL: v = raw[0]; break; case GV_I8: v = (int8_t)raw[0]; break;
            case GV_U16: { uint16_t x; memcpy(&x, raw, 2); v = x; } break; case GV_I16: { int16_t x; memcpy(&x, raw, 2); v = x; } break;
            case GV_U32: { uint32_t x; memcpy(&x, raw, 4); v = x; } break; case GV_I32: { int32_t x; memcpy(&x, raw, 4); v = x; } break;
            case GV_F32: { float x; memcpy(&x, raw, 4); v = x; } break;
            case GV_U64: { uint64_t x; memcpy(&x, raw, 8); v = (double)x; } break; case GV_I64: { int64_t x; memcpy(&x, raw, 8); v = (double)x; } break;
            case GV_F64: memcpy(&v, raw, 8); break;
        }
        return true;
    };
    uint32_t magic = 0; uint64_t n_t = 0, n_kv = 0;
    if (!r.take(&magic, 4)) return "need more";
    if (magic != 0x46554747u) return "not a GGUF file (bad magic)";
    if (!r.take(&out.version, 4)) return "need more";
    if (out.version != 2 && out.version != 3) return "unsupported GGUF version " + std::to_string(out.version);
    if (!r.take(&n_t, 8) || !r.take(&n_kv, 8)) return "need more";
    if (n_t > (1u << 20) || n_kv > (1u << 20)) return "implausible GGUF header counts";
    out.kv.clear(); out.tensors.clear();
    for (uint64_t i = 0; i < n_kv; i++) {
        std::string key; GgufValue v;
        if (!r.str(key)) return r.more ? "need more" : "bad key string";
        if (!r.take(&v.type, 4)) return "need more";
        if (v.type > GV_F64) return "bad value type for key " + key;
        if (v.type == GV_STR) { if (!r.str(v.str)) return r.more ? "need more" : "bad string value"; }
        else if (v.type == GV_ARR) {
            uint64_t cnt; if (!r.take(&v.arr_type, 4) || !r.take(&cnt, 8)) return "need more";
            if (v.arr_type > GV_F64 || v.arr_type == GV_ARR) return "unsupported array element type in " + key;
            if (cnt > (1ull << 28)) return "implausible array length in " + key;
            if (v.arr_type == GV_STR) { v.arr_str.resize(cnt); for (auto& s : v.arr_str) if (!r.str(s)) return r.more ? "need more" : "bad string in array"; }
            else { v.arr_num.resize(cnt); for (auto& d : v.arr_num) if (!scalar(v.arr_type, d)) return "need more"; }
        } else if (!scalar(v.type, v.num)) return "need more";
        out.kv[key] = std::move(v);
    }
    for (uint64_t i = 0; i < n_t; i++) {
        GgufTensor t; uint32_t nd, ty;
        if (!r.str(t.name)) return r.more ? "need more" : "bad tensor name";
        if (!r.take(&nd, 4)) return "need more";
        if (nd > 4) return "tensor " + t.name + " has more than 4 dims";
        t.n_dims = (int)nd;
        for (uint32_t d = 0; d < nd; d++) { uint64_t x; if (!r.take(&x, 8)) return "need more"; t.ne[d] = (int64_t)x; }
        if (!r.take(&ty, 4) || !r.take(&t.offset, 8)) return "need more";
        t.type = (int)ty;
        int be = gguf_block_elems(t.type);
        if (!be) return "tensor " + t.name + " has unsupported ggml type " + std::to_string(ty);
        if (t.ne[0] % be) return "tensor " + t.name + " row size is not a multiple of its block size";
        t.nbytes = (uint64_t)(t.n_elems() / be) * gguf_block_bytes(t.type);
        out.tensors.push_back(std::move(t));
    }
    out.alignment = (uint64_t)out.num("general.alignment", 32);
    if (out.alignment == 0 || (out.alignment & (out.alignment - 1))) return "bad general.alignment";
    uint64_t hdr = (uint64_t)(r.p - buf);
    out.data_offset = (hdr + out.alignment - 1) / out.alignment * out.alignment;
    out.data_bytes = 0;
    for (auto& t : out.tensors) { if (t.offset % out.alignment) return "tensor " + t.name + " is misaligned"; if (t.offset + t.nbytes > out.data_bytes) out.data_bytes = t.offset + t.nbytes; }
    return "";
}

}  // namespace b200
