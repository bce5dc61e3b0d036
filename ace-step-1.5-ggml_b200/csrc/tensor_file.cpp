#include "tensor_file.h"
#include "wpack.h"

#include <cerrno>
#include <cstdlib>
#include <cstring>
#include <fcntl.h>
#include <fstream>
#include <sstream>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace ace {

// ================================================================== JSON
namespace {

struct JsonReader {
    const char* p;
    const char* end;
    std::string err;

    void ws() {
        while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p;
    }
    bool fail(const char* what) {
        if (err.empty()) err = std::string("json: ") + what;
        return false;
    }
    bool parse_string(std::string& out) {
        if (p >= end || *p != '"') return fail("expected string");
        ++p;
        out.clear();
        while (p < end && *p != '"') {
            char c = *p++;
            if (c == '\\') {
                if (p >= end) return fail("bad escape");
                char e = *p++;
                switch (e) {
                    case 'n': out.push_back('\n'); break;
                    case 't': out.push_back('\t'); break;
                    case 'r': out.push_back('\r'); break;
                    case 'b': out.push_back('\b'); break;
                    case 'f': out.push_back('\f'); break;
                    case 'u': {
                        if (end - p < 4) return fail("bad \\u escape");
                        unsigned cp = (unsigned)std::strtoul(std::string(p, p + 4).c_str(), nullptr, 16);
                        p += 4;
                        if (cp < 0x80) out.push_back((char)cp);
                        else if (cp < 0x800) {
                            out.push_back((char)(0xC0 | (cp >> 6)));
                            out.push_back((char)(0x80 | (cp & 0x3F)));
                        } else {
                            out.push_back((char)(0xE0 | (cp >> 12)));
                            out.push_back((char)(0x80 | ((cp >> 6) & 0x3F)));
                            out.push_back((char)(0x80 | (cp & 0x3F)));
                        }
                        break;
                    }
                    default: out.push_back(e);
                }
            } else {
                out.push_back(c);
            }
        }
        if (p >= end) return fail("unterminated string");
        ++p;
        return true;
    }
    bool parse_value(JsonValue& v, int depth) {
        if (depth > 64) return fail("too deep");
        ws();
        if (p >= end) return fail("unexpected end");
        const char c = *p;
        if (c == '{') {
            ++p;
            v.kind = JsonValue::Object;
            ws();
            if (p < end && *p == '}') { ++p; return true; }
            while (true) {
                ws();
                std::string key;
                if (!parse_string(key)) return false;
                ws();
                if (p >= end || *p != ':') return fail("expected ':'");
                ++p;
                JsonValue child;
                if (!parse_value(child, depth + 1)) return false;
                v.obj.emplace_back(std::move(key), std::move(child));
                ws();
                if (p < end && *p == ',') { ++p; continue; }
                if (p < end && *p == '}') { ++p; return true; }
                return fail("expected ',' or '}'");
            }
        }
        if (c == '[') {
            ++p;
            v.kind = JsonValue::Array;
            ws();
            if (p < end && *p == ']') { ++p; return true; }
            while (true) {
                JsonValue child;
                if (!parse_value(child, depth + 1)) return false;
                v.arr.push_back(std::move(child));
                ws();
                if (p < end && *p == ',') { ++p; continue; }
                if (p < end && *p == ']') { ++p; return true; }
                return fail("expected ',' or ']'");
            }
        }
        if (c == '"') {
            v.kind = JsonValue::String;
            return parse_string(v.str);
        }
        if (end - p >= 4 && std::strncmp(p, "true", 4) == 0) { v.kind = JsonValue::Bool; v.b = true; p += 4; return true; }
        if (end - p >= 5 && std::strncmp(p, "false", 5) == 0) { v.kind = JsonValue::Bool; v.b = false; p += 5; return true; }
        if (end - p >= 4 && std::strncmp(p, "null", 4) == 0) { v.kind = JsonValue::Null; p += 4; return true; }
        // number (also accept NaN / Infinity tokens python may emit)
        const char* s = p;
        while (p < end && (std::strchr("+-0123456789.eE", *p) != nullptr)) ++p;
        if (p == s) return fail("unexpected character");
        v.kind = JsonValue::Number;
        v.num = std::strtod(std::string(s, p).c_str(), nullptr);
        return true;
    }
};

}  // namespace

const JsonValue* JsonValue::get(const std::string& key) const {
    if (kind != Object) return nullptr;
    for (const auto& kv : obj)
        if (kv.first == key) return &kv.second;
    return nullptr;
}
bool JsonValue::get_int(const std::string& key, int64_t& out) const {
    const JsonValue* v = get(key);
    if (!v || v->kind != Number) return false;
    out = (int64_t)v->num;
    return true;
}
bool JsonValue::get_num(const std::string& key, double& out) const {
    const JsonValue* v = get(key);
    if (!v || v->kind != Number) return false;
    out = v->num;
    return true;
}
bool JsonValue::get_bool(const std::string& key, bool& out) const {
    const JsonValue* v = get(key);
    if (!v) return false;
    if (v->kind == Bool) { out = v->b; return true; }
    if (v->kind == Number) { out = v->num != 0; return true; }
    return false;
}
bool JsonValue::get_str(const std::string& key, std::string& out) const {
    const JsonValue* v = get(key);
    if (!v || v->kind != String) return false;
    out = v->str;
    return true;
}

bool json_parse(const char* begin, const char* end, JsonValue& out, std::string& err) {
    JsonReader r{begin, end, {}};
    if (!r.parse_value(out, 0)) {
        err = r.err;
        return false;
    }
    return true;
}

bool json_parse_file(const std::string& path, JsonValue& out, std::string& err) {
    std::ifstream f(path, std::ios::binary);
    if (!f) {
        err = "cannot open " + path;
        return false;
    }
    std::stringstream ss;
    ss << f.rdbuf();
    const std::string s = ss.str();
    return json_parse(s.data(), s.data() + s.size(), out, err);
}

bool file_exists(const std::string& path) {
    struct stat st;
    return ::stat(path.c_str(), &st) == 0 && S_ISREG(st.st_mode);
}

// ================================================================== TensorFile
TensorFile::~TensorFile() {
    if (map_ && map_ != MAP_FAILED) ::munmap(map_, map_size_);
}

bool TensorFile::map_file(const std::string& path, std::string& err) {
    const int fd = ::open(path.c_str(), O_RDONLY);
    if (fd < 0) {
        err = "cannot open " + path + ": " + std::strerror(errno);
        return false;
    }
    struct stat st;
    if (::fstat(fd, &st) != 0 || st.st_size <= 0) {
        ::close(fd);
        err = "cannot stat " + path;
        return false;
    }
    map_size_ = (size_t)st.st_size;
    map_ = ::mmap(nullptr, map_size_, PROT_READ, MAP_PRIVATE, fd, 0);
    ::close(fd);
    if (map_ == MAP_FAILED) {
        map_ = nullptr;
        err = "mmap failed for " + path;
        return false;
    }
    return true;
}

const TensorInfo* TensorFile::find(const std::string& name) const {
    auto it = index_.find(name);
    return it == index_.end() ? nullptr : &tensors_[it->second];
}

bool TensorFile::read_f32(const std::string& name, std::vector<float>& out, std::vector<int64_t>* shape,
                          std::string& err) const {
    const TensorInfo* t = find(name);
    if (!t) {
        err = "missing tensor: " + name;
        return false;
    }
    if (shape) *shape = t->shape;
    const int64_t n = t->numel();
    out.resize((size_t)n);
    if (n == 0) return true;
    const int64_t row = t->shape.empty() ? 1 : t->shape.back();
    const size_t rb = ggml_row_bytes(t->ggml_type, row);
    if (rb == 0) {
        err = "unsupported dtype for " + name;
        return false;
    }
    const int64_t rows = n / row;
    for (int64_t r = 0; r < rows; ++r)
        if (!dequant_row(t->ggml_type, t->data + (size_t)r * rb, out.data() + (size_t)r * row, row)) {
            err = "cannot decode " + name;
            return false;
        }
    return true;
}

// every dim in [0, 2^40] and the element count below 2^46 (crafted headers must not wrap size arithmetic)
static bool checked_numel(const std::vector<int64_t>& shape, uint64_t* out) {
    uint64_t n = 1;
    for (int64_t d : shape) {
        if (d < 0 || d > (int64_t)1 << 40) return false;
        if (d != 0 && n > ((uint64_t)1 << 46) / (uint64_t)d) return false;
        n *= (uint64_t)d;
    }
    if (out) *out = n;
    return true;
}

std::unique_ptr<TensorFile> TensorFile::open_safetensors(const std::string& path, std::string& err) {
    std::unique_ptr<TensorFile> f(new TensorFile());
    if (!f->map_file(path, err)) return nullptr;
    const uint8_t* base = static_cast<const uint8_t*>(f->map_);
    if (f->map_size_ < 8) {
        err = "safetensors: file too small";
        return nullptr;
    }
    uint64_t hlen = 0;
    std::memcpy(&hlen, base, 8);
    if (hlen == 0 || hlen > f->map_size_ - 8) {
        err = "safetensors: bad header length";
        return nullptr;
    }
    JsonValue hdr;
    if (!json_parse(reinterpret_cast<const char*>(base + 8), reinterpret_cast<const char*>(base + 8 + hlen), hdr, err)) return nullptr;
    if (hdr.kind != JsonValue::Object) {
        err = "safetensors: header is not an object";
        return nullptr;
    }
    const uint8_t* data0 = base + 8 + hlen;
    const size_t data_size = f->map_size_ - 8 - (size_t)hlen;
    for (const auto& kv : hdr.obj) {
        if (kv.first == "__metadata__") continue;
        const JsonValue& e = kv.second;
        TensorInfo ti;
        ti.name = kv.first;
        std::string dt;
        if (!e.get_str("dtype", dt)) { err = "safetensors: no dtype for " + kv.first; return nullptr; }
        if (dt == "F32") ti.ggml_type = GGML_F32;
        else if (dt == "F16") ti.ggml_type = GGML_F16;
        else if (dt == "BF16") ti.ggml_type = GGML_BF16;
        else { err = "safetensors: unsupported dtype " + dt + " for " + kv.first; return nullptr; }
        const JsonValue* sh = e.get("shape");
        const JsonValue* off = e.get("data_offsets");
        if (!sh || sh->kind != JsonValue::Array || !off || off->kind != JsonValue::Array || off->arr.size() != 2) {
            err = "safetensors: bad entry " + kv.first;
            return nullptr;
        }
        for (const auto& d : sh->arr) ti.shape.push_back((int64_t)d.num);
        if (!checked_numel(ti.shape, nullptr)) { err = "safetensors: bad shape for " + kv.first; return nullptr; }
        const uint64_t o0 = (uint64_t)off->arr[0].num, o1 = (uint64_t)off->arr[1].num;
        if (o1 < o0 || o1 > data_size) { err = "safetensors: offsets out of range for " + kv.first; return nullptr; }
        ti.data = data0 + o0;
        ti.nbytes = (size_t)(o1 - o0);
        const size_t esz = ti.ggml_type == GGML_F32 ? 4 : 2;
        if ((size_t)ti.numel() * esz != ti.nbytes) { err = "safetensors: size mismatch for " + kv.first; return nullptr; }
        f->index_[ti.name] = f->tensors_.size();
        f->tensors_.push_back(std::move(ti));
    }
    return f;
}

namespace {
struct Cursor {
    const uint8_t* p;
    const uint8_t* end;
    bool ok = true;
    template <typename T> T rd() {
        T v{};
        if ((size_t)(end - p) < sizeof(T)) { ok = false; return v; }
        std::memcpy(&v, p, sizeof(T));
        p += sizeof(T);
        return v;
    }
    std::string str() {
        const uint64_t n = rd<uint64_t>();
        if (!ok || (uint64_t)(end - p) < n) { ok = false; return {}; }
        std::string s(reinterpret_cast<const char*>(p), (size_t)n);
        p += n;
        return s;
    }
    void skip(size_t n) {
        if ((size_t)(end - p) < n) ok = false; else p += n;
    }
};

size_t gguf_scalar_size(uint32_t t) {
    switch (t) {
        case 0: case 1: case 7: return 1;
        case 2: case 3: return 2;
        case 4: case 5: case 6: return 4;
        case 10: case 11: case 12: return 8;
    }
    return 0;
}

bool gguf_skip_value(Cursor& c, uint32_t type, uint64_t* u32_out) {
    if (type == 8) { c.str(); return c.ok; }
    if (type == 9) {
        const uint32_t et = c.rd<uint32_t>();
        const uint64_t n = c.rd<uint64_t>();
        if (!c.ok) return false;
        if (et == 8) { for (uint64_t i = 0; i < n && c.ok; ++i) c.str(); return c.ok; }
        const size_t es = gguf_scalar_size(et);
        if (es == 0) return false;
        if (n > (uint64_t)(c.end - c.p) / es) { c.ok = false; return false; }  // es * n must not wrap
        c.skip(es * (size_t)n);
        return c.ok;
    }
    const size_t s = gguf_scalar_size(type);
    if (s == 0) return false;
    if (u32_out && type == 4) { *u32_out = c.rd<uint32_t>(); return c.ok; }
    c.skip(s);
    return c.ok;
}
}  // namespace

std::unique_ptr<TensorFile> TensorFile::open_gguf(const std::string& path, std::string& err) {
    std::unique_ptr<TensorFile> f(new TensorFile());
    if (!f->map_file(path, err)) return nullptr;
    f->gguf_ = true;
    const uint8_t* base = static_cast<const uint8_t*>(f->map_);
    Cursor c{base, base + f->map_size_};
    const uint32_t magic = c.rd<uint32_t>();
    const uint32_t version = c.rd<uint32_t>();
    if (!c.ok || magic != 0x46554747u) { err = "gguf: bad magic in " + path; return nullptr; }
    if (version < 2 || version > 3) { err = "gguf: unsupported version"; return nullptr; }
    const uint64_t n_tensors = c.rd<uint64_t>();
    const uint64_t n_kv = c.rd<uint64_t>();
    if (!c.ok || n_tensors > f->map_size_ / 24 || n_kv > f->map_size_ / 12) { err = "gguf: implausible tensor / KV count"; return nullptr; }
    uint64_t alignment = 32;
    for (uint64_t i = 0; i < n_kv && c.ok; ++i) {
        const std::string key = c.str();
        const uint32_t type = c.rd<uint32_t>();
        uint64_t u = 0;
        const bool want = (key == "general.alignment");
        if (!gguf_skip_value(c, type, want ? &u : nullptr)) { err = "gguf: bad KV " + key; return nullptr; }
        if (want && type == 4 && u > 0 && u <= (1u << 20) && (u & (u - 1)) == 0) alignment = u;
    }
    struct Pending { TensorInfo ti; uint64_t offset; };
    std::vector<Pending> pend;
    for (uint64_t i = 0; i < n_tensors && c.ok; ++i) {
        Pending pd;
        pd.ti.name = c.str();
        const uint32_t nd = c.rd<uint32_t>();
        if (!c.ok || nd > 8) { err = "gguf: bad tensor info"; return nullptr; }
        std::vector<int64_t> ne(nd);
        for (uint32_t d = 0; d < nd; ++d) ne[d] = (int64_t)c.rd<uint64_t>();
        pd.ti.ggml_type = (int)c.rd<uint32_t>();
        pd.offset = c.rd<uint64_t>();
        pd.ti.shape.assign(ne.rbegin(), ne.rend());  // ggml ne[0] is the contiguous dim -> last in PyTorch order
        pend.push_back(std::move(pd));
    }
    if (!c.ok) { err = "gguf: truncated header"; return nullptr; }
    size_t data_off = (size_t)(c.p - base);
    data_off = (data_off + alignment - 1) / alignment * alignment;
    if (data_off > f->map_size_) { err = "gguf: truncated header"; return nullptr; }
    const size_t data_size = f->map_size_ - data_off;
    for (auto& pd : pend) {
        TensorInfo& ti = pd.ti;
        uint64_t numel = 0;
        if (!checked_numel(ti.shape, &numel)) { err = std::string("gguf: bad shape for ") + ti.name; return nullptr; }
        const int64_t row = ti.shape.empty() ? 1 : ti.shape.back();
        const size_t rb = row > 0 ? ggml_row_bytes(ti.ggml_type, row) : 0;
        if (rb == 0) { err = std::string("gguf: unsupported type/shape for ") + ti.name; return nullptr; }
        const uint64_t rows = numel / (uint64_t)row;
        if (rows > data_size / rb) { err = "gguf: tensor out of file range: " + ti.name; return nullptr; }  // rb * rows cannot wrap now
        ti.nbytes = rb * (size_t)rows;
        if (pd.offset > data_size || ti.nbytes > data_size - pd.offset) { err = "gguf: tensor out of file range: " + ti.name; return nullptr; }
        ti.data = base + data_off + pd.offset;
        f->index_[ti.name] = f->tensors_.size();
        f->tensors_.push_back(std::move(ti));
    }
    return f;
}

}  // namespace ace
