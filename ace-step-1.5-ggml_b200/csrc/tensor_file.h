// Weight-file readers: safetensors and GGUF v2/v3, plus the tiny JSON reader both config.json and the
// safetensors header need.  Files are mmap()ed; tensors are exposed as (ggml type, shape, pointer).
//   safetensors format: 8-byte LE header length + JSON header + raw data   (reference reader: cpp/safetensors.cpp:46-169)
//   GGUF format: magic "GGUF", version, tensor count, KV count, KVs, tensor infos, aligned data section
//                (reference: third_party/ggml/src/gguf.cpp; exporter tools/export_safetensors_to_gguf.py)
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace ace {

// ------------------------------------------------------------------ JSON
struct JsonValue {
    enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
    bool b = false;
    double num = 0;
    std::string str;
    std::vector<JsonValue> arr;
    std::vector<std::pair<std::string, JsonValue>> obj;

    const JsonValue* get(const std::string& key) const;
    bool get_int(const std::string& key, int64_t& out) const;
    bool get_num(const std::string& key, double& out) const;
    bool get_bool(const std::string& key, bool& out) const;
    bool get_str(const std::string& key, std::string& out) const;
};
bool json_parse(const char* begin, const char* end, JsonValue& out, std::string& err);
bool json_parse_file(const std::string& path, JsonValue& out, std::string& err);

// ------------------------------------------------------------------ tensor files
struct TensorInfo {
    std::string name;
    int ggml_type = 0;            // GgmlType (wpack.h)
    std::vector<int64_t> shape;   // PyTorch order (outermost first); the LAST dim is the contiguous / quantised one
    const uint8_t* data = nullptr;
    size_t nbytes = 0;
    int64_t numel() const {
        int64_t n = 1;
        for (auto d : shape) n *= d;
        return n;
    }
};

class TensorFile {
public:
    ~TensorFile();
    static std::unique_ptr<TensorFile> open_safetensors(const std::string& path, std::string& err);
    static std::unique_ptr<TensorFile> open_gguf(const std::string& path, std::string& err);

    const TensorInfo* find(const std::string& name) const;
    const std::vector<TensorInfo>& tensors() const { return tensors_; }
    bool is_gguf() const { return gguf_; }
    // whole tensor decoded to fp32 (dequantising block formats row by row)
    bool read_f32(const std::string& name, std::vector<float>& out, std::vector<int64_t>* shape, std::string& err) const;

private:
    bool map_file(const std::string& path, std::string& err);
    void* map_ = nullptr;
    size_t map_size_ = 0;
    bool gguf_ = false;
    std::vector<TensorInfo> tensors_;
    std::map<std::string, size_t> index_;
};

bool file_exists(const std::string& path);

}  // namespace ace
