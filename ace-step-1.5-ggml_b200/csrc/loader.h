// Shared weight-ingest helper of the model loaders (DiT + condition encoders, Qwen3 text encoder): reads tensors from a
// TensorFile, packs 2-D weights into device tiles (wpack.h) and uploads them, recording every allocation in the owner's list.
#pragma once
#include "gemm.h"
#include "tensor_file.h"
#include "wpack.h"

#include <cstring>
#include <string>
#include <vector>

#include <cuda_runtime.h>

namespace ace {

struct WeightLoader {
    const TensorFile* tf;
    std::vector<void*>* allocs;   // owner's allocation list (freed by its release())
    size_t* weight_bytes;
    std::string err;
    int status = 0;
    int online_q = -1;  // W_Q8_0 / W_Q4_K / W_Q6_K when ACE_GGML_*_WEIGHT_QTYPE asks for it (safetensors only)

    bool fail(int st, const std::string& e) {
        if (status == 0) {
            status = st;
            err = e;
        }
        return false;
    }
    void* dev_alloc(size_t bytes) {
        void* p = nullptr;
        if (cudaMalloc(&p, bytes ? bytes : 16) != cudaSuccess) {
            fail(1, "cudaMalloc failed (" + std::to_string(bytes) + " bytes): " + cudaGetErrorString(cudaGetLastError()));
            return nullptr;
        }
        allocs->push_back(p);
        *weight_bytes += bytes;
        return p;
    }
    bool upload(PackedWeight& w, const std::vector<uint8_t>& img) {
        w.dev = dev_alloc(img.size());
        if (!w.dev) return false;
        if (cudaMemcpy(w.dev, img.data(), img.size(), cudaMemcpyHostToDevice) != cudaSuccess)
            return fail(1, std::string("cudaMemcpy failed: ") + cudaGetErrorString(cudaGetLastError()));
        if (w.wtype != W_F16 && predecode_enabled()) {  // decoded F16 twin for the compute-bound regime (gemm.h)
            w.dev_f16 = dev_alloc(gemm_predecode_bytes(w));
            if (!w.dev_f16) return false;
            const cudaError_t e = gemm_predecode(w, w.dev_f16, nullptr);
            if (e != cudaSuccess) return fail(1, std::string("weight predecode failed: ") + cudaGetErrorString(e));
        }
        return true;
    }
    float* upload_f32(const std::vector<float>& v) {
        float* d = static_cast<float*>(dev_alloc(v.size() * 4));
        if (!d) return nullptr;
        if (cudaMemcpy(d, v.data(), v.size() * 4, cudaMemcpyHostToDevice) != cudaSuccess) {
            fail(1, std::string("cudaMemcpy failed: ") + cudaGetErrorString(cudaGetLastError()));
            return nullptr;
        }
        return d;
    }
    // The reference's on-line switch also quantises the [1, rows, row_len] tables it loads as 2-D tensors (scale_shift_table:
    // acestep_dit_model.cpp:279-332) and dequantises them again at run time (ggml_cast); the equivalent here is a round trip.
    void online_roundtrip(std::vector<float>& v, int64_t row_len) const {
        if (tf->is_gguf() || online_q < 0 || row_len <= 0) return;
        const int gt = online_q == W_Q8_0 ? GGML_Q8_0 : (online_q == W_Q4_K ? GGML_Q4_K : GGML_Q6_K);
        if (row_len % ggml_block_elems(gt) != 0) return;
        std::vector<uint8_t> q(ggml_row_bytes(gt, row_len));
        for (size_t r = 0; r + (size_t)row_len <= v.size(); r += (size_t)row_len) {
            if (gt == GGML_Q8_0) quantize_row_q8_0(v.data() + r, q.data(), row_len);
            else if (gt == GGML_Q4_K) quantize_row_q4_k(v.data() + r, q.data(), row_len);
            else quantize_row_q6_k(v.data() + r, q.data(), row_len);
            dequant_row(gt, q.data(), v.data() + r, row_len);
        }
    }
    bool vec(const std::string& name, int64_t expect, float*& out) {
        std::vector<float> v;
        std::string e;
        if (!tf->read_f32(name, v, nullptr, e)) return fail(3, e);
        if ((int64_t)v.size() != expect) return fail(3, "unexpected size for " + name);
        out = upload_f32(v);
        return out != nullptr;
    }

    // Row-concatenation (or pairwise interleave) of 2-D weights [N_i, K] -> one packed matrix.
    bool linear(const std::vector<std::string>& names, bool interleave, int K, PackedWeight& out) {
        std::vector<const TensorInfo*> ti;
        int N = 0;
        for (const auto& n : names) {
            const TensorInfo* t = tf->find(n);
            if (!t) return fail(3, "missing tensor: " + n);
            if (t->shape.size() != 2 || t->shape[1] != K) return fail(3, "invalid tensor shape for " + n);
            ti.push_back(t);
            N += (int)t->shape[0];
        }
        if (interleave && (ti.size() != 2 || ti[0]->shape[0] != ti[1]->shape[0])) return fail(1, "bad interleave request");
        const int gt = ti[0]->ggml_type;
        bool same = true;
        for (auto* t : ti) same = same && t->ggml_type == gt;
        const bool quant_file = same && (gt == GGML_Q8_0 || gt == GGML_Q4_K || gt == GGML_Q6_K) && K % TILE_K == 0;
        std::vector<uint8_t> img;
        std::string e;
        if (quant_file) {
            const size_t rb = ggml_row_bytes(gt, K);
            std::vector<uint8_t> rows((size_t)N * rb);
            if (interleave) {
                const int half = (int)ti[0]->shape[0];
                for (int j = 0; j < half; ++j) {
                    std::memcpy(rows.data() + (size_t)(2 * j) * rb, ti[0]->data + (size_t)j * rb, rb);
                    std::memcpy(rows.data() + (size_t)(2 * j + 1) * rb, ti[1]->data + (size_t)j * rb, rb);
                }
            } else {
                size_t off = 0;
                for (auto* t : ti) {
                    std::memcpy(rows.data() + off, t->data, (size_t)t->shape[0] * rb);
                    off += (size_t)t->shape[0] * rb;
                }
            }
            if (!pack_tiles_ggml(rows.data(), gt, N, K, img, out, e)) return fail(1, e);
        } else {
            std::vector<float> w((size_t)N * K);
            if (interleave) {
                std::vector<float> a, b;
                if (!tf->read_f32(names[0], a, nullptr, e) || !tf->read_f32(names[1], b, nullptr, e)) return fail(3, e);
                const int half = (int)ti[0]->shape[0];
                for (int j = 0; j < half; ++j) {
                    std::memcpy(w.data() + (size_t)(2 * j) * K, a.data() + (size_t)j * K, (size_t)K * 4);
                    std::memcpy(w.data() + (size_t)(2 * j + 1) * K, b.data() + (size_t)j * K, (size_t)K * 4);
                }
            } else {
                size_t off = 0;
                for (const auto& n : names) {
                    std::vector<float> a;
                    if (!tf->read_f32(n, a, nullptr, e)) return fail(3, e);
                    std::memcpy(w.data() + off, a.data(), a.size() * 4);
                    off += a.size();
                }
            }
            if (!pack_f32(w, N, K, img, out)) return false;
        }
        return upload(out, img);
    }
    bool pack_f32(const std::vector<float>& w, int N, int K, std::vector<uint8_t>& img, PackedWeight& out) {
        std::string e;
        // quantise when the input dim is a whole number of blocks, else keep the tensor unquantised (acestep_dit_model.cpp:165-173)
        int wt = W_F16;
        if (!tf->is_gguf() && online_q == W_Q8_0 && K % TILE_K == 0) wt = W_Q8_0;
        if (!tf->is_gguf() && (online_q == W_Q4_K || online_q == W_Q6_K) && K % 256 == 0) wt = online_q;
        if (!pack_tiles_f32(w.data(), N, 1, K, wt, img, out, e)) return fail(1, e);
        return true;
    }
};

}  // namespace ace
