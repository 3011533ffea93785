// layer_driver — runs the reference-shaped C++ layer classes (src/layers/**, written over the
// header-only launch* shim and libb200llama.so) on tensors read from a directory and writes the
// results back, so that tests/test_layers_cpp_gpu.py can compare them with the CPU oracle.
// It plays the role of the reference's examples/cpp/{attention,ffn}/*_example.cpp (which build the
// three layer classes on all-ones data and check nothing) — with real data and a result check.
//
//   layer_driver <dir>
//     <dir>/cfg.txt                   key=value lines (mode, dtype, layers, head_num, kv_head_num,
//                                     head_size, inter_size, batch, max_seq, steps, context, ...)
//     <dir>/model.layers.<i>.*.bin    weights, raw little-endian fp32, reference file naming
//     <dir>/<tensor>.bin              inputs (fp32 / int32); outputs are written next to them
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "src/layers/decoder/context_decoder.h"
#include "src/layers/decoder/self_decoder.h"

namespace {

std::map<std::string, std::string> read_cfg(const std::string& path) {
    std::map<std::string, std::string> m;
    std::ifstream in(path);
    LLM_CHECK_WITH_INFO(in.is_open(), "cannot open " + path);
    std::string line;
    while (std::getline(in, line)) {
        const size_t eq = line.find('=');
        if (eq != std::string::npos) m[line.substr(0, eq)] = line.substr(eq + 1);
    }
    return m;
}

template <typename V>
std::vector<V> read_bin(const std::string& path, size_t n) {
    std::ifstream in(path, std::ios::binary);
    LLM_CHECK_WITH_INFO(in.is_open(), "cannot open " + path);
    std::vector<V> v(n);
    in.read(reinterpret_cast<char*>(v.data()), (std::streamsize)(n * sizeof(V)));
    LLM_CHECK_WITH_INFO((size_t)in.gcount() == n * sizeof(V), path + ": short read");
    return v;
}

void write_bin(const std::string& path, const std::vector<float>& v) {
    std::ofstream out(path, std::ios::binary);
    out.write(reinterpret_cast<const char*>(v.data()), (std::streamsize)(v.size() * sizeof(float)));
}

template <typename T> T to_t(float v);
template <> float to_t<float>(float v) { return v; }
template <> half to_t<half>(float v) { return __float2half(v); }
inline float to_f(float v) { return v; }
inline float to_f(half v) { return __half2float(v); }

template <typename T>
T* to_device(const std::vector<float>& h) {
    std::vector<T> t(h.size());
    for (size_t i = 0; i < h.size(); ++i) t[i] = to_t<T>(h[i]);
    T* d = nullptr;
    CHECK(cudaMalloc((void**)&d, sizeof(T) * (h.size() ? h.size() : 1)));
    CHECK(cudaMemcpy(d, t.data(), sizeof(T) * h.size(), cudaMemcpyHostToDevice));
    return d;
}
int* to_device_i(const std::vector<int>& h) {
    int* d = nullptr;
    CHECK(cudaMalloc((void**)&d, sizeof(int) * (h.size() ? h.size() : 1)));
    CHECK(cudaMemcpy(d, h.data(), sizeof(int) * h.size(), cudaMemcpyHostToDevice));
    return d;
}
template <typename T>
std::vector<float> to_host(const T* d, size_t n) {
    std::vector<T> t(n);
    CHECK(cudaMemcpy(t.data(), d, sizeof(T) * n, cudaMemcpyDeviceToHost));
    std::vector<float> h(n);
    for (size_t i = 0; i < n; ++i) h[i] = to_f(t[i]);
    return h;
}

struct Cfg {
    std::map<std::string, std::string> m;
    int i(const std::string& k, int dflt = 0) const {
        auto it = m.find(k);
        return it == m.end() ? dflt : std::atoi(it->second.c_str());
    }
    float f(const std::string& k, float dflt) const {
        auto it = m.find(k);
        return it == m.end() ? dflt : (float)std::atof(it->second.c_str());
    }
    std::string s(const std::string& k, const std::string& dflt) const {
        auto it = m.find(k);
        return it == m.end() ? dflt : it->second;
    }
};

template <typename T>
int run(const std::string& dir, const Cfg& c) {
    const int L = c.i("layers"), h = c.i("head_num"), hk = c.i("kv_head_num"), d = c.i("head_size");
    const int I = c.i("inter_size"), B = c.i("batch"), max_seq = c.i("max_seq");
    const int H = h * d;
    const float eps = c.f("rms_eps", 1e-5f);
    const std::string mode = c.s("mode", "decode");
    LLaMAAttentionStaticParams sp{c.i("rotary_dim", d), c.f("rotary_base", 10000.f), max_seq, false};
    const DataType dt = getTensorType<T>();
    const WeightType wt = getWeightType<T>();

    CudaAllocator allocator;
    cublasWrapper cublas(nullptr, nullptr);
    if (dt == FP16) cublas.setFP16GemmConfig(); else cublas.setFP32GemmConfig();
    cudaStream_t stream = nullptr;   // the reference's legacy default stream

    std::vector<std::unique_ptr<LlamaLayerWeight<T>>> owned;
    std::vector<LlamaLayerWeight<T>*> weights;
    for (int l = 0; l < L; ++l) {
        owned.emplace_back(new LlamaLayerWeight<T>(h, hk, d, I, wt, false));
        owned.back()->loadWeights(dir + "/model.layers." + std::to_string(l), wt);
        weights.push_back(owned.back().get());
    }
    const size_t cache_elems = (size_t)L * B * hk * max_seq * d;
    T* k_cache = to_device<T>(read_bin<float>(dir + "/k_cache.bin", cache_elems));
    T* v_cache = to_device<T>(read_bin<float>(dir + "/v_cache.bin", cache_elems));
    TensorWrapper<T> all_k(GPU, dt, {L, B, hk, max_seq, d}, k_cache), all_v(GPU, dt, {L, B, hk, max_seq, d}, v_cache);
    int layer0 = 0;
    TensorWrapper<int> layer_id(CPU, INT32, {1}, &layer0);

    if (mode == "decode") {
        // `steps` decode steps starting with `context` tokens in the cache; hidden_in [steps, B, H]
        const int steps = c.i("steps", 1), context = c.i("context");
        std::vector<float> xin = read_bin<float>(dir + "/hidden_in.bin", (size_t)steps * B * H), xout;
        LlamaSelfDecoder<T> dec(h, hk, d, I, L, sp, eps, stream, &cublas, &allocator);
        bool* finished = nullptr;
        CHECK(cudaMalloc((void**)&finished, B));
        CHECK(cudaMemset(finished, 0, B));
        T* out = nullptr;
        CHECK(cudaMalloc((void**)&out, sizeof(T) * (size_t)B * H));
        for (int s = 0; s < steps; ++s) {
            std::vector<float> xs(xin.begin() + (size_t)s * B * H, xin.begin() + (size_t)(s + 1) * B * H);
            T* in = to_device<T>(xs);
            int step = context + s + 1;   // reference `step`: tokens in the cache including the new one
            TensorWrapper<T> din(GPU, dt, {B, H}, in), dout(GPU, dt, {B, H}, out);
            TensorWrapper<int> st(CPU, INT32, {1}, &step);
            TensorWrapper<bool> fin(GPU, BOOL, {B}, finished);
            TensorMap inputs{{"decoder_input", &din}, {"step", &st}, {"finished", &fin}, {"layer_id", &layer_id}};
            TensorMap outputs{{"decoder_output", &dout}, {"all_k_cache", &all_k}, {"all_v_cache", &all_v}};
            LLaMAAttentionDynParams dyn{B, 0, 1, step, L, false};
            dec.forward(inputs, weights, outputs, dyn);
            DeviceSyncAndCheckCudaError();
            std::vector<float> o = to_host(out, (size_t)B * H);
            xout.insert(xout.end(), o.begin(), o.end());
            cudaFree(in);
        }
        write_bin(dir + "/hidden_out.bin", xout);
        cudaFree(finished);
        cudaFree(out);
    } else {
        // prefill: hidden_in [num_tokens, H]; input_length / history_length [B] int32
        std::vector<int> in_len = read_bin<int>(dir + "/input_length.bin", B);
        std::vector<int> hist = read_bin<int>(dir + "/history_length.bin", B);
        std::vector<int> ctx(B);
        int Nt = 0, max_q = 0, max_k = 0;
        for (int b = 0; b < B; ++b) {
            ctx[b] = in_len[b] + hist[b];
            Nt += in_len[b];
            max_q = std::max(max_q, in_len[b]);
            max_k = std::max(max_k, ctx[b]);
        }
        max_q = c.i("max_q_len", max_q);
        T* in = to_device<T>(read_bin<float>(dir + "/hidden_in.bin", (size_t)Nt * H));
        T* out = nullptr;
        CHECK(cudaMalloc((void**)&out, sizeof(T) * (size_t)Nt * H));
        int *d_len = to_device_i(in_len), *d_hist = to_device_i(hist), *d_ctx = to_device_i(ctx);
        LlamaContextDecoder<T> dec(h, hk, d, I, L, sp, eps, stream, &cublas, &allocator);
        dec.setFusedAttention(c.i("fused_attention", 1) != 0);
        TensorWrapper<T> din(GPU, dt, {Nt, H}, in), dout(GPU, dt, {Nt, H}, out);
        TensorWrapper<int> tl(GPU, INT32, {B}, d_len), th(GPU, INT32, {B}, d_hist), tc(GPU, INT32, {B}, d_ctx);
        TensorMap inputs{{"decoder_input", &din}, {"input_length", &tl}, {"history_length", &th},
                         {"context_length", &tc}, {"layer_id", &layer_id}};
        TensorMap outputs{{"decoder_output", &dout}, {"all_k_cache", &all_k}, {"all_v_cache", &all_v}};
        LLaMAAttentionDynParams dyn{B, Nt, max_q, max_k, L, true};
        dec.forward(inputs, weights, outputs, dyn);
        DeviceSyncAndCheckCudaError();
        write_bin(dir + "/hidden_out.bin", to_host(out, (size_t)Nt * H));
        cudaFree(in); cudaFree(out); cudaFree(d_len); cudaFree(d_hist); cudaFree(d_ctx);
    }
    write_bin(dir + "/k_cache_out.bin", to_host(k_cache, cache_elems));
    write_bin(dir + "/v_cache_out.bin", to_host(v_cache, cache_elems));
    cudaFree(k_cache);
    cudaFree(v_cache);
    return 0;
}

}  // namespace

int main(int argc, char** argv) {
    if (argc < 2) {
        std::fprintf(stderr, "usage: %s <dir>\n", argv[0]);
        return 2;
    }
    try {
        int n = 0;
        if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
            std::fprintf(stderr, "[b200][ERROR] no CUDA device: there is no CPU fallback\n");
            return 3;
        }
        Cfg c{read_cfg(std::string(argv[1]) + "/cfg.txt")};
        const int rc = c.s("dtype", "fp32") == "fp16" ? run<half>(argv[1], c) : run<float>(argv[1], c);
        std::printf("layer_driver ok: %llu kernel launches through libb200llama.so\n",
                    (unsigned long long)b200_launch_count());
        return rc;
    } catch (const std::exception& e) {
        std::fprintf(stderr, "%s\n", e.what());
        return 1;
    }
}
