// Tensor carriers at the kernel boundary: Device / DataType enums, Tensor, TensorWrapper<T>,
// TensorMap. Source-compatible with the reference's src/utils/tensor.h:13-304 (same names, same
// member layout, same checks), re-written; size64() is added because Tensor::size() is int
// (the batch-32 x 4096 KV cache has 1.7e10 elements).
#pragma once
#include <cuda_fp16.h>
#include <cstdint>
#include <functional>
#include <initializer_list>
#include <numeric>
#include <string>
#include <type_traits>
#include <unordered_map>
#include <utility>
#include <vector>

#include "src/utils/macro.h"
#include "src/utils/string_utils.h"

enum Device { CPU_PINNED, CPU, GPU };
enum DataType { FP32, FP16, INT8, INT32, BOOL, BYTES, UNSUPPORTED };

template <typename T>
DataType getTensorType() {
    using U = typename std::remove_const<T>::type;
    if (std::is_same<U, float>::value) return FP32;
    if (std::is_same<U, half>::value) return FP16;
    if (std::is_same<U, int>::value) return INT32;
    if (std::is_same<U, int8_t>::value) return INT8;
    if (std::is_same<U, bool>::value) return BOOL;
    if (std::is_same<U, char>::value) return BYTES;
    return UNSUPPORTED;
}

template <typename T>
class TensorWrapper;

struct Tensor {
    Device location;
    DataType dtype;
    std::vector<int> shape;

    Tensor() = default;
    Tensor(Device location_, DataType dtype_, std::vector<int> shape_)
        : location(location_), dtype(dtype_), shape(std::move(shape_)) {}
    virtual ~Tensor() = default;

    virtual int size() const {
        if (shape.empty()) return 0;
        return std::accumulate(shape.begin(), shape.end(), 1, std::multiplies<int>());
    }
    int64_t size64() const {
        if (shape.empty()) return 0;
        int64_t n = 1;
        for (int s : shape) n *= s;
        return n;
    }
    template <typename T>
    TensorWrapper<T>* as() { return static_cast<TensorWrapper<T>*>(this); }

    std::string DeviceString() const {
        return location == CPU ? "CPU" : location == CPU_PINNED ? "CPU_PINNED" : "GPU";
    }
    static const char* TypeString(DataType t) {
        switch (t) {
            case FP32: return "FP32"; case FP16: return "FP16"; case INT8: return "INT8";
            case INT32: return "INT32"; case BOOL: return "BOOL"; case BYTES: return "BYTES";
            default: return "UNSUPPORTED";
        }
    }
    virtual std::string toString() const {
        return fmtstr("Tensor[where=%s, type=%s, shape= %s]", DeviceString().c_str(), TypeString(dtype),
                      vec2str(shape).c_str());
    }
};

template <typename T>
class TensorWrapper : public Tensor {
public:
    T* data = nullptr;

    TensorWrapper(Device location, DataType dtype, std::vector<int> shape) : Tensor(location, dtype, std::move(shape)) {}
    TensorWrapper(Device location, DataType dtype, std::vector<int> shape, T* data_)
        : Tensor(location, dtype, std::move(shape)), data(data_) {
        LLM_CHECK_WITH_INFO(getTensorType<T>() == dtype,
                            "when build TensorWarpper, the passed in data type should be same as dtype in params");
    }

    int size() const override {
        if (data == nullptr || shape.empty()) return 0;
        return std::accumulate(shape.begin(), shape.end(), 1, std::multiplies<int>());
    }
    inline T getVal(int id) const {
        LLM_CHECK(location == CPU);  // host-resident scalars only (step, layer_id)
        return data[id];
    }
    inline T getVal() const {
        LLM_CHECK(location == CPU);
        return getVal(0);
    }
    inline T* getPtr() const { return (T*)data; }
    inline T* getPtrByOffset(int offset) const { return (T*)data + offset; }

    std::string toString() const override {
        return fmtstr("Tensor[where=%s, type=%s, shape=%s, data=%p]", DeviceString().c_str(), TypeString(dtype),
                      vec2str(shape).c_str(), (void*)data);
    }
};

struct TensorMap {
    std::unordered_map<std::string, Tensor*> tensor_map_;

    TensorMap() = default;
    TensorMap(std::initializer_list<std::pair<std::string, Tensor*>> items) {
        for (auto& kv : items) {
            LLM_CHECK_WITH_INFO(isValid(kv.second),
                                fmtstr("%s is not a valid tensor, skipping insert into TensorMap", kv.first.c_str()));
            insert(kv.first, kv.second);
        }
    }
    TensorMap(const std::unordered_map<std::string, Tensor*>& items) {
        for (auto& kv : items) {
            LLM_CHECK_WITH_INFO(isValid(kv.second),
                                fmtstr("%s is not a valid tensor, skipping insert into TensorMap", kv.first.c_str()));
            insert(kv.first, kv.second);
        }
    }
    ~TensorMap() { tensor_map_.clear(); }

    inline size_t size() const { return tensor_map_.size(); }
    inline bool isExist(const std::string& key) const { return tensor_map_.find(key) != tensor_map_.end(); }
    inline bool isValid(const Tensor* tensor) { return tensor != nullptr && tensor->size() > 0; }
    inline void insert(const std::string& key, Tensor* value) { tensor_map_[key] = value; }
    inline void insert(const std::pair<std::string, Tensor*> p) { tensor_map_.insert(p); }

    inline Tensor* at(const std::string& key) {
        LLM_CHECK_WITH_INFO(isExist(key), fmtstr("Cannot find a tensor of name %s in the tensor map (keys: %s)",
                                                 key.c_str(), vec2str(keys()).c_str()));
        return tensor_map_.at(key);
    }
    inline Tensor* operator[](const std::string& key) { return at(key); }

    std::vector<std::string> keys() const {
        std::vector<std::string> names;
        for (auto& kv : tensor_map_) names.push_back(kv.first);
        return names;
    }
    std::string toString() {
        std::string s = "{";
        bool first = true;
        for (auto& kv : tensor_map_) {
            s += (first ? "" : ", ") + kv.first + ": " + kv.second->toString();
            first = false;
        }
        return s + "}";
    }
};
