// Small formatting helpers used by the tensor types and error messages.
// API parity with the reference's src/utils/string_utils.h (fmtstr, vec2str, arr2str).
#pragma once
#include <cstdio>
#include <sstream>
#include <string>
#include <vector>

template <typename... Args>
inline std::string fmtstr(const std::string& format, Args... args) {
    const int n = std::snprintf(nullptr, 0, format.c_str(), args...);
    if (n <= 0) return std::string();
    std::string out(static_cast<size_t>(n) + 1, '\0');
    std::snprintf(&out[0], out.size(), format.c_str(), args...);
    out.resize(static_cast<size_t>(n));
    return out;
}

template <typename T>
inline std::string arr2str(const T* data, size_t count) {
    std::ostringstream os;
    os << "(";
    for (size_t i = 0; i < count; ++i) os << (i ? ", " : "") << data[i];
    os << ")";
    return os.str();
}

template <typename T>
inline std::string vec2str(const std::vector<T>& v) {
    return arr2str(v.data(), v.size());
}
