// Element types of ops::Tensor (API mirror of the reference's core/dtype.h:8-60; fp32 is the compute type,
// int32 carries token ids / labels).
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>

namespace ops {

enum class DType : int8_t { kFloat32 = 0, kFloat16 = 1, kInt32 = 2, kInt64 = 3, kInt8 = 4, kBool = 5, kUInt8 = 6 };

struct DTypeUtils {
    static size_t size_of(DType t) {
        static const size_t k[] = {4, 2, 4, 8, 1, 1, 1};
        int i = static_cast<int>(t);
        return (i >= 0 && i < 7) ? k[i] : 0;
    }
    static std::string to_string(DType t) {
        static const char* k[] = {"float32", "float16", "int32", "int64", "int8", "bool", "uint8"};
        int i = static_cast<int>(t);
        return (i >= 0 && i < 7) ? k[i] : "unknown";
    }
    static bool is_floating_point(DType t) { return t == DType::kFloat32 || t == DType::kFloat16; }
    static bool is_integer(DType t) { return t == DType::kInt32 || t == DType::kInt64 || t == DType::kInt8; }
};

constexpr DType kFloat32 = DType::kFloat32;
constexpr DType kFloat16 = DType::kFloat16;
constexpr DType kInt32 = DType::kInt32;
constexpr DType kInt64 = DType::kInt64;
constexpr DType kInt8 = DType::kInt8;
constexpr DType kBool = DType::kBool;
constexpr DType kUInt8 = DType::kUInt8;

}  // namespace ops
